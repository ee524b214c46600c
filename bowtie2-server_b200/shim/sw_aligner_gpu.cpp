// sw_aligner_gpu.cpp — reference-side shim: SwAligner::align / SwAligner::nextAlignment bodies that run the
// dynamic programming on the GPU through the C ABI (include/bt2gpu.h) instead of aligner_swsse_*.cpp.
//
// Compiled against the reference's own headers (-I<reference tree>) and linked into the unmodified reference in
// place of the two member functions it replaces (aligner_sw.cpp:500-729 align, :737-1146 nextAlignment); every
// other member of SwAligner (initRead, initRef, ungappedAlign, done(), reset(), the metrics) is the reference's.
// SwDriver, AlnSinkWrap, the SAM writer and the client/server front end are untouched, so the SAM the binary
// prints is the parity statement: it must be byte-identical to the stock reference's.
//
// Per call the shim sends ONE problem (the batch-of-one binding; the batched runtime is bt2gpu_hotpath_*).  The
// GPU runs initRef's window + fill + gather + the whole nextAlignment loop at once; nextAlignment() then hands
// the alignments out one per call and replays the reference's RNG protocol (one rnd.nextU32() re-seed per
// attempted backtrace: aligner_sw.cpp:794-795,877-879,974,1062) so the caller's RandomSource is in the same
// state as after the reference's call.
//
// No CPU fallback: anything the CUDA path does not cover (checkpointed reads >= cperMinlen_, ABI errors) throws,
// which bowtie() reports as a fatal error (bt2_search.cpp:5604-5611).
#include <cstdio>
#include <unordered_map>
#include <vector>
#include <limits>
#include "aligner_sw.h"
#include "reference.h"
#include "gpu_shim.h"

#ifndef BT2SHIM_BATCHED
#define BT2SHIM_BATCHED 0      /* 1: compiled for the batched binding (batch_shim.cpp) */
#endif

namespace {

struct ShimState {               // what align() got back for one SwAligner instance
	bt2shim::DpCall call;        // the problem (kept for the raised-minsc replay of nextAlignment()) and its results
	std::vector<uint8_t> seq, qual, win;
	uint32_t next = 0;           // next alignment to hand out
	TAlScore minsc = 0;          // initRef's minsc_
};

// One state per SwAligner instance.  Instances belong to one worker (thread or fiber) and workers never migrate
// between host threads, so the map is per host thread; std::unordered_map keeps references to its elements valid when
// another worker of the same thread inserts, and nothing here is held across a call into the binding except `st`.
thread_local std::unordered_map<const SwAligner*, ShimState> g_st;

const uint32_t kMaxAln = 1024;   // per-call binding: the whole nextAlignment loop in one call
const uint32_t kFirstAln = 16;   // batched binding: first ask; a window with more alignments is asked again in full

} // namespace

/**
 * SwAligner::align (aligner_sw.cpp:500-729): same contract — `best` = best score in the matrix (MIN_I64 when the
 * fill found nothing valid), returns true iff there is at least one backtrace candidate; leaves btncand_ sorted,
 * sse8succ_/sse16succ_ and cural_ as the reference does.
 */
bool SwAligner::align(TAlScore& best) {
	assert(initedRef() && initedRead());
	state_ = STATE_ALIGNED;
	btncand_.clear();
	btncanddone_.clear();
	btncanddoneSucc_ = btncanddoneFail_ = 0;
	best = std::numeric_limits<TAlScore>::min();
	sse8succ_ = sse16succ_ = false;
	cural_ = 0;
	const size_t rdlen = rdf_ - rdi_;
	const size_t ncol = (size_t)(rff_ - rfi_);
	if(rdlen >= cperMinlen_) {
		fprintf(stderr, "bt2gpu shim: checkpointed DP (read of %zu rows >= %zu) is not on the CUDA path\n", rdlen, cperMinlen_);
		throw 1;
	}
	ShimState& st = g_st[this];
	// the oriented read, rows rdi_..rdf_-1 (bytes 0..4 + Phred+33), and the window incl. the extra column of initRef
	st.seq.resize(rdlen); st.qual.resize(rdlen); st.win.resize(ncol + 1);
	for(size_t i = 0; i < rdlen; i++) { st.seq[i] = (uint8_t)(*rd_)[rdi_ + i]; st.qual[i] = (uint8_t)(*qu_)[rdi_ + i]; }
	for(size_t i = 0; i <= ncol; i++) st.win[i] = (uint8_t)rf_[rfi_ + i];
	bt2shim::DpCall& c = st.call;
	c.sc = sc_; c.seq = st.seq.data(); c.qual = st.qual.data(); c.win = st.win.data();
	bt2gpu_dp_job& job = c.job;
	job.rd_off = 0; job.rd_len = (uint32_t)rdlen; job.rf_off = 0; job.ncol = (uint32_t)ncol;
	job.minsc = (int32_t)minsc_; job.nceil = nceil_;
	job.refl = rect_->refl; job.triml = (uint32_t)rect_->triml;
	job.corel = (uint32_t)rect_->corel; job.corer = (uint32_t)rect_->corer;
	job.fw = fw_ ? 1 : 0; job.enable8 = (enable8_ && !readSse16_) ? 1 : 0;
	job.reserved[0] = job.reserved[1] = 0;
	c.max_aln = BT2SHIM_BATCHED ? kFirstAln : kMaxAln;
	bt2shim::dpAlign(c);
	if(c.res.naln >= c.max_aln && c.max_aln < kMaxAln) { c.max_aln = kMaxAln; bt2shim::dpAlign(c); }
	if(c.res.naln >= kMaxAln) { fprintf(stderr, "bt2gpu shim: more than %u alignments in one window\n", kMaxAln); throw 1; }
	const bt2gpu_dp_result& r = c.res;
	st.next = 0; st.minsc = minsc_;
	sse8succ_ = r.width == 8; sse16succ_ = r.width == 16;
	best = r.best;                           // INT64_MIN == MIN_I64
	// SSEMetrics (aligner_swsse.h:32-78) as the reference's fill and gather routines keep them (ee_u8.cpp:791,1109-1144,
	// 1197 and twins; NWORDS_PER_REG of the default SSE2 build: 16 / 8).  Which variants ran follows from align()'s own
	// selection (aligner_sw.cpp:516-621): end-to-end runs u8 OR i16; local runs u8 and, when that saturated (the answer
	// then came from the 16-bit fill), i16 as well.  Not kept: `fixup` (the lazy-F loop is an artefact of the striped
	// layout), the columns of an 8-bit fill that saturated, and of the backtrace counters btcell, corerej and nrej (the
	// kernel reports attempts and successes, not the reason of a failure: every failed attempt is booked as btfail); see
	// nextAlignment below.  Checked against the reference's own counters by tests/test_shim_dp_replay.py.
	{
		const bool try8 = sc_->monotone ? (enable8_ && !readSse16_ && minsc_ >= -254) : (enable8_ && !readSse16_);
		const bool ran16 = !try8 || (!sc_->monotone && r.width == 16);
		SSEMetrics& m8 = extend_ ? sseU8ExtendMet_ : sseU8MateMet_;
		SSEMetrics& m16 = extend_ ? sseI16ExtendMet_ : sseI16MateMet_;
		if(try8) {
			const size_t ninner = ncol * ((rdlen + 15) / 16);
			m8.dp++;
			// a saturated 8-bit fill stops at the column where it saturates: how far it got is not known here, and its
			// columns are left out rather than overstated
			if(!ran16) { m8.col += ncol; m8.cell += ninner * 16; m8.inner += ninner; }
			if(ran16) m8.dpsat++;
			else if(r.width == 8) m8.dpsucc++;
			else m8.dpfail++;
		}
		if(ran16) {
			const size_t ninner = ncol * ((rdlen + 7) / 8);
			m16.dp++; m16.col += ncol; m16.cell += ninner * 8; m16.inner += ninner;
			if(r.width == 16) m16.dpsucc++; else m16.dpfail++;
		}
		if(r.aligned || r.ncand > 0) (r.width == 16 ? m16 : m8).gathsol += r.ncand;
	}
	if(best == MIN_I64 || best < minsc_) {   // aligner_sw.cpp:681-686
		if(dpLog_ != NULL) (*dpLog_) << ",0,0";
		return false;
	}
	if(dpLog_ != NULL) (*dpLog_) << ",1," << best;
	for(uint32_t i = 0; i < r.ncand; i++) {
		const bt2gpu_dp_cand& k = c.cands[i];
		btncand_.push_back(DpBtCandidate(k.row, k.col, k.score));
	}
	return r.aligned != 0;
}

/**
 * SwAligner::nextAlignment (aligner_sw.cpp:737-1146): hands out the alignments of the candidate loop the GPU
 * already ran, in order, leaving `res`, cural_ and `rnd` exactly as the reference's call would.
 */
bool SwAligner::nextAlignment(SwResult& res, TAlScore minsc, RandomSource& rnd) {
	assert(initedRead() && initedRef());
	if(done()) { res.reset(); return false; }
	ShimState& st = g_st[this];
	const bt2gpu_dp_result& r0 = st.call.res;
	if(minsc < st.minsc) { fprintf(stderr, "bt2gpu shim: nextAlignment minsc below initRef's\n"); throw 1; }
	const bt2gpu_dp_aln *a = (st.next < r0.naln) ? &st.call.alns[st.next] : NULL;
	uint32_t tries = a ? a->tries : r0.tail_tries;
	if(minsc > st.minsc) {
		// The driver tightened minsc after reporting an alignment (aligner_sw_driver.cpp:1255-1284).  nextAlignment only
		// uses its minsc argument to drop candidates (aligner_sw.cpp:762); btncand_ is sorted by score, so from the
		// first candidate below the new minsc on, everything is dropped without a backtrace or an RNG draw.
		if(btncand_[cural_].score < minsc) { res.reset(); cural_ = btncand_.size(); return false; }
		if(a == NULL || btncand_[a->cural - 1].score < minsc) {
			// The GPU loop's next success (or its tail) lies beyond that point: the backtraces attempted before it are
			// those of the same problem with minsc raised from the start (candidates >= the new minsc come first and
			// see the same history) -- replay it and take its tail count.
			bt2shim::DpCall c2;
			c2.sc = sc_; c2.seq = st.seq.data(); c2.qual = st.qual.data(); c2.win = st.win.data();
			c2.job = st.call.job;
			c2.job.minsc = (int32_t)minsc; c2.job.enable8 = sse8succ_ ? 1 : 0;
			c2.max_aln = st.next + 1 < kFirstAln ? kFirstAln : kMaxAln;     // the replay stops after st.next alignments
			bt2shim::dpAlign(c2);
			const bt2gpu_dp_result& r2 = c2.res;
			if(r2.naln != st.next || (r2.width != 0 && r2.width != r0.width)) {
				fprintf(stderr, "bt2gpu shim: raised-minsc replay diverged (%u alignments, expected %u)\n", r2.naln, st.next);
				throw 1;
			}
			a = NULL; tries = r2.tail_tries;
		}
	}
	for(uint32_t k = 0; k < tries; k++) {    // u8: init(reseed) ... init(reseed+1); i16: ... init(reseed)
		uint32_t reseed = rnd.nextU32() + 1;
		rnd.init(sse8succ_ ? reseed + 1 : reseed);
	}
	{	// one backtrace per draw (met.bt++, ee_u8.cpp:1296), one success when an alignment is handed out (:1875)
		SSEMetrics& m = sse8succ_ ? (extend_ ? sseU8ExtendMet_ : sseU8MateMet_) : (extend_ ? sseI16ExtendMet_ : sseI16MateMet_);
		m.bt += tries;
		if(a != NULL) m.btsucc++;
		m.btfail += tries - (a != NULL ? 1 : 0);
	}
	res.reset();
	if(a == NULL) { cural_ = btncand_.size(); return false; }
	const size_t rdlen = rdf_ - rdi_;
	const size_t trimBeg = fw_ ? a->trim5 : a->trim3;
	EList<Edit>& ned = res.alres.ned();
	for(uint32_t i = 0; i < a->nedits; i++) {
		const bt2gpu_edit& e = st.call.edits[a->edit_off + i];
		ned.push_back(Edit(e.pos, (int)e.chr, (int)e.qchr, (int)e.type));
	}
	// back to the backtrace's frame (w.r.t. the upstream end, before AlnRes::setShape's trim shift), so that
	// setShape and invertEdits below run exactly as in the reference (ee_u8.cpp:1819-1847, aligner_sw.cpp:1137)
	if(!fw_) Edit::invertPoss(ned, rdlen - a->trim5 - a->trim3, false);
	for(size_t i = 0; i < ned.size(); i++) ned[i].pos += (uint32_t)trimBeg;
	AlnScore score;
	score.score_ = a->score; score.ns_ = a->ns; score.gaps_ = a->gaps;
	score.edits_ = (int)a->nedits;
	score.basesAligned_ = (int)(rdlen - a->trim5 - a->trim3 - a->nedits);
	res.alres.setScore(score);
	res.alres.setShape(refidx_, a->refoff, reflen_, fw_, rdlen, true, 0, 0, true, a->trim5, a->trim3);
	res.alres.setRefNs(a->refns);
	if(!fw_) res.alres.invertEdits();
	cural_ = a->cural;
	st.next++;
	bt2shim::note(bt2shim::C_ALN, 1);
	return true;
}

/**
 * SwAligner::ungappedAlign (aligner_sw.cpp:286-494): the single-diagonal alignment SwDriver tries when no gap is
 * affordable.  The window is cut out of the BitPairReference here (Ns off the ends, as :326-360); the overhang
 * rules, the scoring scan, the soft trims and the edit list come from bt2gpu_ungapped_batch.
 */
int SwAligner::ungappedAlign(
	const BTDnaString& rd, const BTString& qu, const Coord& coord, const BitPairReference& refs, size_t reflen,
	const Scoring& sc, bool ohang, TAlScore minsc, SwResult& res)
{
	const size_t len = rd.length();
	std::vector<uint8_t> seq(len), qual(len), win(len);
	const TRefOff rfi = coord.off();
	for(size_t i = 0; i < len; i++) {
		seq[i] = (uint8_t)rd[i]; qual[i] = (uint8_t)qu[i];
		const TRefOff p = rfi + (TRefOff)i;
		win[i] = (p < 0 || (size_t)p >= reflen) ? 4 : (uint8_t)refs.getBase(coord.ref(), (size_t)p);
	}
	bt2shim::UngapCall uc;
	uc.sc = &sc; uc.seq = seq.data(); uc.qual = qual.data(); uc.win = win.data();
	bt2gpu_ungap_job& j = uc.job;
	j.rd_off = 0; j.rd_len = (uint32_t)len; j.rf_off = 0; j.minsc = (int32_t)minsc;
	j.nceil = sc.nCeil.f<int>((double)len);
	j.fw = coord.fw() ? 1 : 0; j.ohang = ohang ? 1 : 0; j.pad[0] = j.pad[1] = 0;
	j.off = rfi; j.reflen = (int64_t)reflen;
	bt2shim::ungapped(uc);
	const bt2gpu_ungap_result& r = uc.res;
	if(r.ret != 1) {
		// the reference has reset res.alres by the time it scores (:386); before that it returns with res untouched
		bool early = false;
		if(rfi < 0 && !ohang) early = true;
		if(rfi + (TRefOff)len > (TRefOff)reflen && !ohang) early = true;
		size_t leftNs = rfi < 0 ? (size_t)(-rfi) : 0, rightNs = rfi + (TRefOff)len > (TRefOff)reflen ? (size_t)(rfi + (TRefOff)len - (TRefOff)reflen) : 0;
		if(leftNs + rightNs > (size_t)j.nceil) early = true;
		if(!early) res.alres.reset();
		return r.ret;
	}
	res.alres.reset();
	const bool fw = coord.fw();
	const size_t rowi = fw ? r.trim5 : r.trim3, trimEnd = fw ? r.trim3 : r.trim5;
	EList<Edit>& ned = res.alres.ned();
	for(uint32_t i = 0; i < r.nedits; i++) {
		const bt2gpu_edit& e = uc.edits[r.edit_off + i];
		ned.push_back(Edit(e.pos, (int)e.chr, (int)e.qchr, (int)e.type));
	}
	// back to row coordinates so that setShape / invertEdits below run as in the reference (:456-490)
	if(!fw) Edit::invertPoss(ned, len - rowi - trimEnd, false);
	for(size_t i = 0; i < ned.size(); i++) ned[i].pos += (uint32_t)rowi;
	res.alres.setScore(AlnScore(r.score, (int)(len - ned.size()), (int)ned.size(), r.ns, 0));
	res.alres.setShape(coord.ref(), r.refoff, reflen, fw, len, true, 0, 0, true, r.trim5, r.trim3);
	res.alres.setRefNs(r.refns);
	if(!fw) res.alres.invertEdits();
	return 1;
}
