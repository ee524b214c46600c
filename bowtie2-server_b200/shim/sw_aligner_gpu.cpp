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

namespace {

struct ShimState {               // what align() got back for one SwAligner instance
	bt2gpu_dp_result res;
	std::vector<bt2gpu_dp_aln> alns;
	std::vector<bt2gpu_edit> edits;
	uint32_t next = 0;           // next alignment to hand out
	TAlScore minsc = 0;          // initRef's minsc_
	bt2gpu_dp_job job;           // the problem, kept for the raised-minsc replay of nextAlignment()
	std::vector<uint8_t> seq, qual, win;
};

struct ShimThread {
	std::unordered_map<const SwAligner*, ShimState> st;
	std::vector<bt2gpu_dp_cand> cands;
	std::vector<bt2gpu_dp_aln> alns2;
	std::vector<bt2gpu_edit> edits2;
	std::vector<uint8_t> useq, uqual, uwin;   // ungappedAlign
	std::vector<bt2gpu_edit> uedits;
};
thread_local ShimThread g_t;

const uint32_t kMaxAln = 1024;
const size_t kEditCap = (size_t)1 << 18;

size_t candCap(const bt2gpu_dp_job& j, bool monotone) {
	return monotone ? (size_t)j.ncol + 64 : (size_t)j.rd_len * j.ncol + 64;
}

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
	bt2shim::bindScoring(*sc_);
	ShimThread& t = g_t;
	ShimState& st = t.st[this];
	// the oriented read, rows rdi_..rdf_-1 (bytes 0..4 + Phred+33), and the window incl. the extra column of initRef
	st.seq.resize(rdlen); st.qual.resize(rdlen); st.win.resize(ncol + 1);
	for(size_t i = 0; i < rdlen; i++) { st.seq[i] = (uint8_t)(*rd_)[rdi_ + i]; st.qual[i] = (uint8_t)(*qu_)[rdi_ + i]; }
	for(size_t i = 0; i <= ncol; i++) st.win[i] = (uint8_t)rf_[rfi_ + i];
	bt2gpu_dp_job& job = st.job;
	job.rd_off = 0; job.rd_len = (uint32_t)rdlen; job.rf_off = 0; job.ncol = (uint32_t)ncol;
	job.minsc = (int32_t)minsc_; job.nceil = nceil_;
	job.refl = rect_->refl; job.triml = (uint32_t)rect_->triml;
	job.corel = (uint32_t)rect_->corel; job.corer = (uint32_t)rect_->corer;
	job.fw = fw_ ? 1 : 0; job.enable8 = (enable8_ && !readSse16_) ? 1 : 0;
	job.reserved[0] = job.reserved[1] = 0;
	const size_t cc = candCap(job, sc_->monotone);
	t.cands.resize(cc); st.alns.resize(kMaxAln); st.edits.resize(kEditCap);
	if(bt2gpu_dp_align_batch(bt2shim::ctx(), st.seq.data(), st.qual.data(), rdlen, st.win.data(), ncol + 1, &job, 1, kMaxAln,
	                         &st.res, t.cands.data(), cc, st.alns.data(), st.edits.data(), kEditCap))
		bt2shim::fail("bt2gpu_dp_align_batch");
	if(st.res.naln >= kMaxAln) { fprintf(stderr, "bt2gpu shim: more than %u alignments in one window\n", kMaxAln); throw 1; }
	bt2shim::note(bt2shim::C_ALIGN, 1);
	st.next = 0; st.minsc = minsc_;
	sse8succ_ = st.res.width == 8; sse16succ_ = st.res.width == 16;
	best = st.res.best;                      // INT64_MIN == MIN_I64
	if(best == MIN_I64 || best < minsc_) {   // aligner_sw.cpp:681-686
		if(dpLog_ != NULL) (*dpLog_) << ",0,0";
		return false;
	}
	if(dpLog_ != NULL) (*dpLog_) << ",1," << best;
	for(uint32_t i = 0; i < st.res.ncand; i++) {
		const bt2gpu_dp_cand& c = t.cands[st.res.cand_off + i];
		btncand_.push_back(DpBtCandidate(c.row, c.col, c.score));
	}
	return st.res.aligned != 0;
}

/**
 * SwAligner::nextAlignment (aligner_sw.cpp:737-1146): hands out the alignments of the candidate loop the GPU
 * already ran, in order, leaving `res`, cural_ and `rnd` exactly as the reference's call would.
 */
bool SwAligner::nextAlignment(SwResult& res, TAlScore minsc, RandomSource& rnd) {
	assert(initedRead() && initedRef());
	if(done()) { res.reset(); return false; }
	ShimState& st = g_t.st[this];
	if(minsc < st.minsc) { fprintf(stderr, "bt2gpu shim: nextAlignment minsc below initRef's\n"); throw 1; }
	const bt2gpu_dp_aln *a = (st.next < st.res.naln) ? &st.alns[st.res.aln_off + st.next] : NULL;
	uint32_t tries = a ? a->tries : st.res.tail_tries;
	if(minsc > st.minsc) {
		// The driver tightened minsc after reporting an alignment (aligner_sw_driver.cpp:1255-1284).  nextAlignment only
		// uses its minsc argument to drop candidates (aligner_sw.cpp:762); btncand_ is sorted by score, so from the
		// first candidate below the new minsc on, everything is dropped without a backtrace or an RNG draw.
		if(btncand_[cural_].score < minsc) { res.reset(); cural_ = btncand_.size(); return false; }
		if(a == NULL || btncand_[a->cural - 1].score < minsc) {
			// The GPU loop's next success (or its tail) lies beyond that point: the backtraces attempted before it are
			// those of the same problem with minsc raised from the start (candidates >= the new minsc come first and
			// see the same history) -- replay it and take its tail count.
			ShimThread& t = g_t;
			bt2gpu_dp_job j2 = st.job; bt2gpu_dp_result r2;
			j2.minsc = (int32_t)minsc; j2.enable8 = sse8succ_ ? 1 : 0;
			const size_t cc = candCap(j2, sc_->monotone);
			t.cands.resize(cc); t.alns2.resize(kMaxAln); t.edits2.resize(kEditCap);
			bt2shim::bindScoring(*sc_);
			if(bt2gpu_dp_align_batch(bt2shim::ctx(), st.seq.data(), st.qual.data(), j2.rd_len, st.win.data(), j2.ncol + 1, &j2, 1,
			                         kMaxAln, &r2, t.cands.data(), cc, t.alns2.data(), t.edits2.data(), kEditCap))
				bt2shim::fail("bt2gpu_dp_align_batch (raised minsc)");
			bt2shim::note(bt2shim::C_ALIGN, 1);
			if(r2.naln != st.next || (r2.width != 0 && r2.width != st.res.width)) {
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
	res.reset();
	if(a == NULL) { cural_ = btncand_.size(); return false; }
	const size_t rdlen = rdf_ - rdi_;
	const size_t trimBeg = fw_ ? a->trim5 : a->trim3;
	EList<Edit>& ned = res.alres.ned();
	for(uint32_t i = 0; i < a->nedits; i++) {
		const bt2gpu_edit& e = st.edits[a->edit_off + i];
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
	bt2shim::bindScoring(sc);
	ShimThread& t = g_t;
	std::vector<uint8_t>& seq = t.useq; std::vector<uint8_t>& qual = t.uqual; std::vector<uint8_t>& win = t.uwin;
	seq.resize(len); qual.resize(len); win.resize(len);
	const TRefOff rfi = coord.off();
	for(size_t i = 0; i < len; i++) {
		seq[i] = (uint8_t)rd[i]; qual[i] = (uint8_t)qu[i];
		const TRefOff p = rfi + (TRefOff)i;
		win[i] = (p < 0 || (size_t)p >= reflen) ? 4 : (uint8_t)refs.getBase(coord.ref(), (size_t)p);
	}
	bt2gpu_ungap_job j;
	j.rd_off = 0; j.rd_len = (uint32_t)len; j.rf_off = 0; j.minsc = (int32_t)minsc;
	j.nceil = sc.nCeil.f<int>((double)len);
	j.fw = coord.fw() ? 1 : 0; j.ohang = ohang ? 1 : 0; j.pad[0] = j.pad[1] = 0;
	j.off = rfi; j.reflen = (int64_t)reflen;
	bt2gpu_ungap_result r; size_t ne = 0;
	t.uedits.resize(len + 1);
	if(bt2gpu_ungapped_batch(bt2shim::ctx(), seq.data(), qual.data(), len, win.data(), len, &j, 1, &r, t.uedits.data(),
	                         len + 1, &ne))
		bt2shim::fail("bt2gpu_ungapped_batch");
	bt2shim::note(bt2shim::C_UNGAPPED, 1);
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
		const bt2gpu_edit& e = t.uedits[r.edit_off + i];
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
