// hotpath.cu — batch runtime: one chained pass of the hot path (include/bt2gpu.h,
// bt2gpu_hotpath_round0_batch), resident in HBM from the staged reads to the alignments.
// The arithmetic runs in the kernels of fm.cu and dp.cu; what is here is the glue the reference
// executes per read between those calls -- seed instantiation (aligner_seed.cpp:498-587),
// choosing a seed hit to extend, framing the window (dp_framer.cpp:81-129) and fetching it as
// masks from the 2-bit reference (aligner_sw.cpp:155-271, reference.cpp getStretch) -- written
// as small kernels over the whole batch, so that no intermediate (seed jobs, BW ranges, rows,
// offsets, windows, DP job table) ever crosses PCIe.  The host reads back three scalars on the
// way (seed count, DP job count, arena totals) to size the next launches.
#include "common.cuh"
#include "scan.cuh"
#include "frame.cuh"
#include <algorithm>
#include <cmath>
#include <chrono>

// 2-bit reference text on the device (BitPairReference::buf_, reference.h:180; `.4.bt2` layout:
// base i in bits 2*(i&15).. of word i>>4)
// plus the unambiguous stretches of the `.3.bt2` RefRecords (reference.h:59-103, reference.cpp:122-167): per record the
// sequence coordinate where its bases start (after its leading Ns), its length and its offset in the joined text; per
// sequence its record range and its length (Ns included) -- what BitPairReference::getStretch (reference.cpp:425-600)
// and Ebwt::joinedToTextOff (bt2_idx.cpp:54-124: rstarts triples, plen) work from.
struct bt2gpu_ref {
	int device = 0;
	uint64_t len = 0;
	bt2::DevBuf packed;
	uint32_t nrec = 0, nseq = 0, nfrag = 0;
	bt2::DevBuf rec_start, rec_len, rec_buf, seq_rec0, plen, rstarts;
};

namespace bt2 {

typedef unsigned long long u64;

struct HpState {
	DevBuf cnt, first, spart, tot, hitlist, sjobs, shits, xjobs, xres, pick, val, ppos, pk_read, pk_seed, rows, rdoff, offs, steps,
	       minsc, nceil, job_read, job_start, job_fw, job_tidx, stats;
	HostBuf h_tot, h_stats;
	cudaEvent_t ev[12] = {};
	// paired-end pass (bt2gpu_hotpath_pairs_batch): what pass 1 left behind, and the mate-pass scratch
	FmReadsDev rv = {nullptr, nullptr, nullptr};
	size_t last_npick = 0; uint32_t last_max_aln = 0;
	DevBuf anc, mtmp, mval, mpos, wval, wpos, minfo, mstats;
	HostBuf h_mstats;
	cudaEvent_t mev[6] = {};
};
// device scratch reused across calls; owned by the context (freed by bt2gpu_ctx_destroy), so it lives on the
// context's device and the paired pass finds what pass 1 left behind whichever host thread drives the context
static HpState &hp_state(bt2gpu_ctx *ctx) { if(!ctx->hp) ctx->hp = new HpState(); return *ctx->hp; }
void hp_state_free(HpState *st) {
	if(!st) return;
	DevBuf *bufs[] = { &st->cnt, &st->first, &st->spart, &st->tot, &st->hitlist, &st->sjobs, &st->shits, &st->xjobs, &st->xres, &st->pick,
	                   &st->val, &st->ppos, &st->pk_read, &st->pk_seed, &st->rows, &st->rdoff, &st->offs, &st->steps, &st->minsc, &st->nceil,
	                   &st->job_read, &st->job_start, &st->job_fw, &st->job_tidx, &st->stats, &st->anc, &st->mtmp, &st->mval, &st->mpos, &st->wval,
	                   &st->wpos, &st->minfo, &st->mstats };
	for(DevBuf *b : bufs) b->release();
	st->h_tot.release(); st->h_stats.release(); st->h_mstats.release();
	for(int i = 0; i < 12; i++) if(st->ev[i]) cudaEventDestroy(st->ev[i]);
	for(int i = 0; i < 6; i++) if(st->mev[i]) cudaEventDestroy(st->mev[i]);
	delete st;
}

// what the kernels need of a bt2gpu_ref, by value
struct RefView {
	const uint32_t *text2; u64 len;
	const uint32_t *rec_start, *rec_len, *rec_buf, *seq_rec0, *plen, *rstarts;
	uint32_t nfrag;
};
static RefView ref_view(const bt2gpu_ref *r) {
	RefView v;
	v.text2 = r->packed.as<uint32_t>(); v.len = r->len;
	v.rec_start = r->rec_start.as<uint32_t>(); v.rec_len = r->rec_len.as<uint32_t>(); v.rec_buf = r->rec_buf.as<uint32_t>();
	v.seq_rec0 = r->seq_rec0.as<uint32_t>(); v.plen = r->plen.as<uint32_t>(); v.rstarts = r->rstarts.as<uint32_t>();
	v.nfrag = r->nfrag;
	return v;
}

// BitPairReference::getStretch for one position (reference.cpp:425-600): base 0..3 of sequence `tidx` at offset p, 4 (N)
// in the gaps between its unambiguous stretches, after the last one and off either end.  Binary search over the
// sequence's records for the last stretch starting at or before p.
__device__ __forceinline__ uint32_t ref_code(const RefView &v, uint32_t tidx, long long p) {
	if(p < 0) return 4u;
	uint32_t lo = v.seq_rec0[tidx], hi = v.seq_rec0[tidx + 1];
	if(hi == lo || (u64)p < v.rec_start[lo]) return 4u;
	while(hi - lo > 1) {
		const uint32_t mid = lo + ((hi - lo) >> 1);
		if((u64)v.rec_start[mid] <= (u64)p) lo = mid; else hi = mid;
	}
	const u64 d = (u64)p - v.rec_start[lo];
	if(d >= v.rec_len[lo]) return 4u;
	const u64 q = (u64)v.rec_buf[lo] + d;
	return (__ldg(v.text2 + (q >> 4)) >> (2 * (q & 15))) & 3u;
}

enum { ST_SEED_BWOPS = 0, ST_SEED_SIDES, ST_SEED_HITS, ST_EXT_FMOPS, ST_LOC_STEPS, ST_ALIGNED, ST_SWEEP_BWOPS, ST_N };

// msIval.f<int>(len) for SIMPLE_FUNC_SQRT (simple_func.h:89-112), then max(.,1) (bt2_search.cpp:3396);
// explicit round-to-nearest steps so that the product is not contracted into an fma
// boost: both mates of a pair are searched -> (int)(interval * 1.2 + 0.5) (bt2_search.cpp:3392-3395)
__device__ __forceinline__ int seed_interval(double c, double l, uint32_t rdlen, int boost) {
	double v = __dadd_rn(c, __dmul_rn(l, sqrt((double)rdlen)));
	if(v < 0.0) v = 0.0;
	int iv = (int)v;
	if(boost) iv = (int)__dadd_rn(__dmul_rn((double)iv, 1.2), 0.5);
	return iv < 1 ? 1 : iv;
}

// ---- seed instantiation: count, then emit (fw seeds of a read, then its rc seeds) ----------
template<bool EMIT> __global__ void __launch_bounds__(256)
k_hp_seeds(const uint8_t *__restrict__ seq, const bt2gpu_read *__restrict__ reads, uint32_t nreads, int L, double ic, double il, int boost,
           uint32_t *__restrict__ cnt, const uint32_t *__restrict__ first, bt2gpu_seed_job *__restrict__ out)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= nreads) return;
	const uint32_t rdlen = reads[i].rd_len;
	const uint8_t *rd = seq + reads[i].rd_off;
	const int per = seed_interval(ic, il, rdlen, boost);
	const int sl = (int)min((uint32_t)L, rdlen);
	int nseeds = 1;
	if((int)rdlen > L) nseeds += ((int)rdlen - L) / per;
	uint32_t c = 0;
	bt2gpu_seed_job *o = EMIT ? out + first[i] : nullptr;
	const uint32_t half = EMIT ? cnt[i] / 2 : 0;
	int lastN = -1, scanned = 0;                      // position of the last N seen in rd[0, scanned)
	for(int k = 0; k < nseeds; k++) {
		const int depth = k * per, end = depth + sl;
		for(; scanned < end; scanned++) if(rd[scanned] > 3) lastN = scanned;
		if(lastN >= depth) continue;                  // window holds an N: the reference drops it (:556-566)
		if(EMIT) {
			bt2gpu_seed_job j; j.read = i; j.off = (uint16_t)depth; j.len = (uint8_t)sl; j.fw = 1;
			o[c] = j; j.fw = 0; o[half + c] = j;
		}
		c++;
	}
	if(!EMIT) cnt[i] = 2 * c;
}

// ---- every seed becomes an extend job (k_extend skips the empty ranges) --------------------
__global__ void __launch_bounds__(256)
k_hp_mkextend(const bt2gpu_seed_job *__restrict__ sj, const bt2gpu_seed_hit *__restrict__ sh, uint32_t n,
              bt2gpu_extend_job *__restrict__ xj, uint32_t *__restrict__ hitlist, uint32_t *__restrict__ nhit, u64 *__restrict__ stats)
{
	const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
	u64 bw = 0, sd = 0, hit = 0;
	if(k < n) {
		const bt2gpu_seed_job j = sj[k]; const bt2gpu_seed_hit h = sh[k];
		bt2gpu_extend_job x; memset(&x, 0, sizeof(x));
		x.read = j.read; x.topf = h.topf; x.botf = h.botf; x.topb = h.topb; x.botb = h.botb; x.off = j.off; x.len = j.len; x.fw = j.fw;
		xj[k] = x;
		bw = h.bwops; sd = h.sides; hit = h.botf > h.topf;
	}
	{   // the hits, compacted (order is irrelevant: results land at the seed's own slot)
		const unsigned bal = __ballot_sync(0xffffffffu, hit != 0);
		uint32_t base = 0;
		if((threadIdx.x & 31) == 0 && bal) base = atomicAdd(nhit, (uint32_t)__popc(bal));
		base = __shfl_sync(0xffffffffu, base, 0);
		if(hit) hitlist[base + __popc(bal & ((1u << (threadIdx.x & 31)) - 1u))] = k;
	}
	for(int o = 16; o; o >>= 1) {
		bw += __shfl_xor_sync(0xffffffffu, bw, o); sd += __shfl_xor_sync(0xffffffffu, sd, o); hit += __shfl_xor_sync(0xffffffffu, hit, o);
	}
	if((threadIdx.x & 31) == 0) {
		if(bw) atomicAdd(stats + ST_SEED_BWOPS, bw);
		if(sd) atomicAdd(stats + ST_SEED_SIDES, sd);
		if(hit) atomicAdd(stats + ST_SEED_HITS, hit);
	}
}

// ---- one hit per read: narrowest range, then longest exact extension, then first ----------
// val[i] = (1 << 32) | rdlen for a read that goes on to DP, 0 otherwise (scanned next)
__global__ void __launch_bounds__(256)
k_hp_pick(const bt2gpu_read *__restrict__ reads, uint32_t nreads, const uint32_t *__restrict__ first, const uint32_t *__restrict__ cnt,
          const bt2gpu_extend_job *__restrict__ xj, const bt2gpu_extend_result *__restrict__ xr,
          uint32_t *__restrict__ pick, u64 *__restrict__ val, u64 *__restrict__ stats)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	u64 fm = 0;
	if(i < nreads) {
		const uint32_t lo = first[i], hi = lo + cnt[i];
		uint32_t best = 0xffffffffu, bsz = 0, bex = 0;
		for(uint32_t e = lo; e < hi; e++) {
			const uint32_t sz = xj[e].botf - xj[e].topf;
			if(xj[e].botf <= xj[e].topf) continue;
			const bt2gpu_extend_result r = xr[e];
			fm += r.fmops;
			const uint32_t ex = (uint32_t)r.nlex + r.nrex;
			if(best == 0xffffffffu || sz < bsz || (sz == bsz && ex > bex)) { best = e; bsz = sz; bex = ex; }
		}
		pick[i] = best;
		val[i] = best == 0xffffffffu ? 0ull : ((1ull << 32) | reads[i].rd_len);
	}
	for(int o = 16; o; o >>= 1) fm += __shfl_xor_sync(0xffffffffu, fm, o);
	if((threadIdx.x & 31) == 0 && fm) atomicAdd(stats + ST_EXT_FMOPS, fm);
}

__global__ void __launch_bounds__(256)
k_hp_compact(uint32_t nreads, const uint32_t *__restrict__ pick, const u64 *__restrict__ pos, const bt2gpu_extend_job *__restrict__ xj,
             uint32_t *__restrict__ pk_read, uint32_t *__restrict__ pk_seed, uint32_t *__restrict__ rows, uint32_t *__restrict__ rdoff)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= nreads || pick[i] == 0xffffffffu) return;
	const uint32_t k = (uint32_t)(pos[i] >> 32);
	pk_read[k] = i; pk_seed[k] = pick[i]; rows[k] = xj[pick[i]].topf; rdoff[k] = (uint32_t)pos[i];
}

// ---- frame the window around the located seed and fetch it as masks; orient the read --------
// One warp per DP problem.  DynProgFramer::frameSeedExtensionRect with maxgap on both sides
// (dp_framer.cpp:81-129, frame.cuh): the oriented read is laid at text offset `start`, the window spans
// [start - 2*maxgap, start + rdlen + 2*maxgap] trimmed to the text, the core diagonals are [maxgap, 3*maxgap].
// Window bytes are 1 << base, 16 outside the text, plus one pad column (aligner_sw.cpp:155-271).
__global__ void __launch_bounds__(256)
k_hp_frame(const uint8_t *__restrict__ seq, const uint8_t *__restrict__ qual, const bt2gpu_read *__restrict__ reads,
           RefView rv, uint32_t npick, int mg,
           const uint32_t *__restrict__ pk_read, const uint32_t *__restrict__ pk_seed, const uint32_t *__restrict__ rdoff,
           const uint32_t *__restrict__ offs, const uint32_t *__restrict__ steps, const bt2gpu_extend_job *__restrict__ xj,
           const int32_t *__restrict__ minsc, const int32_t *__restrict__ nceil,
           uint8_t *__restrict__ dseq, uint8_t *__restrict__ dqual, uint8_t *__restrict__ dwin, bt2gpu_dp_job *__restrict__ dj,
           uint32_t *__restrict__ job_read, long long *__restrict__ job_start, uint8_t *__restrict__ job_fw,
           uint32_t *__restrict__ job_tidx, u64 *__restrict__ stats)
{
	const int lane = threadIdx.x & 31;
	const uint32_t k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	if(k >= npick) return;
	const bt2gpu_extend_job x = xj[pk_seed[k]];
	const uint32_t ri = pk_read[k], rdlen = reads[ri].rd_len;
	const uint8_t *rd = seq + reads[ri].rd_off, *qu = qual + reads[ri].rd_off;
	const uint32_t ro = rdoff[k], wo = ro + k * (uint32_t)(4 * mg + 1);
	uint8_t *os = dseq + ro, *oq = dqual + ro, *ow = dwin + wo;
	if(x.fw) for(uint32_t i = lane; i < rdlen; i += 32) { os[i] = rd[i]; oq[i] = qu[i]; }
	else for(uint32_t i = lane; i < rdlen; i += 32) { const uint8_t c = rd[rdlen - 1 - i]; os[i] = c > 3 ? 4 : (uint8_t)(3 - c); oq[i] = qu[rdlen - 1 - i]; }
	const long long seedoff = x.fw ? (long long)x.off : (long long)rdlen - x.off - x.len;   // in the oriented read
	// Ebwt::joinedToTextOff of the located seed hit (aligner_sw_driver.cpp:940; seed hits that straddle a fragment end
	// are kept outside eeMode): sequence, offset inside it, its length
	uint32_t tidx, toff, tlen; bool straddled;
	joined_to_text(rv.rstarts, rv.nfrag, rv.plen, (uint32_t)rv.len, true, x.len, offs[k], false, tidx, toff, tlen, straddled);
	const long long start = (long long)toff - seedoff;
	bt2gpu_rect rect;         // trimToRef_ = !gReportOverhangs = true (bt2_search.cpp:188); the window holds the located seed, so it is never empty
	frame_seed_ext(start, rdlen, (long long)tlen, (u64)mg, (u64)mg, nceil[ri], (u64)mg, true, rect);
	const long long refl = rect.refl;
	const uint32_t ncol = (uint32_t)(rect.refr - rect.refl + 1);
	for(uint32_t c = lane; c <= ncol; c += 32)
		ow[c] = (uint8_t)(1u << ref_code(rv, tidx, refl + c));         // 16 = N: sequence gaps and beyond its ends
	if(lane == 0) {
		bt2gpu_dp_job j; memset(&j, 0, sizeof(j));
		j.rd_off = ro; j.rd_len = rdlen; j.rf_off = wo; j.ncol = ncol;
		j.minsc = minsc[ri]; j.nceil = nceil[ri]; j.refl = refl; j.triml = (uint32_t)rect.triml;
		j.corel = (uint32_t)rect.corel; j.corer = (uint32_t)rect.corer; j.fw = x.fw; j.enable8 = 1;
		dj[k] = j;
		job_read[k] = ri; job_start[k] = start; job_fw[k] = x.fw; job_tidx[k] = tidx;
		atomicAdd(stats + ST_LOC_STEPS, (u64)steps[k]);
	}
}

__global__ void __launch_bounds__(256)
k_hp_final(const bt2gpu_dp_result *__restrict__ res, uint32_t npick, const bt2gpu_sweep_result *__restrict__ sweep, uint32_t nreads,
           u64 *__restrict__ stats)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	u64 al = 0, sw = 0;
	if(i < npick) al = res[i].naln > 0;
	if(i < nreads) sw = sweep[i].bwops;
	for(int o = 16; o; o >>= 1) { al += __shfl_xor_sync(0xffffffffu, al, o); sw += __shfl_xor_sync(0xffffffffu, sw, o); }
	if((threadIdx.x & 31) == 0) {
		if(al) atomicAdd(stats + ST_ALIGNED, al);
		if(sw) atomicAdd(stats + ST_SWEEP_BWOPS, sw);
	}
}

// 16 text bytes (values 0..3) -> one 2-bit word
__global__ void __launch_bounds__(256)
k_ref_pack(const uint8_t *__restrict__ bytes, u64 n, uint32_t *__restrict__ out, int *__restrict__ bad) {
	const u64 w = (u64)blockIdx.x * blockDim.x + threadIdx.x;
	if(w * 16 >= n) return;
	uint32_t v = 0;
	for(int i = 0; i < 16; i++) {
		const u64 p = w * 16 + i;
		if(p < n) { const uint8_t c = bytes[p]; if(c > 3) *bad = 1; v |= (uint32_t)(c & 3) << (2 * i); }
	}
	out[w] = v;
}

} // namespace bt2

using namespace bt2;

static int ref_create(int device, const uint8_t *text, const uint8_t *bitpairs, uint64_t text_len,
                      const bt2gpu_ref_record *recs, uint64_t nrec, bt2gpu_ref **out)
{
	if((!text && !bitpairs) || !out || !recs) BT2_FAIL("null argument");
	*out = nullptr;
	int n = 0;
	BT2_CUDA(cudaGetDeviceCount(&n));
	if(n <= 0) BT2_FAIL("no CUDA device: libbt2gpu has no CPU fallback");
	if(device < 0 || device >= n) BT2_FAIL("bad device ordinal");
	if(text_len == 0 || text_len > 0xfffffff0ull) BT2_FAIL("reference text length must be in 1..2^32-16 (small index)");
	if(nrec == 0 || nrec > 0x7fffffffull) BT2_FAIL("bad record count");
	if(!recs[0].first) BT2_FAIL("first reference record is not marked 'first' (reference.cpp:146-149)");
	// the tables BitPairReference's constructor derives from the records (reference.cpp:122-167) and the fragment
	// list Ebwt carries as rstarts/plen (bt2_io.cpp:240-322)
	std::vector<uint32_t> rstart(nrec), rlen(nrec), rbuf(nrec), seq0, plen, rst;
	uint64_t cumsz = 0, cumlen = 0;
	for(uint64_t i = 0; i < nrec; i++) {
		if(recs[i].first) { if(!seq0.empty()) plen.push_back((uint32_t)cumlen); seq0.push_back((uint32_t)i); cumlen = 0; }
		cumlen += recs[i].off;
		if(cumlen + recs[i].len > 0xffffffffull) BT2_FAIL("reference sequence longer than 2^32-1");
		rstart[i] = (uint32_t)cumlen; rlen[i] = recs[i].len; rbuf[i] = (uint32_t)cumsz;
		if(recs[i].len) { rst.push_back((uint32_t)cumsz); rst.push_back((uint32_t)seq0.size() - 1); rst.push_back((uint32_t)cumlen); }
		cumsz += recs[i].len; cumlen += recs[i].len;
	}
	plen.push_back((uint32_t)cumlen);
	seq0.push_back((uint32_t)nrec);
	if(cumsz != text_len) BT2_FAIL("reference records do not add up to the text length");
	BT2_CUDA(cudaSetDevice(device));
	bt2gpu_ref *r = new bt2gpu_ref();
	r->device = device; r->len = text_len;
	r->nrec = (uint32_t)nrec; r->nseq = (uint32_t)plen.size(); r->nfrag = (uint32_t)(rst.size() / 3);
	const uint64_t nwords = (text_len + 15) / 16;
	DevBuf stage, flag;
	const uint64_t CH = 64ull << 20;                     // bytes per staged chunk (a multiple of 16)
	int rc = 0, bad = 0;
	auto up = [&](DevBuf &b, const std::vector<uint32_t> &v) {
		if(b.reserve(v.size() * 4 + 16)) return 1;
		if(cudaMemcpy(b.p, v.data(), v.size() * 4, cudaMemcpyHostToDevice) != cudaSuccess) return fail("cudaMemcpy(reference records)", __FILE__, __LINE__);
		return 0;
	};
	if(up(r->rec_start, rstart) || up(r->rec_len, rlen) || up(r->rec_buf, rbuf) || up(r->seq_rec0, seq0) || up(r->plen, plen) ||
	   up(r->rstarts, rst)) rc = 1;
	if(!rc && r->packed.reserve(nwords * 4 + 16)) rc = 1;
	if(!rc && bitpairs) {                                  // `.4.bt2` bytes: base i at bits 2*(i&3) of byte i>>2 == our words
		if(cudaMemcpy(r->packed.p, bitpairs, (text_len + 3) / 4, cudaMemcpyHostToDevice) != cudaSuccess)
			rc = fail("cudaMemcpy(reference bit pairs)", __FILE__, __LINE__);
	} else if(!rc) {
		if(stage.reserve(std::min<uint64_t>(CH, text_len) + 16) || flag.reserve(16)) rc = 1;
		if(!rc && cudaMemset(flag.p, 0, 4) != cudaSuccess) rc = fail("cudaMemset", __FILE__, __LINE__);
		for(uint64_t off = 0; !rc && off < text_len; off += CH) {
			const uint64_t nb = std::min<uint64_t>(CH, text_len - off);
			if(cudaMemcpy(stage.p, text + off, nb, cudaMemcpyHostToDevice) != cudaSuccess) { rc = fail("cudaMemcpy(reference text)", __FILE__, __LINE__); break; }
			const uint64_t nw = (nb + 15) / 16;
			k_ref_pack<<<(unsigned)((nw + 255) / 256), 256>>>(stage.as<uint8_t>(), nb, r->packed.as<uint32_t>() + off / 16, flag.as<int>());
			if(cudaDeviceSynchronize() != cudaSuccess) { rc = fail("k_ref_pack", __FILE__, __LINE__); break; }
		}
		if(!rc && cudaMemcpy(&bad, flag.p, 4, cudaMemcpyDeviceToHost) != cudaSuccess) rc = fail("cudaMemcpy", __FILE__, __LINE__);
		stage.release(); flag.release();
		if(!rc && bad) rc = fail("reference text holds a byte outside 0..3", __FILE__, __LINE__);
	}
	if(rc) { bt2gpu_ref_destroy(r); return 1; }
	*out = r;
	return 0;
}

extern "C" int bt2gpu_ref_create(int device, const uint8_t *text, uint64_t text_len, bt2gpu_ref **out) {
	bt2gpu_ref_record one; memset(&one, 0, sizeof(one));
	one.off = 0; one.len = (uint32_t)text_len; one.first = 1;
	return ref_create(device, text, nullptr, text_len, &one, 1, out);
}

extern "C" int bt2gpu_ref_create_records(int device, const uint8_t *text, uint64_t text_len,
                                         const bt2gpu_ref_record *recs, uint64_t nrec, bt2gpu_ref **out) {
	return ref_create(device, text, nullptr, text_len, recs, nrec, out);
}

extern "C" int bt2gpu_ref_create_bitpairs(int device, const uint8_t *bitpairs, uint64_t text_len,
                                          const bt2gpu_ref_record *recs, uint64_t nrec, bt2gpu_ref **out) {
	return ref_create(device, nullptr, bitpairs, text_len, recs, nrec, out);
}

extern "C" void bt2gpu_ref_destroy(bt2gpu_ref *r) {
	if(!r) return;
	cudaSetDevice(r->device);
	r->packed.release();
	r->rec_start.release(); r->rec_len.release(); r->rec_buf.release(); r->seq_rec0.release(); r->plen.release(); r->rstarts.release();
	delete r;
}

extern "C" uint64_t bt2gpu_ref_bytes(const bt2gpu_ref *r) { return r ? ((r->len + 15) / 16) * 4 : 0; }

namespace bt2 { const bt2gpu_dp_result *dp_dev_results(bt2gpu_ctx *ctx); }

static int hotpath_round0(
	bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw, const bt2gpu_ref *ref,
	const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const bt2gpu_read *reads, size_t nreads, const int32_t *minsc, const int32_t *nceil,
	const bt2gpu_hotpath_params *P,
	bt2gpu_sweep_result *sweep, uint32_t *job_read, int64_t *job_start, uint8_t *job_fw, size_t *njobs_out,
	bt2gpu_dp_result *results, bt2gpu_dp_aln *alns, bt2gpu_edit *edits, size_t edit_cap, size_t *nedits_out,
	bt2gpu_hotpath_stats *S)
{
	if(!ctx || !fw || !bw || !ref || !seq_pool || !qual_pool || !reads || !minsc || !nceil || !P || !sweep ||
	   !job_read || !job_start || !job_fw || !njobs_out || !results || !alns || !edits || !nedits_out || !S)
		BT2_FAIL("null argument");
	if(P->seed_len < 1 || P->seed_len > 32) BT2_FAIL("seed length must be 1..32");
	if(P->maxgap < 0 || P->maxgap > 1000) BT2_FAIL("bad maxgap");
	if(!ctx->have_sc) BT2_FAIL("bt2gpu_scoring_set must be called first");
	if(ref->device != ctx->device) BT2_FAIL("reference text lives on another device");
	if(nreads > 0x7ffffff0u) BT2_FAIL("too many reads");
	memset(S, 0, sizeof(*S));
	*njobs_out = 0; *nedits_out = 0;
	if(nreads == 0) return 0;
	BT2_CUDA(cudaSetDevice(ctx->device));
	HpState &st = hp_state(ctx);
	st.last_npick = 0;
	cudaStream_t s = ctx->stream;
	for(int i = 0; i < 12; i++) if(!st.ev[i]) BT2_CUDA(cudaEventCreate(&st.ev[i]));
	auto tick = std::chrono::steady_clock::now();
	auto lap = [&](float &dst) {
		const auto now = std::chrono::steady_clock::now();
		dst = std::chrono::duration<float, std::milli>(now - tick).count(); tick = now;
	};
	const uint32_t max_aln = P->max_aln ? P->max_aln : 1;
	const uint32_t n32 = (uint32_t)nreads;
	const unsigned nbR = (n32 + 255) / 256;

	// ---- 1. reads to HBM, exactSweep, seed instantiation (count) ----
	FmReadsDev rv;
	cudaStream_t cs = ctx->cstream;
	if(fm_hp_reads_stage(ctx, seq_pool, qual_pool, pool_len, reads, nreads, &rv, cs)) return 1;
	uint64_t sum_len = 0;
	for(size_t i = 0; i < nreads; i++) sum_len += reads[i].rd_len;
	if(sum_len + (uint64_t)nreads * (4 * P->maxgap + 1) > 0x70000000ull) BT2_FAIL("DP pools exceed 32-bit offsets: split the batch");
	if(st.minsc.reserve(nreads * 4) || st.nceil.reserve(nreads * 4) || st.cnt.reserve(nreads * 4 + 16) || st.first.reserve(nreads * 4 + 16) ||
	   st.spart.reserve((scan_parts(nreads) + 2) * 8) || st.tot.reserve(64) || st.h_tot.reserve(64) || st.stats.reserve(ST_N * 8) ||
	   st.h_stats.reserve(ST_N * 8) || st.pick.reserve(nreads * 4) || st.val.reserve(nreads * 8) || st.ppos.reserve(nreads * 8)) return 1;
	BT2_CUDA(cudaMemcpyAsync(st.minsc.p, minsc, nreads * 4, cudaMemcpyHostToDevice, cs));
	BT2_CUDA(cudaMemcpyAsync(st.nceil.p, nceil, nreads * 4, cudaMemcpyHostToDevice, cs));
	BT2_CUDA(cudaEventRecord(ctx->cev[0], cs));        // qualities, minsc, nceil in HBM
	BT2_CUDA(cudaMemsetAsync(st.stats.p, 0, ST_N * 8, s));
	S->h2d_bytes += 2 * pool_len + nreads * (sizeof(bt2gpu_read) + 8);
	u64 *stats = st.stats.as<u64>();
	BT2_CUDA(cudaEventRecord(st.ev[0], s));
	bt2gpu_sweep_result *dsweep = nullptr;
	if(fm_hp_sweep(ctx, fw, P->mine_max ? P->mine_max : 2, &dsweep)) return 1;
	BT2_CUDA(cudaEventRecord(st.ev[1], s));
	uint32_t *cnt = st.cnt.as<uint32_t>(), *first = st.first.as<uint32_t>();
	k_hp_seeds<false><<<nbR, 256, 0, s>>>(rv.seq, rv.reads, n32, P->seed_len, P->ival_const, P->ival_coeff, P->pair_boost, cnt, nullptr, nullptr);
	ctx->launches++;
	if(dev_exclusive_scan<uint32_t>(ctx, cnt, first, st.spart.as<uint32_t>(), st.tot.as<uint32_t>(), nreads)) return 1;
	BT2_CUDA(cudaMemcpyAsync(st.h_tot.p, st.tot.p, 4, cudaMemcpyDeviceToHost, s));
	BT2_CUDA(cudaStreamWaitEvent(cs, st.ev[1], 0));       // sweep results go home under the kernels that follow
	BT2_CUDA(cudaMemcpyAsync(sweep, dsweep, nreads * sizeof(bt2gpu_sweep_result), cudaMemcpyDeviceToHost, cs));
	S->d2h_bytes += nreads * sizeof(bt2gpu_sweep_result);
	BT2_CUDA(cudaStreamSynchronize(s));
	const size_t nseed = *st.h_tot.as<uint32_t>();
	S->seeds = nseed;
	lap(S->wall_sweep);

	// ---- 2. seeds (emit), exact search, extend every hit, pick one per read ----
	size_t npick = 0, tot_rd = 0;
	if(nseed) {
		if(st.sjobs.reserve(nseed * sizeof(bt2gpu_seed_job)) || st.shits.reserve(nseed * sizeof(bt2gpu_seed_hit)) ||
		   st.xjobs.reserve(nseed * sizeof(bt2gpu_extend_job)) || st.xres.reserve(nseed * sizeof(bt2gpu_extend_result)) ||
		   st.hitlist.reserve(nseed * 4)) return 1;
		BT2_CUDA(cudaMemsetAsync(st.tot.as<uint32_t>() + 8, 0, 4, s));
		bt2gpu_seed_job *sj = st.sjobs.as<bt2gpu_seed_job>(); bt2gpu_seed_hit *sh = st.shits.as<bt2gpu_seed_hit>();
		bt2gpu_extend_job *xj = st.xjobs.as<bt2gpu_extend_job>(); bt2gpu_extend_result *xr = st.xres.as<bt2gpu_extend_result>();
		k_hp_seeds<true><<<nbR, 256, 0, s>>>(rv.seq, rv.reads, n32, P->seed_len, P->ival_const, P->ival_coeff, P->pair_boost, cnt, first, sj);
		BT2_CUDA(cudaEventRecord(st.ev[2], s));
		if(fm_hp_seed(ctx, fw, bw, sj, nseed, sh)) return 1;
		BT2_CUDA(cudaEventRecord(st.ev[3], s));
		k_hp_mkextend<<<(unsigned)((nseed + 255) / 256), 256, 0, s>>>(sj, sh, (uint32_t)nseed, xj, st.hitlist.as<uint32_t>(),
			st.tot.as<uint32_t>() + 8, stats);
		BT2_CUDA(cudaEventRecord(st.ev[4], s));
		if(fm_hp_extend(ctx, fw, bw, xj, nseed, xr, st.hitlist.as<uint32_t>(), st.tot.as<uint32_t>() + 8)) return 1;
		BT2_CUDA(cudaEventRecord(st.ev[5], s));
		k_hp_pick<<<nbR, 256, 0, s>>>(rv.reads, n32, first, cnt, xj, xr, st.pick.as<uint32_t>(), st.val.as<u64>(), stats);
		ctx->launches += 3;
		BT2_CUDA(cudaGetLastError());
		if(dev_exclusive_scan<u64>(ctx, st.val.as<u64>(), st.ppos.as<u64>(), st.spart.as<u64>(), st.tot.as<u64>() + 1, nreads)) return 1;
		BT2_CUDA(cudaMemcpyAsync(st.h_tot.as<u64>() + 1, st.tot.as<u64>() + 1, 8, cudaMemcpyDeviceToHost, s));
		BT2_CUDA(cudaStreamSynchronize(s));
		const u64 t = st.h_tot.as<u64>()[1];
		npick = (size_t)(t >> 32); tot_rd = (size_t)(t & 0xffffffffu);
	}
	lap(S->wall_seed);

	// ---- 3. locate, frame, DP ----
	if(npick) {
		const int mg = P->maxgap;
		const size_t tot_win = tot_rd + npick * (size_t)(4 * mg + 1);
		if(st.pk_read.reserve(npick * 4) || st.pk_seed.reserve(npick * 4) || st.rows.reserve(npick * 4) || st.rdoff.reserve(npick * 4) ||
		   st.offs.reserve(npick * 4) || st.steps.reserve(npick * 4) || st.job_read.reserve(npick * 4) || st.job_start.reserve(npick * 8) ||
		   st.job_fw.reserve(npick) || st.job_tidx.reserve(npick * 4)) return 1;
		bt2gpu_extend_job *xj = st.xjobs.as<bt2gpu_extend_job>();
		k_hp_compact<<<nbR, 256, 0, s>>>(n32, st.pick.as<uint32_t>(), st.ppos.as<u64>(), xj, st.pk_read.as<uint32_t>(),
			st.pk_seed.as<uint32_t>(), st.rows.as<uint32_t>(), st.rdoff.as<uint32_t>());
		BT2_CUDA(cudaEventRecord(st.ev[6], s));
		if(fm_hp_locate(ctx, fw, st.rows.as<uint32_t>(), npick, st.offs.as<uint32_t>(), st.steps.as<uint32_t>())) return 1;
		BT2_CUDA(cudaEventRecord(st.ev[7], s));
		DpDevPools pools;
		const size_t cpj = P->cand_per_job ? P->cand_per_job : (ctx->sc.monotone ? 256 : 4096);
		if(dp_dev_begin(ctx, npick, tot_rd, tot_win, max_aln, cpj * npick, edit_cap, &pools)) return 1;
		BT2_CUDA(cudaStreamWaitEvent(s, ctx->cev[0], 0));
		k_hp_frame<<<(unsigned)((npick * 32 + 255) / 256), 256, 0, s>>>(rv.seq, rv.qual, rv.reads, ref_view(ref),
			(uint32_t)npick, mg, st.pk_read.as<uint32_t>(), st.pk_seed.as<uint32_t>(), st.rdoff.as<uint32_t>(), st.offs.as<uint32_t>(),
			st.steps.as<uint32_t>(), xj, st.minsc.as<int32_t>(), st.nceil.as<int32_t>(), pools.seq, pools.qual, pools.win, pools.jobs,
			st.job_read.as<uint32_t>(), st.job_start.as<long long>(), st.job_fw.as<uint8_t>(), st.job_tidx.as<uint32_t>(), stats);
		ctx->launches += 2;
		BT2_CUDA(cudaGetLastError());
		if(dp_dev_finish(ctx)) return 1;
		lap(S->wall_frame);
		BT2_CUDA(cudaEventRecord(st.ev[8], s));
		if(bt2gpu_dp_run(ctx, 7)) return 1;
		BT2_CUDA(cudaEventRecord(st.ev[9], s));
	}
	k_hp_final<<<(unsigned)((std::max(nreads, npick) + 255) / 256), 256, 0, s>>>(npick ? dp_dev_results(ctx) : nullptr, (uint32_t)npick,
		dsweep, n32, stats);
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	BT2_CUDA(cudaEventRecord(st.ev[10], s));
	BT2_CUDA(cudaMemcpyAsync(st.h_stats.p, stats, ST_N * 8, cudaMemcpyDeviceToHost, s));
	size_t ne = 0, nc_total = 0;
	if(npick) {
		BT2_CUDA(cudaMemcpyAsync(job_read, st.job_read.p, npick * 4, cudaMemcpyDeviceToHost, s));
		BT2_CUDA(cudaMemcpyAsync(job_start, st.job_start.p, npick * 8, cudaMemcpyDeviceToHost, s));
		BT2_CUDA(cudaMemcpyAsync(job_fw, st.job_fw.p, npick, cudaMemcpyDeviceToHost, s));
		if(bt2gpu_dp_fetch(ctx, results, nullptr, alns, edits, &nc_total, &ne)) return 1;
		S->d2h_bytes += npick * (13 + sizeof(bt2gpu_dp_result) + max_aln * sizeof(bt2gpu_dp_aln)) + ne * sizeof(bt2gpu_edit);
	} else BT2_CUDA(cudaStreamSynchronize(s));
	BT2_CUDA(cudaStreamSynchronize(cs));
	lap(S->wall_dp);

	// ---- 4. counters and device times ----
	const u64 *hs = st.h_stats.as<u64>();
	S->seed_bwops = hs[ST_SEED_BWOPS]; S->seed_sides = hs[ST_SEED_SIDES]; S->seed_hits = hs[ST_SEED_HITS];
	S->extend_fmops = hs[ST_EXT_FMOPS]; S->locate_steps = hs[ST_LOC_STEPS]; S->aligned = hs[ST_ALIGNED];
	S->sweep_bwops = hs[ST_SWEEP_BWOPS];
	S->locate_rows = npick; S->dp_jobs = npick;
	S->dp_cells = npick ? bt2gpu_dp_staged_cells(ctx) : 0;
	S->bt_cells = npick ? bt2gpu_dp_walked_cells(ctx) : 0;
	S->dp_cands = nc_total;
	auto span = [&](int a, int b) -> float { float ms = 0.f; cudaEventElapsedTime(&ms, st.ev[a], st.ev[b]); return ms; };
	S->ms_sweep = span(0, 1);
	if(nseed) { S->ms_seed = span(2, 3); S->ms_extend = span(4, 5); }
	if(npick) {
		S->ms_locate = span(6, 7); S->ms_dp = span(8, 9);
		float ph[3] = {0, 0, 0};
		if(bt2gpu_dp_phase_ms(ctx, ph) == 0) { S->ms_dp_fill = ph[0]; S->ms_dp_gather = ph[1]; S->ms_dp_backtrace = ph[2]; }
	}
	S->ms_total = span(0, 10);
	S->ms_glue = S->ms_total - S->ms_sweep - S->ms_seed - S->ms_extend - S->ms_locate - S->ms_dp;
	*njobs_out = npick; *nedits_out = ne;
	st.rv = rv; st.last_npick = npick; st.last_max_aln = max_aln;
	return 0;
}

extern "C" int bt2gpu_hotpath_job_tidx(bt2gpu_ctx *ctx, uint32_t *job_tidx) {
	if(!ctx || !job_tidx) BT2_FAIL("null argument");
	if(!ctx->hp) BT2_FAIL("no batch has run on this context");
	const size_t n = ctx->hp->last_npick;
	if(n == 0) return 0;
	BT2_CUDA(cudaSetDevice(ctx->device));
	BT2_CUDA(cudaMemcpyAsync(job_tidx, ctx->hp->job_tidx.p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	BT2_CUDA(cudaStreamSynchronize(ctx->stream));
	return 0;
}

extern "C" int bt2gpu_hotpath_round0_batch(
	bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw, const bt2gpu_ref *ref,
	const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const bt2gpu_read *reads, size_t nreads, const int32_t *minsc, const int32_t *nceil,
	const bt2gpu_hotpath_params *P,
	bt2gpu_sweep_result *sweep, uint32_t *job_read, int64_t *job_start, uint8_t *job_fw, size_t *njobs_out,
	bt2gpu_dp_result *results, bt2gpu_dp_aln *alns, bt2gpu_edit *edits, size_t edit_cap, size_t *nedits_out,
	bt2gpu_hotpath_stats *S)
{
	const int rc = hotpath_round0(ctx, fw, bw, ref, seq_pool, qual_pool, pool_len, reads, nreads, minsc, nceil, P, sweep,
	                              job_read, job_start, job_fw, njobs_out, results, alns, edits, edit_cap, nedits_out, S);
	if(rc && ctx) {          // failed part-way: nothing may still be copying from / into the caller's buffers
		cudaStreamSynchronize(ctx->stream);
		if(ctx->cstream) cudaStreamSynchronize(ctx->cstream);
	}
	return rc;
}

// =======================================================================================
// Paired-end pass: anchor alignments -> mate windows -> DP -> pair classes
// =======================================================================================
namespace bt2 {
const bt2gpu_dp_aln *dp_dev_alns(bt2gpu_ctx *ctx);
const bt2gpu_edit *dp_dev_edits(bt2gpu_ctx *ctx);
int check_pe_policy(const bt2gpu_pe_policy *pol);

enum { MS_ALIGNED = 0, MS_CONCORDANT, MS_BADWIN, MS_N };

struct AnchorSum { long long refoff; uint32_t extent; int32_t score; };

// AlnRes::refExtent(): read characters aligned, minus those opposite a reference gap, plus the
// reference characters opposite a read gap
__device__ __forceinline__ uint32_t ref_extent(uint32_t rdlen, const bt2gpu_dp_aln &a, const bt2gpu_edit *__restrict__ edits) {
	int ext = (int)rdlen - (int)a.trim5 - (int)a.trim3;
	for(uint32_t e = 0; e < a.nedits; e++) {
		const uint8_t t = edits[a.edit_off + e].type;
		if(t == BT2GPU_EDIT_REF_GAP) ext--;
		else if(t == BT2GPU_EDIT_READ_GAP) ext++;
	}
	return (uint32_t)ext;
}

// One thread per pass-1 DP job: the first alignment of the job is the anchor; compute where the opposite
// mate may lie (otherMate) and the rectangle to search (frameFindMateRect).
// mval[k] = (1 << 32) | opposite read length, wval[k] = window bytes (ncol + 1), both 0 without a window.
__global__ void __launch_bounds__(256)
k_hp_anchor(const bt2gpu_dp_result *__restrict__ res, const bt2gpu_dp_aln *__restrict__ alns, const bt2gpu_edit *__restrict__ edits,
            uint32_t max_aln, uint32_t njobs, const uint32_t *__restrict__ job_read, const uint8_t *__restrict__ job_fw,
            const bt2gpu_read *__restrict__ reads, const int32_t *__restrict__ minsc, const int32_t *__restrict__ nceil,
            bt2gpu_pe_policy pol, uint32_t maxhalf, const uint32_t *__restrict__ job_tidx, const uint32_t *__restrict__ plen,
            int match, int rdo, int rde, int rfo, int rfe,
            AnchorSum *__restrict__ anc, bt2gpu_mate_frame *__restrict__ mtmp, u64 *__restrict__ mval, u64 *__restrict__ wval,
            u64 *__restrict__ mstats)
{
	const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
	if(k >= njobs) return;
	u64 mv = 0, wv = 0;
	if(res[k].naln > 0) {
		const bt2gpu_dp_aln a = alns[(size_t)k * max_aln];
		const uint32_t ri = job_read[k], mate = ri ^ 1u;
		const bool is1 = (ri & 1u) == 0, fw = job_fw[k] != 0;
		const uint32_t len_a = reads[ri].rd_len, len_o = reads[mate].rd_len;
		AnchorSum s; s.refoff = a.refoff; s.extent = ref_extent(len_a, a, edits); s.score = a.score;
		anc[k] = s;
		const long long ominsc = minsc[mate];
		const long long reflen = (long long)plen[job_tidx[k]];      // the mate is searched in the anchor's sequence
		const int ordg = max_gaps(ominsc, len_o, match, rdo, rde, false), orfg = max_gaps(ominsc, len_o, match, rfo, rfe, true);
		bt2gpu_mate_frame f; memset(&f, 0, sizeof(f));
		MateWin w;
		bool ok = len_o > 0 && ordg >= 0 && orfg >= 0 &&
		          other_mate(pol, is1, fw, a.refoff, (long long)len_o + ordg, is1 ? len_a : len_o, is1 ? len_o : len_a, w);
		if(ok) {
			f.found = 1; f.oleft = w.oleft; f.ofw = w.ofw; f.oll = w.oll; f.olr = w.olr; f.orl = w.orl; f.orr = w.orr;
			ok = frame_mate(!w.oleft, w.oll, w.olr, w.orl, w.orr, len_o, reflen, (u64)ordg, (u64)orfg, nceil[mate], maxhalf,
			                pol.trim_to_ref != 0, f.rect);
		}
		if(ok) {
			const long long ncol = f.rect.refr - f.rect.refl + 1;
			if(ncol > 60000 || f.rect.corer > 0xffffffffull) { atomicAdd(mstats + MS_BADWIN, 1ull); ok = false; }
			else { f.ok = 1; mv = (1ull << 32) | len_o; wv = (u64)ncol + 1; }
		}
		mtmp[k] = f;
	}
	mval[k] = mv; wval[k] = wv;
}

// One warp per anchor job that has a window: orient the opposite mate, fetch the window as masks
// (16 outside the text, plus the pad column, aligner_sw.cpp:155-271), write the DP job and its description.
__global__ void __launch_bounds__(256)
k_hp_mateframe(const uint8_t *__restrict__ seq, const uint8_t *__restrict__ qual, const bt2gpu_read *__restrict__ reads,
               RefView rv, uint32_t njobs, const uint32_t *__restrict__ job_read, const uint32_t *__restrict__ job_tidx,
               const int32_t *__restrict__ minsc, const int32_t *__restrict__ nceil, const AnchorSum *__restrict__ anc,
               const bt2gpu_mate_frame *__restrict__ mtmp, const u64 *__restrict__ mval, const u64 *__restrict__ mpos,
               const u64 *__restrict__ wpos, uint8_t *__restrict__ dseq, uint8_t *__restrict__ dqual, uint8_t *__restrict__ dwin,
               bt2gpu_dp_job *__restrict__ dj, bt2gpu_mate_info *__restrict__ info)
{
	const int lane = threadIdx.x & 31;
	const uint32_t k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	if(k >= njobs || mval[k] == 0) return;
	const uint32_t m = (uint32_t)(mpos[k] >> 32), ro = (uint32_t)mpos[k], wo = (uint32_t)wpos[k];
	const uint32_t mate = job_read[k] ^ 1u, rdlen = reads[mate].rd_len;
	const bt2gpu_mate_frame f = mtmp[k];
	const uint8_t *rd = seq + reads[mate].rd_off, *qu = qual + reads[mate].rd_off;
	uint8_t *os = dseq + ro, *oq = dqual + ro, *ow = dwin + wo;
	if(f.ofw) for(uint32_t i = lane; i < rdlen; i += 32) { os[i] = rd[i]; oq[i] = qu[i]; }
	else for(uint32_t i = lane; i < rdlen; i += 32) { const uint8_t c = rd[rdlen - 1 - i]; os[i] = c > 3 ? 4 : (uint8_t)(3 - c); oq[i] = qu[rdlen - 1 - i]; }
	const long long refl = f.rect.refl;
	const uint32_t ncol = (uint32_t)(f.rect.refr - f.rect.refl + 1);
	const uint32_t tidx = job_tidx[k];
	for(uint32_t c = lane; c <= ncol; c += 32)
		ow[c] = (uint8_t)(1u << ref_code(rv, tidx, refl + c));
	if(lane == 0) {
		bt2gpu_dp_job j; memset(&j, 0, sizeof(j));
		j.rd_off = ro; j.rd_len = rdlen; j.rf_off = wo; j.ncol = ncol;
		j.minsc = minsc[mate]; j.nceil = nceil[mate]; j.refl = refl; j.triml = (uint32_t)f.rect.triml;
		j.corel = (uint32_t)f.rect.corel; j.corer = (uint32_t)f.rect.corer; j.fw = f.ofw; j.enable8 = 1;
		dj[m] = j;
		bt2gpu_mate_info o; memset(&o, 0, sizeof(o));
		o.anchor_job = k; o.read = mate; o.ofw = f.ofw; o.oleft = f.oleft;
		o.anchor_refoff = anc[k].refoff; o.anchor_extent = anc[k].extent; o.rect = f.rect;
		info[m] = o;
	}
}

// One thread per mate job: peClassifyPair of the anchor alignment and the first mate alignment.
__global__ void __launch_bounds__(256)
k_hp_pairs(const bt2gpu_dp_result *__restrict__ res, const bt2gpu_dp_aln *__restrict__ alns, const bt2gpu_edit *__restrict__ edits,
           uint32_t max_aln, uint32_t nmates, const bt2gpu_read *__restrict__ reads, const uint8_t *__restrict__ job_fw,
           bt2gpu_pe_policy pol, bt2gpu_mate_info *__restrict__ info, u64 *__restrict__ mstats)
{
	const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
	u64 al = 0, cc = 0;
	if(m < nmates && res[m].naln > 0) {
		bt2gpu_mate_info o = info[m];
		const bt2gpu_dp_aln a = alns[(size_t)m * max_aln];
		const uint32_t ext = ref_extent(reads[o.read].rd_len, a, edits);
		const bool mate_is1 = (o.read & 1u) == 0, afw = job_fw[o.anchor_job] != 0;
		const long long off1 = mate_is1 ? a.refoff : o.anchor_refoff, off2 = mate_is1 ? o.anchor_refoff : a.refoff;
		const uint32_t len1 = mate_is1 ? ext : o.anchor_extent, len2 = mate_is1 ? o.anchor_extent : ext;
		const bool fw1 = mate_is1 ? (o.ofw != 0) : afw, fw2 = mate_is1 ? afw : (o.ofw != 0);
		const int cls = pe_classify(pol, off1, len1, fw1, off2, len2, fw2);
		info[m].cls = cls; info[m].mate_extent = ext;
		al = 1; cc = cls != BT2GPU_PE_ALS_DISCORD;
	}
	for(int o = 16; o; o >>= 1) { al += __shfl_xor_sync(0xffffffffu, al, o); cc += __shfl_xor_sync(0xffffffffu, cc, o); }
	if((threadIdx.x & 31) == 0) {
		if(al) atomicAdd(mstats + MS_ALIGNED, al);
		if(cc) atomicAdd(mstats + MS_CONCORDANT, cc);
	}
}

} // namespace bt2

static int hotpath_mates(
	bt2gpu_ctx *ctx, const bt2gpu_ref *ref, size_t nreads, const bt2gpu_hotpath_params *P, const bt2gpu_pe_policy *pol, uint32_t maxhalf,
	bt2gpu_mate_info *mates, size_t *nmates_out, bt2gpu_dp_result *mresults, bt2gpu_dp_aln *malns, bt2gpu_edit *medits,
	size_t medit_cap, size_t *nmedits_out, bt2gpu_pairs_stats *S)
{
	HpState &st = hp_state(ctx);
	cudaStream_t s = ctx->stream;
	const size_t npick = st.last_npick;
	const uint32_t max_aln = st.last_max_aln;
	if(npick == 0) return 0;
	for(int i = 0; i < 6; i++) if(!st.mev[i]) BT2_CUDA(cudaEventCreate(&st.mev[i]));
	if(st.anc.reserve(npick * sizeof(AnchorSum)) || st.mtmp.reserve(npick * sizeof(bt2gpu_mate_frame)) || st.mval.reserve(npick * 8) ||
	   st.mpos.reserve(npick * 8) || st.wval.reserve(npick * 8) || st.wpos.reserve(npick * 8) || st.minfo.reserve(npick * sizeof(bt2gpu_mate_info)) ||
	   st.mstats.reserve(MS_N * 8) || st.h_mstats.reserve(MS_N * 8) || st.spart.reserve((scan_parts(std::max(npick, nreads)) + 2) * 8)) return 1;
	u64 *mstats = st.mstats.as<u64>();
	BT2_CUDA(cudaMemsetAsync(mstats, 0, MS_N * 8, s));
	BT2_CUDA(cudaEventRecord(st.mev[0], s));
	const bt2gpu_scoring &sc = ctx->sc;
	const unsigned nbJ = (unsigned)((npick + 255) / 256);
	k_hp_anchor<<<nbJ, 256, 0, s>>>(dp_dev_results(ctx), dp_dev_alns(ctx), dp_dev_edits(ctx), max_aln, (uint32_t)npick,
		st.job_read.as<uint32_t>(), st.job_fw.as<uint8_t>(), st.rv.reads, st.minsc.as<int32_t>(), st.nceil.as<int32_t>(), *pol, maxhalf,
		st.job_tidx.as<uint32_t>(), ref->plen.as<uint32_t>(), sc.match_bonus, sc.rd_gap_open, sc.rd_gap_ext, sc.rf_gap_open, sc.rf_gap_ext,
		st.anc.as<AnchorSum>(), st.mtmp.as<bt2gpu_mate_frame>(), st.mval.as<u64>(), st.wval.as<u64>(), mstats);
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	if(dev_exclusive_scan<u64>(ctx, st.mval.as<u64>(), st.mpos.as<u64>(), st.spart.as<u64>(), st.tot.as<u64>() + 2, npick)) return 1;
	if(dev_exclusive_scan<u64>(ctx, st.wval.as<u64>(), st.wpos.as<u64>(), st.spart.as<u64>(), st.tot.as<u64>() + 3, npick)) return 1;
	BT2_CUDA(cudaMemcpyAsync(st.h_tot.as<u64>() + 2, st.tot.as<u64>() + 2, 16, cudaMemcpyDeviceToHost, s));
	BT2_CUDA(cudaMemcpyAsync(st.h_mstats.p, mstats, MS_N * 8, cudaMemcpyDeviceToHost, s));
	BT2_CUDA(cudaStreamSynchronize(s));
	if(st.h_mstats.as<u64>()[MS_BADWIN]) BT2_FAIL("mate window wider than 60000 columns: lower the maximum fragment length");
	const u64 t = st.h_tot.as<u64>()[2];
	const size_t nm = (size_t)(t >> 32), tot_rd = (size_t)(t & 0xffffffffu), tot_win = (size_t)st.h_tot.as<u64>()[3];
	*nmates_out = nm;
	S->mate_jobs = nm;
	if(nm == 0) return 0;
	DpDevPools pools;
	const size_t cpj = std::max<size_t>(P->cand_per_job ? P->cand_per_job : (ctx->sc.monotone ? 256 : 4096), ctx->sc.monotone ? 1024 : 4096);
	if(dp_dev_begin(ctx, nm, tot_rd, tot_win, max_aln, cpj * nm, medit_cap, &pools)) return 1;
	k_hp_mateframe<<<(unsigned)((npick * 32 + 255) / 256), 256, 0, s>>>(st.rv.seq, st.rv.qual, st.rv.reads, ref_view(ref),
		(uint32_t)npick, st.job_read.as<uint32_t>(), st.job_tidx.as<uint32_t>(), st.minsc.as<int32_t>(), st.nceil.as<int32_t>(), st.anc.as<AnchorSum>(),
		st.mtmp.as<bt2gpu_mate_frame>(), st.mval.as<u64>(), st.mpos.as<u64>(), st.wpos.as<u64>(), pools.seq, pools.qual, pools.win, pools.jobs,
		st.minfo.as<bt2gpu_mate_info>());
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	BT2_CUDA(cudaEventRecord(st.mev[1], s));
	if(dp_dev_finish(ctx)) return 1;
	BT2_CUDA(cudaEventRecord(st.mev[2], s));
	if(bt2gpu_dp_run(ctx, 7)) return 1;
	BT2_CUDA(cudaEventRecord(st.mev[3], s));
	k_hp_pairs<<<(unsigned)((nm + 255) / 256), 256, 0, s>>>(dp_dev_results(ctx), dp_dev_alns(ctx), dp_dev_edits(ctx), max_aln, (uint32_t)nm,
		st.rv.reads, st.job_fw.as<uint8_t>(), *pol, st.minfo.as<bt2gpu_mate_info>(), mstats);
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	BT2_CUDA(cudaEventRecord(st.mev[4], s));
	BT2_CUDA(cudaMemcpyAsync(st.h_mstats.p, mstats, MS_N * 8, cudaMemcpyDeviceToHost, s));
	BT2_CUDA(cudaMemcpyAsync(mates, st.minfo.p, nm * sizeof(bt2gpu_mate_info), cudaMemcpyDeviceToHost, s));
	size_t ne = 0, nc_total = 0;
	if(bt2gpu_dp_fetch(ctx, mresults, nullptr, malns, medits, &nc_total, &ne)) return 1;
	*nmedits_out = ne;
	const u64 *hs = st.h_mstats.as<u64>();
	S->mate_aligned = hs[MS_ALIGNED]; S->concordant = hs[MS_CONCORDANT];
	S->mate_cells = bt2gpu_dp_staged_cells(ctx); S->bt_cells = bt2gpu_dp_walked_cells(ctx); S->dp_cands = nc_total;
	S->d2h_bytes = nm * (sizeof(bt2gpu_mate_info) + sizeof(bt2gpu_dp_result) + max_aln * sizeof(bt2gpu_dp_aln)) + ne * sizeof(bt2gpu_edit);
	auto span = [&](int a, int b) -> float { float ms = 0.f; cudaEventElapsedTime(&ms, st.mev[a], st.mev[b]); return ms; };
	S->ms_frame = span(0, 2); S->ms_dp = span(2, 3); S->ms_total = span(0, 4);
	float ph[3] = {0, 0, 0};
	if(bt2gpu_dp_phase_ms(ctx, ph) == 0) { S->ms_dp_fill = ph[0]; S->ms_dp_gather = ph[1]; S->ms_dp_backtrace = ph[2]; }
	return 0;
}

extern "C" int bt2gpu_hotpath_pairs_batch(
	bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw, const bt2gpu_ref *ref,
	const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const bt2gpu_read *reads, size_t nreads, const int32_t *minsc, const int32_t *nceil,
	const bt2gpu_hotpath_params *P, const bt2gpu_pe_policy *pol, uint32_t maxhalf,
	bt2gpu_sweep_result *sweep, uint32_t *job_read, int64_t *job_start, uint8_t *job_fw, size_t *njobs_out,
	bt2gpu_dp_result *results, bt2gpu_dp_aln *alns, bt2gpu_edit *edits, size_t edit_cap, size_t *nedits_out,
	bt2gpu_hotpath_stats *S,
	bt2gpu_mate_info *mates, size_t *nmates_out,
	bt2gpu_dp_result *mresults, bt2gpu_dp_aln *malns, bt2gpu_edit *medits, size_t medit_cap, size_t *nmedits_out,
	bt2gpu_pairs_stats *PS)
{
	if(!mates || !nmates_out || !mresults || !malns || !medits || !nmedits_out || !PS) BT2_FAIL("null argument");
	if(nreads & 1) BT2_FAIL("paired batch needs an even number of reads (mate 1, mate 2, mate 1, ...)");
	if(check_pe_policy(pol)) return 1;
	memset(PS, 0, sizeof(*PS));
	*nmates_out = 0; *nmedits_out = 0;
	int rc = hotpath_round0(ctx, fw, bw, ref, seq_pool, qual_pool, pool_len, reads, nreads, minsc, nceil, P, sweep,
	                        job_read, job_start, job_fw, njobs_out, results, alns, edits, edit_cap, nedits_out, S);
	if(!rc) rc = hotpath_mates(ctx, ref, nreads, P, pol, maxhalf, mates, nmates_out, mresults, malns, medits, medit_cap, nmedits_out, PS);
	if(rc && ctx) {
		cudaStreamSynchronize(ctx->stream);
		if(ctx->cstream) cudaStreamSynchronize(ctx->cstream);
	}
	return rc;
}
