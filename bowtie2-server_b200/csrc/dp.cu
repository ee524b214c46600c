// dp.cu — Smith-Waterman fill / gather / backtrace kernels (sm_100a) and their C ABI.
//
// Replaces, behind include/bt2gpu.h, the reference's SwAligner::align + nextAlignment
// (aligner_sw.cpp:500-1146) and its four striped-SSE back ends
// (aligner_swsse_{ee,loc}_{u8,i16}.cpp).  The reference vectorises INSIDE one problem
// (Farrar stripes + lazy-F loop); this design is a B200 one:
//
//  * k_dp_fill: a group of 8 lanes owns one read x window problem; each lane owns two
//    "virtual lanes" packed into the low/high half of s16x2 registers (rows
//    [2l*S, 2l*S+S) and [(2l+1)*S, (2l+1)*S+S) of a 16*S-row stripe).  The 16 virtual lanes
//    run a skewed wavefront (virtual lane v is at column t-v at step t), so the vertical
//    F/H dependency never crosses the two halves of a register in the same step.  All
//    recurrences are DPX: VIADDMNMX.S16x2 (add + clamp-at-floor) and VIMNMX3.S16x2.  The
//    gap barrier (--gbar) is folded into per-row gap constants, the query profile lives in
//    shared memory.  Only H of every cell is streamed to HBM (8-byte stores in wavefront order: one byte
//    per cell on the u8 end-to-end path, a halfword otherwise): the backtrace never needs the VALUE of
//    an E or F cell, only whether it EQUALS a number it already knows, and that follows from H alone
//    (BtJob::e_is / f_is below).
//  * k_dp_gather_*: candidate cells + best/flag/width with the exact u8/i16 decision rules
//    of aligner_sw.cpp:516-621 (including the stripe-padding rows that leak into the local
//    column maxima), two passes + device scan, then a per-job bitonic sort in the order of
//    DpBtCandidate::operator< (aligner_sw_nuc.h:134).
//  * k_dp_backtrace: one warp per problem replays the nextAlignment() candidate loop and
//    the deterministic-priority walk of backtraceNucleotides* (ee_u8.cpp:1283-1877), 32
//    speculated diagonal cells per round trip to the matrices.
//
// Values are kept in the SCORE domain (the reference's raw value minus its offset) with the
// variant's floor: -255 for the u8 end-to-end path (exact saturation emulation), 0 for
// local, and -16000 for the i16 end-to-end path (the reference's floor of -65535 is
// unreachable for the supported read lengths; checked in stage()).
#include "common.cuh"
#include "scan.cuh"
#include <algorithm>

namespace bt2 {

static constexpr int   GL      = 8;       // lanes per problem group
static constexpr int   VL      = 16;      // virtual lanes per group
static constexpr int   BIG     = 16000;   // veto / padding penalty
static constexpr int   FL_I16  = -16000;  // floor of the 16-bit end-to-end domain
static constexpr int   NCH     = 6;       // profile rows: A C G T N pad

__constant__ int32_t c_mmpens[256];
__constant__ int32_t c_npens[256];
struct ScConst { int32_t match, rdo, rde, rfo, rfe, gapbar, monotone; };
__constant__ ScConst c_sc;

// Per-job data prepared on the host by stage().
struct DpAux {
	uint64_t mat_off;     // word offset of the E/F/H stream in the matrix arena
	uint64_t rep_off;     // byte offset of the reportedThrough bitmap
	uint32_t bound_off;   // word offset of the 2 stripe-boundary rows (2*ncol words)
	uint32_t escr_off;    // offset (edits) of the per-job edit scratch
	uint32_t S;           // packed registers per lane (stripe = 16*S rows)
	uint32_t nstripes;
	uint32_t T;           // steps per stripe = ncol + 16
	uint32_t RW;          // words per lane-step record (H only): roundup2(S), or roundup2(ceil(S/2)) for byte records
	int32_t  fl;          // floor of the score domain
	uint32_t pad;
};

__device__ __forceinline__ uint32_t pack2(int lo, int hi) {
	return (uint32_t)(lo & 0xffff) | ((uint32_t)(hi & 0xffff) << 16);
}
__device__ __forceinline__ int lo16(uint32_t w) { return (int)(int16_t)(w & 0xffff); }
__device__ __forceinline__ int hi16(uint32_t w) { return (int)(int16_t)(w >> 16); }

// Scoring::score (scoring.h:241) for profile row c (0..4 = A,C,G,T,N).
__device__ __forceinline__ int dev_score(int rdc, int c, int q) {
	q = q < 0 ? 0 : (q > 255 ? 255 : q);
	if(rdc > 3 || c > 3) return -c_npens[q];
	return rdc == c ? c_sc.match : -c_mmpens[q];
}
__device__ __forceinline__ int dev_score_mask(int rdc, int refm, int q) {
	q = q < 0 ? 0 : (q > 255 ? 255 : q);
	if(rdc > 3 || refm > 15) return -c_npens[q];
	return ((refm >> rdc) & 1) ? c_sc.match : -c_mmpens[q];
}
__device__ __forceinline__ int dev_mm(int rdc, int refm, int q) {
	q = q < 0 ? 0 : (q > 255 ? 255 : q);
	return (rdc > 3 || refm > 15) ? c_npens[q] : c_mmpens[q];
}

// ---------------------------------------------------------------------------------------
// Fill.  One 8-lane group per job; jobs of one S class per launch, taken in `order`.
// ---------------------------------------------------------------------------------------
// a[idx] for a register array (idx is not a compile-time constant): a tree of selects on the bits of idx
template<int N> __device__ __forceinline__ uint32_t pick_reg(const uint32_t *a, int idx) {
	uint32_t t[N];
#pragma unroll
	for(int i = 0; i < N; i++) t[i] = a[i];
#pragma unroll
	for(int bit = 0, n = N; n > 1; bit++, n = (n + 1) / 2) {
		const bool b = (idx >> bit) & 1;
#pragma unroll
		for(int i = 0; i < (n + 1) / 2; i++) t[i] = (b && 2 * i + 1 < n) ? t[2 * i + 1] : t[2 * i];
	}
	return t[0];
}

template<int S>
__global__ void __launch_bounds__(128)
k_dp_fill(const uint8_t *__restrict__ seq, const uint8_t *__restrict__ qual,
          const uint8_t *__restrict__ win, const bt2gpu_dp_job *__restrict__ jobs,
          const DpAux *__restrict__ aux, const uint32_t *__restrict__ order, uint32_t norder,
          uint32_t *__restrict__ mat, uint32_t *__restrict__ bound)
{
	constexpr int SP = (S + 3) & ~3;          // profile row stride (words)
	constexpr int RW = (S + 1) & ~1;          // record words (s16 H values)
	extern __shared__ uint32_t smem[];
	const int lane  = threadIdx.x & 31;
	const int l     = lane & (GL - 1);
	const int gInW  = lane >> 3;
	const int gInB  = (threadIdx.x >> 5) * 4 + gInW;
	const unsigned gmask = 0xffu << (gInW * 8);
	uint32_t *prof = smem + (size_t)gInB * (NCH * GL * SP);
	const uint32_t ngroups = gridDim.x * (blockDim.x / GL);
	for(uint32_t oi = blockIdx.x * (blockDim.x / GL) + gInB; oi < norder; oi += ngroups) {
		const uint32_t ji = order[oi];
		const bt2gpu_dp_job jb = jobs[ji];
		const DpAux ax = aux[ji];
		const int nrow = (int)jb.rd_len, ncol = (int)jb.ncol;
		const uint8_t *rd = seq + jb.rd_off, *qu = qual + jb.rd_off, *rf = win + jb.rf_off;
		const int fl = ax.fl;
		const uint32_t FLv = pack2(fl, fl);
		const uint32_t nrde = pack2(-c_sc.rde, -c_sc.rde);
		const int T = (int)ax.T;
		uint32_t *bnd0 = bound + ax.bound_off, *bnd1 = bnd0 + ncol;
		for(int st = 0; st < (int)ax.nstripes; st++) {
			const int row0 = st * (VL * S) + (2 * l) * S;     // first row of my low half
			// --- per-row constants and query profile of this stripe ---
			uint32_t nrfo[S], nrfe[S], nrdo[S];
			__syncwarp(gmask);
#pragma unroll
			for(int k = 0; k < S; k++) {
				int rl = row0 + k, rh = row0 + S + k;
				bool bl = rl < nrow && (rl < c_sc.gapbar || (nrow - rl - 1) < c_sc.gapbar);
				bool bh = rh < nrow && (rh < c_sc.gapbar || (nrow - rh - 1) < c_sc.gapbar);
				nrfo[k] = pack2(bl ? -BIG : -c_sc.rfo, bh ? -BIG : -c_sc.rfo);
				nrfe[k] = pack2(bl ? -BIG : -c_sc.rfe, bh ? -BIG : -c_sc.rfe);
				nrdo[k] = pack2(bl ? -BIG : -c_sc.rdo, bh ? -BIG : -c_sc.rdo);
				int cl = rl < nrow ? rd[rl] : 4, ch = rh < nrow ? rd[rh] : 4;
				int ql = rl < nrow ? (int)qu[rl] - 33 : 0, qh = rh < nrow ? (int)qu[rh] - 33 : 0;
#pragma unroll
				for(int c = 0; c < 5; c++) {
					int sl = rl < nrow ? dev_score(cl, c, ql) : 0;
					int sh = rh < nrow ? dev_score(ch, c, qh) : 0;
					prof[(c * GL + l) * SP + k] = pack2(sl, sh);
				}
				prof[(5 * GL + l) * SP + k] = pack2(-BIG, -BIG);
			}
			__syncwarp(gmask);
			// --- state ---
			uint32_t Hprev[S], E[S], Pprev[S];
#pragma unroll
			for(int k = 0; k < S; k++) { Hprev[k] = FLv; E[k] = FLv; Pprev[k] = pack2(-BIG, -BIG); }
			uint32_t Hbot = FLv, Fbot = FLv, HtopPrev = FLv;
			const uint32_t *bin = (st & 1) ? bnd1 : bnd0;      // written by stripe st-1
			uint32_t *bout = (st & 1) ? bnd0 : bnd1;
			const bool lastStripe = (st + 1 == (int)ax.nstripes);
			// the matrix's last row (end-to-end candidates, gatherCellsNucleotidesEnd2End*) lives in one half of one
			// lane of the last stripe; that lane copies its H values, column by column, into the boundary buffer the
			// last stripe leaves free, so the gather reads one contiguous row instead of one sector per column
			const int rlast = nrow - 1 - st * (VL * S);
			const int vlast = lastStripe ? rlast / S : -1, klast = lastStripe ? rlast - vlast * S : 0;
			const bool ownLast = lastStripe && (vlast >> 1) == l;
			// top boundary of the matrix: H(-1,j) = 0 in both modes (0xff / 0x7fff raw end-to-end,
			// 0 local), F(-1,j) = floor  (ee_u8.cpp:920-927, loc_u8.cpp:1066-1075)
			const uint32_t top0 = pack2(0, fl);
			if(l == 0) HtopPrev = (st == 0) ? pack2(0, fl) : FLv;
			// u8 end-to-end jobs (floor -255) store E/F/H as bytes: every value is in [-255, 0], so its low
			// byte identifies it (0 -> 0, b -> b - 256); halves the matrix stream of the headline path
			const bool rec8 = (fl == -255);
			const int rwj = (int)ax.RW;
			uint32_t *rec_base = mat + ax.mat_off + (size_t)st * T * GL * rwj + l * 2;
			for(int t = 0; t < T; t++) {
				const int col_lo = t - 2 * l;
				int c = 5;
				if(col_lo >= 0 && col_lo < ncol) c = __ffs((int)rf[col_lo]) - 1;
				uint32_t P[S];
				{
					const uint4 *pp = reinterpret_cast<const uint4*>(prof + (c * GL + l) * SP);
#pragma unroll
					for(int q = 0; q < SP / 4; q++) {
						uint4 v = pp[q];
						if(4 * q + 0 < S) P[4 * q + 0] = v.x;
						if(4 * q + 1 < S) P[4 * q + 1] = v.y;
						if(4 * q + 2 < S) P[4 * q + 2] = v.z;
						if(4 * q + 3 < S) P[4 * q + 3] = v.w;
					}
				}
				// bottom row of the previous step: high halves go down to lane l+1 / next stripe
				const uint32_t X = __byte_perm(Hbot, Fbot, 0x7632);
				uint32_t recv = __shfl_up_sync(gmask, X, 1, GL);
				if(l == GL - 1 && !lastStripe) {
					int cb = t - 2 * (GL - 1) - 2;           // column X belongs to
					if(cb >= 0 && cb < ncol) bout[cb] = X;
				}
				if(l == 0) {
					if(st == 0) recv = top0;
					else recv = (col_lo >= 0 && col_lo < ncol) ? __ldcg(bin + col_lo) : FLv;
				}
				const uint32_t Htop = __byte_perm(recv, Hbot, 0x5410);
				const uint32_t Ftop = __byte_perm(recv, Fbot, 0x5432);
				uint32_t hdiag = HtopPrev; HtopPrev = Htop;
				uint32_t Hup = Htop, Fup = Ftop;
#pragma unroll
				for(int k = 0; k < S; k++) {
					const uint32_t sck = __byte_perm(P[k], Pprev[k], 0x7610);
					Pprev[k] = P[k];
					const uint32_t F = __viaddmax_s16x2(Fup, nrfe[k], __viaddmax_s16x2(Hup, nrfo[k], FLv));
					const uint32_t d = __viaddmax_s16x2(hdiag, sck, FLv);
					const uint32_t e = E[k];
					const uint32_t H = __vimax3_s16x2(d, e, F);
					E[k] = __viaddmax_s16x2(e, nrde, __viaddmax_s16x2(H, nrdo[k], FLv));
					hdiag = Hprev[k]; Hprev[k] = H; Hup = H; Fup = F;
				}
				Hbot = Hup; Fbot = Fup;
				if(ownLast) {
					const int cl = t - vlast;
					if(cl >= 0 && cl < ncol) {
						const uint32_t hw = pick_reg<S>(Hprev, klast);
						bout[cl] = (uint32_t)((vlast & 1) ? hi16(hw) : lo16(hw));
					}
				}
				// H of this step (now in Hprev[]) goes to the matrix stream, 8 bytes per store: chunk q of the 8 lanes
				// is one contiguous 64-byte run (two full sectors)
				if(rec8) {
					// byte records: word k / 2 holds the low bytes of [lo(k), hi(k), lo(k+1), hi(k+1)] (one PRMT per word)
					constexpr int SH = (S + 1) / 2;
					constexpr int RW8 = (SH + 1) & ~1;
					uint32_t r8[RW8];
#pragma unroll
					for(int kp = 0; kp < SH; kp++) {
						const uint32_t a = Hprev[2 * kp];
						const uint32_t b = (2 * kp + 1 < S) ? Hprev[2 * kp + 1] : 0u;
						r8[kp] = __byte_perm(a, b, 0x6420);
					}
#pragma unroll
					for(int k = SH; k < RW8; k++) r8[k] = 0;
					uint2 *dst = reinterpret_cast<uint2*>(rec_base + (size_t)t * GL * RW8);
#pragma unroll
					for(int q = 0; q < RW8 / 2; q++) dst[q * GL] = make_uint2(r8[2 * q], r8[2 * q + 1]);
				} else {
					uint2 *dst = reinterpret_cast<uint2*>(rec_base + (size_t)t * GL * RW);
#pragma unroll
					for(int q = 0; q < RW / 2; q++)
						dst[q * GL] = make_uint2(Hprev[2 * q], (2 * q + 1 < S) ? Hprev[2 * q + 1] : 0u);
				}
			}
			if(!lastStripe) __threadfence();   // boundary row visible to lane 0 of the next stripe
		}
	}
}

// ---------------------------------------------------------------------------------------
// Matrix access for gather / backtrace: H(row, col).
// ---------------------------------------------------------------------------------------
// A DP row resolved once: its step offset (stripe * T + virtual lane) and the byte offset of its
// H element inside a lane-step record; a cell is then one byte/halfword load at
// base + (tv + col) * record bytes + off.
struct RowRef { int tv; int off; };
struct MatView {
	const uint32_t *base; int S, T, RW;
	uint32_t invS, invRS; int stride;     // ceil(2^32 / S), ceil(2^32 / (16 S)), record bytes of one step (GL * RW * 4)
	bool rec8;                            // byte records (u8 end-to-end path)
	__device__ __forceinline__ void init(const uint32_t *b, int S_, int T_, int RW_, bool rec8_) {
		base = b; S = S_; T = T_; RW = RW_; rec8 = rec8_;
		invS = 0xffffffffu / (uint32_t)S + 1u; invRS = 0xffffffffu / (uint32_t)(VL * S) + 1u;
		stride = GL * RW * 4;
	}
	// row -> (stripe, virtual lane, register) by multiply-high: exact for rows < 2^16
	__device__ __forceinline__ RowRef rowref(int row) const {
		const int st = (int)__umulhi((uint32_t)row, invRS), r = row - st * (VL * S);
		const int v = (int)__umulhi((uint32_t)r, invS), k = r - v * S;
		RowRef rr; rr.tv = st * T + v;
		// word idx of the lane's record, in 8-byte chunks interleaved over the 8 lanes
		const int idx = rec8 ? (k >> 1) : k;
		const int sub = rec8 ? (((k & 1) << 1) | (v & 1)) : ((v & 1) << 1);
		rr.off = (idx >> 1) * (GL * 8) + (v >> 1) * 8 + (idx & 1) * 4 + sub;
		return rr;
	}
	// H(row, col)
	__device__ __forceinline__ int at(const RowRef &rr, int col) const {
		// a job's matrix is far below 4 GiB (dp_stage bounds rows and columns): 32-bit offset, one IMAD
		const unsigned char *p = reinterpret_cast<const unsigned char*>(base) + (uint32_t)((rr.tv + col) * stride + rr.off);
		if(rec8) { const int b = (int)__ldcg(p); return b ? b - 256 : 0; }
		return (int)__ldcg(reinterpret_cast<const short*>(p));
	}
	__device__ __forceinline__ int get(int row, int col) const { return at(rowref(row), col); }
};

__device__ __forceinline__ uint64_t cand_key(int score, int row, int col) {
	// ascending order of this key == DpBtCandidate::operator< (score desc, row desc, col desc)
	uint64_t k = ((uint64_t)(uint32_t)(score + 32768) << 32) | ((uint64_t)row << 16) | (uint64_t)col;
	return ~k;
}

// ---------------------------------------------------------------------------------------
// Gather.  One warp per job.  EMIT=false: best/flag/width/ncand (SwAligner::align,
// aligner_sw.cpp:500-729 + gatherCells*); EMIT=true: write candidate keys.
// ---------------------------------------------------------------------------------------
template<bool EMIT>
__global__ void __launch_bounds__(128)
k_dp_gather(const uint8_t *__restrict__ seq, const uint8_t *__restrict__ qual,
            const uint8_t *__restrict__ win, const bt2gpu_dp_job *__restrict__ jobs,
            const DpAux *__restrict__ aux, uint32_t job0, uint32_t njobs,
            const uint32_t *__restrict__ mat, bt2gpu_dp_result *__restrict__ res,
            uint32_t *__restrict__ counts, const uint32_t *__restrict__ cand_off,
            uint64_t *__restrict__ keys, uint32_t *__restrict__ fillptr, int maxcol)
{
	extern __shared__ int gsm[];
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const uint32_t ji = job0 + blockIdx.x * (blockDim.x >> 5) + wib;
	if(ji >= job0 + njobs) return;
	const bt2gpu_dp_job jb = jobs[ji];
	const DpAux ax = aux[ji];
	const int nrow = (int)jb.rd_len, ncol = (int)jb.ncol;
	const int minsc = jb.minsc;
	MatView mv; mv.init(mat + ax.mat_off, (int)ax.S, (int)ax.T, (int)ax.RW, ax.fl == -255);
	const uint8_t *rd = seq + jb.rd_off, *qu = qual + jb.rd_off, *rf = win + jb.rf_off;
	const bool local = !c_sc.monotone;
	if(EMIT && res[ji].ncand == 0) return;
	const uint32_t coff = EMIT ? cand_off[ji] : 0;
	if(!local) {
		// gatherCellsNucleotidesEnd2EndSse* (ee_u8.cpp:1176-1208): last row, score >= minsc
		int best = INT32_MIN, cnt = 0;
		for(int j = lane; j < ncol; j += 32) {
			int h = mv.get(nrow - 1, j);
			best = max(best, h);
			if(h >= minsc) {
				if(EMIT) { uint32_t p = atomicAdd(fillptr + ji, 1u); keys[coff + p] = cand_key(h, nrow - 1, j); }
				cnt++;
			}
		}
		if(EMIT) return;
		for(int o = 16; o; o >>= 1) {
			best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
			cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
		}
		if(lane == 0) {
			bt2gpu_dp_result r; memset(&r, 0, sizeof(r));
			const bool w8 = (ax.fl == -255);
			int64_t b = best; bool ok = true;
			if(ncol <= 0) { b = INT64_MIN; ok = false; }                 // !found
			else if(best < minsc) ok = false;                            // flag -1, returns score
			else if(w8 && best == -255) { b = INT64_MIN; ok = false; }   // flag -2 (ee_u8.cpp:1137)
			r.best = b; r.width = ok ? (w8 ? 8 : 16) : 0;
			r.ncand = ok ? (uint32_t)cnt : 0; r.aligned = (ok && cnt > 0) ? 1 : 0;
			res[ji] = r; counts[ji - job0] = r.ncand;
		}
		return;
	}
	// ---- local mode ----
	const int P8 = ((nrow + 7) >> 3) << 3, P16 = ((nrow + 15) >> 4) << 4;
	int *cmA = gsm + (size_t)wib * 2 * maxcol, *cmB = cmA + maxcol;
	if(!EMIT) { for(int j = lane; j < ncol; j += 32) { cmA[j] = 0; cmB[j] = 0; } __syncwarp(); }
	const uint64_t bonus = (uint64_t)c_sc.match;
	const uint64_t minrow64 = (((uint64_t)(int64_t)minsc + bonus - 1) / bonus) - 1;   // loc_u8.cpp:1408
	const int S = mv.S, T = mv.T, RS = VL * S;
	const int nrec = (int)ax.nstripes * T * GL;
	int cnt = 0;
	for(int rix = lane; rix < nrec; rix += 32) {
		const int l = rix & (GL - 1), t = (rix >> 3) % T, st = (rix >> 3) / T;
		const uint32_t *rp = mv.base + (size_t)(rix >> 3) * GL * mv.RW;   // whole records, word-addressed (s16 layout: local mode)
		// the lane's H words 0..S-1, an 8-byte chunk (two words) per load: the 8 lanes' chunks of one index are one
		// contiguous 64-byte run
		for(int ch = 0; ch < (S + 1) >> 1; ch++) {
			const uint2 q2 = __ldcg(reinterpret_cast<const uint2*>(rp + (ch * GL + l) * 2));
			const uint32_t qw[2] = {q2.x, q2.y};
#pragma unroll
			for(int jw = 0; jw < 2; jw++) {
				const int k = 2 * ch + jw;
				if(k >= S) continue;
				const uint32_t w = qw[jw];
#pragma unroll
				for(int half = 0; half < 2; half++) {
					const int v = 2 * l + half;
					const int row = st * RS + v * S + k, col = t - v;
					if(col < 0 || col >= ncol || row >= P16) continue;
					const int h = half ? hi16(w) : lo16(w);
					if(!EMIT) { if(h > 0) atomicMax(row < P8 ? &cmA[col] : &cmB[col], h); }
					if(row < nrow && h >= minsc && (uint64_t)row >= minrow64) {
						const int readc = rd[row], refc = rf[col];
						const bool match = (refc & (1 << readc)) != 0;
						bool matchSucc = false;
						if(row < nrow - 1) matchSucc = ((int)rf[col + 1] & (1 << (int)rd[row + 1])) != 0;
						if(match && !matchSucc) {
							if(EMIT) { uint32_t p = atomicAdd(fillptr + ji, 1u); keys[coff + p] = cand_key(h, row, col); }
							cnt++;
						}
					}
				}
			}
		}
	}
	if(EMIT) return;
	// bias of the u8 profile (loc_u8.cpp:98-110): largest penalty over the read
	int bias = 0;
	for(int i = lane; i < nrow; i += 32) {
		int q = (int)qu[i] - 33; q = q < 0 ? 0 : (q > 255 ? 255 : q);
		int pen = c_npens[q];
		if(rd[i] < 4) pen = max(pen, c_mmpens[q]);
		bias = max(bias, pen);
	}
	for(int o = 16; o; o >>= 1) {
		cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
		bias = max(bias, __shfl_xor_sync(0xffffffffu, bias, o));
	}
	__syncwarp();
	if(lane == 0) {
		bt2gpu_dp_result r; memset(&r, 0, sizeof(r));
		int flag = -2; int64_t best = INT64_MIN; int width = 0;
		if(jb.enable8) {                                   // alignNucleotidesLocalSseU8
			int vmax = 0; bool sat = false;
			for(int j = 0; j < ncol; j++) {
				const int cm = max(cmA[j], cmB[j]);
				vmax = max(vmax, cm);
				if(cm + bias >= 255) { sat = true; break; }
				if(cm < minsc && (int64_t)cm + (int64_t)(ncol - j - 1) * c_sc.match < minsc) break;
			}
			if(sat || vmax + bias >= 255) flag = -2;
			else if(vmax == 0 || vmax < minsc) { flag = -1; best = vmax; }
			else { flag = 0; best = vmax; width = 8; }
		}
		if(flag == -2) {                                   // alignNucleotidesLocalSseI16
			int vmax = 0;
			for(int j = 0; j < ncol; j++) {
				const int cm = cmA[j];
				vmax = max(vmax, cm);
				if(cm < minsc && (int64_t)cm + (int64_t)(ncol - j - 1) * c_sc.match < minsc) break;
			}
			if(vmax == 0) { flag = -1; best = INT64_MIN; }
			else if(vmax < minsc) { flag = -1; best = vmax; }
			else { flag = 0; best = vmax; width = 16; }
		}
		const bool ok = (flag == 0) && best != INT64_MIN && best >= minsc;
		r.best = best; r.width = width;
		r.ncand = ok ? (uint32_t)cnt : 0; r.aligned = (ok && cnt > 0) ? 1 : 0;
		res[ji] = r; counts[ji - job0] = r.ncand;
	}
}

// Exclusive scan of per-job candidate counts (one block); also resets the fill pointers.
__global__ void k_dp_scan(const uint32_t *__restrict__ counts, uint32_t n, uint32_t job0,
                          uint32_t *__restrict__ cand_off, uint32_t *__restrict__ fillptr,
                          bt2gpu_dp_result *__restrict__ res, unsigned long long *__restrict__ total,
                          unsigned long long cap, int *__restrict__ errflag)
{
	__shared__ unsigned long long warpsum[32];
	__shared__ unsigned long long carry;
	if(threadIdx.x == 0) carry = *total;
	__syncthreads();
	for(uint32_t base = 0; base < n; base += blockDim.x) {
		uint32_t i = base + threadIdx.x;
		unsigned long long v = i < n ? counts[i] : 0, x = v;
		for(int o = 1; o < 32; o <<= 1) { unsigned long long y = __shfl_up_sync(0xffffffffu, x, o); if((threadIdx.x & 31) >= o) x += y; }
		if((threadIdx.x & 31) == 31) warpsum[threadIdx.x >> 5] = x;
		__syncthreads();
		if(threadIdx.x < 32) {
			unsigned long long w = threadIdx.x < (blockDim.x >> 5) ? warpsum[threadIdx.x] : 0, z = w;
			for(int o = 1; o < 32; o <<= 1) { unsigned long long y = __shfl_up_sync(0xffffffffu, z, o); if(threadIdx.x >= o) z += y; }
			warpsum[threadIdx.x] = z - w;
		}
		__syncthreads();
		unsigned long long off = carry + warpsum[threadIdx.x >> 5] + x - v;
		if(i < n) {
			// pool overflow: the call fails in bt2gpu_dp_fetch; until then the job must look empty to the emit, sort and
			// backtrace passes (they index keys/cands/done by cand_off + [0, ncand)), as k_dp_gather_e2e does
			if(off + v > cap) { *errflag = 1; off = 0; res[job0 + i].ncand = 0; res[job0 + i].aligned = 0; }
			cand_off[job0 + i] = (uint32_t)off; fillptr[job0 + i] = 0; res[job0 + i].cand_off = (uint32_t)off;
		}
		__syncthreads();
		if(threadIdx.x == blockDim.x - 1) carry = off + v;
		__syncthreads();
	}
	if(threadIdx.x == 0) *total = carry;
}

// Per-job bitonic sort of the candidate keys (one CTA per job) + conversion to bt2gpu_dp_cand.
__global__ void __launch_bounds__(256)
k_dp_sort(uint32_t job0, const bt2gpu_dp_result *__restrict__ res, uint64_t *__restrict__ keys,
          bt2gpu_dp_cand *__restrict__ cands)
{
	extern __shared__ uint64_t skeys[];
	const uint32_t ji = job0 + blockIdx.x;
	const uint32_t n = res[ji].ncand;
	if(n == 0) return;
	uint64_t *g = keys + res[ji].cand_off;
	const bool insm = n <= 4096;
	uint64_t *a = insm ? skeys : g;
	if(insm) { for(uint32_t i = threadIdx.x; i < n; i += blockDim.x) skeys[i] = g[i]; }
	__syncthreads();
	uint32_t np2 = 1; while(np2 < n) np2 <<= 1;
	for(uint32_t size = 2; size <= np2; size <<= 1) {
		for(uint32_t i = threadIdx.x; i < np2 / 2; i += blockDim.x) {     // flip step
			uint32_t blk = i / (size / 2), off = i % (size / 2);
			uint32_t x = blk * size + off, y = blk * size + size - 1 - off;
			if(y < n) { uint64_t kx = a[x], ky = a[y]; if(kx > ky) { a[x] = ky; a[y] = kx; } }
		}
		__syncthreads();
		for(uint32_t stride = size / 4; stride > 0; stride >>= 1) {
			for(uint32_t i = threadIdx.x; i < np2 / 2; i += blockDim.x) {
				uint32_t x = (i / stride) * stride * 2 + (i % stride), y = x + stride;
				if(y < n) { uint64_t kx = a[x], ky = a[y]; if(kx > ky) { a[x] = ky; a[y] = kx; } }
			}
			__syncthreads();
		}
	}
	bt2gpu_dp_cand *out = cands + res[ji].cand_off;
	for(uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
		uint64_t k = ~a[i];
		if(insm) g[i] = a[i];
		bt2gpu_dp_cand c; c.score = (int32_t)(uint32_t)(k >> 32) - 32768; c.row = (uint32_t)((k >> 16) & 0xffff); c.col = (uint32_t)(k & 0xffff);
		out[i] = c;
	}
}

// End-to-end windows of at most E2E_CAP columns: count, place, sort and convert in one launch (one
// warp per job).  The candidates are the last-row cells with H >= minsc (at most ncol of them), so
// a warp sorts them in its slice of shared memory and takes its slot in the pool with one atomic.
static constexpr int E2E_CAP = 1024;   // mate windows of -X 500 are ~700 columns wide; seed-extension windows use the 512 instance
template<int CAP> __global__ void __launch_bounds__(128)
k_dp_gather_e2e(const bt2gpu_dp_job *__restrict__ jobs, const DpAux *__restrict__ aux, uint32_t job0, uint32_t njobs,
                const uint32_t *__restrict__ bound, bt2gpu_dp_result *__restrict__ res, bt2gpu_dp_cand *__restrict__ cands,
                unsigned long long *__restrict__ total, unsigned long long cap, int *__restrict__ errflag)
{
	__shared__ uint64_t s_keys[4][CAP];
	const unsigned FULL = 0xffffffffu;
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const uint32_t ji = job0 + blockIdx.x * (blockDim.x >> 5) + wib;
	if(ji >= job0 + njobs) return;
	const bt2gpu_dp_job jb = jobs[ji];
	const DpAux ax = aux[ji];
	const int nrow = (int)jb.rd_len, ncol = (int)jb.ncol, minsc = jb.minsc;
	uint64_t *a = s_keys[wib];
	// last row of H, left by the fill in the boundary buffer its last stripe does not use
	const int32_t *lastrow = reinterpret_cast<const int32_t*>(bound + ax.bound_off + (((ax.nstripes - 1) & 1) ? 0 : ncol));
	int best = INT32_MIN; uint32_t cnt = 0;
	for(int j0 = 0; j0 < ncol; j0 += 32) {
		const int j = j0 + lane;
		int h = INT32_MIN;
		if(j < ncol) h = __ldcg(lastrow + j);
		best = max(best, h);
		const bool in = j < ncol && h >= minsc;
		const unsigned bal = __ballot_sync(FULL, in);
		if(in) a[cnt + __popc(bal & ((1u << lane) - 1u))] = cand_key(h, nrow - 1, j);
		cnt += __popc(bal);
	}
	for(int o = 16; o; o >>= 1) best = max(best, __shfl_xor_sync(FULL, best, o));
	bt2gpu_dp_result r; memset(&r, 0, sizeof(r));
	const bool w8 = (ax.fl == -255);
	int64_t b = best; bool ok = true;
	if(ncol <= 0) { b = INT64_MIN; ok = false; }                 // !found
	else if(best < minsc) ok = false;                            // flag -1, returns score
	else if(w8 && best == -255) { b = INT64_MIN; ok = false; }   // flag -2 (ee_u8.cpp:1137)
	r.best = b; r.width = ok ? (w8 ? 8 : 16) : 0;
	r.ncand = ok ? cnt : 0; r.aligned = (ok && cnt > 0) ? 1 : 0;
	const uint32_t n = r.ncand;
	if(n) {
		// bitonic sort of the warp's keys (ascending key == DpBtCandidate::operator<)
		uint32_t np2 = 1; while(np2 < n) np2 <<= 1;
		__syncwarp();
		for(uint32_t size = 2; size <= np2; size <<= 1) {
			for(uint32_t i = lane; i < np2 / 2; i += 32) {
				const uint32_t blk = i / (size / 2), off = i % (size / 2);
				const uint32_t x = blk * size + off, y = blk * size + size - 1 - off;
				if(y < n) { const uint64_t kx = a[x], ky = a[y]; if(kx > ky) { a[x] = ky; a[y] = kx; } }
			}
			__syncwarp();
			for(uint32_t stride = size / 4; stride > 0; stride >>= 1) {
				for(uint32_t i = lane; i < np2 / 2; i += 32) {
					const uint32_t x = (i / stride) * stride * 2 + (i % stride), y = x + stride;
					if(y < n) { const uint64_t kx = a[x], ky = a[y]; if(kx > ky) { a[x] = ky; a[y] = kx; } }
				}
				__syncwarp();
			}
		}
		unsigned long long off = 0;
		if(lane == 0) off = atomicAdd(total, (unsigned long long)n);
		off = __shfl_sync(FULL, off, 0);
		if(off + n > cap) { if(lane == 0) { *errflag = 1; r.ncand = 0; r.aligned = 0; res[ji] = r; } return; }
		r.cand_off = (uint32_t)off;
		for(uint32_t i = lane; i < n; i += 32) {
			const uint64_t k = ~a[i];
			bt2gpu_dp_cand c; c.score = (int32_t)(uint32_t)(k >> 32) - 32768; c.row = (uint32_t)((k >> 16) & 0xffff); c.col = (uint32_t)(k & 0xffff);
			cands[off + i] = c;
		}
	}
	if(lane == 0) res[ji] = r;
}

// ---------------------------------------------------------------------------------------
// Backtrace.  One WARP per job replays SwAligner::nextAlignment (aligner_sw.cpp:737-1146)
// over the sorted candidates.  The walk is backtraceNucleotides* (ee_u8.cpp:1283-1877).
// In the reference a cell is marked reportedThrough the moment it is visited, and a popped
// branch frame re-enters its (already marked) cell, so every dead end unwinds the whole
// branch stack and fails: the masks_/btnstack_ state never changes an outcome.  The kernel
// therefore keeps only the reportedThrough bit and fails at the first blocked cell
// (equivalence is exercised by the goldens; the CPU oracle keeps the full state).
//
// The reference walks one cell per step, each step a dependent gather from the E/F/H
// matrices.  What a cell decides in the H state depends only on the stored values around it,
// and >95% of the steps are diagonal, so the warp SPECULATES the diagonal: lane i evaluates
// cell (row-i, col-i) as if the path arrived there in the H state; a ballot finds the first
// lane whose cell is not a plain diagonal move; lanes before it are committed together
// (reportedThrough bits, core-diagonal test, mismatch edits, score), that lane's event is
// handled, and the E/F (gap extension) states run as uniform scalar code.  One round trip to
// the matrices then covers up to 32 cells of the path.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint8_t dev_mask2dna(int m) { return (uint8_t)"?ACMGRSVTWYHKDBNN"[m & 31 ? (m > 16 ? 16 : m) : 0]; }

struct BtJob {
	MatView mv; const uint8_t *rd, *qu, *rf; uint32_t *rp32;
	int nrow, ncol, gapbar; bool local;
	long long triml; unsigned long long corel, corer;
	__device__ __forceinline__ bool reported(int row, int col) const {
		const size_t b = (size_t)row * ncol + col;
		return (__ldcg(rp32 + (b >> 5)) >> (b & 31)) & 1u;
	}
	__device__ __forceinline__ void mark(int row, int col) const {
		const size_t b = (size_t)row * ncol + col;
		atomicOr(rp32 + (b >> 5), 1u << (b & 31));
	}
	__device__ __forceinline__ bool on_core(int row, int col) const {
		const long long d = (long long)col - row + triml;
		return d >= 0 && (unsigned long long)d >= corel && (unsigned long long)d <= corer;
	}
	__device__ __forceinline__ bool in_barrier(int row) const { return row < gapbar || (nrow - row - 1) < gapbar; }
	// E(row, col) == T ?  The matrix stream holds H only.  The fill computes, along a row,
	//   E(r,c) = max(floor, E(r,c-1) - rde, H(r,c-1) - rdo)     (E(r,0) = floor; no read gap opens inside the gap barrier)
	// so E(r,c) = T (T above the floor) iff the chain "E(r,c) extends E(r,c-1) extends ..." reaches a column c' where
	// H(r,c'-1) - rdo equals what E(r,c') has to be.  The chain stops as soon as the required value exceeds what the
	// cell can hold: E(r,c') <= H(r,c') always (H is the maximum of its three sources), and H(r,c'-1) - rdo <= E(r,c').
	// One H load per link; a gap of length g costs g loads, and nearly every test ends at the first.
	__device__ __forceinline__ bool e_is(int row, int col, int T) const {
		if(in_barrier(row)) return false;
		const RowRef rr = mv.rowref(row);
		int need = T;
		for(int c = col; c > 0; c--) {
			const int h = mv.at(rr, c - 1);
			const int o = h - c_sc.rdo;
			if(o == need) return true;
			if(o > need) return false;         // E(row,c) >= o: it cannot be `need`
			need += c_sc.rde;                  // then E(row,c) extends E(row,c-1), which has to be need + rde ...
			if(need > h) return false;         // ... and cannot exceed H(row,c-1)
		}
		return false;
	}
	// F(row, col) == T ?  Down a column: F(r,c) = max(floor, F(r-1,c) - rfe, H(r-1,c) - rfo), no reference gap inside the
	// gap barrier (F is the floor there, and in row 0).
	__device__ __forceinline__ bool f_is(int row, int col, int T) const {
		int need = T;
		for(int r = row; r > 0; r--) {
			if(in_barrier(r)) return false;
			const int h = mv.get(r - 1, col);
			const int o = h - c_sc.rfo;
			if(o == need) return true;
			if(o > need) return false;
			need += c_sc.rfe;
			if(need > h) return false;
		}
		return false;
	}
	// move out of an H cell: 0 = none (empty mask), 1 diag, 2 ref-open, 5 ref-ext, 3 read-open, 4 read-ext.
	// sc_cur = H(row, col): the value of the gap cell the walk enters on move 4 (E(row,col-1) = sc_cur + rde) or 5
	// (F(row-1,col) = sc_cur + rfe) follows from it.
	__device__ __forceinline__ int h_move(int row, int col, int &sc_cur) const {
		const bool gapsAllowed = !in_barrier(row);
		const RowRef rc = mv.rowref(row), ru = mv.rowref(row > 0 ? row - 1 : 0);   // row 0: callers discard the result
		// The diagonal wins whenever it is legal (ee_u8.cpp:1510-1520), and most cells of a walk are diagonal moves:
		// its two cells are fetched first, the gap neighbours only when the diagonal fails.
		sc_cur = mv.at(rc, col);
		const int cl = col > 0 ? col - 1 : 0;
		if(col > 0) {
			const int hd = mv.at(ru, cl);
			const int sd = dev_score_mask(rd[row], rf[col], (int)qu[row] - 33);
			if((!local || hd > 0) && sc_cur == hd + sd) return 1;
		}
		if(!gapsAllowed) return 0;
		// (diag >) ref open > ref ext > read open > read ext; "x > 0" of the local rules is a test on the value the
		// neighbour would have to hold
		const int hu = mv.at(ru, col);
		if((!local || hu > 0) && sc_cur == hu - c_sc.rfo) return 2;
		if((!local || sc_cur + c_sc.rfe > 0) && f_is(row - 1, col, sc_cur + c_sc.rfe)) return 5;
		if(col > 0) {
			const int hl = mv.at(rc, cl);
			if((!local || hl > 0) && sc_cur == hl - c_sc.rdo) return 3;
			if((!local || sc_cur + c_sc.rde > 0) && e_is(row, cl, sc_cur + c_sc.rde)) return 4;
		}
		return 0;
	}
	// move out of the E (ct 0) or F (ct 1) cell (row, col) whose value is gv: 3/2 open, 4/5 extend, 0 none
	__device__ __forceinline__ int g_move(int ct, int row, int col, int gv) const {
		if(ct == 0 && col == 0) return 0;
		const int pr = max(row - ct, 0), pc = max(col - (1 - ct), 0);      // callers discard row 0 results
		const int hp = mv.get(pr, pc);
		const int open = ct ? c_sc.rfo : c_sc.rdo, ext = ct ? c_sc.rfe : c_sc.rde;
		if((!local || hp > 0) && hp - open == gv) return ct ? 2 : 3;
		if((!local || gv + ext > 0) && (ct ? f_is(pr, pc, gv + ext) : e_is(pr, pc, gv + ext))) return ct ? 5 : 4;
		return 0;
	}
};

// Trial group (end-to-end mode).  After the first alignment of a window, almost every remaining
// candidate is a neighbour of its end cell whose walk runs a few cells (the gap barrier's rows, then
// a gap back to the reported path) and dies on a cell that an earlier walk has reported.  The
// reference tries them one after another; here 32 consecutive candidates walk at once, one per
// lane, against the bitmap as it stands, each recording its cells.  The reference's order is then
// replayed over the recorded walks -- candidate q stops at the first of its cells that an earlier
// candidate of the group has reported -- with a small shared-memory bitmap of the window the walks
// live in, which is exactly the sequential outcome.  Candidates up to the first one whose walk
// neither gets skipped nor blocked are committed (bitmap bits, try count); that one goes to the
// full walker below.
static constexpr int TR_LMAX = 32;                // recorded cells per trial walk = rows of the window

__global__ void __launch_bounds__(128, 8)
k_dp_backtrace(const uint8_t *__restrict__ seq, const uint8_t *__restrict__ qual,
               const uint8_t *__restrict__ win, const bt2gpu_dp_job *__restrict__ jobs,
               const DpAux *__restrict__ aux, uint32_t job0, uint32_t njobs,
               const uint32_t *__restrict__ mat, bt2gpu_dp_result *__restrict__ res,
               const bt2gpu_dp_cand *__restrict__ cands, uint8_t *__restrict__ rep,
               uint32_t *__restrict__ donelist, bt2gpu_dp_aln *__restrict__ alns,
               bt2gpu_edit *__restrict__ escr, bt2gpu_edit *__restrict__ edits,
               unsigned long long *__restrict__ nedits, unsigned long long edit_cap,
               uint32_t max_aln, int *__restrict__ errflag)
{
	const unsigned FULL = 0xffffffffu;
	const int lane = threadIdx.x & 31;
	const uint32_t ji = job0 + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if(ji >= job0 + njobs) return;
	bt2gpu_dp_result r = res[ji];
	r.naln = 0; r.aln_off = ji * max_aln; r.tail_tries = 0; r.cural = 0;
	if(!r.aligned) { if(lane == 0) res[ji] = r; return; }
	const bt2gpu_dp_job jb = jobs[ji];
	const DpAux ax = aux[ji];
	BtJob J;
	J.nrow = (int)jb.rd_len; J.ncol = (int)jb.ncol;
	J.mv.init(mat + ax.mat_off, (int)ax.S, (int)ax.T, (int)ax.RW, ax.fl == -255);
	J.rd = seq + jb.rd_off; J.qu = qual + jb.rd_off; J.rf = win + jb.rf_off;
	J.rp32 = reinterpret_cast<uint32_t*>(rep + ax.rep_off);
	J.local = !c_sc.monotone; J.gapbar = c_sc.gapbar;
	J.triml = (long long)jb.triml; J.corel = jb.corel; J.corer = jb.corer;
	const int nrow = J.nrow;
	const uint8_t *rd = J.rd, *qu = J.qu, *rf = J.rf;
	const bool local = J.local;
	const bt2gpu_dp_cand *cd = cands + r.cand_off;
	uint32_t *done = donelist + r.cand_off;
	bt2gpu_edit *ed = escr + ax.escr_off;
	const uint32_t ncand = r.ncand;
	uint32_t SQ = (uint32_t)nrow >> 4; if(SQ == 0) SQ = 1;
	uint32_t cural = 0, ndone = 0, tries = 0, cbase = 0xffffffffu;
	uint32_t btcells = 0;                // matrix cells the walks visited (the reference's SSEMetrics.btcell)
	__shared__ uint16_t s_tcells[4][TR_LMAX * 32];   // recorded cells of the trial walks: (rows above the last) << 8 | window column
	__shared__ uint32_t s_twin[4][TR_LMAX * 8];      // cells reported by the group so far: TR_LMAX rows x 256 columns
	uint16_t *tcells = s_tcells[threadIdx.x >> 5]; uint32_t *twin = s_twin[threadIdx.x >> 5];
	const bool trial_ok = !local;
	bool walked = false;                 // a full walk has happened: the bitmap is no longer empty
	uint32_t trial_at = 0xffffffffu;     // candidate the last trial group stopped at (walked in full next)
	bt2gpu_dp_cand cc; cc.row = cc.col = 0; cc.score = 0;
	while(r.naln < max_aln && cural < ncand) {
		bool got = false;
		bt2gpu_dp_aln out;
		while(cural < ncand) {
			if(trial_ok && walked && cural != trial_at) {
				// ---- trial group: lanes = candidates cural .. cural+31 ----
				const uint32_t ci = cural + lane;
				enum { T_RUN = 0, T_INACTIVE, T_SKIP, T_BLOCKED, T_END, T_HORIZON };
				int status = T_INACTIVE, row = 0, col = 0, ct = 2, len = 0, gv = 0;
				int cmin = INT32_MAX, cmax = -1;
				if(ci < ncand) {
					const bt2gpu_dp_cand c = cd[ci];
					row = (int)c.row; col = (int)c.col;
					status = c.score < jb.minsc ? T_SKIP : T_RUN;
					cmin = cmax = col;
					if(row != nrow - 1) cmax = INT32_MAX;          // not an end-to-end candidate: no trial
				}
				for(int o = 16; o; o >>= 1) {
					cmin = min(cmin, __shfl_xor_sync(FULL, cmin, o));
					cmax = max(cmax, __shfl_xor_sync(FULL, cmax, o));
				}
				const int wbase = cmin - TR_LMAX;                   // window: TR_LMAX rows x 256 columns from wbase
				if(cmax == INT32_MAX || cmax - wbase >= 256) { trial_at = cural; }
				else {
					for(int t = 0; t < TR_LMAX; t++) {
						if(!__any_sync(FULL, status == T_RUN)) break;
						if(status == T_RUN) {
							const bool rep = J.reported(row, col);
							int m = 0, hv = 0;
							if(!rep && row > 0) m = ct == 2 ? J.h_move(row, col, hv) : J.g_move(ct, row, col, gv);
							tcells[t * 32 + lane] = (uint16_t)(((nrow - 1 - row) << 8) | (col - wbase));
							len = t + 1;
							if(rep) status = t == 0 ? T_SKIP : T_BLOCKED;
							else if(row == 0 || m == 0) status = T_END;
							else {
								// the value of the gap cell entered by an extension: one extension above the cell left
								const int from = ct == 2 ? hv : gv;
								if(m == 1) { row--; col--; ct = 2; }
								else if(m == 2) { row--; ct = 2; }
								else if(m == 5) { row--; ct = 1; gv = from + c_sc.rfe; }
								else if(m == 3) { col--; ct = 2; }
								else { col--; ct = 0; gv = from + c_sc.rde; }
							}
						}
					}
					if(status == T_RUN) status = T_HORIZON;
					for(int i = lane; i < TR_LMAX * 8; i += 32) twin[i] = 0;
					__syncwarp();
					// the reference's order, replayed over the recorded walks: candidate q stops at the first of
					// its cells that an earlier candidate of the group has reported
					int nacc = 0; uint32_t nfail = 0;
					for(; nacc < 32; nacc++) {
						const int st = __shfl_sync(FULL, status, nacc), ln = __shfl_sync(FULL, len, nacc);
						if(st == T_INACTIVE) break;
						if(st == T_SKIP) continue;
						uint32_t cell = 0; bool set = false;
						if(lane < ln) { cell = tcells[lane * 32 + nacc]; set = (twin[cell >> 5] >> (cell & 31)) & 1u; }
						const int first = __ffs(__ballot_sync(FULL, set)) - 1;
						int nmark;
						if(first < 0) {
							if(st != T_BLOCKED) break;                     // runs to its end or past the horizon: full walk
							nmark = ln;
						} else if(first == 0) nmark = 0;                   // start cell already reported: skipped, not a try
						else nmark = first;                                // cell `first` is set already
						if(nmark) nfail++;
						btcells += nmark;
						if(lane < nmark) {
							atomicOr(&twin[cell >> 5], 1u << (cell & 31));
							const uint32_t b = (uint32_t)(nrow - 1 - (int)(cell >> 8)) * (uint32_t)J.ncol + (uint32_t)(wbase + (int)(cell & 255u));
							atomicOr(J.rp32 + (b >> 5), 1u << (b & 31));
						}
						__syncwarp();
					}
					tries += nfail;
					cural += nacc;
					trial_at = cural;
					__syncwarp();
					if(cural >= ncand) break;
				}
			}
			if(local) {
				// Skip scan (local mode: thousands of candidates per window, nearly all dropped untried).  The next 32
				// candidates are tested at once for the three reasons nextAlignment drops one without a walk -- score
				// below minsc, start cell already reported, within SQ rows and columns of a tried one -- and, as
				// dropping a candidate changes no state, everything before the first survivor goes in one step.
				const uint32_t ci = cural + lane;
				bool keepc = false;
				if(ci < ncand) {
					const bt2gpu_dp_cand c = cd[ci];
					keepc = c.score >= jb.minsc && !J.reported((int)c.row, (int)c.col);
					// |row - dr| <= SQ && |col - dc| <= SQ as two unsigned range tests (x - d + SQ <= 2 SQ)
					const uint32_t rs = c.row + SQ, cs = c.col + SQ, SQ2 = 2 * SQ;
					for(uint32_t i = 0; keepc && i < ndone; i++) {
						const uint32_t dv = done[i];
						if(rs - (dv >> 16) <= SQ2 && cs - (dv & 0xffff) <= SQ2) keepc = false;
					}
				}
				const unsigned keep = __ballot_sync(FULL, keepc);
				if(keep == 0) { cural += min(32u, ncand - cural); continue; }
				cural += (uint32_t)(__ffs(keep) - 1);
			}
			// candidates are read 32 at a time (one round trip per 32 of them)
			if(cural < cbase || cural >= cbase + 32) {
				cbase = cural;
				if(cbase + lane < ncand) cc = cd[cbase + lane];
			}
			bt2gpu_dp_cand c0;
			c0.row = __shfl_sync(FULL, cc.row, cural - cbase);
			c0.col = __shfl_sync(FULL, cc.col, cural - cbase);
			c0.score = __shfl_sync(FULL, cc.score, cural - cbase);
			if(c0.score < jb.minsc) { cural++; continue; }
			int row = (int)c0.row, col = (int)c0.col;
			// the "start cell already reported" test (aligner_sw.cpp:774-789) rides on the first
			// speculative batch below (lane 0 is the start cell) instead of its own round trip
			bool fresh = true, skipped = false;
			// ---- walk ----
			int score = 0, ns = 0, gaps = 0, ned = 0, ct = 2;
			int gval = 0;                          // value of the E / F cell the walk is in (ct 0 / 1)
			const int origCol = col;
			const int trimEnd = nrow - row - 1;
			int trimBeg = 0;
			bool ok = false, overlapped = false, blocked = false;
			while(row >= 0) {
				if(ct == 2) {
					// ---- speculative diagonal batch: lane i looks at (row-i, col-i) in the H state ----
					const int ri = row - lane, ci = col - lane;
					const bool valid = ri >= 0 && ci >= 0;
					bool rep_i = false; int mv_i = 0, hv_i = 0;
					if(valid) {
						rep_i = J.reported(ri, ci);
						if(!rep_i && ri > 0) mv_i = J.h_move(ri, ci, hv_i);
					}
					const bool plain = valid && !rep_i && ri > 0 && mv_i == 1;
					const unsigned pb = __ballot_sync(FULL, plain);
					const int k = __ffs(~pb) - 1;             // first lane that is not a plain diagonal step (32 if none)
					const int nk = k < 0 ? 32 : k;
					if(fresh) {
						fresh = false;
						if(__shfl_sync(FULL, (int)rep_i, 0)) { skipped = true; break; }
						if(local) {
							bool dom = false;
							for(uint32_t i = lane; i < ndone; i += 32) {
								const uint32_t dv = done[i];
								uint32_t dr = dv >> 16, dc = dv & 0xffff;
								uint32_t rh = max((uint32_t)row, dr), rl = min((uint32_t)row, dr);
								uint32_t ch = max((uint32_t)col, dc), cl = min((uint32_t)col, dc);
								if(ch - cl <= SQ && rh - rl <= SQ) dom = true;
							}
							if(__any_sync(FULL, dom)) { skipped = true; break; }
						}
						tries++; walked = true;
					}
					// commit lanes [0, nk): diagonal moves
					{
						const bool mine = lane < nk;
						int mm = 1, pen = 0;
						if(mine) {
							J.mark(ri, ci);
							const int rc = rd[ri], rm = rf[ci];
							mm = (rm >= 16 || rc > 3) ? -1 : (((1 << rc) & rm) ? 1 : 0);
							pen = (mm != 1) ? -dev_mm(rc, rm, (int)qu[ri] - 33) : c_sc.match;
						}
						const bool isEd = mine && mm != 1;
						const unsigned eb = __ballot_sync(FULL, isEd);
						if(isEd) {
							bt2gpu_edit e; e.pad = 0;
							e.pos = (uint32_t)ri; e.chr = dev_mask2dna(rf[ci]); e.qchr = (uint8_t)"ACGTN"[rd[ri]]; e.type = BT2GPU_EDIT_MM;
							ed[ned + __popc(eb & ((1u << lane) - 1u))] = e;
						}
						ned += __popc(eb);
						score += __reduce_add_sync(FULL, mine ? pen : 0);
						ns += __popc(__ballot_sync(FULL, mine && mm == -1));
						if(__any_sync(FULL, mine && J.on_core(ri, ci))) overlapped = true;
					}
					row -= nk; col -= nk; btcells += nk;
					if(nk == 32) continue;
					// ---- event at lane nk: cell (row, col), still in the H state ----
					if(col < 0) {
						// ran off the left edge on a diagonal: the scalar walk would be at (row, -1); cannot
						// happen (a diagonal move needs col > 0), kept as a guard
						blocked = true; break;
					}
					const bool e_rep = __shfl_sync(FULL, (int)rep_i, nk) != 0;
					const int e_mv = __shfl_sync(FULL, mv_i, nk), e_hv = __shfl_sync(FULL, hv_i, nk);
					J.mark(row, col); btcells++;
					if(e_rep) { blocked = true; break; }
					if(J.on_core(row, col)) overlapped = true;
					if(row == 0 || e_mv == 0) { trimBeg = row; ok = true; break; }
					bt2gpu_edit e; e.pad = 0;
					if(e_mv == 2 || e_mv == 5) {
						e.pos = (uint32_t)row; e.chr = '-'; e.qchr = (uint8_t)"ACGTN"[rd[row]]; e.type = BT2GPU_EDIT_REF_GAP;
						if(lane == 0) ed[ned] = e;
						ned++; row--;
						if(e_mv == 2) { ct = 2; score -= c_sc.rfo; } else { ct = 1; score -= c_sc.rfe; gval = e_hv + c_sc.rfe; }
						gaps++;
					} else {
						e.pos = (uint32_t)row + 1; e.chr = dev_mask2dna(rf[col]); e.qchr = '-'; e.type = BT2GPU_EDIT_READ_GAP;
						if(lane == 0) ed[ned] = e;
						ned++; col--;
						if(e_mv == 3) { ct = 2; score -= c_sc.rdo; } else { ct = 0; score -= c_sc.rde; gval = e_hv + c_sc.rde; }
						gaps++;
					}
					continue;
				}
				// ---- E / F states: uniform scalar steps (every lane computes the same thing) ----
				const bool reported = J.reported(row, col);
				bool empty = false;
				int cur = 0;
				if(!reported && row > 0) {
					cur = J.g_move(ct, row, col, gval);          // 3 / 2 open (back to the H state), 4 / 5 extend
					empty = cur == 0;
				}
				__syncwarp();
				J.mark(row, col); btcells++;
				if(reported) { blocked = true; break; }
				if(J.on_core(row, col)) overlapped = true;
				if(empty || row == 0) { trimBeg = row; ok = true; break; }
				bt2gpu_edit e; e.pad = 0;
				if(cur == 2 || cur == 5) {
					e.pos = (uint32_t)row; e.chr = '-'; e.qchr = (uint8_t)"ACGTN"[rd[row]]; e.type = BT2GPU_EDIT_REF_GAP;
					if(lane == 0) ed[ned] = e;
					ned++; row--;
					if(cur == 2) { ct = 2; score -= c_sc.rfo; } else { ct = 1; score -= c_sc.rfe; gval += c_sc.rfe; }
					gaps++;
				} else {
					e.pos = (uint32_t)row + 1; e.chr = dev_mask2dna(rf[col]); e.qchr = '-'; e.type = BT2GPU_EDIT_READ_GAP;
					if(lane == 0) ed[ned] = e;
					ned++; col--;
					if(cur == 3) { ct = 2; score -= c_sc.rdo; } else { ct = 0; score -= c_sc.rde; gval += c_sc.rde; }
					gaps++;
				}
			}
			__syncwarp();
			if(skipped) { cural++; continue; }
			if(ok && !blocked) {
				ok = overlapped;                                           // core-diagonal test
				if(ok) {
					const int rc = rd[row], rm = rf[col];
					const int mm = (rm >= 16 || rc > 3) ? -1 : (((1 << rc) & rm) ? 1 : 0);
					if(mm != 1) {
						bt2gpu_edit e; e.pad = 0;
						e.pos = (uint32_t)row; e.chr = dev_mask2dna(rm); e.qchr = (uint8_t)"ACGTN"[rc]; e.type = BT2GPU_EDIT_MM;
						if(lane == 0) ed[ned] = e;
						ned++;
						score -= dev_mm(rc, rm, (int)qu[row] - 33);
					} else score += c_sc.match;
					if(mm == -1) ns++;
					if(ns > jb.nceil) ok = false;
				}
			} else ok = false;
			if(local) { if(lane == 0) done[ndone] = ((uint32_t)c0.row << 16) | (uint32_t)c0.col; ndone++; }
			__syncwarp();
			if(ok) {
				out.score = score; out.ns = ns; out.gaps = gaps;
				out.refoff = (int64_t)col + jb.refl;
				out.trim5 = (uint32_t)(jb.fw ? trimBeg : trimEnd);
				out.trim3 = (uint32_t)(jb.fw ? trimEnd : trimBeg);
				int refns = 0;
				for(int i = col + lane; i <= origCol; i += 32) if(rf[i] > 15) refns++;
				out.refns = __reduce_add_sync(FULL, refns);
				unsigned long long eo = 0;
				if(lane == 0) eo = atomicAdd(nedits, (unsigned long long)ned);
				eo = __shfl_sync(FULL, eo, 0);
				if(eo + ned > edit_cap) { if(lane == 0) { *errflag = 2; res[ji] = r; } return; }
				out.edit_off = (uint32_t)eo; out.nedits = (uint32_t)ned;
				// scratch holds edits in walk order (descending row); reference order after
				// res.reverse() is ascending; '-' strand then inverts again (edit.cpp:50-78)
				const int rdex = nrow - trimBeg - trimEnd;
				for(int i = lane; i < ned; i += 32) {
					bt2gpu_edit e = ed[jb.fw ? (ned - 1 - i) : i];
					e.pos -= (uint32_t)trimBeg;
					if(!jb.fw) e.pos = (uint32_t)(rdex - (int)e.pos - (e.type == BT2GPU_EDIT_READ_GAP ? 0 : 1));
					edits[eo + i] = e;
				}
				__syncwarp();
				got = true;
				break;
			}
			cural++;
		}
		if(!got) break;
		cural++;
		out.cural = cural; out.tries = tries; tries = 0;
		if(lane == 0) alns[(size_t)ji * max_aln + r.naln] = out;
		r.naln++;
	}
	r.tail_tries = tries; r.cural = cural;
	if(lane == 0) { res[ji] = r; atomicAdd(nedits + 2, (unsigned long long)btcells); }   // scal[3]
}

// ---------------------------------------------------------------------------------------
// Host side: staging, chunking, launches.
// ---------------------------------------------------------------------------------------
static const int kSClasses[] = {3, 4, 5, 6, 8, 10, 12, 16};
static constexpr int kNClasses = sizeof(kSClasses) / sizeof(int);

struct DpChunk {
	uint32_t job0, njobs;
	uint64_t mat_words, rep_bytes;
	uint32_t bound_words, escr_edits;
	uint32_t cls_off[kNClasses + 1];   // ranges into `order` (absolute)
	int maxcol;
};

struct DpState {
	size_t njobs = 0; uint32_t max_aln = 1;
	size_t cand_cap = 0, edit_cap = 0;
	uint64_t cells = 0;
	std::vector<DpChunk> chunks;
	DevBuf seq, qual, win, jobs, aux, order, res, counts, cand_off, fillptr, keys, cands, done,
	       alns, edits, escr, mat, bound, rep, scal;   // scal: [0]=total cands, [1]=total edits, [2]=errflag
	HostBuf h_scal;
	DevBuf cls, sizes, soffs, spart, dsum;   // device-side staging scratch
	HostBuf h_dsum;
	size_t dev_pool_len = 0, dev_win_len = 0;
	int maxcol_all = 0;
	cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};   // fill | gather | backtrace boundaries (single-chunk runs)
	uint32_t ev_phases = 0;
};

template<int S> static int launch_fill(bt2gpu_ctx *ctx, DpState *d, const uint32_t *order, uint32_t n) {
	if(n == 0) return 0;
	constexpr int SP = (S + 3) & ~3;
	const size_t smem = (size_t)16 * NCH * GL * SP * 4;
	static bool attr = false;
	if(!attr) {
		BT2_CUDA(cudaFuncSetAttribute(k_dp_fill<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		attr = true;
	}
	uint32_t blocks = (n + 15) / 16;
	uint32_t maxb = (uint32_t)ctx->sm_count * 8;
	if(blocks > maxb) blocks = maxb;
	k_dp_fill<S><<<blocks, 128, smem, ctx->stream>>>(
		d->seq.as<uint8_t>(), d->qual.as<uint8_t>(), d->win.as<uint8_t>(), d->jobs.as<bt2gpu_dp_job>(),
		d->aux.as<DpAux>(), order, n, d->mat.as<uint32_t>(), d->bound.as<uint32_t>());
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	return 0;
}

static int dispatch_fill(bt2gpu_ctx *ctx, DpState *d, int cls, const uint32_t *order, uint32_t n) {
	switch(kSClasses[cls]) {
	case 3:  return launch_fill<3>(ctx, d, order, n);
	case 4:  return launch_fill<4>(ctx, d, order, n);
	case 5:  return launch_fill<5>(ctx, d, order, n);
	case 6:  return launch_fill<6>(ctx, d, order, n);
	case 8:  return launch_fill<8>(ctx, d, order, n);
	case 10: return launch_fill<10>(ctx, d, order, n);
	case 12: return launch_fill<12>(ctx, d, order, n);
	case 16: return launch_fill<16>(ctx, d, order, n);
	}
	BT2_FAIL("bad S class");
}


// ---------------------------------------------------------------------------------------
// Device-side staging (chained batch runtime): the job table, reads and windows were written
// by a kernel, so the per-job layout (class, arena offsets, launch order) is computed here too.
// ---------------------------------------------------------------------------------------
struct DpSizes { unsigned long long mat, rep; uint32_t bound, escr; };
__host__ __device__ __forceinline__ DpSizes operator+(const DpSizes &a, const DpSizes &b) {
	DpSizes r; r.mat = a.mat + b.mat; r.rep = a.rep + b.rep; r.bound = a.bound + b.bound; r.escr = a.escr + b.escr; return r;
}
// summary read back by the host: [0..7] jobs per class, [8] max ncol, [9] error code, then cells (u64 at word 10)
static constexpr int DPSUM_WORDS = 24;    // [16..23]: per-class fill counters of k_dp_order
__constant__ int c_sclasses[8] = {3, 4, 5, 6, 8, 10, 12, 16};

__global__ void __launch_bounds__(256)
k_dp_aux(const bt2gpu_dp_job *__restrict__ jobs, uint32_t njobs, unsigned long long pool_len, unsigned long long win_len,
         int local, DpAux *__restrict__ aux, uint8_t *__restrict__ cls, DpSizes *__restrict__ sizes, uint32_t *__restrict__ sum)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	unsigned long long cells = 0;
	int mycls = -1; uint32_t mycol = 0;
	if(i < njobs) {
		const bt2gpu_dp_job j = jobs[i];
		int err = 0;
		if(j.rd_len == 0 || j.ncol == 0) err = 1;
		else if((unsigned long long)j.rd_off + j.rd_len > pool_len || (unsigned long long)j.rf_off + j.ncol + 1 > win_len) err = 2;
		else if(j.ncol > 60000 || j.rd_len > 60000) err = 3;
		else if(!local && j.rd_len > 2270) err = 4;
		else if(local && j.rd_len > 16000) err = 5;
		DpAux a; memset(&a, 0, sizeof(a));
		int best = -1; long long bestRows = 0, bestStripes = 0;            // choose_class
		for(int c = 0; c < 8; c++) {
			const long long rs = 16LL * c_sclasses[c], ns = ((long long)j.rd_len + rs - 1) / rs, rows = ns * rs;
			if(best < 0 || rows < bestRows || (rows == bestRows && ns < bestStripes)) { best = c; bestRows = rows; bestStripes = ns; }
		}
		a.S = (uint32_t)c_sclasses[best];
		a.nstripes = (j.rd_len + 16 * a.S - 1) / (16 * a.S);
		a.T = j.ncol + 16;
		if(local) a.fl = 0;
		else if(j.enable8 && j.minsc >= -254) a.fl = -255;
		else { a.fl = FL_I16; if(j.minsc < -15000) err = 6; }
		a.RW = (a.fl == -255) ? ((((a.S + 1) / 2) + 1) & ~1u) : ((a.S + 1) & ~1u);
		DpSizes z;
		z.mat = (unsigned long long)a.nstripes * a.T * GL * a.RW;
		z.rep = (((unsigned long long)j.rd_len * j.ncol + 31) / 32) * 4;
		z.bound = 2 * j.ncol; z.escr = j.rd_len + j.ncol + 4;
		if(err) { atomicMax(&sum[9], (uint32_t)err); z.mat = z.rep = 0; z.bound = z.escr = 0; }
		aux[i] = a; cls[i] = (uint8_t)best; sizes[i] = z;
		mycls = best; mycol = j.ncol;
		cells = (unsigned long long)j.rd_len * j.ncol;
	}
	// one atomic per warp and class (a batch is mostly one class: 65 us of same-address atomics otherwise)
	const int lane = threadIdx.x & 31;
#pragma unroll
	for(int c = 0; c < 8; c++) {
		const unsigned bal = __ballot_sync(0xffffffffu, mycls == c);
		if(lane == 0 && bal) atomicAdd(&sum[c], (uint32_t)__popc(bal));
	}
	for(int o = 16; o; o >>= 1) {
		cells += __shfl_xor_sync(0xffffffffu, cells, o);
		mycol = max(mycol, __shfl_xor_sync(0xffffffffu, mycol, o));
	}
	if(lane == 0 && cells) { atomicAdd(reinterpret_cast<unsigned long long*>(sum + 10), cells); atomicMax(&sum[8], mycol); }
}

__global__ void __launch_bounds__(256)
k_dp_aux_apply(DpAux *__restrict__ aux, const DpSizes *__restrict__ offs, uint32_t njobs) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= njobs) return;
	const DpSizes o = offs[i];
	aux[i].mat_off = o.mat; aux[i].rep_off = o.rep; aux[i].bound_off = o.bound; aux[i].escr_off = o.escr;
}

// launch order: by class; inside a class in (nearly) job order -- a warp takes its slots with one atomic per
// class.  `order` only schedules the fill, results are addressed by job.  fillc[8] must be zero on entry.
__global__ void __launch_bounds__(256)
k_dp_order(const uint8_t *__restrict__ cls, uint32_t njobs, const uint32_t *__restrict__ sum, uint32_t *__restrict__ fillc,
           uint32_t *__restrict__ order) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const int lane = threadIdx.x & 31;
	const int c = i < njobs ? (int)cls[i] : -1;
#pragma unroll
	for(int k = 0; k < 8; k++) {
		const unsigned bal = __ballot_sync(0xffffffffu, c == k);
		if(!bal) continue;
		uint32_t base = 0;
		if(lane == 0) base = atomicAdd(&fillc[k], (uint32_t)__popc(bal));
		base = __shfl_sync(0xffffffffu, base, 0);
		if(c == k) {
			uint32_t cb = 0;
			for(int q = 0; q < k; q++) cb += sum[q];
			order[cb + base + __popc(bal & ((1u << lane) - 1u))] = i;
		}
	}
}

static int choose_class(int nrow) {
	int best = -1; long bestRows = 0, bestStripes = 0;
	for(int c = 0; c < kNClasses; c++) {
		long rs = 16L * kSClasses[c];
		long ns = (nrow + rs - 1) / rs;
		long rows = ns * rs;
		if(best < 0 || rows < bestRows || (rows == bestRows && ns < bestStripes)) { best = c; bestRows = rows; bestStripes = ns; }
	}
	return best;
}

int fm_set_scoring(bt2gpu_ctx *ctx, const bt2gpu_scoring *sc);   // fm.cu

} // namespace bt2

using namespace bt2;

namespace bt2 { int scoring_acquire(bt2gpu_ctx *ctx, const bt2gpu_scoring *sc, bool *upload); }   // api.cu

extern "C" int bt2gpu_scoring_set(bt2gpu_ctx *ctx, const bt2gpu_scoring *sc) {
	if(!ctx || !sc) BT2_FAIL("null argument");
	BT2_CUDA(cudaSetDevice(ctx->device));
	if(sc->gapbar < 1) BT2_FAIL("gapbar must be >= 1 (ee_u8.cpp:781)");
	if(sc->rd_gap_ext > sc->rd_gap_open || sc->rf_gap_ext > sc->rf_gap_open) BT2_FAIL("gap extend > gap open");
	// The kernels read the scheme from __constant__ symbols, which every context of a device shares: the registry
	// admits one scheme per device at a time and refuses a second, different one while another context holds the first.
	bool upload = false;
	if(scoring_acquire(ctx, sc, &upload)) return 1;
	ctx->sc = *sc; ctx->have_sc = true;
	if(!upload) return 0;
	ScConst c = { sc->match_bonus, sc->rd_gap_open, sc->rd_gap_ext, sc->rf_gap_open, sc->rf_gap_ext, sc->gapbar, sc->monotone };
	BT2_CUDA(cudaMemcpyToSymbolAsync(c_mmpens, sc->mmpens, sizeof(c_mmpens), 0, cudaMemcpyHostToDevice, ctx->stream));
	BT2_CUDA(cudaMemcpyToSymbolAsync(c_npens, sc->npens, sizeof(c_npens), 0, cudaMemcpyHostToDevice, ctx->stream));
	BT2_CUDA(cudaMemcpyToSymbolAsync(c_sc, &c, sizeof(c), 0, cudaMemcpyHostToDevice, ctx->stream));
	if(fm_set_scoring(ctx, sc)) return 1;
	BT2_CUDA(cudaStreamSynchronize(ctx->stream));
	return 0;
}

extern "C" int bt2gpu_dp_stage(
	bt2gpu_ctx *ctx, const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const uint8_t *win_pool, size_t win_len, const bt2gpu_dp_job *jobs, size_t njobs,
	uint32_t max_aln, size_t cand_cap, size_t edit_cap)
{
	if(!ctx) BT2_FAIL("null ctx");
	if(!ctx->have_sc) BT2_FAIL("bt2gpu_scoring_set must be called first");
	BT2_CUDA(cudaSetDevice(ctx->device));
	if(!ctx->dp) ctx->dp = new DpState();
	DpState *d = ctx->dp;
	d->njobs = njobs; d->max_aln = max_aln ? max_aln : 1; d->cand_cap = cand_cap; d->edit_cap = edit_cap;
	d->chunks.clear(); d->cells = 0; d->maxcol_all = 0;
	if(njobs == 0) return 0;
	if(njobs > 0x7fffffffu / d->max_aln) BT2_FAIL("too many jobs");
	const bool local = !ctx->sc.monotone;
	// cudaMemGetInfo is slow (and serialises the device's contexts): the arena budget is only asked for when the batch
	// does not fit the arena this context already holds (first pass with no limit, second with the real one)
	uint64_t mat_budget_words = ~0ull;
	std::vector<DpAux> aux(njobs);
	std::vector<uint32_t> order(njobs);
	std::vector<uint8_t> cls(njobs);
	DpChunk ch;
	std::vector<std::pair<uint64_t, uint32_t>> tmp;
	auto close_chunk = [&](uint32_t end) {
		ch.njobs = end - ch.job0;
		// order: by class, then by work descending (keeps the 4 groups of a warp in step)
		uint32_t pos = ch.job0;
		for(int c = 0; c < kNClasses; c++) {
			ch.cls_off[c] = pos;
			tmp.clear();
			for(uint32_t i = ch.job0; i < end; i++) if(cls[i] == c) tmp.emplace_back((uint64_t)aux[i].nstripes * aux[i].T, i);
			std::sort(tmp.begin(), tmp.end(), [](const std::pair<uint64_t, uint32_t> &a, const std::pair<uint64_t, uint32_t> &b) {
				return a.first != b.first ? a.first > b.first : a.second < b.second; });
			for(auto &p : tmp) order[pos++] = p.second;
		}
		ch.cls_off[kNClasses] = pos;
		d->chunks.push_back(ch);
	};
	for(int pass = 0; pass < 2; pass++) {
	d->chunks.clear(); d->cells = 0; d->maxcol_all = 0;
	memset(&ch, 0, sizeof(ch)); ch.job0 = 0;
	for(size_t i = 0; i < njobs; i++) {
		const bt2gpu_dp_job &j = jobs[i];
		if(j.rd_len == 0 || j.ncol == 0) BT2_FAIL("empty DP problem");
		if((size_t)j.rd_off + j.rd_len > pool_len || (size_t)j.rf_off + j.ncol + 1 > win_len) BT2_FAIL("job outside pool");
		if(j.ncol > 60000 || j.rd_len > 60000) BT2_FAIL("DP problem too large");
		if(!local && j.rd_len > 2270) BT2_FAIL("end-to-end read longer than 2270 rows: checkpointed path (rdlen >= cminlen, aligner_sw.cpp:514) not implemented");
		if(local && j.rd_len > 16000) BT2_FAIL("local read longer than 16000 rows not supported by the s16 domain");
		DpAux a; memset(&a, 0, sizeof(a));
		const int c = choose_class((int)j.rd_len);
		cls[i] = (uint8_t)c;
		a.S = (uint32_t)kSClasses[c];
		a.nstripes = (j.rd_len + 16 * a.S - 1) / (16 * a.S);
		a.T = j.ncol + 16;
		if(local) a.fl = 0;
		else if(j.enable8 && j.minsc >= -254) a.fl = -255;            // aligner_sw.cpp:518
		else { a.fl = FL_I16; if(j.minsc < -15000) BT2_FAIL("minsc below the 16-bit end-to-end domain"); }
		a.RW = (a.fl == -255) ? ((((a.S + 1) / 2) + 1) & ~1u) : ((a.S + 1) & ~1u);   // H only; byte records on the u8 path
		const uint64_t mw = (uint64_t)a.nstripes * a.T * GL * a.RW;
		if(mw > mat_budget_words) BT2_FAIL("single DP problem exceeds the matrix arena");
		if(ch.mat_words + mw > mat_budget_words || ch.bound_words > 0x70000000u || ch.escr_edits > 0x70000000u) {
			close_chunk((uint32_t)i);
			memset(&ch, 0, sizeof(ch)); ch.job0 = (uint32_t)i;
		}
		a.mat_off = ch.mat_words; ch.mat_words += mw;
		a.rep_off = ch.rep_bytes; ch.rep_bytes += (((uint64_t)j.rd_len * j.ncol + 31) / 32) * 4;   // u32 words (atomicOr)
		a.bound_off = ch.bound_words; ch.bound_words += 2 * j.ncol;
		a.escr_off = ch.escr_edits; ch.escr_edits += j.rd_len + j.ncol + 4;
		if((int)j.ncol > ch.maxcol) ch.maxcol = (int)j.ncol;
		aux[i] = a;
		d->cells += (uint64_t)j.rd_len * j.ncol;
	}
	if(pass == 0 && d->chunks.empty() && ch.mat_words * 4 <= d->mat.cap) break;       // fits what is already there
	if(pass == 0) {
		size_t freeB = 0, totalB = 0;
		BT2_CUDA(cudaMemGetInfo(&freeB, &totalB));
		mat_budget_words = (uint64_t)((freeB + d->mat.cap) * 0.55) / 4;
		if(d->chunks.empty() && ch.mat_words <= mat_budget_words) break;               // one chunk, and it may grow
	}
	}
	close_chunk((uint32_t)njobs);
	uint64_t maxMat = 0, maxRep = 0; uint32_t maxBound = 0, maxEscr = 0;
	for(auto &c : d->chunks) {
		maxMat = std::max(maxMat, c.mat_words); maxRep = std::max(maxRep, c.rep_bytes);
		maxBound = std::max(maxBound, c.bound_words); maxEscr = std::max(maxEscr, c.escr_edits);
		d->maxcol_all = std::max(d->maxcol_all, c.maxcol);
	}
	if(d->seq.reserve(pool_len) || d->qual.reserve(pool_len) || d->win.reserve(win_len) ||
	   d->jobs.reserve(njobs * sizeof(bt2gpu_dp_job)) || d->aux.reserve(njobs * sizeof(DpAux)) ||
	   d->order.reserve(njobs * 4) || d->res.reserve(njobs * sizeof(bt2gpu_dp_result)) ||
	   d->counts.reserve(njobs * 4) || d->cand_off.reserve(njobs * 4) || d->fillptr.reserve(njobs * 4) ||
	   d->keys.reserve(cand_cap * 8) || d->cands.reserve(cand_cap * sizeof(bt2gpu_dp_cand)) ||
	   d->done.reserve(cand_cap * 4) || d->alns.reserve(njobs * d->max_aln * sizeof(bt2gpu_dp_aln)) ||
	   d->edits.reserve(edit_cap * sizeof(bt2gpu_edit)) || d->escr.reserve((size_t)maxEscr * sizeof(bt2gpu_edit)) ||
	   d->mat.reserve(maxMat * 4) || d->bound.reserve((size_t)maxBound * 4) || d->rep.reserve(maxRep) ||
	   d->scal.reserve(64) || d->h_scal.reserve(64)) return 1;
	cudaStream_t s = ctx->stream;
	BT2_CUDA(cudaMemcpyAsync(d->seq.p, seq_pool, pool_len, cudaMemcpyHostToDevice, s));
	BT2_CUDA(cudaMemcpyAsync(d->qual.p, qual_pool, pool_len, cudaMemcpyHostToDevice, s));
	BT2_CUDA(cudaMemcpyAsync(d->win.p, win_pool, win_len, cudaMemcpyHostToDevice, s));
	BT2_CUDA(cudaMemcpyAsync(d->jobs.p, jobs, njobs * sizeof(bt2gpu_dp_job), cudaMemcpyHostToDevice, s));
	BT2_CUDA(cudaMemcpyAsync(d->aux.p, aux.data(), njobs * sizeof(DpAux), cudaMemcpyHostToDevice, s));
	BT2_CUDA(cudaMemcpyAsync(d->order.p, order.data(), njobs * 4, cudaMemcpyHostToDevice, s));
	// aux/order are stack-owned host vectors.  Pageable sources have been copied out by the driver when cudaMemcpyAsync
	// returns; the caller's pools, if pinned, must outlive the copies: inside bt2gpu_dp_align_batch they do (it returns after
	// its fetch), a separate bt2gpu_dp_stage waits here
	if(!ctx->in_batch) BT2_CUDA(bt2::wait_stream(ctx, s));
	return 0;
}

extern "C" int bt2gpu_dp_run(bt2gpu_ctx *ctx, uint32_t phases) {
	if(!ctx || !ctx->dp) BT2_FAIL("nothing staged");
	BT2_CUDA(cudaSetDevice(ctx->device));
	DpState *d = ctx->dp;
	if(d->njobs == 0) return 0;
	cudaStream_t s = ctx->stream;
	if(d->chunks.size() > 1 && phases != 7) BT2_FAIL("partial phases need a single-chunk batch");
	if(ctx->kstart()) return 1;
	BT2_CUDA(cudaMemsetAsync(d->scal.p, 0, 64, s));
	unsigned long long *scal = d->scal.as<unsigned long long>();
	int *errflag = reinterpret_cast<int*>(scal + 2);
	const bool local = !ctx->sc.monotone;
	const bool timed = d->chunks.size() == 1 && ctx->timing;
	if(timed) {
		for(int i = 0; i < 4; i++) if(!d->ev[i]) BT2_CUDA(cudaEventCreate(&d->ev[i]));
		d->ev_phases = phases;
		BT2_CUDA(cudaEventRecord(d->ev[0], s));
	} else d->ev_phases = 0;
	for(const DpChunk &c : d->chunks) {
		if(phases & 1) {
			for(int k = 0; k < kNClasses; k++) {
				uint32_t n = c.cls_off[k + 1] - c.cls_off[k];
				if(dispatch_fill(ctx, d, k, d->order.as<uint32_t>() + c.cls_off[k], n)) return 1;
			}
		}
		if(timed) BT2_CUDA(cudaEventRecord(d->ev[1], s));
		if((phases & 2) && !local && c.maxcol <= E2E_CAP) {
			if(c.maxcol <= 512)
				k_dp_gather_e2e<512><<<(c.njobs + 3) / 4, 128, 0, s>>>(d->jobs.as<bt2gpu_dp_job>(), d->aux.as<DpAux>(), c.job0, c.njobs,
					d->bound.as<uint32_t>(), d->res.as<bt2gpu_dp_result>(), d->cands.as<bt2gpu_dp_cand>(), scal,
					(unsigned long long)d->cand_cap, errflag);
			else
				k_dp_gather_e2e<E2E_CAP><<<(c.njobs + 3) / 4, 128, 0, s>>>(d->jobs.as<bt2gpu_dp_job>(), d->aux.as<DpAux>(), c.job0, c.njobs,
					d->bound.as<uint32_t>(), d->res.as<bt2gpu_dp_result>(), d->cands.as<bt2gpu_dp_cand>(), scal,
					(unsigned long long)d->cand_cap, errflag);
			ctx->launches++;
			BT2_CUDA(cudaGetLastError());
		} else if(phases & 2) {
			const uint32_t gb = (c.njobs + 3) / 4;
			const size_t gsm = local ? (size_t)4 * 2 * c.maxcol * sizeof(int) : 0;
			if(gsm > 48 * 1024) {
				BT2_CUDA(cudaFuncSetAttribute(k_dp_gather<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsm));
				BT2_CUDA(cudaFuncSetAttribute(k_dp_gather<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsm));
			}
			k_dp_gather<false><<<gb, 128, gsm, s>>>(
				d->seq.as<uint8_t>(), d->qual.as<uint8_t>(), d->win.as<uint8_t>(), d->jobs.as<bt2gpu_dp_job>(),
				d->aux.as<DpAux>(), c.job0, c.njobs, d->mat.as<uint32_t>(), d->res.as<bt2gpu_dp_result>(),
				d->counts.as<uint32_t>(), d->cand_off.as<uint32_t>(), d->keys.as<uint64_t>(), d->fillptr.as<uint32_t>(), c.maxcol);
			k_dp_scan<<<1, 1024, 0, s>>>(d->counts.as<uint32_t>(), c.njobs, c.job0, d->cand_off.as<uint32_t>(),
				d->fillptr.as<uint32_t>(), d->res.as<bt2gpu_dp_result>(), scal, (unsigned long long)d->cand_cap, errflag);
			k_dp_gather<true><<<gb, 128, gsm, s>>>(
				d->seq.as<uint8_t>(), d->qual.as<uint8_t>(), d->win.as<uint8_t>(), d->jobs.as<bt2gpu_dp_job>(),
				d->aux.as<DpAux>(), c.job0, c.njobs, d->mat.as<uint32_t>(), d->res.as<bt2gpu_dp_result>(),
				d->counts.as<uint32_t>(), d->cand_off.as<uint32_t>(), d->keys.as<uint64_t>(), d->fillptr.as<uint32_t>(), c.maxcol);
			k_dp_sort<<<c.njobs, 256, 4096 * 8, s>>>(c.job0, d->res.as<bt2gpu_dp_result>(), d->keys.as<uint64_t>(), d->cands.as<bt2gpu_dp_cand>());
			ctx->launches += 4;
			BT2_CUDA(cudaGetLastError());
		}
		if(timed) BT2_CUDA(cudaEventRecord(d->ev[2], s));
		if(phases & 4) {
			BT2_CUDA(cudaMemsetAsync(d->rep.p, 0, c.rep_bytes, s));
			k_dp_backtrace<<<(c.njobs + 3) / 4, 128, 0, s>>>(
				d->seq.as<uint8_t>(), d->qual.as<uint8_t>(), d->win.as<uint8_t>(), d->jobs.as<bt2gpu_dp_job>(),
				d->aux.as<DpAux>(), c.job0, c.njobs, d->mat.as<uint32_t>(), d->res.as<bt2gpu_dp_result>(),
				d->cands.as<bt2gpu_dp_cand>(), d->rep.as<uint8_t>(), d->done.as<uint32_t>(), d->alns.as<bt2gpu_dp_aln>(),
				d->escr.as<bt2gpu_edit>(), d->edits.as<bt2gpu_edit>(), scal + 1, (unsigned long long)d->edit_cap,
				d->max_aln, errflag);
			ctx->launches++;
			BT2_CUDA(cudaGetLastError());
		}
		if(timed) BT2_CUDA(cudaEventRecord(d->ev[3], s));
	}
	if(ctx->kend()) return 1;
	return 0;
}

extern "C" int bt2gpu_dp_phase_ms(bt2gpu_ctx *ctx, float *ms3) {
	if(!ctx || !ctx->dp || !ms3) BT2_FAIL("nothing staged");
	DpState *d = ctx->dp;
	ms3[0] = ms3[1] = ms3[2] = 0.f;
	if(!d->ev_phases) BT2_FAIL("phase times are recorded for single-chunk runs only");
	BT2_CUDA(cudaEventSynchronize(d->ev[3]));
	for(int i = 0; i < 3; i++) BT2_CUDA(cudaEventElapsedTime(&ms3[i], d->ev[i], d->ev[i + 1]));
	return 0;
}

extern "C" int bt2gpu_dp_fetch(
	bt2gpu_ctx *ctx, bt2gpu_dp_result *results, bt2gpu_dp_cand *cands, bt2gpu_dp_aln *alns,
	bt2gpu_edit *edits, size_t *ncands, size_t *nedits)
{
	if(!ctx || !ctx->dp) BT2_FAIL("nothing staged");
	BT2_CUDA(cudaSetDevice(ctx->device));
	DpState *d = ctx->dp;
	cudaStream_t s = ctx->stream;
	if(ncands) *ncands = 0;
	if(nedits) *nedits = 0;
	if(d->njobs == 0) return 0;
	BT2_CUDA(bt2::settle(ctx, s));
	BT2_CUDA(cudaMemcpyAsync(d->h_scal.p, d->scal.p, 64, cudaMemcpyDeviceToHost, s));
	if(results) BT2_CUDA(cudaMemcpyAsync(results, d->res.p, d->njobs * sizeof(bt2gpu_dp_result), cudaMemcpyDeviceToHost, s));
	if(alns) BT2_CUDA(cudaMemcpyAsync(alns, d->alns.p, d->njobs * d->max_aln * sizeof(bt2gpu_dp_aln), cudaMemcpyDeviceToHost, s));
	BT2_CUDA(bt2::wait_stream(ctx, s));
	const unsigned long long *hs = d->h_scal.as<unsigned long long>();
	const int err = *reinterpret_cast<const int*>(hs + 2);
	if(err == 1) BT2_FAIL("candidate pool overflow (raise cand_cap)");
	if(err == 2) BT2_FAIL("edit pool overflow (raise edit_cap)");
	if(cands && hs[0]) BT2_CUDA(cudaMemcpyAsync(cands, d->cands.p, hs[0] * sizeof(bt2gpu_dp_cand), cudaMemcpyDeviceToHost, s));
	if(edits && hs[1]) BT2_CUDA(cudaMemcpyAsync(edits, d->edits.p, hs[1] * sizeof(bt2gpu_edit), cudaMemcpyDeviceToHost, s));
	BT2_CUDA(bt2::wait_stream(ctx, s));
	if(ncands) *ncands = (size_t)hs[0];
	if(nedits) *nedits = (size_t)hs[1];
	return 0;
}

extern "C" uint64_t bt2gpu_dp_staged_cells(bt2gpu_ctx *ctx) { return (ctx && ctx->dp) ? ctx->dp->cells : 0; }
/* matrix cells the nextAlignment walks visited in the last fetched run (the reference's SSEMetrics.btcell) */
extern "C" uint64_t bt2gpu_dp_walked_cells(bt2gpu_ctx *ctx) { return (ctx && ctx->dp && ctx->dp->h_scal.p) ? ctx->dp->h_scal.as<unsigned long long>()[3] : 0; }

extern "C" int bt2gpu_dp_align_batch(
	bt2gpu_ctx *ctx, const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const uint8_t *win_pool, size_t win_len, const bt2gpu_dp_job *jobs, size_t njobs, uint32_t max_aln,
	bt2gpu_dp_result *results, bt2gpu_dp_cand *cands, size_t cand_cap, bt2gpu_dp_aln *alns,
	bt2gpu_edit *edits, size_t edit_cap)
{
	if(!ctx) BT2_FAIL("null ctx");
	ctx->in_batch = true;
	int rc = bt2gpu_dp_stage(ctx, seq_pool, qual_pool, pool_len, win_pool, win_len, jobs, njobs, max_aln, cand_cap, edit_cap);
	ctx->in_batch = false;
	if(rc) return 1;
	if(bt2gpu_dp_run(ctx, 7)) return 1;
	return bt2gpu_dp_fetch(ctx, results, cands, alns, edits, nullptr, nullptr);
}

namespace bt2 {

static int dp_reserve_common(DpState *d, size_t njobs, size_t pool_len, size_t win_len) {
	if(d->seq.reserve(pool_len) || d->qual.reserve(pool_len) || d->win.reserve(win_len) ||
	   d->jobs.reserve(njobs * sizeof(bt2gpu_dp_job)) || d->aux.reserve(njobs * sizeof(DpAux)) ||
	   d->order.reserve(njobs * 4) || d->res.reserve(njobs * sizeof(bt2gpu_dp_result)) ||
	   d->counts.reserve(njobs * 4) || d->cand_off.reserve(njobs * 4) || d->fillptr.reserve(njobs * 4) ||
	   d->keys.reserve(d->cand_cap * 8) || d->cands.reserve(d->cand_cap * sizeof(bt2gpu_dp_cand)) ||
	   d->done.reserve(d->cand_cap * 4) || d->alns.reserve(njobs * d->max_aln * sizeof(bt2gpu_dp_aln)) ||
	   d->edits.reserve(d->edit_cap * sizeof(bt2gpu_edit)) || d->scal.reserve(64) || d->h_scal.reserve(64)) return 1;
	return 0;
}

int dp_dev_begin(bt2gpu_ctx *ctx, size_t njobs, size_t pool_len, size_t win_len, uint32_t max_aln, size_t cand_cap,
                 size_t edit_cap, DpDevPools *pools)
{
	if(!ctx->have_sc) BT2_FAIL("bt2gpu_scoring_set must be called first");
	BT2_CUDA(cudaSetDevice(ctx->device));
	if(!ctx->dp) ctx->dp = new DpState();
	DpState *d = ctx->dp;
	d->njobs = njobs; d->max_aln = max_aln ? max_aln : 1; d->cand_cap = cand_cap; d->edit_cap = edit_cap;
	d->chunks.clear(); d->cells = 0; d->maxcol_all = 0;
	d->dev_pool_len = pool_len; d->dev_win_len = win_len;
	if(njobs > 0x7fffffffu / d->max_aln) BT2_FAIL("too many jobs");
	if(2 * (uint64_t)win_len > 0x70000000u || (uint64_t)pool_len + win_len + 4 * njobs > 0x70000000u)
		BT2_FAIL("DP batch too large for 32-bit scratch offsets: split the batch");
	if(dp_reserve_common(d, njobs, pool_len + 16, win_len + 16)) return 1;
	if(d->cls.reserve(njobs + 16) || d->sizes.reserve((njobs + 1) * sizeof(DpSizes)) || d->soffs.reserve((njobs + 1) * sizeof(DpSizes)) ||
	   d->spart.reserve((scan_parts(njobs) + 2) * sizeof(DpSizes)) || d->dsum.reserve(DPSUM_WORDS * 4 + sizeof(DpSizes)) ||
	   d->h_dsum.reserve(DPSUM_WORDS * 4 + sizeof(DpSizes))) return 1;
	pools->seq = d->seq.as<uint8_t>(); pools->qual = d->qual.as<uint8_t>(); pools->win = d->win.as<uint8_t>();
	pools->jobs = d->jobs.as<bt2gpu_dp_job>();
	return 0;
}

int dp_dev_finish(bt2gpu_ctx *ctx) {
	DpState *d = ctx->dp;
	if(!d) BT2_FAIL("nothing staged");
	const size_t njobs = d->njobs;
	if(njobs == 0) return 0;
	cudaStream_t s = ctx->stream;
	const bool local = !ctx->sc.monotone;
	uint32_t *sum = d->dsum.as<uint32_t>();
	DpSizes *total = reinterpret_cast<DpSizes*>(sum + DPSUM_WORDS);
	BT2_CUDA(cudaMemsetAsync(sum, 0, DPSUM_WORDS * 4, s));
	const unsigned nb = (unsigned)((njobs + 255) / 256);
	k_dp_aux<<<nb, 256, 0, s>>>(d->jobs.as<bt2gpu_dp_job>(), (uint32_t)njobs, d->dev_pool_len, d->dev_win_len, local ? 1 : 0,
		d->aux.as<DpAux>(), d->cls.as<uint8_t>(), d->sizes.as<DpSizes>(), sum);
	if(dev_exclusive_scan<DpSizes>(ctx, d->sizes.as<DpSizes>(), d->soffs.as<DpSizes>(), d->spart.as<DpSizes>(), total, njobs)) return 1;
	k_dp_aux_apply<<<nb, 256, 0, s>>>(d->aux.as<DpAux>(), d->soffs.as<DpSizes>(), (uint32_t)njobs);
	k_dp_order<<<nb, 256, 0, s>>>(d->cls.as<uint8_t>(), (uint32_t)njobs, sum, sum + 16, d->order.as<uint32_t>());
	ctx->launches += 3;
	BT2_CUDA(cudaGetLastError());
	BT2_CUDA(cudaMemcpyAsync(d->h_dsum.p, sum, DPSUM_WORDS * 4 + sizeof(DpSizes), cudaMemcpyDeviceToHost, s));
	BT2_CUDA(cudaStreamSynchronize(s));
	const uint32_t *hs = d->h_dsum.as<uint32_t>();
	switch(hs[9]) {
	case 0: break;
	case 1: BT2_FAIL("empty DP problem");
	case 2: BT2_FAIL("job outside pool");
	case 3: BT2_FAIL("DP problem too large");
	case 4: BT2_FAIL("end-to-end read longer than 2270 rows: checkpointed path (rdlen >= cminlen, aligner_sw.cpp:514) not implemented");
	case 5: BT2_FAIL("local read longer than 16000 rows not supported by the s16 domain");
	default: BT2_FAIL("minsc below the 16-bit end-to-end domain");
	}
	DpSizes tot; memcpy(&tot, hs + DPSUM_WORDS, sizeof(tot));
	if(tot.mat * 4 > d->mat.cap) {          // the arena has to grow: check the budget first (cudaMemGetInfo is slow, so only then)
		size_t freeB = 0, totalB = 0;
		BT2_CUDA(cudaMemGetInfo(&freeB, &totalB));
		const uint64_t mat_budget_words = (uint64_t)((freeB + d->mat.cap) * 0.55) / 4;
		if(tot.mat > mat_budget_words) BT2_FAIL("DP batch exceeds the matrix arena: split the batch");
	}
	DpChunk ch; memset(&ch, 0, sizeof(ch));
	ch.job0 = 0; ch.njobs = (uint32_t)njobs; ch.mat_words = tot.mat; ch.rep_bytes = tot.rep;
	ch.bound_words = tot.bound; ch.escr_edits = tot.escr; ch.maxcol = (int)hs[8];
	uint32_t pos = 0;
	for(int c = 0; c < kNClasses; c++) { ch.cls_off[c] = pos; pos += hs[c]; }
	ch.cls_off[kNClasses] = pos;
	if(pos != njobs) BT2_FAIL("internal: class histogram does not cover the batch");
	d->chunks.push_back(ch);
	d->maxcol_all = ch.maxcol;
	memcpy(&d->cells, hs + 10, 8);
	if(d->escr.reserve((size_t)tot.escr * sizeof(bt2gpu_edit)) || d->mat.reserve(tot.mat * 4) ||
	   d->bound.reserve((size_t)tot.bound * 4) || d->rep.reserve(tot.rep)) return 1;
	return 0;
}

} // namespace bt2

namespace bt2 {
const bt2gpu_dp_result *dp_dev_results(bt2gpu_ctx *ctx) { return ctx->dp ? ctx->dp->res.as<bt2gpu_dp_result>() : nullptr; }
const bt2gpu_dp_aln *dp_dev_alns(bt2gpu_ctx *ctx) { return ctx->dp ? ctx->dp->alns.as<bt2gpu_dp_aln>() : nullptr; }
const bt2gpu_edit *dp_dev_edits(bt2gpu_ctx *ctx) { return ctx->dp ? ctx->dp->edits.as<bt2gpu_edit>() : nullptr; }
void dp_state_free(DpState *d) {
	if(!d) return;
	DevBuf *bufs[] = { &d->seq, &d->qual, &d->win, &d->jobs, &d->aux, &d->order, &d->res, &d->counts, &d->cand_off,
	                   &d->fillptr, &d->keys, &d->cands, &d->done, &d->alns, &d->edits, &d->escr, &d->mat, &d->bound,
	                   &d->rep, &d->scal, &d->cls, &d->sizes, &d->soffs, &d->spart, &d->dsum };
	for(DevBuf *b : bufs) b->release();
	d->h_scal.release(); d->h_dsum.release();
	for(int i = 0; i < 4; i++) if(d->ev[i]) cudaEventDestroy(d->ev[i]);
	delete d;
}
}

// ---------------------------------------------------------------------------------------
// SwAligner::ungappedAlign (aligner_sw.cpp:286-494): the read laid on one diagonal.  A warp per job.
//   pass 1  lanes score 32 positions at a time (coalesced byte loads of read, quality, window).  End-to-end: every
//           term is <= 0 and the N count only grows, so the reference's per-prefix bail-out (score < minsc || ns >
//           nceil, :391-396) is the test on the totals -- two warp reductions.  Local: the running score with its
//           floor at 0 and the (scoreMax, rowf, rowi, sols) bookkeeping (:399-424) is a sequential state machine;
//           the lanes' terms are broadcast one by one and every lane steps the same scalar state (no divergence).
//   pass 2  mismatch edits of rows rowi..rowf, compacted with ballots, written in their final form (after
//           AlnRes::setShape's trim shift and, for the '-' strand, invertEdits) into a range of the edit pool
//           reserved with one atomic per job.
__global__ void __launch_bounds__(128)
k_ungapped(const uint8_t *__restrict__ seq, const uint8_t *__restrict__ qual, const uint8_t *__restrict__ win,
           const bt2gpu_ungap_job *__restrict__ jobs, uint32_t njobs, bt2gpu_ungap_result *__restrict__ res,
           bt2gpu_edit *__restrict__ edits, unsigned long long *__restrict__ nedits, unsigned long long edit_cap,
           int *__restrict__ errflag)
{
	const uint32_t FULL = 0xffffffffu;
	const int lane = threadIdx.x & 31;
	const uint32_t ji = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if(ji >= njobs) return;
	const bt2gpu_ungap_job j = jobs[ji];
	bt2gpu_ungap_result r;
	r.ret = 0; r.score = r.ns = r.refns = 0; r.refoff = 0; r.trim5 = r.trim3 = r.nedits = r.edit_off = 0;
	const int len = (int)j.rd_len;
	const uint8_t *rd = seq + j.rd_off, *qu = qual + j.rd_off, *rf = win + j.rf_off;
	const long long rfi = j.off, rff = j.off + len;
	long long leftNs = 0, rightNs = 0;
	bool ok = true;
	if(rfi < 0) { if(j.ohang) leftNs = -rfi; else ok = false; }
	if(rff > j.reflen) { if(j.ohang) rightNs = rff - j.reflen; else ok = false; }
	if(leftNs + rightNs > (long long)j.nceil) ok = false;
	int rowi = 0, rowf = len - 1, score = 0, ns = 0;
	if(ok) {
		if(c_sc.monotone) {
			int tot = 0;
			for(int i = lane; i < len; i += 32) {
				const int rdc = rd[i], c = rf[i];
				ns += (rdc > 3 || c > 3);
				tot += dev_score(rdc, c, (int)qu[i] - 33);
			}
			for(int o = 16; o; o >>= 1) { tot += __shfl_xor_sync(FULL, tot, o); ns += __shfl_xor_sync(FULL, ns, o); }
			score = tot;
			ok = !(score < j.minsc || ns > j.nceil);
		} else {
			int scoreMax = 0, lastfloor = 0, sols = 0, run = 0;
			rowi = -1;
			for(int base = 0; base < len; base += 32) {
				const int i = base + lane;
				int s = 0, isn = 0;
				if(i < len) {
					const int rdc = rd[i], c = rf[i];
					isn = (rdc > 3 || c > 3);
					s = dev_score(rdc, c, (int)qu[i] - 33);
				}
				ns += __popc(__ballot_sync(FULL, isn));
				const int cnt = min(32, len - base);
				for(int k = 0; k < cnt; k++) {
					run += __shfl_sync(FULL, s, k);
					if(run >= j.minsc && run >= scoreMax) {
						scoreMax = run; rowf = base + k;
						if(rowi != lastfloor) { rowi = lastfloor; sols++; }
					}
					if(run <= 0) { run = 0; lastfloor = base + k + 1; }
				}
			}
			if(ns > j.nceil || scoreMax < j.minsc || rowi < 0) ok = false;
			else if(sols > 1) { ok = false; r.ret = -1; }
			score = scoreMax;
		}
	}
	if(ok) {
		// count, reserve, write
		uint32_t n = 0, refns = 0;
		for(int base = rowi; base <= rowf; base += 32) {
			const int i = base + lane;
			bool e = false, rn = false;
			if(i <= rowf) { const int c = rf[i]; rn = c > 3; e = rn || rd[i] != c; }
			n += __popc(__ballot_sync(FULL, e));
			refns += __popc(__ballot_sync(FULL, rn));
		}
		unsigned long long e0 = 0;
		if(lane == 0 && n) e0 = atomicAdd(nedits, (unsigned long long)n);
		e0 = __shfl_sync(FULL, e0, 0);
		if(e0 + n > edit_cap) { if(lane == 0) { *errflag = 2; res[ji] = r; } return; }
		uint32_t k0 = 0;
		for(int base = rowi; base <= rowf; base += 32) {
			const int i = base + lane;
			bool e = false; int c = 0, rdc = 0;
			if(i <= rowf) { c = rf[i]; rdc = rd[i]; e = c > 3 || rdc != c; }
			const uint32_t m = __ballot_sync(FULL, e);
			if(e) {
				const uint32_t k = k0 + __popc(m & ((1u << lane) - 1u));
				bt2gpu_edit ed;
				ed.pos = j.fw ? (uint32_t)(i - rowi) : (uint32_t)(rowf - i);
				ed.chr = (uint8_t)"ACGTN"[c]; ed.qchr = (uint8_t)"ACGTN"[rdc];
				ed.type = BT2GPU_EDIT_MM; ed.pad = 0;
				edits[e0 + (j.fw ? k : n - 1 - k)] = ed;
			}
			k0 += __popc(m);
		}
		const int trimEnd = (len - 1) - rowf;
		r.ret = 1; r.score = score; r.ns = ns; r.refns = (int32_t)refns;
		r.refoff = j.off + rowi;
		r.trim5 = (uint32_t)(j.fw ? rowi : trimEnd); r.trim3 = (uint32_t)(j.fw ? trimEnd : rowi);
		r.nedits = n; r.edit_off = (uint32_t)e0;
	}
	if(lane == 0) res[ji] = r;
}

extern "C" int bt2gpu_ungapped_batch(
	bt2gpu_ctx *ctx, const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const uint8_t *win_pool, size_t win_len, const bt2gpu_ungap_job *jobs, size_t njobs,
	bt2gpu_ungap_result *results, bt2gpu_edit *edits, size_t edit_cap, size_t *nedits)
{
	using namespace bt2;
	if(!ctx) BT2_FAIL("null context");
	if(!ctx->have_sc) BT2_FAIL("bt2gpu_scoring_set has not been called");
	if(nedits) *nedits = 0;
	if(njobs == 0) return 0;
	if(!seq_pool || !qual_pool || !win_pool || !jobs || !results || (!edits && edit_cap)) BT2_FAIL("null argument");
	if(njobs > 0x7fffffffu) BT2_FAIL("too many jobs");
	for(size_t i = 0; i < njobs; i++) {
		const bt2gpu_ungap_job& j = jobs[i];
		if(j.rd_len == 0) BT2_FAIL("empty read");
		if((size_t)j.rd_off + j.rd_len > pool_len || (size_t)j.rf_off + j.rd_len > win_len) BT2_FAIL("job outside its pool");
	}
	BT2_CUDA(cudaSetDevice(ctx->device));
	DevBuf buf;
	auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
	const size_t o_seq = 0, o_qual = o_seq + al(pool_len), o_win = o_qual + al(pool_len), o_jobs = o_win + al(win_len),
	             o_res = o_jobs + al(njobs * sizeof(bt2gpu_ungap_job)), o_ed = o_res + al(njobs * sizeof(bt2gpu_ungap_result)),
	             o_scal = o_ed + al(edit_cap * sizeof(bt2gpu_edit)), total = o_scal + 256;
	if(buf.reserve(total)) return 1;
	uint8_t *b = buf.as<uint8_t>();
	int rc = 0;
	do {
		cudaStream_t s = ctx->stream;
		cudaError_t e = cudaMemcpyAsync(b + o_seq, seq_pool, pool_len, cudaMemcpyHostToDevice, s);
		if(e == cudaSuccess) e = cudaMemcpyAsync(b + o_qual, qual_pool, pool_len, cudaMemcpyHostToDevice, s);
		if(e == cudaSuccess) e = cudaMemcpyAsync(b + o_win, win_pool, win_len, cudaMemcpyHostToDevice, s);
		if(e == cudaSuccess) e = cudaMemcpyAsync(b + o_jobs, jobs, njobs * sizeof(bt2gpu_ungap_job), cudaMemcpyHostToDevice, s);
		if(e == cudaSuccess) e = cudaMemsetAsync(b + o_scal, 0, 256, s);
		if(e != cudaSuccess) { rc = fail("ungapped H2D", __FILE__, __LINE__, e); break; }
		ctx->kstart();
		const unsigned wpb = 4;
		k_ungapped<<<(unsigned)((njobs + wpb - 1) / wpb), wpb * 32, 0, s>>>(
			b + o_seq, b + o_qual, b + o_win, reinterpret_cast<const bt2gpu_ungap_job*>(b + o_jobs), (uint32_t)njobs,
			reinterpret_cast<bt2gpu_ungap_result*>(b + o_res), reinterpret_cast<bt2gpu_edit*>(b + o_ed),
			reinterpret_cast<unsigned long long*>(b + o_scal), (unsigned long long)edit_cap,
			reinterpret_cast<int*>(b + o_scal + 8));
		ctx->launches++;
		ctx->kend();
		unsigned long long scal[2] = {0, 0};
		e = cudaMemcpyAsync(results, b + o_res, njobs * sizeof(bt2gpu_ungap_result), cudaMemcpyDeviceToHost, s);
		if(e == cudaSuccess) e = cudaMemcpyAsync(scal, b + o_scal, sizeof scal, cudaMemcpyDeviceToHost, s);
		if(e == cudaSuccess) e = cudaStreamSynchronize(s);
		if(e != cudaSuccess) { rc = fail("ungapped", __FILE__, __LINE__, e); break; }
		if((int)(scal[1] & 0xffffffffu) == 2 || scal[0] > edit_cap) { rc = fail("edit pool overflow (raise edit_cap)", __FILE__, __LINE__); break; }
		if(scal[0]) {
			e = cudaMemcpy(edits, b + o_ed, (size_t)scal[0] * sizeof(bt2gpu_edit), cudaMemcpyDeviceToHost);
			if(e != cudaSuccess) { rc = fail("ungapped D2H", __FILE__, __LINE__, e); break; }
		}
		if(nedits) *nedits = (size_t)scal[0];
	} while(0);
	buf.release();
	return rc;
}

// ---------------------------------------------------------------------------------------
// bt2gpu_ungapped_batch_noalloc: the same kernel and the same host-buffer signature as bt2gpu_ungapped_batch, for callers
// that may not allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  No counterpart in the reference.
//
// bt2gpu_ungapped_batch cudaMallocs its scratch arena per call (staged pools, job table, results, edit pool, and the
// counter block the kernel reserves edit ranges from) and frees it afterwards.  Here the arena is a file-scope device
// array of BT2GPU_NOALLOC_SCRATCH_BYTES; every GPU that loads the library holds it until the process exits.  The call
// makes no cudaMalloc / cudaFree and no allocating library call; a batch whose arena does not fit is rejected.
//
// In HOST code the array's name is only a host-side stand-in for the symbol, not a device address: its device address
// is asked for once per device, with that device current (cudaGetSymbolAddress), and that pointer is what the kernel
// and the copies get.  All calls on a device share the array and it keeps the previous call's contents: the counter block
// is cleared on the stream before every launch (everything else is overwritten before it is read), and two calls that
// overlap on the GPU (different contexts / streams of one device) RACE -- serialise them.
// ---------------------------------------------------------------------------------------
#include <mutex>
static __device__ __align__(256) unsigned char g_noalloc_scratch[BT2GPU_NOALLOC_SCRATCH_BYTES];

namespace bt2 {
static std::mutex g_noalloc_mu;
static unsigned char *g_noalloc_ptr[64];     // device address of g_noalloc_scratch, per device ordinal

// set-up, once per device, with that device current
static int noalloc_scratch(int device, unsigned char **out) {
	if(device < 0 || device >= 64) BT2_FAIL("device ordinal out of range");
	std::lock_guard<std::mutex> lk(g_noalloc_mu);
	if(!g_noalloc_ptr[device]) {
		BT2_CUDA(cudaSetDevice(device));
		void *p = nullptr;
		BT2_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
		if(!p) BT2_FAIL("cudaGetSymbolAddress returned a null scratch address");
		g_noalloc_ptr[device] = static_cast<unsigned char*>(p);
	}
	*out = g_noalloc_ptr[device];
	return 0;
}
}

extern "C" uint64_t bt2gpu_noalloc_scratch_bytes(void) { return BT2GPU_NOALLOC_SCRATCH_BYTES; }

extern "C" int bt2gpu_ungapped_batch_noalloc(
	bt2gpu_ctx *ctx, const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const uint8_t *win_pool, size_t win_len, const bt2gpu_ungap_job *jobs, size_t njobs,
	bt2gpu_ungap_result *results, bt2gpu_edit *edits, size_t edit_cap, size_t *nedits)
{
	using namespace bt2;
	if(!ctx) BT2_FAIL("null context");
	if(!ctx->have_sc) BT2_FAIL("bt2gpu_scoring_set has not been called");
	if(nedits) *nedits = 0;
	if(njobs == 0) return 0;
	if(!seq_pool || !qual_pool || !win_pool || !jobs || !results || (!edits && edit_cap)) BT2_FAIL("null argument");
	if(njobs > 0x7fffffffu) BT2_FAIL("too many jobs");
	for(size_t i = 0; i < njobs; i++) {
		const bt2gpu_ungap_job& j = jobs[i];
		if(j.rd_len == 0) BT2_FAIL("empty read");
		if((size_t)j.rd_off + j.rd_len > pool_len || (size_t)j.rf_off + j.rd_len > win_len) BT2_FAIL("job outside its pool");
	}
	// the arena of bt2gpu_ungapped_batch, laid out in the static array: the counter block first
	auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
	const size_t cap = BT2GPU_NOALLOC_SCRATCH_BYTES;
	if(pool_len > cap || win_len > cap || njobs > cap / sizeof(bt2gpu_ungap_job) || edit_cap > cap / sizeof(bt2gpu_edit))
		BT2_FAIL("batch exceeds the no-allocation scratch (BT2GPU_NOALLOC_SCRATCH_BYTES): split it or use bt2gpu_ungapped_batch");
	const size_t o_scal = 0, o_seq = 256, o_qual = o_seq + al(pool_len), o_win = o_qual + al(pool_len), o_jobs = o_win + al(win_len),
	             o_res = o_jobs + al(njobs * sizeof(bt2gpu_ungap_job)), o_ed = o_res + al(njobs * sizeof(bt2gpu_ungap_result)),
	             total = o_ed + al(edit_cap * sizeof(bt2gpu_edit));
	if(total > cap)
		BT2_FAIL("batch exceeds the no-allocation scratch (BT2GPU_NOALLOC_SCRATCH_BYTES): split it or use bt2gpu_ungapped_batch");
	unsigned char *b = nullptr;
	if(noalloc_scratch(ctx->device, &b)) return 1;
	BT2_CUDA(cudaSetDevice(ctx->device));
	cudaStream_t s = ctx->stream;
	BT2_CUDA(cudaMemsetAsync(b + o_scal, 0, 256, s));          // edit counter + error flag must start at zero
	BT2_CUDA(cudaMemcpyAsync(b + o_seq, seq_pool, pool_len, cudaMemcpyHostToDevice, s));
	BT2_CUDA(cudaMemcpyAsync(b + o_qual, qual_pool, pool_len, cudaMemcpyHostToDevice, s));
	BT2_CUDA(cudaMemcpyAsync(b + o_win, win_pool, win_len, cudaMemcpyHostToDevice, s));
	BT2_CUDA(cudaMemcpyAsync(b + o_jobs, jobs, njobs * sizeof(bt2gpu_ungap_job), cudaMemcpyHostToDevice, s));
	const unsigned wpb = 4;
	k_ungapped<<<(unsigned)((njobs + wpb - 1) / wpb), wpb * 32, 0, s>>>(
		b + o_seq, b + o_qual, b + o_win, reinterpret_cast<const bt2gpu_ungap_job*>(b + o_jobs), (uint32_t)njobs,
		reinterpret_cast<bt2gpu_ungap_result*>(b + o_res), reinterpret_cast<bt2gpu_edit*>(b + o_ed),
		reinterpret_cast<unsigned long long*>(b + o_scal), (unsigned long long)edit_cap,
		reinterpret_cast<int*>(b + o_scal + 8));
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	unsigned long long scal[2] = {0, 0};
	BT2_CUDA(cudaMemcpyAsync(results, b + o_res, njobs * sizeof(bt2gpu_ungap_result), cudaMemcpyDeviceToHost, s));
	BT2_CUDA(cudaMemcpyAsync(scal, b + o_scal, sizeof scal, cudaMemcpyDeviceToHost, s));
	BT2_CUDA(cudaStreamSynchronize(s));
	if((int)(scal[1] & 0xffffffffu) == 2 || scal[0] > edit_cap) BT2_FAIL("edit pool overflow (raise edit_cap)");
	if(scal[0]) {
		BT2_CUDA(cudaMemcpyAsync(edits, b + o_ed, (size_t)scal[0] * sizeof(bt2gpu_edit), cudaMemcpyDeviceToHost, s));
		BT2_CUDA(cudaStreamSynchronize(s));
	}
	if(nedits) *nedits = (size_t)scal[0];
	return 0;
}
