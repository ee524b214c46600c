// fm.cu — FM-index kernels (sm_100a) and their C ABI: exact seed search, exactSweep,
// oneMmSearch, seed-hit extension and SA-offset location over the reference's .bt2 layout.
//
// Replaces, behind include/bt2gpu.h, the LF primitives of Ebwt (countBt2SideEx / mapBiLFEx /
// mapLF1, bt2_idx.h:1758-2472) and their callers on the hot path: SeedAligner::searchAllSeeds
// (exact seeds), ::exactSweep, ::oneMmSearch (aligner_seed.cpp:597-1323), SwDriver::extend
// (aligner_sw_driver.cpp:299-484) and Ebwt::getOffset (bt2_idx.cpp:145-171).
//
// The reference walks one query at a time and hides DRAM latency with software prefetch
// (8 seeds in flight per thread, aligner_seed.cpp:628).  Here the unit is a QUAD of lanes per
// BWT side: the 64-byte side (48 B of 2-bit BWT + u32 occ[4]) is fetched as four coalesced
// 16-byte loads, every lane pop-counts its 64 bases for all four symbols, the packed counts
// are summed with two warp shuffles and the lane holding the occ[] words finishes the rank.
// An OCTET (two quads) resolves the top and bottom of a BW range in one step, so a warp
// advances four independent searches per instruction stream and each SM keeps hundreds of
// dependent 64-byte gathers in flight.  The index stays resident in HBM in the file's own
// layout; ftab/eftab/fchr are read through the same path (L2-resident in practice).
#include "common.cuh"
#include <algorithm>

namespace bt2 {

static constexpr uint32_t OFFMASK = 0xffffffffu;

struct IndexDev {
	const uint4    *sides;     // numSides x 4 x 16 bytes
	const uint32_t *offs;      // SA sample (forward index only)
	const uint32_t *ftab, *eftab;
	uint32_t fchr[5];
	uint32_t zoff, zside, zco; // '$' row, its side and char offset inside the side
	uint32_t len;
	uint32_t off_mask;         // OFF_MASK << offRate
	int off_rate, ftab_chars, is_fw;
};

} // namespace bt2

struct bt2gpu_index {
	int device = 0;
	bt2::DevBuf ebwt, offs, ftab, eftab;
	bt2::IndexDev dev;
	uint64_t bytes = 0;
	uint64_t nsides = 0;
};

namespace bt2 {

struct ScFm { int32_t match; };
__constant__ int32_t c_fm_mmpens[256];
__constant__ int32_t c_fm_npens[256];
__constant__ ScFm c_fm_sc;

// ---------------------------------------------------------------------------------------
// Rank primitive.  quad_lf: the 4 lanes of a quad (q = lane & 3) cooperatively evaluate
// countBt2SideEx (bt2_idx.h:1887-1919) for `row`: val[c] = fchr[c] + occ_side[c] +
// #c among the first (row % 192) bases - ('$' fix for c == A).  Also returns the BWT
// character at `row` (rowL, bt2_idx.h:2247).  All 4 lanes get all results.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pos_mask16(int m) {       // low m (0..16) base slots of a word
	return m >= 16 ? 0x55555555u : (((1u << (2 * m)) - 1u) & 0x55555555u);
}

__device__ __forceinline__ void quad_lf(const IndexDev &ix, uint32_t row, unsigned gmask, int q,
                                        uint32_t val[4], int &rowc)
{
	const uint32_t side = row / 192u;
	const int co = (int)(row - side * 192u);
	const uint4 v = __ldg(ix.sides + (size_t)side * 4 + q);
	uint32_t pk = 0;
	int myc = 0;
	if(q < 3) {
		int m = co - 64 * q;
		m = m < 0 ? 0 : (m > 64 ? 64 : m);
		const uint32_t w[4] = {v.x, v.y, v.z, v.w};
		uint32_t nC = 0, nG = 0, nT = 0;
#pragma unroll
		for(int k = 0; k < 4; k++) {
			int mw = m - 16 * k;
			mw = mw < 0 ? 0 : mw;
			const uint32_t pm = pos_mask16(mw);
			const uint32_t lo = w[k] & pm, hi = (w[k] >> 1) & pm;
			nT += __popc(lo & hi);
			nG += __popc(hi & ~lo);
			nC += __popc(lo & ~hi);
		}
		const uint32_t nA = (uint32_t)m - (nC + nG + nT);
		pk = nA | (nC << 8) | (nG << 16) | (nT << 24);
		const int ci = co - 64 * q;                               // my slot of the row's own char
		if(ci >= 0 && ci < 64) myc = (int)((w[ci >> 4] >> (2 * (ci & 15))) & 3u);
	}
	pk += __shfl_xor_sync(gmask, pk, 1);
	pk += __shfl_xor_sync(gmask, pk, 2);
	const uint32_t oA = __shfl_sync(gmask, v.x, 3, 4), oC = __shfl_sync(gmask, v.y, 3, 4);
	const uint32_t oG = __shfl_sync(gmask, v.z, 3, 4), oT = __shfl_sync(gmask, v.w, 3, 4);
	rowc = __shfl_sync(gmask, myc, co >> 6, 4);
	const uint32_t fix = (side == ix.zside && (uint32_t)co > ix.zco) ? 1u : 0u;
	val[0] = ix.fchr[0] + oA + (pk & 0xff) - fix;
	val[1] = ix.fchr[1] + oC + ((pk >> 8) & 0xff);
	val[2] = ix.fchr[2] + oG + ((pk >> 16) & 0xff);
	val[3] = ix.fchr[3] + oT + (pk >> 24);
}

// octet_lf: lanes 0-3 of an octet rank `top`, lanes 4-7 rank `bot` (mapLFEx / mapBiLFEx's two
// countBt2SideEx calls, bt2_idx.h:2372-2413); every lane gets t[4], b[4] and rowL(top).
__device__ __forceinline__ void octet_lf(const IndexDev &ix, uint32_t top, uint32_t bot, unsigned gmask,
                                         int l8, uint32_t t[4], uint32_t b[4], int &topc)
{
	const int h = l8 >> 2;
	uint32_t mine[4]; int c;
	quad_lf(ix, h ? bot : top, gmask, l8 & 3, mine, c);
#pragma unroll
	for(int k = 0; k < 4; k++) {
		const uint32_t other = __shfl_xor_sync(gmask, mine[k], 4);
		t[k] = h ? other : mine[k];
		b[k] = h ? mine[k] : other;
	}
	const int oc = __shfl_xor_sync(gmask, c, 4);
	topc = h ? oc : c;
}

// A search that follows one known symbol and has no mirror range to carry (exactSweep) does not need all four
// ranks (the LT form, which also serves mapBiLFEx's mirror update, measured slower than the four-symbol rank in the
// seed kernel -- 0.57 -> 0.61 ms -- and is kept for reference only).
// quad_lf_c: LF(row, c) alone -- one pop-count per word instead of three -- and, when LT, the number of
// BWT symbols below c before `row` plus their side totals (lt = sum over j < c of occ_side[j] + #j), which is all
// mapBiLFEx's mirror-range update needs: topb' = topb + sum_{j<c} (b[j] - t[j]), where fchr cancels
// (bt2_idx.h:2395-2411).  The '$' correction (stored as A, bt2_idx.h:1891-1899) applies to A's count.
template<bool LT>
__device__ __forceinline__ void quad_lf_c(const IndexDev &ix, uint32_t row, int c, unsigned gmask, int q,
                                          uint32_t &val, uint32_t &lt, int &rowc)
{
	const uint32_t side = row / 192u;
	const int co = (int)(row - side * 192u);
	const uint4 v = __ldg(ix.sides + (size_t)side * 4 + q);
	uint32_t pk = 0;
	int myc = 0;
	if(q < 3) {
		int m = co - 64 * q;
		m = m < 0 ? 0 : (m > 64 ? 64 : m);
		const uint32_t w[4] = {v.x, v.y, v.z, v.w};
		const uint32_t pat = (uint32_t)c * 0x55555555u;
		uint32_t nc = 0, nlt = 0;
#pragma unroll
		for(int k = 0; k < 4; k++) {
			int mw = m - 16 * k;
			mw = mw < 0 ? 0 : mw;
			const uint32_t pm = pos_mask16(mw);
			const uint32_t x = ~(w[k] ^ pat);
			nc += __popc(x & (x >> 1) & pm);
			if(LT) {
				const uint32_t lo = w[k], hi = w[k] >> 1;
				// symbol < c:  c=1: A;  c=2: high bit clear;  c=3: not T
				const uint32_t ltm = c == 1 ? ~(lo | hi) : (c == 2 ? ~hi : ~(lo & hi));
				nlt += __popc(ltm & pm);
			}
		}
		if(LT && c == 0) nlt = 0;
		pk = nc | (nlt << 16);
		const int ci = co - 64 * q;                               // my slot of the row's own char
		if(ci >= 0 && ci < 64) myc = (int)((w[ci >> 4] >> (2 * (ci & 15))) & 3u);
	}
	pk += __shfl_xor_sync(gmask, pk, 1);
	pk += __shfl_xor_sync(gmask, pk, 2);
	// lane 3 of the quad holds occ[A,C,G,T] of the side
	const uint32_t oc_mine = c == 0 ? v.x : (c == 1 ? v.y : (c == 2 ? v.z : v.w));
	const uint32_t oc = __shfl_sync(gmask, oc_mine, 3, 4);
	rowc = __shfl_sync(gmask, myc, co >> 6, 4);
	const uint32_t fix = (side == ix.zside && (uint32_t)co > ix.zco) ? 1u : 0u;
	val = ix.fchr[c] + oc + (pk & 0xffffu) - (c == 0 ? fix : 0u);
	if(LT) {
		const uint32_t ol_mine = (c > 0 ? v.x : 0u) + (c > 1 ? v.y : 0u) + (c > 2 ? v.z : 0u);
		const uint32_t ol = __shfl_sync(gmask, ol_mine, 3, 4);
		lt = ol + (pk >> 16) - (c > 0 ? fix : 0u);
	} else lt = 0;
}

// octet_lf_c: lanes 0-3 rank `top`, lanes 4-7 rank `bot` for the one symbol c; every lane gets LF(top,c), LF(bot,c),
// rowL(top) and, when LT, below = sum_{j<c} (LF(bot,j) - LF(top,j)).
template<bool LT>
__device__ __forceinline__ void octet_lf_c(const IndexDev &ix, uint32_t top, uint32_t bot, int c, unsigned gmask, int l8,
                                           uint32_t &tc, uint32_t &bc, uint32_t &below, int &topc)
{
	const int h = l8 >> 2;
	uint32_t mine, lt; int rc;
	quad_lf_c<LT>(ix, h ? bot : top, c, gmask, l8 & 3, mine, lt, rc);
	const uint32_t other = __shfl_xor_sync(gmask, mine, 4);
	tc = h ? other : mine; bc = h ? mine : other;
	if(LT) {
		const uint32_t olt = __shfl_xor_sync(gmask, lt, 4);
		below = h ? lt - olt : olt - lt;
	} else below = 0;
	const int oc = __shfl_xor_sync(gmask, rc, 4);
	topc = h ? oc : rc;
}

// v[c] without dynamic register indexing (keeps t[]/b[] out of local memory)
__device__ __forceinline__ uint32_t pick4(const uint32_t v[4], int c) {
	return c == 0 ? v[0] : (c == 1 ? v[1] : (c == 2 ? v[2] : v[3]));
}

// ftabHi / ftabLo (bt2_idx.h:1425-1554)
__device__ __forceinline__ uint32_t ftab_hi(const IndexDev &ix, uint32_t i) {
	const uint32_t v = __ldg(ix.ftab + i);
	return v <= ix.len ? v : __ldg(ix.eftab + (size_t)(v ^ OFFMASK) * 2 + 1);
}
__device__ __forceinline__ uint32_t ftab_lo(const IndexDev &ix, uint32_t i) {
	const uint32_t v = __ldg(ix.ftab + i);
	return v <= ix.len ? v : __ldg(ix.eftab + (size_t)(v ^ OFFMASK) * 2);
}

__device__ __forceinline__ int comp_dna(int c) { return c > 3 ? 4 : 3 - c; }

// One strand-view of a staged read: base i of patFw / patRc / patFwRev / patRcRev (read.h:125-131).
struct ReadView {
	const uint8_t *rd; int len; bool rc, rev;
	__device__ __forceinline__ int at(int i) const {
		const int j = (rc != rev) ? len - 1 - i : i;
		const int c = rd[j];
		return rc ? comp_dna(c) : c;
	}
};

// ftabSeqToInt (bt2_idx.h:1374-1398) over a ReadView; OFFMASK if an N is in the window.
__device__ __forceinline__ uint32_t ftab_key(const IndexDev &ix, const ReadView &s, int off, bool rev) {
	const int fc = ix.ftab_chars;
	bool fwex = ix.is_fw != 0;
	if(rev) fwex = !fwex;
	uint32_t v = 0;
	for(int i = 0; i < fc; i++) {
		const int c = fwex ? s.at(off + i) : s.at(off + fc - i - 1);
		if(c > 3) return OFFMASK;
		v = (v << 2) | (uint32_t)c;
	}
	return v;
}

} // namespace bt2

namespace bt2 {

// ---------------------------------------------------------------------------------------
// Kernel 1a: exact seeds (searchSeedBi for SEED_TYPE_EXACT, aligner_seed.cpp:1854-2033).
// A CTA stages a tile of seed jobs in shared memory (each seed window packed to 2 bits per
// base, already oriented to the Watson strand as instantiateSeq does, :471-491), then every
// octet walks its seeds right-to-left over the forward index with the mirror range riding
// along (mapBiLFEx's prefix sums).
// ---------------------------------------------------------------------------------------
static constexpr int SEED_BLOCK = 256;
static constexpr int SEED_TILE  = 256;

__global__ void __launch_bounds__(SEED_BLOCK)
k_seed_exact(const IndexDev fw, const IndexDev bw, const uint8_t *__restrict__ seq,
             const bt2gpu_read *__restrict__ reads, const bt2gpu_seed_job *__restrict__ jobs,
             uint32_t njobs, uint32_t tile_sz, bt2gpu_seed_hit *__restrict__ out)
{
	__shared__ unsigned long long s_seq[SEED_TILE];
	__shared__ int s_len[SEED_TILE];
	const int lane = threadIdx.x & 31, l8 = lane & 7;
	const unsigned gmask = 0xffu << (lane & 24);
	const int oct = threadIdx.x >> 3;                        // octet in block
	constexpr int NOCT = SEED_BLOCK / 8;
	for(uint32_t tile = blockIdx.x * tile_sz; tile < njobs; tile += gridDim.x * tile_sz) {
		__syncthreads();
		if(threadIdx.x < tile_sz) {   // stage: one job per thread
			const uint32_t ji = tile + threadIdx.x;
			unsigned long long p = 0; int len = 0;
			if(ji < njobs) {
				const bt2gpu_seed_job jb = jobs[ji];
				const uint8_t *rd = seq + reads[jb.read].rd_off + jb.off;
				len = jb.len;
				for(int i = 0; i < len; i++) {
					int c = jb.fw ? rd[i] : 3 - rd[len - 1 - i];
					p |= (unsigned long long)(c & 3) << (2 * i);
				}
			}
			s_seq[threadIdx.x] = p; s_len[threadIdx.x] = len;
		}
		__syncthreads();
		for(int k = oct; k < (int)tile_sz; k += NOCT) {
			const uint32_t ji = tile + k;
			if(ji >= njobs) break;
			const unsigned long long p = s_seq[k];
			const int len = s_len[k];
			uint32_t topf, botf, topb, botb, bwops = 0, nsides = 0;
			int step;
			const int fl = fw.ftab_chars;
			bool alive = true;
			if(fl > 1 && fl <= len) {
				// CacheAndSeed (aligner_seed.cpp:84-112): the mirror key consumes the window
				// left-to-right = the packed bits as they are; the forward key is its reversal
				const uint32_t kb = (uint32_t)(p >> (2 * (len - fl))) & ((1u << (2 * fl)) - 1u);
				uint32_t r = __brev(kb) >> (32 - 2 * fl);
				const uint32_t kf = ((r & 0x55555555u) << 1) | ((r >> 1) & 0x55555555u);
				topf = ftab_hi(fw, kf); botf = ftab_lo(fw, kf + 1);
				topb = 0; botb = 0;
				if(botf == topf) alive = false;
				else { topb = ftab_hi(bw, kb); botb = topb + (botf - topf); }
				step = fl;
			} else {
				const int c = (int)(p >> (2 * (len - 1))) & 3;
				topf = topb = fw.fchr[c]; botf = botb = fw.fchr[c + 1];
				if(botf == topf) alive = false;
				step = 1;
			}
			for(; alive && step < len; step++) {
				const int c = (int)(p >> (2 * (len - 1 - step))) & 3;
				uint32_t t[4], b[4]; int topc;
				const bool single = (botf - topf == 1);
				octet_lf(fw, topf, single ? topf : botf, gmask, l8, t, b, topc);
				bwops++;
				nsides += (single || topf / 192u == botf / 192u) ? 1 : 2;
				if(single) {                                    // mapLF1(row, l, c), bt2_idx.h:2420
					if(topc != c || topf == fw.zoff) { alive = false; break; }
					topf = pick4(t, c); botf = topf + 1;        // mirror range unchanged
				} else {                                        // mapBiLFEx, bt2_idx.h:2372
					uint32_t below = 0;
#pragma unroll
					for(int j = 0; j < 3; j++) if(j < c) below += b[j] - t[j];
					topf = pick4(t, c); botf = pick4(b, c);
					if(botf == topf) { alive = false; break; }
					topb += below; botb = topb + (botf - topf);
				}
			}
			if(l8 == 0) {
				bt2gpu_seed_hit h;
				if(alive) { h.topf = topf; h.botf = botf; h.topb = topb; h.botb = botb; }
				else { h.topf = h.botf = h.topb = h.botb = 0; }
				h.bwops = bwops; h.sides = nsides;
				out[ji] = h;
			}
		}
	}
}

// ---------------------------------------------------------------------------------------
// Kernel 1b: SeedAligner::exactSweep (aligner_seed.cpp:750-968).  One octet per
// (read, strand); the two strands are independent in the reference (it interleaves them
// only to overlap prefetches), so they run as separate octets.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_exact_sweep(const IndexDev ix, const uint8_t *__restrict__ seq, const bt2gpu_read *__restrict__ reads,
              uint32_t nreads, int nofw, int norc, uint32_t mine_max, bt2gpu_sweep_result *__restrict__ out)
{
	const int lane = threadIdx.x & 31, l8 = lane & 7;
	const unsigned gmask = 0xffu << (lane & 24);
	const uint32_t task = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
	if(task >= 2 * nreads) return;
	const uint32_t ri = task >> 1;
	const int fwi = (int)(task & 1);
	const bool skip = fwi == 0 ? (nofw != 0) : (norc != 0);
	ReadView s; s.rd = seq + reads[ri].rd_off; s.len = (int)reads[ri].rd_len; s.rc = (fwi == 1); s.rev = false;
	const int len = s.len, fl = ix.ftab_chars;
	uint32_t top = 0, bot = 0, bwops = 0, nedit = 0;
	int dep = 0;
	bool done = skip, do_init = true;
	while(dep < len && !done) {
		if(do_init) {                                           // exactSweepInit :750-790
			top = bot = 0;
			const int left = len - dep;
			bool do_ftab = fl > 1 && left >= fl;
			uint32_t key = 0;
			if(do_ftab) { key = ftab_key(ix, s, left - fl, false); if(key == OFFMASK) do_ftab = false; }
			if(do_ftab) { top = ftab_hi(ix, key); bot = ftab_lo(ix, key + 1); dep += fl; }
			else {
				const int c = s.at(len - dep - 1);
				if(c < 4) { top = ix.fchr[c]; bot = ix.fchr[c + 1]; }
				dep++;
			}
			if(bot <= top) {                                    // exactSweepStep :827-846
				nedit++;
				if(nedit >= mine_max) done = true;
				continue;
			}
			do_init = false;
		}
		if(dep < len) {                                         // exactSweepMapLF :792-824
			const int c = s.at(len - dep - 1);
			if(c > 3) top = bot = 0;
			else {
				uint32_t tc, bc, below; int topc;
				const bool single = (bot - top == 1);
				octet_lf_c<false>(ix, top, single ? top : bot, c, gmask, l8, tc, bc, below, topc);
				if(single) {
					bwops++;
					if(topc != c || top == ix.zoff) top = bot = 0;
					else { top = tc; bot = top + 1; }
				} else { bwops += 2; top = tc; bot = bc; }
			}
			if(bot <= top) {
				nedit++;
				if(nedit >= mine_max) done = true;
				do_init = true;
			}
			dep++;
		}
	}
	if(l8 == 0) {
		bt2gpu_sweep_result *o = out + ri;
		o->top[fwi] = top; o->bot[fwi] = bot;
		o->nedit[fwi] = (uint16_t)nedit; o->done[fwi] = done ? 1 : 0;
		atomicAdd(&o->bwops, bwops);
	}
}

// ---------------------------------------------------------------------------------------
// Kernel 1c: Ebwt::getOffset (bt2_idx.cpp:145-171) = what GroupWalk2S::advanceElement
// resolves per element (group_walk.h:1105-1216).  One quad per row.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_locate(const IndexDev ix, const uint32_t *__restrict__ rows, uint32_t n,
         uint32_t *__restrict__ offs_out, uint32_t *__restrict__ steps_out)
{
	const int lane = threadIdx.x & 31, q = lane & 3;
	const unsigned gmask = 0xfu << (lane & 28);
	const uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 2;
	if(i >= n) return;
	uint32_t row = rows[i], jumps = 0, res;
	for(;;) {
		if(row > ix.len) { res = OFFMASK; break; }             // not a BW row
		if(row == ix.zoff) { res = jumps; break; }
		if((row & ix.off_mask) == row) { res = jumps + __ldg(ix.offs + (row >> ix.off_rate)); break; }
		uint32_t val[4]; int c;
		quad_lf(ix, row, gmask, q, val, c);                      // mapLF(l), bt2_idx.h:2313
		row = pick4(val, c);
		jumps++;
	}
	if(q == 0) { offs_out[i] = res; if(steps_out) steps_out[i] = jumps; }
}

// ---------------------------------------------------------------------------------------
// Kernel 1d: SwDriver::extend (aligner_sw_driver.cpp:299-484).  One octet per seed hit.
// ---------------------------------------------------------------------------------------
// One step of SwDriver::extend's loop in one direction (:332-404 left over the forward index,
// :405-484 right over the mirror): false when the extension stops before this character.
__device__ __forceinline__ bool extend_step(const IndexDev &ix, const ReadView &s, uint32_t &top, uint32_t &bot,
                                            int i, unsigned gmask, int l8)
{
	const int rdc = s.at(i);
	uint32_t t[4], b[4]; int topc;
	const bool single = (bot - top == 1);
	octet_lf(ix, top, single ? top : bot, gmask, l8, t, b, topc);
	if(!single) {
		int nonz = -1; bool abort = false;
		const uint32_t orig = bot - top;
#pragma unroll
		for(int j = 0; j < 4; j++) {
			if(!abort && b[j] > t[j]) {
				if(nonz >= 0) abort = true;
				else { nonz = j; top = t[j]; bot = b[j]; }
			}
		}
		if(abort || (nonz != rdc && rdc <= 3) || bot - top < orig) return false;
	} else {
		int c = -1;                                         // mapLF1(row&, l), bt2_idx.h:2451
		if(top != ix.zoff) { c = topc; top = pick4(t, c); }
		if(c != rdc && rdc <= 3) return false;
		bot = top + 1;
	}
	return true;
}

__global__ void __launch_bounds__(256)
k_extend(const IndexDev fw, const IndexDev bw, const uint8_t *__restrict__ seq,
         const bt2gpu_read *__restrict__ reads, const bt2gpu_extend_job *__restrict__ jobs, uint32_t njobs,
         bt2gpu_extend_result *__restrict__ out, const uint32_t *__restrict__ list, const uint32_t *__restrict__ nlist)
{
	// Persistent octets: extensions run 0..255 characters each way, so an octet that is done takes the next
	// hit at once (one flat loop, one rank step per iteration; octets diverge freely -- the walk is
	// latency-bound and independently scheduled octets hide more of it than a converged warp does).
	// list mode (chained batch runtime): the jobs to run are list[0 .. *nlist), a count only the device knows.
	const int lane = threadIdx.x & 31, l8 = lane & 7;
	const unsigned gmask = 0xffu << (lane & 24);
	const uint32_t n = list ? *nlist : njobs;
	const uint32_t noct = (gridDim.x * blockDim.x) >> 3;
	uint32_t next = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
	bt2gpu_extend_job jb; memset(&jb, 0, sizeof(jb));
	ReadView s; s.rd = seq; s.len = 0; s.rc = false; s.rev = false;
	uint32_t ji = 0, top = 0, bot = 0, fmops = 0, nex = 0, nlex = 0;
	int phase = 2, ii = 0, lim = 0;          // 0 = left over fw, 1 = right over bw, 2 = no job
	for(;;) {
		if(phase == 2) {
			if(next >= n) break;
			ji = list ? list[next] : next; next += noct;
			jb = jobs[ji];
			s.rd = seq + reads[jb.read].rd_off; s.len = (int)reads[jb.read].rd_len; s.rc = !jb.fw;
			fmops = 0; nlex = 0;
			phase = 0; ii = 0; nex = 0; top = jb.topf; bot = jb.botf;
			lim = jb.fw ? (int)jb.off : s.len - (int)jb.len - (int)jb.off;
		}
		if(ii < lim) {
			const int rdlen = s.len, off = jb.off, len = jb.len;
			int i;
			if(phase == 0) i = jb.fw ? off - ii - 1 : rdlen - off - len - 1 - ii;
			else           i = jb.fw ? ii + len + off : rdlen - off + ii;
			fmops++;
			if(extend_step(phase == 0 ? fw : bw, s, top, bot, i, gmask, l8)) { ii++; if(++nex == 255) ii = lim; }
			else ii = lim;
		} else if(phase == 0) {
			nlex = nex;
			phase = 1; ii = 0; nex = 0; top = jb.topb; bot = jb.botb;
			lim = jb.fw ? s.len - (int)jb.len - (int)jb.off : (int)jb.off;
		} else {
			if(l8 == 0) { out[ji].nlex = (uint16_t)nlex; out[ji].nrex = (uint16_t)nex; out[ji].fmops = fmops; }
			phase = 2;
		}
	}
}

// ---------------------------------------------------------------------------------------
// Kernel 1e: SeedAligner::oneMmSearch (aligner_seed.cpp:973-1323).  One octet per
// (read, strand, index) task -- the reference's two nested loops (fw/rc x BWT/BWT') are
// independent.  Hits go to a pool tagged (task, ordinal); the host orders them as the
// reference's hits.add1mmEe / addExactEe* call sequence.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ int fm_score(int rdc, int refm, int q) {    // Scoring::score, scoring.h:241
	q = q < 0 ? 0 : (q > 255 ? 255 : q);
	if(rdc > 3 || refm > 15) return -c_fm_npens[q];
	return ((refm >> rdc) & 1) ? c_fm_sc.match : -c_fm_mmpens[q];
}

__global__ void __launch_bounds__(256)
k_onemm(const IndexDev ifw, const IndexDev ibw, const uint8_t *__restrict__ seq, const uint8_t *__restrict__ qual,
        const bt2gpu_read *__restrict__ reads, uint32_t nreads, const bt2gpu_onemm_job *__restrict__ jobs,
        int local, int repex, int rep1mm, bt2gpu_onemm_result *__restrict__ res,
        bt2gpu_ee_hit *__restrict__ hits, unsigned long long *__restrict__ keys,
        unsigned long long hit_cap, unsigned long long *__restrict__ nhits)
{
	const int lane = threadIdx.x & 31, l8 = lane & 7;
	const unsigned gmask = 0xffu << (lane & 24);
	const uint32_t task = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
	if(task >= 4 * nreads) return;
	const uint32_t ri = task >> 2;
	const bool fw = ((task >> 1) & 1) == 0, ebwtfw = (task & 1) == 0;
	const bt2gpu_onemm_job jb = jobs[ri];
	if((fw && jb.nofw) || (!fw && jb.norc)) return;
	if(!ebwtfw && !rep1mm) return;
	const uint8_t *rd = seq + reads[ri].rd_off, *qu = qual + reads[ri].rd_off;
	const int len = (int)reads[ri].rd_len;
	int ns = 0;
	for(int i = 0; i < len; i++) ns += rd[i] > 3;
	if(ns > 1 || (ns == 1 && !rep1mm)) return;
	const IndexDev &ebwt = ebwtfw ? ifw : ibw;
	const IndexDev &ebwtp = ebwtfw ? ibw : ifw;
	ReadView s; s.rd = rd; s.len = len; s.rc = !fw; s.rev = !ebwtfw;
	const bool qrev = (fw != ebwtfw);
	const int fl = ebwt.ftab_chars;
	const int nea = ebwtfw ? (len >> 1) : ((len >> 1) + (len & 1));
	for(int d = 0; d < nea; d++) if(s.at(len - d - 1) > 3) return;
	uint32_t top = 0, bot = 0, topp = 0, botp = 0, bwops = 0, ord = 0;
	int dep;
	if(fl > 1 && fl <= nea) {
		const bool rev = !ebwtfw;
		const uint32_t k1 = ftab_key(ebwt, s, len - fl, rev);
		top = ftab_hi(ebwt, k1); bot = ftab_lo(ebwt, k1 + 1);
		if(rep1mm) { const uint32_t k2 = ftab_key(ebwtp, s, len - fl, rev); topp = ftab_hi(ebwtp, k2); botp = ftab_lo(ebwtp, k2 + 1); }
		if(bot - top == 0) return;
		dep = fl;
	} else {
		const int c = s.at(len - 1);
		top = topp = ebwt.fchr[c]; bot = botp = ebwt.fchr[c + 1];
		if(bot - top == 0) return;
		dep = 1;
	}
	bool dead = false;
	for(; dep < nea; dep++) {                                   // near half :1078-1116
		const int rdc = s.at(len - dep - 1);
		uint32_t t[4], b[4]; int topc;
		const bool single = (bot - top == 1);
		octet_lf(ebwt, top, single ? top : bot, gmask, l8, t, b, topc);
		bwops++;
		if(!single) {
			uint32_t below = 0;
#pragma unroll
			for(int j = 0; j < 3; j++) if(j < rdc) below += b[j] - t[j];
			top = t[rdc]; bot = b[rdc];
			if(bot <= top) { dead = true; break; }
			topp += below; botp = topp + (bot - top);
		} else {
			if(topc != rdc || top == ebwt.zoff) { dead = true; break; }
			top = t[rdc]; bot = top + 1;
		}
	}
	for(; !dead && dep < len; dep++) {                          // far half :1121-1318
		const int rdc = s.at(len - dep - 1);
		const int quc = qu[qrev ? dep : len - dep - 1];
		if(rdc > 3 && jb.nceil == 0) break;
		uint32_t t[4], b[4], tp[4], bp[4]; int topc;
		const bool single = (bot - top == 1);
		int clo = 0, chi = 3; bool match = true;
		const uint32_t otop = top;
		octet_lf(ebwt, top, single ? top : bot, gmask, l8, t, b, topc);
		bwops++;
		if(!single) {
			uint32_t acc = topp;
#pragma unroll
			for(int j = 0; j < 4; j++) { tp[j] = acc; acc += b[j] - t[j]; bp[j] = acc; }
			match = rdc < 4;
			if(rdc < 4) { top = t[rdc]; bot = b[rdc]; topp = tp[rdc]; botp = bp[rdc]; }
		} else {
			if(otop == ebwt.zoff) break;                        // mapLF1 hit the '$'
			clo = chi = topc;
			match = (clo == rdc);
			top = t[clo]; bot = top + 1;
#pragma unroll
			for(int j = 0; j < 4; j++) { t[j] = top; b[j] = bot; tp[j] = topp; bp[j] = botp; }
		}
		if(rep1mm && (ns == 0 || rdc > 3)) {
			for(int j = clo; j <= chi; j++) {
				if(j == rdc || b[j] == t[j]) continue;
				int depm = dep + 1;
				uint32_t topm = t[j], botm = b[j], topmp = tp[j], botmp = bp[j];
				for(; depm < len; depm++) {
					const int rdcm = s.at(len - depm - 1);
					uint32_t tm[4], bm[4]; int tc;
					const bool sm = (botm - topm == 1);
					octet_lf(ebwt, topm, sm ? topm : botm, gmask, l8, tm, bm, tc);
					bwops++;
					if(!sm) {
						uint32_t below = 0;
#pragma unroll
						for(int k = 0; k < 3; k++) if(k < rdcm) below += bm[k] - tm[k];
						topm = tm[rdcm]; botm = bm[rdcm];
						topmp += below; botmp = topmp + (botm - topm);
						if(botm <= topm) break;
					} else {
						if(tc != rdcm || topm == ebwt.zoff) break;
						topm = tm[rdcm]; botm = topm + 1;
					}
				}
				if(depm == len) {
					int off5p = dep;
					if(fw == ebwtfw) off5p = len - off5p - 1;
					long long score = (long long)(len - 1) * c_fm_sc.match;
					const int pen = fm_score(rdc, 1 << j, quc - 33);
					score += pen;
					bool valid = true;
					if(local) {
						long long lf = 0, lb = 0;
						for(int i = 0; i < len; i++) {
							if(i == dep) { if(lf + pen <= 0) { valid = false; break; } lf += pen; }
							else lf += c_fm_sc.match;
							if(len - i - 1 == dep) { if(lb + pen <= 0) { valid = false; break; } lb += pen; }
							else lb += c_fm_sc.match;
						}
					}
					if(valid) valid = score >= (long long)jb.minsc;
					if(valid) {
						if(l8 == 0) {
							const unsigned long long slot = atomicAdd(nhits, 1ull);
							if(slot < hit_cap) {
								bt2gpu_ee_hit h;
								h.top = ebwtfw ? topm : topmp; h.bot = ebwtfw ? botm : botmp;
								h.score = (int32_t)score; h.pos = (uint16_t)off5p;
								h.chr = (uint8_t)"ACGT"[j]; h.qchr = (uint8_t)"ACGTN"[rdc]; h.fw = fw ? 1 : 0;
								h.pad[0] = h.pad[1] = h.pad[2] = 0;
								hits[slot] = h;
								keys[slot] = ((unsigned long long)task << 24) | ord;
							}
						}
						ord++;
					}
				}
			}
		}
		if(bot > top && match) {
			if(dep == len - 1) {
				if(ebwtfw && repex) {
					if(l8 == 0) {
						const unsigned long long slot = atomicAdd(nhits, 1ull);
						if(slot < hit_cap) {
							bt2gpu_ee_hit h;
							h.top = top; h.bot = bot; h.score = (int32_t)((long long)len * c_fm_sc.match);
							h.pos = 0xffff; h.chr = 0; h.qchr = 0; h.fw = fw ? 1 : 0;
							h.pad[0] = h.pad[1] = h.pad[2] = 0;
							hits[slot] = h;
							keys[slot] = ((unsigned long long)task << 24) | ord;
						}
					}
					ord++;
				}
				break;
			}
		} else break;
	}
	if(l8 == 0 && bwops) atomicAdd(&res[ri].bwops, bwops);
}

} // namespace bt2

// ---------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------
namespace bt2 {

struct FmState {
	DevBuf seq, qual, reads;            // staged read batch
	size_t nreads = 0, pool_len = 0;
	std::vector<uint32_t> rdlen;        // host copy of the staged read lengths (job validation)
	DevBuf jobs, out, aux, keys, scal;  // per-call scratch
	HostBuf h_scal, h_keys;
	size_t n_seed_jobs = 0, n_rows = 0;
	DevBuf sjobs, sout, lrows, loffs, lsteps;   // staged (device-resident) seed / locate batches
};

void fm_state_free(FmState *d) {
	if(!d) return;
	DevBuf *bufs[] = { &d->seq, &d->qual, &d->reads, &d->jobs, &d->out, &d->aux, &d->keys, &d->scal,
	                   &d->sjobs, &d->sout, &d->lrows, &d->loffs, &d->lsteps };
	for(DevBuf *b : bufs) b->release();
	d->h_scal.release(); d->h_keys.release();
	delete d;
}

int fm_set_scoring(bt2gpu_ctx *ctx, const bt2gpu_scoring *sc) {
	ScFm c = { sc->match_bonus };
	BT2_CUDA(cudaMemcpyToSymbolAsync(c_fm_mmpens, sc->mmpens, sizeof(c_fm_mmpens), 0, cudaMemcpyHostToDevice, ctx->stream));
	BT2_CUDA(cudaMemcpyToSymbolAsync(c_fm_npens, sc->npens, sizeof(c_fm_npens), 0, cudaMemcpyHostToDevice, ctx->stream));
	BT2_CUDA(cudaMemcpyToSymbolAsync(c_fm_sc, &c, sizeof(c), 0, cudaMemcpyHostToDevice, ctx->stream));
	return 0;
}

static FmState *fm_state(bt2gpu_ctx *ctx) {
	if(!ctx->fm) ctx->fm = new FmState();
	return ctx->fm;
}

static int check_idx(bt2gpu_ctx *ctx, const bt2gpu_index *ix, bool need_offs) {
	if(!ix) BT2_FAIL("null index");
	if(ix->device != ctx->device) BT2_FAIL("index lives on another device");
	if(need_offs && !ix->dev.offs) BT2_FAIL("index has no SA sample (offs)");
	return 0;
}

static inline uint32_t blocks_for(size_t tasks, int lanes_per_task, int block) {
	return (uint32_t)((tasks * lanes_per_task + block - 1) / block);
}

} // namespace bt2

using namespace bt2;

extern "C" int bt2gpu_index_create(int device, const bt2gpu_index_desc *d, bt2gpu_index **out) {
	if(!d || !out) BT2_FAIL("null argument");
	*out = nullptr;
	int n = 0;
	BT2_CUDA(cudaGetDeviceCount(&n));
	if(n <= 0) BT2_FAIL("no CUDA device: libbt2gpu has no CPU fallback");
	if(device < 0 || device >= n) BT2_FAIL("bad device ordinal");
	if(!d->ebwt || !d->ftab || !d->eftab) BT2_FAIL("index arrays missing");
	if(d->ebwt_bytes % 64) BT2_FAIL("ebwt size is not a multiple of the 64-byte side");
	if(d->ftab_chars < 1 || d->ftab_chars > 15) BT2_FAIL("bad ftabChars");
	if((uint64_t)d->zoff > d->len) BT2_FAIL("bad zOff");
	BT2_CUDA(cudaSetDevice(device));
	bt2gpu_index *ix = new bt2gpu_index();
	ix->device = device;
	ix->nsides = d->ebwt_bytes / 64;
	const size_t ftab_n = ((size_t)1 << (2 * d->ftab_chars)) + 1, eftab_n = (size_t)2 * d->ftab_chars;
	auto up = [&](DevBuf &b, const void *src, size_t bytes) -> int {
		if(b.reserve(bytes)) return 1;
		BT2_CUDA(cudaMemcpy(b.p, src, bytes, cudaMemcpyHostToDevice));
		ix->bytes += bytes;
		return 0;
	};
	if(up(ix->ebwt, d->ebwt, d->ebwt_bytes) || up(ix->ftab, d->ftab, ftab_n * 4) || up(ix->eftab, d->eftab, eftab_n * 4) ||
	   (d->offs && d->n_offs && up(ix->offs, d->offs, d->n_offs * 4))) {
		bt2gpu_index_destroy(ix);
		return 1;
	}
	IndexDev &v = ix->dev;
	v.sides = ix->ebwt.as<uint4>();
	v.offs = (d->offs && d->n_offs) ? ix->offs.as<uint32_t>() : nullptr;
	v.ftab = ix->ftab.as<uint32_t>(); v.eftab = ix->eftab.as<uint32_t>();
	for(int i = 0; i < 5; i++) v.fchr[i] = d->fchr[i];
	v.zoff = d->zoff; v.zside = d->zoff / 192u; v.zco = d->zoff % 192u;
	v.len = d->len; v.off_rate = d->off_rate; v.off_mask = OFFMASK << d->off_rate;
	v.ftab_chars = d->ftab_chars; v.is_fw = d->is_fw;
	*out = ix;
	return 0;
}

extern "C" void bt2gpu_index_destroy(bt2gpu_index *ix) {
	if(!ix) return;
	cudaSetDevice(ix->device);
	ix->ebwt.release(); ix->offs.release(); ix->ftab.release(); ix->eftab.release();
	delete ix;
}

extern "C" uint64_t bt2gpu_index_bytes(const bt2gpu_index *ix) { return ix ? ix->bytes : 0; }

extern "C" int bt2gpu_reads_stage(bt2gpu_ctx *ctx, const uint8_t *seq_pool, const uint8_t *qual_pool,
                                  size_t pool_len, const bt2gpu_read *reads, size_t nreads)
{
	if(!ctx) BT2_FAIL("null ctx");
	BT2_CUDA(cudaSetDevice(ctx->device));
	FmState *f = fm_state(ctx);
	f->nreads = 0;
	for(size_t i = 0; i < nreads; i++) {
		if((size_t)reads[i].rd_off + reads[i].rd_len > pool_len) BT2_FAIL("read outside pool");
		if(reads[i].rd_len == 0 || reads[i].rd_len > 65000) BT2_FAIL("read length must be in 1..65000");
	}
	if(f->seq.reserve(pool_len + 16) || f->qual.reserve(pool_len + 16) || f->reads.reserve(nreads * sizeof(bt2gpu_read) + 16)) return 1;
	cudaStream_t s = ctx->stream;
	if(pool_len) {
		BT2_CUDA(cudaMemcpyAsync(f->seq.p, seq_pool, pool_len, cudaMemcpyHostToDevice, s));
		if(qual_pool) BT2_CUDA(cudaMemcpyAsync(f->qual.p, qual_pool, pool_len, cudaMemcpyHostToDevice, s));
	}
	if(nreads) BT2_CUDA(cudaMemcpyAsync(f->reads.p, reads, nreads * sizeof(bt2gpu_read), cudaMemcpyHostToDevice, s));
	BT2_CUDA(bt2::wait_stream(ctx, s));
	f->nreads = nreads; f->pool_len = pool_len;
	f->rdlen.resize(nreads);
	for(size_t i = 0; i < nreads; i++) f->rdlen[i] = reads[i].rd_len;
	return 0;
}


// ---- device-resident launches for the chained batch runtime (hotpath.cu) -----------------
// Same kernels as the entry points below; inputs and outputs stay in HBM, nothing synchronises.
namespace bt2 {

int fm_hp_reads_stage(bt2gpu_ctx *ctx, const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
                      const bt2gpu_read *reads, size_t nreads, FmReadsDev *view, cudaStream_t qual_stream)
{
	BT2_CUDA(cudaSetDevice(ctx->device));
	FmState *f = fm_state(ctx);
	f->nreads = 0;
	if(f->seq.reserve(pool_len + 16) || f->qual.reserve(pool_len + 16) || f->reads.reserve(nreads * sizeof(bt2gpu_read) + 16)) return 1;
	cudaStream_t s = ctx->stream;
	BT2_CUDA(cudaMemcpyAsync(f->reads.p, reads, nreads * sizeof(bt2gpu_read), cudaMemcpyHostToDevice, s));
	BT2_CUDA(cudaMemcpyAsync(f->seq.p, seq_pool, pool_len, cudaMemcpyHostToDevice, s));
	// qualities are first read by the DP framing: their copy rides a second stream under the FM kernels
	BT2_CUDA(cudaMemcpyAsync(f->qual.p, qual_pool, pool_len, cudaMemcpyHostToDevice, qual_stream ? qual_stream : s));
	// validation runs while the copies are in flight
	f->rdlen.resize(nreads);
	for(size_t i = 0; i < nreads; i++) {
		if((size_t)reads[i].rd_off + reads[i].rd_len > pool_len) BT2_FAIL("read outside pool");
		if(reads[i].rd_len == 0 || reads[i].rd_len > 65000) BT2_FAIL("read length must be in 1..65000");
		f->rdlen[i] = reads[i].rd_len;
	}
	f->nreads = nreads; f->pool_len = pool_len;
	view->seq = f->seq.as<uint8_t>(); view->qual = f->qual.as<uint8_t>(); view->reads = f->reads.as<bt2gpu_read>();
	return 0;
}

int fm_hp_sweep(bt2gpu_ctx *ctx, const bt2gpu_index *fw, uint32_t mine_max, bt2gpu_sweep_result **dev_out) {
	if(check_idx(ctx, fw, false)) return 1;
	FmState *f = ctx->fm;
	const size_t n = f->nreads;
	if(f->out.reserve(n * sizeof(bt2gpu_sweep_result))) return 1;
	cudaStream_t s = ctx->stream;
	BT2_CUDA(cudaMemsetAsync(f->out.p, 0, n * sizeof(bt2gpu_sweep_result), s));
	k_exact_sweep<<<blocks_for(2 * n, 8, 256), 256, 0, s>>>(fw->dev, f->seq.as<uint8_t>(), f->reads.as<bt2gpu_read>(),
		(uint32_t)n, 0, 0, mine_max, f->out.as<bt2gpu_sweep_result>());
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	*dev_out = f->out.as<bt2gpu_sweep_result>();
	return 0;
}

int fm_hp_seed(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw, const bt2gpu_seed_job *djobs,
               size_t njobs, bt2gpu_seed_hit *dout)
{
	if(check_idx(ctx, fw, false) || check_idx(ctx, bw, false)) return 1;
	if(fw->dev.ftab_chars != bw->dev.ftab_chars) BT2_FAIL("forward and mirror index disagree on ftabChars");
	if(njobs == 0) return 0;
	FmState *f = ctx->fm;
	const uint64_t resident = (uint64_t)ctx->sm_count * 8 * (SEED_BLOCK / 8);
	uint32_t per_oct = (uint32_t)((njobs + resident - 1) / resident);
	if(per_oct < 1) per_oct = 1;
	if(per_oct > SEED_TILE / (SEED_BLOCK / 8)) per_oct = SEED_TILE / (SEED_BLOCK / 8);
	const uint32_t tile_sz = per_oct * (SEED_BLOCK / 8);
	uint32_t blocks = (uint32_t)((njobs + tile_sz - 1) / tile_sz);
	const uint32_t maxb = (uint32_t)ctx->sm_count * 8;
	if(blocks > maxb && per_oct == SEED_TILE / (SEED_BLOCK / 8)) blocks = maxb;
	k_seed_exact<<<blocks, SEED_BLOCK, 0, ctx->stream>>>(fw->dev, bw->dev, f->seq.as<uint8_t>(), f->reads.as<bt2gpu_read>(),
		djobs, (uint32_t)njobs, tile_sz, dout);
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	return 0;
}

int fm_hp_extend(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw, const bt2gpu_extend_job *djobs,
                 size_t njobs, bt2gpu_extend_result *dout, const uint32_t *dlist, const uint32_t *dnlist)
{
	if(njobs == 0) return 0;
	FmState *f = ctx->fm;
	uint32_t blocks = blocks_for(njobs, 8, 256);
	const uint32_t maxb = (uint32_t)ctx->sm_count * 8;
	if(blocks > maxb) blocks = maxb;
	k_extend<<<blocks, 256, 0, ctx->stream>>>(fw->dev, bw->dev, f->seq.as<uint8_t>(), f->reads.as<bt2gpu_read>(),
		djobs, (uint32_t)njobs, dout, dlist, dnlist);
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	return 0;
}

int fm_hp_locate(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const uint32_t *drows, size_t n, uint32_t *doffs, uint32_t *dsteps) {
	if(check_idx(ctx, fw, true)) return 1;
	if(n == 0) return 0;
	k_locate<<<blocks_for(n, 4, 256), 256, 0, ctx->stream>>>(fw->dev, drows, (uint32_t)n, doffs, dsteps);
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	return 0;
}

} // namespace bt2

// ---- exact seeds ------------------------------------------------------------------------
extern "C" int bt2gpu_seed_exact_stage(bt2gpu_ctx *ctx, const bt2gpu_seed_job *jobs, size_t njobs) {
	if(!ctx || !ctx->fm) BT2_FAIL("bt2gpu_reads_stage must be called first");
	BT2_CUDA(cudaSetDevice(ctx->device));
	FmState *f = ctx->fm;
	f->n_seed_jobs = 0;
	if(njobs > 0xfffffff0u) BT2_FAIL("too many seed jobs");
	for(size_t i = 0; i < njobs; i++) {
		const bt2gpu_seed_job &j = jobs[i];
		if(j.read >= f->nreads) BT2_FAIL("seed job names a read that is not staged");
		if(j.len < 1 || j.len > 32) BT2_FAIL("seed length must be 1..32");
		if((uint32_t)j.off + j.len > f->rdlen[j.read]) BT2_FAIL("seed window outside the read");
	}
	if(f->sjobs.reserve(njobs * sizeof(bt2gpu_seed_job) + 16) || f->sout.reserve(njobs * sizeof(bt2gpu_seed_hit) + 16)) return 1;
	if(njobs) BT2_CUDA(cudaMemcpyAsync(f->sjobs.p, jobs, njobs * sizeof(bt2gpu_seed_job), cudaMemcpyHostToDevice, ctx->stream));
	if(!ctx->in_batch) BT2_CUDA(bt2::wait_stream(ctx, ctx->stream));   // inside bt2gpu_seed_exact_batch the fetch waits
	f->n_seed_jobs = njobs;
	return 0;
}

extern "C" int bt2gpu_seed_exact_run(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw) {
	if(!ctx || !ctx->fm) BT2_FAIL("nothing staged");
	if(check_idx(ctx, fw, false) || check_idx(ctx, bw, false)) return 1;
	if(fw->dev.ftab_chars != bw->dev.ftab_chars) BT2_FAIL("forward and mirror index disagree on ftabChars");
	BT2_CUDA(cudaSetDevice(ctx->device));
	FmState *f = ctx->fm;
	if(f->n_seed_jobs == 0) return 0;
	// tile = seeds staged per CTA pass: as small as keeps every resident octet busy (latency-bound walk:
	// parallel seeds beat sequential ones), at most SEED_TILE
	const uint64_t resident = (uint64_t)ctx->sm_count * 8 * (SEED_BLOCK / 8);   // octets resident on the device
	uint32_t per_oct = (uint32_t)((f->n_seed_jobs + resident - 1) / resident);
	if(per_oct < 1) per_oct = 1;
	if(per_oct > SEED_TILE / (SEED_BLOCK / 8)) per_oct = SEED_TILE / (SEED_BLOCK / 8);
	const uint32_t tile_sz = per_oct * (SEED_BLOCK / 8);
	uint32_t blocks = (uint32_t)((f->n_seed_jobs + tile_sz - 1) / tile_sz);
	const uint32_t maxb = (uint32_t)ctx->sm_count * 8;          // 8 resident CTAs of 256 threads per SM
	if(blocks > maxb && per_oct == SEED_TILE / (SEED_BLOCK / 8)) blocks = maxb;
	if(ctx->kstart()) return 1;
	k_seed_exact<<<blocks, SEED_BLOCK, 0, ctx->stream>>>(fw->dev, bw->dev, f->seq.as<uint8_t>(), f->reads.as<bt2gpu_read>(),
		f->sjobs.as<bt2gpu_seed_job>(), (uint32_t)f->n_seed_jobs, tile_sz, f->sout.as<bt2gpu_seed_hit>());
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	if(ctx->kend()) return 1;
	return 0;
}

extern "C" int bt2gpu_seed_exact_fetch(bt2gpu_ctx *ctx, bt2gpu_seed_hit *out) {
	if(!ctx || !ctx->fm) BT2_FAIL("nothing staged");
	BT2_CUDA(cudaSetDevice(ctx->device));
	FmState *f = ctx->fm;
	BT2_CUDA(bt2::settle(ctx, ctx->stream));
	if(f->n_seed_jobs && out)
		BT2_CUDA(cudaMemcpyAsync(out, f->sout.p, f->n_seed_jobs * sizeof(bt2gpu_seed_hit), cudaMemcpyDeviceToHost, ctx->stream));
	BT2_CUDA(bt2::wait_stream(ctx, ctx->stream));
	return 0;
}

extern "C" int bt2gpu_seed_exact_batch(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw,
                                       const bt2gpu_seed_job *jobs, size_t njobs, bt2gpu_seed_hit *out) {
	if(!ctx) BT2_FAIL("null ctx");
	ctx->in_batch = true;
	const int rc = bt2gpu_seed_exact_stage(ctx, jobs, njobs);
	ctx->in_batch = false;
	if(rc) return 1;
	if(bt2gpu_seed_exact_run(ctx, fw, bw)) return 1;
	return bt2gpu_seed_exact_fetch(ctx, out);
}

// ---- exactSweep -------------------------------------------------------------------------
extern "C" int bt2gpu_exact_sweep_batch(bt2gpu_ctx *ctx, const bt2gpu_index *fw, int nofw, int norc,
                                        uint32_t mine_max, bt2gpu_sweep_result *out) {
	if(!ctx || !ctx->fm) BT2_FAIL("bt2gpu_reads_stage must be called first");
	if(check_idx(ctx, fw, false)) return 1;
	if(mine_max == 0) BT2_FAIL("mine_max must be > 0 (aligner_seed.cpp:867)");
	BT2_CUDA(cudaSetDevice(ctx->device));
	FmState *f = ctx->fm;
	const size_t n = f->nreads;
	if(n == 0) return 0;
	if(f->out.reserve(n * sizeof(bt2gpu_sweep_result))) return 1;
	cudaStream_t s = ctx->stream;
	BT2_CUDA(cudaMemsetAsync(f->out.p, 0, n * sizeof(bt2gpu_sweep_result), s));
	if(ctx->kstart()) return 1;
	k_exact_sweep<<<blocks_for(2 * n, 8, 256), 256, 0, s>>>(fw->dev, f->seq.as<uint8_t>(), f->reads.as<bt2gpu_read>(),
		(uint32_t)n, nofw, norc, mine_max, f->out.as<bt2gpu_sweep_result>());
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	if(ctx->kend()) return 1;
	BT2_CUDA(bt2::settle(ctx, s));
	BT2_CUDA(cudaMemcpyAsync(out, f->out.p, n * sizeof(bt2gpu_sweep_result), cudaMemcpyDeviceToHost, s));
	BT2_CUDA(bt2::wait_stream(ctx, s));
	return 0;
}

// ---- oneMmSearch ------------------------------------------------------------------------
extern "C" int bt2gpu_onemm_batch(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw,
                                  const bt2gpu_onemm_job *jobs, int local, int repex, int rep1mm,
                                  bt2gpu_onemm_result *out, bt2gpu_ee_hit *hits, size_t hit_cap, size_t *nhits) {
	if(!ctx || !ctx->fm) BT2_FAIL("bt2gpu_reads_stage must be called first");
	if(!ctx->have_sc) BT2_FAIL("bt2gpu_scoring_set must be called first");
	if(check_idx(ctx, fw, false)) return 1;
	if(rep1mm && check_idx(ctx, bw, false)) return 1;
	if(!jobs || !out || !nhits) BT2_FAIL("null argument");
	BT2_CUDA(cudaSetDevice(ctx->device));
	FmState *f = ctx->fm;
	const size_t n = f->nreads;
	*nhits = 0;
	if(n == 0) return 0;
	if(n > 0x3fffffffu) BT2_FAIL("too many reads");
	if(f->jobs.reserve(n * sizeof(bt2gpu_onemm_job)) || f->out.reserve(n * sizeof(bt2gpu_onemm_result)) ||
	   f->aux.reserve(hit_cap * sizeof(bt2gpu_ee_hit) + 16) || f->keys.reserve(hit_cap * 8 + 16) ||
	   f->scal.reserve(64) || f->h_scal.reserve(64)) return 1;
	cudaStream_t s = ctx->stream;
	BT2_CUDA(cudaMemcpyAsync(f->jobs.p, jobs, n * sizeof(bt2gpu_onemm_job), cudaMemcpyHostToDevice, s));
	BT2_CUDA(cudaMemsetAsync(f->out.p, 0, n * sizeof(bt2gpu_onemm_result), s));
	BT2_CUDA(cudaMemsetAsync(f->scal.p, 0, 64, s));
	if(ctx->kstart()) return 1;
	k_onemm<<<blocks_for(4 * n, 8, 256), 256, 0, s>>>(fw->dev, rep1mm ? bw->dev : fw->dev, f->seq.as<uint8_t>(),
		f->qual.as<uint8_t>(), f->reads.as<bt2gpu_read>(), (uint32_t)n, f->jobs.as<bt2gpu_onemm_job>(), local, repex, rep1mm,
		f->out.as<bt2gpu_onemm_result>(), f->aux.as<bt2gpu_ee_hit>(), f->keys.as<unsigned long long>(),
		(unsigned long long)hit_cap, f->scal.as<unsigned long long>());
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	if(ctx->kend()) return 1;
	BT2_CUDA(cudaMemcpyAsync(out, f->out.p, n * sizeof(bt2gpu_onemm_result), cudaMemcpyDeviceToHost, s));
	BT2_CUDA(cudaMemcpyAsync(f->h_scal.p, f->scal.p, 8, cudaMemcpyDeviceToHost, s));
	BT2_CUDA(cudaStreamSynchronize(s));
	const unsigned long long nh = *f->h_scal.as<unsigned long long>();
	if(nh > hit_cap) BT2_FAIL("1-mismatch hit pool overflow (raise hit_cap)");
	// order the pool as the reference's call sequence: by task (read, strand, index), then ordinal
	std::vector<bt2gpu_ee_hit> tmp(nh);
	std::vector<unsigned long long> keys(nh);
	if(nh) {
		BT2_CUDA(cudaMemcpyAsync(tmp.data(), f->aux.p, nh * sizeof(bt2gpu_ee_hit), cudaMemcpyDeviceToHost, s));
		BT2_CUDA(cudaMemcpyAsync(keys.data(), f->keys.p, nh * 8, cudaMemcpyDeviceToHost, s));
		BT2_CUDA(cudaStreamSynchronize(s));
	}
	std::vector<uint32_t> perm(nh);
	for(size_t i = 0; i < nh; i++) perm[i] = (uint32_t)i;
	std::sort(perm.begin(), perm.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
	size_t k = 0;
	for(size_t r = 0; r < n; r++) {
		out[r].hit_off = (uint32_t)k;
		while(k < nh && (keys[perm[k]] >> 26) == r) { hits[k] = tmp[perm[k]]; k++; }
		out[r].nhits = (uint32_t)k - out[r].hit_off;
	}
	*nhits = (size_t)nh;
	return 0;
}

// ---- extend -----------------------------------------------------------------------------
extern "C" int bt2gpu_extend_batch(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw,
                                   const bt2gpu_extend_job *jobs, size_t njobs, bt2gpu_extend_result *out) {
	if(!ctx || !ctx->fm) BT2_FAIL("bt2gpu_reads_stage must be called first");
	if(check_idx(ctx, fw, false) || check_idx(ctx, bw, false)) return 1;
	BT2_CUDA(cudaSetDevice(ctx->device));
	FmState *f = ctx->fm;
	if(njobs == 0) return 0;
	for(size_t i = 0; i < njobs; i++) {
		const bt2gpu_extend_job &j = jobs[i];
		if(j.read >= f->nreads) BT2_FAIL("extend job names a read that is not staged");
		if((uint32_t)j.off + j.len > f->rdlen[j.read]) BT2_FAIL("seed window outside the read");
		if(j.botf <= j.topf || j.botb <= j.topb) BT2_FAIL("empty BW range");
	}
	if(f->jobs.reserve(njobs * sizeof(bt2gpu_extend_job)) || f->out.reserve(njobs * sizeof(bt2gpu_extend_result))) return 1;
	cudaStream_t s = ctx->stream;
	BT2_CUDA(cudaMemcpyAsync(f->jobs.p, jobs, njobs * sizeof(bt2gpu_extend_job), cudaMemcpyHostToDevice, s));
	if(ctx->kstart()) return 1;
	k_extend<<<blocks_for(njobs, 8, 256), 256, 0, s>>>(fw->dev, bw->dev, f->seq.as<uint8_t>(), f->reads.as<bt2gpu_read>(),
		f->jobs.as<bt2gpu_extend_job>(), (uint32_t)njobs, f->out.as<bt2gpu_extend_result>(), nullptr, nullptr);
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	if(ctx->kend()) return 1;
	BT2_CUDA(bt2::settle(ctx, s));
	BT2_CUDA(cudaMemcpyAsync(out, f->out.p, njobs * sizeof(bt2gpu_extend_result), cudaMemcpyDeviceToHost, s));
	BT2_CUDA(bt2::wait_stream(ctx, s));
	return 0;
}

// ---- locate -----------------------------------------------------------------------------
extern "C" int bt2gpu_locate_stage(bt2gpu_ctx *ctx, const uint32_t *rows, size_t n) {
	if(!ctx) BT2_FAIL("null ctx");
	BT2_CUDA(cudaSetDevice(ctx->device));
	FmState *f = fm_state(ctx);
	f->n_rows = 0;
	if(n > 0x3ffffff0u) BT2_FAIL("too many rows");
	if(f->lrows.reserve(n * 4 + 16) || f->loffs.reserve(n * 4 + 16) || f->lsteps.reserve(n * 4 + 16)) return 1;
	if(n) BT2_CUDA(cudaMemcpyAsync(f->lrows.p, rows, n * 4, cudaMemcpyHostToDevice, ctx->stream));
	if(!ctx->in_batch) BT2_CUDA(bt2::wait_stream(ctx, ctx->stream));   // inside bt2gpu_locate_batch the fetch waits
	f->n_rows = n;
	return 0;
}

extern "C" int bt2gpu_locate_run(bt2gpu_ctx *ctx, const bt2gpu_index *fw) {
	if(!ctx || !ctx->fm) BT2_FAIL("nothing staged");
	if(check_idx(ctx, fw, true)) return 1;
	BT2_CUDA(cudaSetDevice(ctx->device));
	FmState *f = ctx->fm;
	if(f->n_rows == 0) return 0;
	if(ctx->kstart()) return 1;
	k_locate<<<blocks_for(f->n_rows, 4, 256), 256, 0, ctx->stream>>>(fw->dev, f->lrows.as<uint32_t>(), (uint32_t)f->n_rows,
		f->loffs.as<uint32_t>(), f->lsteps.as<uint32_t>());
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	if(ctx->kend()) return 1;
	return 0;
}

extern "C" int bt2gpu_locate_fetch(bt2gpu_ctx *ctx, uint32_t *offs_out, uint32_t *steps_out) {
	if(!ctx || !ctx->fm) BT2_FAIL("nothing staged");
	BT2_CUDA(cudaSetDevice(ctx->device));
	FmState *f = ctx->fm;
	if(f->n_rows) {
		BT2_CUDA(bt2::settle(ctx, ctx->stream));
		if(offs_out) BT2_CUDA(cudaMemcpyAsync(offs_out, f->loffs.p, f->n_rows * 4, cudaMemcpyDeviceToHost, ctx->stream));
		if(steps_out) BT2_CUDA(cudaMemcpyAsync(steps_out, f->lsteps.p, f->n_rows * 4, cudaMemcpyDeviceToHost, ctx->stream));
	}
	BT2_CUDA(bt2::wait_stream(ctx, ctx->stream));
	return 0;
}

extern "C" int bt2gpu_locate_batch(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const uint32_t *rows, size_t n,
                                   uint32_t *offs_out, uint32_t *steps_out) {
	if(!ctx) BT2_FAIL("null ctx");
	ctx->in_batch = true;
	const int rc = bt2gpu_locate_stage(ctx, rows, n);
	ctx->in_batch = false;
	if(rc) return 1;
	if(bt2gpu_locate_run(ctx, fw)) return 1;
	return bt2gpu_locate_fetch(ctx, offs_out, steps_out);
}


// ---- multi-zone seeds: -N 1 / -N 2 ----------------------------------------------------------------------------
// Seed::mmSeeds (aligner_seed.cpp:365-448), Seed::instantiate (:213-358), startSearchSeedBi (:1633-1714), nextLocsBi
// (:1364-1432), searchSeedBi with its mismatch branches (:1854-2033) and reportHit (:1575-1630), one THREAD per seed
// offset of a read strand: the seed types of an offset (2 for -N 1, 3 for -N 2) advance in lock step, a mismatch
// branch is followed to its end before anything else moves -- the order in which the reference's hits reach the
// offset's SeedSearchCache, which the driver's RNG-driven choices see.  First version: scalar rank (one thread reads
// the whole 64-byte side), correctness before speed; the branching is compile-time bounded (at most two edits).
namespace bt2 {

#define MM_MAXI 0x7fffffff
#define MM_MAXLEN 32
struct MmCons { int edits, mms, ins, dels, penalty, editsCeil, mmsCeil, insCeil, delsCeil, penaltyCeil; };
__device__ __forceinline__ MmCons mmc_init() { MmCons c = { MM_MAXI, MM_MAXI, MM_MAXI, MM_MAXI, MM_MAXI, MM_MAXI, MM_MAXI, MM_MAXI, MM_MAXI, MM_MAXI }; return c; }
__device__ __forceinline__ MmCons mmc_exact() { MmCons c = mmc_init(); c.edits = c.mms = c.ins = c.dels = c.penalty = 0; return c; }
__device__ __forceinline__ MmCons mmc_mm(int n) { MmCons c = mmc_init(); c.mms = n; c.edits = c.dels = c.ins = 0; return c; }
__device__ __forceinline__ bool mmc_must_match(const MmCons &c) { return (c.mms == 0 && c.edits == 0) || c.penalty == 0 || (c.mms == 0 && c.dels == 0 && c.ins == 0); }
__device__ __forceinline__ bool mmc_can(const MmCons &c, int pen) { return (c.mms > 0 || c.edits > 0) && c.penalty >= pen; }
__device__ __forceinline__ void mmc_charge(MmCons &c, int pen) { if(c.mms == 0) c.edits--; else c.mms--; c.penalty -= pen; }
__device__ __forceinline__ bool mmc_ok(const MmCons &c) {
	return c.edits <= c.editsCeil && c.mms <= c.mmsCeil && c.ins <= c.insCeil && c.dels <= c.delsCeil && c.penalty <= c.penaltyCeil;
}
__device__ __forceinline__ bool mmc_can_gap(const MmCons &c) { return ((c.ins > 0 || c.dels > 0) || c.edits > 0) && c.penalty > 0; }

// countBt2SideEx + rowL by one thread (the quad version above, without the shuffles)
__device__ __forceinline__ void thread_lf(const IndexDev &ix, uint32_t row, uint32_t val[4], int &rowc) {
	const uint32_t side = row / 192u;
	const int co = (int)(row - side * 192u);
	const uint4 *sp = ix.sides + (size_t)side * 4;
	const uint4 v0 = __ldg(sp), v1 = __ldg(sp + 1), v2 = __ldg(sp + 2), v3 = __ldg(sp + 3);
	const uint32_t w[12] = { v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w, v2.x, v2.y, v2.z, v2.w };
	uint32_t nC = 0, nG = 0, nT = 0;
#pragma unroll
	for(int k = 0; k < 12; k++) {
		int mw = co - 16 * k;
		mw = mw < 0 ? 0 : mw;
		const uint32_t pm = pos_mask16(mw);
		const uint32_t lo = w[k] & pm, hi = (w[k] >> 1) & pm;
		nT += __popc(lo & hi);
		nG += __popc(hi & ~lo);
		nC += __popc(lo & ~hi);
	}
	const uint32_t nA = (uint32_t)co - (nC + nG + nT);
	uint32_t wc = 0;
#pragma unroll
	for(int k = 0; k < 12; k++) if(k == (co >> 4)) wc = w[k];
	rowc = (int)((wc >> (2 * (co & 15))) & 3u);
	const uint32_t fix = (side == ix.zside && (uint32_t)co > ix.zco) ? 1u : 0u;
	val[0] = ix.fchr[0] + v3.x + nA - fix;
	val[1] = ix.fchr[1] + v3.y + nC;
	val[2] = ix.fchr[2] + v3.z + nG;
	val[3] = ix.fchr[3] + v3.w + nT;
}

struct MmSeed {                       // InstantiatedSeed: steps, zones, constraints after the Ns were charged, maxjump
	int n; int8_t steps[MM_MAXLEN], z1[MM_MAXLEN], z2[MM_MAXLEN];
	MmCons cons[3], overall;
	int maxjump;
};
struct MmParam {                      // SeedAlignerSearchParams; the loci are functions of their rows
	int seed;                         // which MmSeed
	int step;
	uint32_t topf, botf, topb, botb;
	uint32_t trow, brow; bool bvalid; // SideLocus tloc / bloc
	MmCons cv[3], overall;
	int nedits; uint8_t epos[2], echr[2];
	bool hasi0; uint32_t fwi0, bwi0;
};
struct MmCtx {
	const IndexDev *fw, *bw;
	const MmSeed *seeds;
	const uint8_t *seq, *qual;        // the seed string and its qualities
	bt2gpu_mmseed_hit *hits; uint32_t cap, n;
	uint32_t bwops, bwedits;
	int err;                          // 1 = hit slice full, 2 = a gap would be legal (the reference throws), 3 = deeper branching than two edits
};

// Seed::instantiate: false when the Ns of the window disqualify the seed
__device__ bool mm_instantiate(int type, const MmCons zones[3], int seedlen, const uint8_t *seq, const uint8_t *qual, MmSeed &is) {
	is.n = seedlen;
	if(type == BT2GPU_SEED_EXACT) {
		for(int k = 0; k < seedlen; k++) { is.steps[k] = (int8_t)-(seedlen - k); is.z1[k] = is.z2[k] = 0; }
	} else if(type == BT2GPU_SEED_LEFT_TO_RIGHT) {
		for(int k = 0; k < seedlen; k++) { is.steps[k] = (int8_t)(k + 1); is.z1[k] = is.z2[k] = (k < (seedlen + 1) / 2) ? 0 : 1; }
		is.z1[seedlen - 1] = is.z2[seedlen - 1] = -1;
	} else if(type == BT2GPU_SEED_RIGHT_TO_LEFT) {
		for(int k = 0; k < seedlen; k++) {
			is.steps[k] = (int8_t)-(seedlen - k);
			is.z1[k] = (k < seedlen / 2) ? 0 : 1;
			is.z2[k] = (k < (seedlen + 1) / 2 + 1) ? 0 : 1;
		}
		is.z1[seedlen - 1] = is.z2[seedlen - 1] = -1;
	} else {
		int step = 0;
		for(int k = (seedlen + 3) / 4; k < seedlen - (seedlen / 4); k++) { is.z1[step] = is.z2[step] = 0; is.steps[step++] = (int8_t)(k + 1); }
		for(int k = seedlen - (seedlen / 4); k < seedlen; k++) { is.z1[step] = is.z2[step] = 1; is.steps[step++] = (int8_t)(k + 1); }
		is.z1[step - 1] = is.z2[step - 1] = -1;
		for(int k = ((seedlen + 3) / 4) - 1; k >= 0; k--) { is.z1[step] = is.z2[step] = 2; is.steps[step++] = (int8_t)-(k + 1); }
		is.z1[step - 1] = is.z2[step - 1] = -2;
	}
	for(int i = 0; i < 3; i++) is.cons[i] = zones[i];
	is.overall = mmc_init();
	bool streak = true;
	const bool ltr = is.steps[0] > 0;
	is.maxjump = 0;
	for(int i = 0; i < is.n; i++) {
		const int off = abs((int)is.steps[i]) - 1;
		MmCons &cons = is.cons[abs((int)is.z1[i])];
		const int c = seq[off], q = qual[off];
		if(ltr != (is.steps[i] > 0) || is.z1[i] != 0 || is.z2[i] != 0) streak = false;
		if(c == 4) {
			if(mmc_can(cons, c_fm_npens[q])) mmc_charge(cons, c_fm_npens[q]);
			else return false;
		}
		if(streak) is.maxjump++;
	}
	return true;
}

__device__ void mm_report(MmCtx &x, const MmParam &p) {
	const MmSeed &s = x.seeds[p.seed];
	unsigned long long key = 0;
	for(int i = 0; i < s.n; i++) {
		int c = x.seq[i];
		for(int e = 0; e < p.nedits; e++) if(p.epos[e] == i) c = p.echr[e];
		key = (key << 2) | (unsigned long long)(c & 3);
	}
	if(x.n >= x.cap) { x.err = 1; x.n++; return; }
	bt2gpu_mmseed_hit h;
	h.topf = p.topf; h.botf = p.botf; h.topb = p.topb; h.botb = p.botb; h.key = key;
	x.hits[x.n++] = h;
}

__device__ __forceinline__ void mm_next_locs(const MmSeed &s, MmParam &p, int step) {      // nextLocsBi
	if(step == s.n) return;
	const bool ltr = s.steps[step] > 0;
	const uint32_t top = ltr ? p.topb : p.topf, bot = ltr ? p.botb : p.botf;
	p.trow = top;
	p.bvalid = (bot - top != 1);
	p.brow = bot;
}

__device__ __forceinline__ uint32_t mm_ftab_key(const IndexDev &ix, const uint8_t *seq, int off) {     // ftabSeqToInt(seq, off, false)
	const int fc = ix.ftab_chars;
	uint32_t v = 0;
	for(int i = 0; i < fc; i++) {
		const int c = ix.is_fw ? seq[off + i] : seq[off + fc - i - 1];
		if(c > 3) return OFFMASK;
		v = (v << 2) | (uint32_t)c;
	}
	return v;
}

__device__ void mm_param_init(MmCtx &x, MmParam &p, int seed) {
	const MmSeed &s = x.seeds[seed];
	p.seed = seed; p.step = 0;
	p.topf = p.botf = p.topb = p.botb = 0; p.trow = p.brow = 0; p.bvalid = false;
	for(int i = 0; i < 3; i++) p.cv[i] = s.cons[i];
	p.overall = s.overall;
	p.nedits = 0;
	int off = abs((int)s.steps[0]) - 1;
	const bool ltr = s.steps[0] > 0;
	const int ftab_len = x.fw->ftab_chars;
	p.hasi0 = (ftab_len > 1 && ftab_len <= s.maxjump);
	p.fwi0 = p.bwi0 = 0;
	if(p.hasi0) {
		if(!ltr) off = off - ftab_len + 1;
		p.fwi0 = mm_ftab_key(*x.fw, x.seq, off);
		p.bwi0 = mm_ftab_key(*x.bw, x.seq, off);
	}
}

__device__ bool mm_start(MmCtx &x, MmParam &p) {               // startSearchSeedBi: true = already done
	const MmSeed &s = x.seeds[p.seed];
	if(p.step == s.n) { mm_report(x, p); return true; }
	if(p.step == 0) {
		const int off = abs((int)s.steps[0]) - 1;
		if(p.hasi0) {
			p.topf = ftab_hi(*x.fw, p.fwi0); p.botf = ftab_lo(*x.fw, p.fwi0 + 1);
			if(p.botf - p.topf == 0) return true;
			p.topb = ftab_hi(*x.bw, p.bwi0); p.botb = p.topb + (p.botf - p.topf);
			p.step += x.fw->ftab_chars;
		} else if(s.maxjump > 0) {
			const int c = x.seq[off];
			p.topf = p.topb = x.fw->fchr[c];
			p.botf = p.botb = x.fw->fchr[c + 1];
			if(p.botf - p.topf == 0) return true;
			p.step++;
		} else {
			p.topf = p.topb = 0;
			p.botf = p.botb = x.fw->fchr[4];
		}
		if(p.step == s.n) { mm_report(x, p); return true; }
		mm_next_locs(s, p, p.step);
	}
	return false;
}

template<int D> __device__ void mm_run(MmCtx &x, MmParam &p);
template<> __device__ void mm_run<3>(MmCtx &x, MmParam &p);     // never reached: two edits at most

// one step of one search (the body of searchSeedBi's inner loop); true = this search is done
template<int D> __device__ bool mm_step(MmCtx &x, MmParam &p) {
	const MmSeed &s = x.seeds[p.seed];
	if(p.step >= s.n) return true;
	const int i = p.step++;
	uint32_t t[4] = {0, 0, 0, 0}, b[4] = {0, 0, 0, 0}, tp[4], bp[4];
	const bool ltr = s.steps[i] > 0;
	const int off = abs((int)s.steps[i]) - 1;
	const IndexDev &ebwt = ltr ? *x.bw : *x.fw;
	uint32_t ntop = ltr ? p.topb : p.topf;
#pragma unroll
	for(int k = 0; k < 4; k++) { tp[k] = ltr ? p.topf : p.topb; bp[k] = ltr ? p.botf : p.botb; }
	// tf/bf = forward-index ranges, tb/bb = mirror ranges: (t, b) belong to the index of this step, (tp, bp) to the other
	int tlc = 0;                                                  // rowL(tloc)
	bool have_tl = false;
	uint32_t tlv[4];
	if(p.bvalid) {
		x.bwops++;
		int dummy;
		thread_lf(ebwt, p.trow, t, tlc);
		thread_lf(ebwt, p.brow, b, dummy);
		bp[0] = tp[0] + (b[0] - t[0]); tp[1] = bp[0];
		bp[1] = tp[1] + (b[1] - t[1]); tp[2] = bp[1];
		bp[2] = tp[2] + (b[2] - t[2]); tp[3] = bp[2];
		bp[3] = tp[3] + (b[3] - t[3]);
	}
	const int c = x.seq[off];
	const bool leave_zone = s.z1[i] < 0;
	MmCons &cons = p.cv[abs((int)s.z1[i])];
	if(!(mmc_must_match(cons) && !mmc_must_match(p.overall)) || c == 4) {
		bool bail = false;
		if(!p.bvalid) {
			x.bwops++;
			thread_lf(ebwt, p.trow, tlv, tlc); have_tl = true;
			if(ntop == ebwt.zoff) bail = true;                    // mapLF1(row&, l): -1 at '$'
			else { ntop = pick4(tlv, tlc); t[tlc] = ntop; b[tlc] = ntop + 1; }
		}
		if(!bail) {
			const int q = x.qual[off], pen = c_fm_mmpens[q];          // the raw quality character indexes the table (:1934)
			if((mmc_can(cons, pen) && mmc_can(p.overall, pen)) || c == 4) {
				const MmCons save_c = cons, save_o = p.overall;
				if(c != 4) { mmc_charge(cons, pen); mmc_charge(p.overall, pen); }
				if(!leave_zone || (mmc_ok(cons) && mmc_ok(p.overall))) {
					for(int j = 0; j < 4; j++) {
						if(j == c || b[j] == t[j]) continue;
						if(D >= 2 || p.nedits >= 2) { x.err = 3; continue; }
						MmParam p2 = p;
						p2.topf = ltr ? tp[j] : t[j]; p2.botf = ltr ? bp[j] : b[j];
						p2.topb = ltr ? t[j] : tp[j]; p2.botb = ltr ? b[j] : bp[j];
						p2.epos[p2.nedits] = (uint8_t)off; p2.echr[p2.nedits] = (uint8_t)j; p2.nedits++;
						mm_next_locs(s, p2, i + 1);
						x.bwedits++;
						p2.step = i + 1;
						mm_run<D + 1>(x, p2);
					}
				}
				cons = save_c; p.overall = save_o;
			}
			if(mmc_can_gap(cons) && mmc_can_gap(p.overall)) { x.err = 2; return true; }
		}
	}
	if(c == 4) return true;
	if(leave_zone && (!mmc_ok(cons) || !mmc_ok(p.overall))) return true;
	if(!p.bvalid) {
		x.bwops++;
		if(!have_tl) { thread_lf(ebwt, p.trow, tlv, tlc); }
		// mapLF1(row, l, c): the row tested against '$' is the one the mismatch branch above may have advanced
		if(tlc != c || ntop == ebwt.zoff) return true;
		t[c] = pick4(tlv, c);
		b[c] = t[c] + 1;
	}
	if(pick4(b, c) == pick4(t, c)) return true;
	p.topf = ltr ? pick4(tp, c) : pick4(t, c); p.botf = ltr ? pick4(bp, c) : pick4(b, c);
	p.topb = ltr ? pick4(t, c) : pick4(tp, c); p.botb = ltr ? pick4(b, c) : pick4(bp, c);
	if(i + 1 == s.n) { mm_report(x, p); return true; }
	mm_next_locs(s, p, i + 1);
	return false;
}

template<int D> __device__ void mm_run(MmCtx &x, MmParam &p) {   // searchSeedBi(1, &p) of a mismatch branch
	if(mm_start(x, p)) return;
	while(!mm_step<D>(x, p)) { }
}
template<> __device__ void mm_run<3>(MmCtx &x, MmParam &p) { x.err = 3; }

__global__ void __launch_bounds__(64)
k_seed_mm(const IndexDev fw, const IndexDev bw, const uint8_t *__restrict__ seqs, const uint8_t *__restrict__ quals,
          const bt2gpu_read *__restrict__ reads, const bt2gpu_seed_job *__restrict__ jobs, uint32_t njobs, int mms,
          uint32_t cap_per_job, bt2gpu_mmseed_result *__restrict__ out, bt2gpu_mmseed_hit *__restrict__ hits, int *__restrict__ errflag)
{
	const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
	if(j >= njobs) return;
	const bt2gpu_seed_job jb = jobs[j];
	const bt2gpu_read rd = reads[jb.read];
	const int len = jb.len;
	uint8_t seq[MM_MAXLEN], qual[MM_MAXLEN];
	for(int i = 0; i < len; i++) {                               // instantiateSeq: the window as it aligns to the Watson strand
		const uint32_t o = rd.rd_off + jb.off + (jb.fw ? i : len - 1 - i);
		const int c = seqs[o];
		seq[i] = (uint8_t)(jb.fw ? c : comp_dna(c));
		qual[i] = quals[o];
	}
	// Seed::mmSeeds
	int types[3]; MmCons zones[3][3]; int npol = 0;
	if(mms == 0) {
		types[0] = BT2GPU_SEED_EXACT; zones[0][0] = zones[0][1] = zones[0][2] = mmc_exact(); npol = 1;
	} else if(mms == 1) {
		types[0] = BT2GPU_SEED_LEFT_TO_RIGHT; zones[0][0] = mmc_exact(); zones[0][1] = mmc_mm(1); zones[0][2] = mmc_exact();
		types[1] = BT2GPU_SEED_RIGHT_TO_LEFT; zones[1][0] = mmc_exact(); zones[1][1] = mmc_mm(1); zones[1][1].mmsCeil = 0; zones[1][2] = mmc_exact();
		npol = 2;
	} else {
		types[0] = BT2GPU_SEED_LEFT_TO_RIGHT; zones[0][0] = mmc_exact(); zones[0][1] = mmc_mm(2); zones[0][2] = mmc_exact();
		types[1] = BT2GPU_SEED_RIGHT_TO_LEFT; zones[1][0] = mmc_exact(); zones[1][1] = mmc_mm(2); zones[1][1].mmsCeil = 1; zones[1][2] = mmc_exact();
		types[2] = BT2GPU_SEED_INSIDE_OUT; zones[2][0] = mmc_exact(); zones[2][1] = mmc_mm(1); zones[2][1].mmsCeil = 0;
		zones[2][2] = mmc_mm(1); zones[2][2].mmsCeil = 0;
		npol = 3;
	}
	MmSeed seeds[3]; MmParam pv[3]; bool done[3];
	MmCtx x;
	x.fw = &fw; x.bw = &bw; x.seeds = seeds; x.seq = seq; x.qual = qual;
	x.hits = hits + (size_t)j * cap_per_job; x.cap = cap_per_job; x.n = 0; x.bwops = x.bwedits = 0; x.err = 0;
	int np = 0;
	for(int k = 0; k < npol; k++)
		if(mm_instantiate(types[k], zones[k], len, seq, qual, seeds[np])) { mm_param_init(x, pv[np], np); np++; }
	int nleft = np;
	for(int n = 0; n < np; n++) { done[n] = mm_start(x, pv[n]); if(done[n]) nleft--; }
	while(nleft > 0) {
		for(int n = 0; n < np; n++) {
			if(done[n]) continue;
			if(mm_step<0>(x, pv[n])) { done[n] = true; nleft--; }
		}
	}
	bt2gpu_mmseed_result r;
	r.hit_off = 0; r.nhits = x.n; r.bwops = x.bwops; r.bwedits = x.bwedits; r.ntypes = (uint32_t)np; r.pad = 0;
	out[j] = r;
	if(x.err) atomicMax(errflag, x.err);
}

} // namespace bt2

extern "C" int bt2gpu_seed_mm_batch(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw,
                                    const bt2gpu_seed_job *jobs, size_t njobs, int mms, bt2gpu_mmseed_result *out,
                                    bt2gpu_mmseed_hit *hits, size_t hit_cap, size_t *nhits)
{
	if(!ctx || !ctx->fm) BT2_FAIL("bt2gpu_reads_stage must be called first");
	if(!ctx->have_sc) BT2_FAIL("bt2gpu_scoring_set must be called first");
	if(!out || !hits || !nhits) BT2_FAIL("null argument");
	if(mms < 0 || mms > 2) BT2_FAIL("seed mismatches must be 0, 1 or 2 (-N)");
	if(check_idx(ctx, fw, false) || check_idx(ctx, bw, false)) return 1;
	if(fw->dev.ftab_chars != bw->dev.ftab_chars) BT2_FAIL("forward and mirror index disagree on ftabChars");
	BT2_CUDA(cudaSetDevice(ctx->device));
	FmState *f = ctx->fm;
	*nhits = 0;
	if(njobs == 0) return 0;
	if(njobs > 0x0fffffffu) BT2_FAIL("too many seed jobs");
	for(size_t i = 0; i < njobs; i++) {
		const bt2gpu_seed_job &j = jobs[i];
		if(j.read >= f->nreads) BT2_FAIL("seed job names a read that is not staged");
		if(j.len < 1 || j.len > MM_MAXLEN) BT2_FAIL("seed length must be 1..32");
		if((uint32_t)j.off + j.len > f->rdlen[j.read]) BT2_FAIL("seed window outside the read");
	}
	size_t per = hit_cap / njobs;
	if(per < 1) BT2_FAIL("seed hit pool overflow (raise hit_cap)");
	if(per > 4096) per = 4096;
	if(f->jobs.reserve(njobs * sizeof(bt2gpu_seed_job)) || f->out.reserve(njobs * sizeof(bt2gpu_mmseed_result)) ||
	   f->aux.reserve(njobs * per * sizeof(bt2gpu_mmseed_hit)) || f->scal.reserve(64) || f->h_scal.reserve(64)) return 1;
	cudaStream_t s = ctx->stream;
	BT2_CUDA(cudaMemcpyAsync(f->jobs.p, jobs, njobs * sizeof(bt2gpu_seed_job), cudaMemcpyHostToDevice, s));
	BT2_CUDA(cudaMemsetAsync(f->scal.p, 0, 64, s));
	if(ctx->kstart()) return 1;
	k_seed_mm<<<(uint32_t)((njobs + 63) / 64), 64, 0, s>>>(fw->dev, bw->dev, f->seq.as<uint8_t>(), f->qual.as<uint8_t>(),
		f->reads.as<bt2gpu_read>(), f->jobs.as<bt2gpu_seed_job>(), (uint32_t)njobs, mms, (uint32_t)per,
		f->out.as<bt2gpu_mmseed_result>(), f->aux.as<bt2gpu_mmseed_hit>(), f->scal.as<int>());
	ctx->launches++;
	BT2_CUDA(cudaGetLastError());
	if(ctx->kend()) return 1;
	BT2_CUDA(bt2::settle(ctx, s));
	BT2_CUDA(cudaMemcpyAsync(out, f->out.p, njobs * sizeof(bt2gpu_mmseed_result), cudaMemcpyDeviceToHost, s));
	BT2_CUDA(cudaMemcpyAsync(f->h_scal.p, f->scal.p, 64, cudaMemcpyDeviceToHost, s));
	BT2_CUDA(bt2::wait_stream(ctx, s));
	const int err = *f->h_scal.as<int>();
	if(err == 1) BT2_FAIL("seed hit pool overflow (raise hit_cap)");
	if(err == 2) BT2_FAIL("a seed zone would allow a gap: not implemented by the reference either (aligner_seed.cpp:1967)");
	if(err) BT2_FAIL("seed search branched deeper than two edits");
	// compact the per-job slices into the caller's pool, job order = hit order
	std::vector<bt2gpu_mmseed_hit> tmp(njobs * per);
	size_t tot = 0;
	for(size_t i = 0; i < njobs; i++) tot += out[i].nhits;
	if(tot > hit_cap) BT2_FAIL("seed hit pool overflow (raise hit_cap)");
	if(tot) {
		BT2_CUDA(cudaMemcpyAsync(tmp.data(), f->aux.p, njobs * per * sizeof(bt2gpu_mmseed_hit), cudaMemcpyDeviceToHost, s));
		BT2_CUDA(bt2::wait_stream(ctx, s));
	}
	size_t o = 0;
	for(size_t i = 0; i < njobs; i++) {
		out[i].hit_off = (uint32_t)o;
		for(uint32_t k = 0; k < out[i].nhits; k++) hits[o++] = tmp[i * per + k];
	}
	*nhits = tot;
	return 0;
}

// ---- roofline probe ---------------------------------------------------------------------
// Independent (not dependent-chain) random 64-byte side gathers over the index itself at full
// occupancy: the HBM random-access ceiling the seed kernels are measured against
// ("measured 64-B-sector random-gather GB/s on the same B200").
namespace bt2 {
__global__ void __launch_bounds__(256)
k_gather_probe(const uint4 *__restrict__ sides, uint32_t nsides, uint32_t per_quad, uint32_t *__restrict__ sink)
{
	const uint32_t gq = (blockIdx.x * blockDim.x + threadIdx.x) >> 2;
	const int q = threadIdx.x & 3;
	uint32_t x = gq * 2654435761u + 12345u, acc = 0;
#pragma unroll 4
	for(uint32_t i = 0; i < per_quad; i++) {
		x = x * 1664525u + 1013904223u;
		const uint32_t side = (uint32_t)(((unsigned long long)(x ^ (x >> 15)) * nsides) >> 32);
		const uint4 v = __ldg(sides + (size_t)side * 4 + q);
		acc ^= v.x ^ v.y ^ v.z ^ v.w;
	}
	if(acc == 0x9e3779b9u) sink[0] = acc;
}
}

extern "C" int bt2gpu_index_gather_probe(bt2gpu_ctx *ctx, const bt2gpu_index *ix, uint64_t nloads, double *gbs) {
	if(!ctx || !gbs) BT2_FAIL("null argument");
	if(check_idx(ctx, ix, false)) return 1;
	BT2_CUDA(cudaSetDevice(ctx->device));
	FmState *f = fm_state(ctx);
	if(f->scal.reserve(64)) return 1;
	const uint32_t nsides = (uint32_t)ix->nsides;
	if(nsides == 0) BT2_FAIL("empty index");
	const uint32_t blocks = (uint32_t)ctx->sm_count * 8, quads = blocks * 64;
	uint32_t per = (uint32_t)std::max<uint64_t>(1, nloads / quads);
	for(int rep = 0; rep < 2; rep++) {                          // first pass warms up
		if(ctx->kstart()) return 1;
		k_gather_probe<<<blocks, 256, 0, ctx->stream>>>(ix->dev.sides, nsides, per, f->scal.as<uint32_t>());
		ctx->launches++;
		BT2_CUDA(cudaGetLastError());
		if(ctx->kend()) return 1;
	}
	float ms = 0.f;
	if(bt2gpu_ctx_kernel_ms(ctx, &ms)) return 1;
	*gbs = (double)quads * per * 64.0 / (ms * 1e-3) / 1e9;
	return 0;
}
