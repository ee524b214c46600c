// hotpath.cu — batch runtime: one chained pass of the hot path (include/bt2gpu.h,
// bt2gpu_hotpath_round0_batch).  Host C++ only: the arithmetic runs in the kernels of fm.cu and
// dp.cu through the same C-ABI entry points a caller would use; what is here is the glue the
// reference executes per read between those calls -- seed instantiation
// (aligner_seed.cpp:498-587), choosing a seed hit to extend, framing the window
// (dp_framer.cpp:81-129) and fetching it as masks (aligner_sw.cpp:155-271) -- done per batch on
// several host threads into pinned staging buffers.
#include "common.cuh"
#include <algorithm>
#include <cmath>
#include <thread>
#include <functional>

namespace bt2 {

struct HpState {
	HostBuf sjobs, shits, xjobs, xres, rows, offs, steps, dseq, dqual, dwin, djobs, pickidx;
	std::vector<uint32_t> seed_count, seed_first;      // per read
	std::vector<uint32_t> hit_idx;                     // seed-job index of each extend job
};

template<typename F> static void parallel_for(size_t n, int nthreads, F f) {
	if(nthreads <= 1 || n < 2048) { f(0, n); return; }
	std::vector<std::thread> th;
	const size_t chunk = (n + nthreads - 1) / nthreads;
	for(int t = 0; t < nthreads; t++) {
		const size_t lo = t * chunk, hi = std::min(n, lo + chunk);
		if(lo >= hi) break;
		th.emplace_back([=]() { f(lo, hi); });
	}
	for(auto &x : th) x.join();
}

// msIval.f<int>(len) for SIMPLE_FUNC_SQRT (simple_func.h:89-112), then max(.,1) (bt2_search.cpp:3396)
static inline int seed_interval(double c, double l, uint32_t rdlen) {
	double v = c + l * std::sqrt((double)rdlen);
	if(v < 0.0) v = 0.0;
	int iv = (int)v;
	return iv < 1 ? 1 : iv;
}

} // namespace bt2

using namespace bt2;

extern "C" int bt2gpu_hotpath_round0_batch(
	bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw,
	const uint8_t *text, uint64_t text_len,
	const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const bt2gpu_read *reads, size_t nreads, const int32_t *minsc, const int32_t *nceil,
	const bt2gpu_hotpath_params *P,
	bt2gpu_sweep_result *sweep, uint32_t *job_read, int64_t *job_start, uint8_t *job_fw, size_t *njobs_out,
	bt2gpu_dp_result *results, bt2gpu_dp_aln *alns, bt2gpu_edit *edits, size_t edit_cap, size_t *nedits_out,
	bt2gpu_hotpath_stats *S)
{
	if(!ctx || !fw || !bw || !text || !seq_pool || !qual_pool || !reads || !minsc || !nceil || !P || !sweep ||
	   !job_read || !job_start || !job_fw || !njobs_out || !results || !alns || !edits || !nedits_out || !S)
		BT2_FAIL("null argument");
	if(P->seed_len < 1 || P->seed_len > 32) BT2_FAIL("seed_len must be 1..32");
	if(P->maxgap < 0 || P->maxgap > 1000) BT2_FAIL("bad maxgap");
	if(!ctx->have_sc) BT2_FAIL("bt2gpu_scoring_set must be called first");
	memset(S, 0, sizeof(*S));
	*njobs_out = 0; *nedits_out = 0;
	if(nreads == 0) return 0;
	static thread_local HpState tl_st;         // staging reused across calls of this host thread
	HpState &st = tl_st;                       // (a local reference: worker-thread lambdas must not re-resolve the thread_local)
	int nth = P->nthreads > 0 ? P->nthreads : (int)std::thread::hardware_concurrency();
	if(nth < 1) nth = 1;
	if(nth > 32) nth = 32;
	float ms = 0.f;
	const uint32_t max_aln = P->max_aln ? P->max_aln : 1;

	// ---- 1. stage reads, exactSweep ----
	if(bt2gpu_reads_stage(ctx, seq_pool, qual_pool, pool_len, reads, nreads)) return 1;
	S->h2d_bytes += 2 * pool_len + nreads * sizeof(bt2gpu_read);
	if(bt2gpu_exact_sweep_batch(ctx, fw, 0, 0, P->mine_max ? P->mine_max : 2, sweep)) return 1;
	if(bt2gpu_ctx_kernel_ms(ctx, &ms)) return 1;
	S->ms_sweep = ms;
	S->d2h_bytes += nreads * sizeof(bt2gpu_sweep_result);

	// ---- 2. seed instantiation (windows without N; fw seeds of a read, then its rc seeds) ----
	st.seed_count.assign(nreads, 0);
	st.seed_first.assign(nreads + 1, 0);
	const int L = P->seed_len;
	auto count_or_emit = [&](size_t lo, size_t hi, bt2gpu_seed_job *out) {
		for(size_t i = lo; i < hi; i++) {
			const uint32_t rdlen = reads[i].rd_len;
			const uint8_t *rd = seq_pool + reads[i].rd_off;
			const int per = seed_interval(P->ival_const, P->ival_coeff, rdlen);
			const int sl = (int)std::min<uint32_t>((uint32_t)L, rdlen);
			int nseeds = 1;
			if((int)rdlen > L) nseeds += ((int)rdlen - L) / per;
			uint32_t cnt = 0;
			bt2gpu_seed_job *o = out ? out + st.seed_first[i] : nullptr;
			for(int k = 0; k < nseeds; k++) {
				const int depth = k * per;
				bool clean = true;
				for(int j = 0; j < sl; j++) if(rd[depth + j] > 3) { clean = false; break; }
				if(!clean) continue;
				if(o) { o[cnt].read = (uint32_t)i; o[cnt].off = (uint16_t)depth; o[cnt].len = (uint8_t)sl; o[cnt].fw = 1; }
				cnt++;
			}
			if(o) for(uint32_t c = 0; c < cnt; c++) { o[cnt + c] = o[c]; o[cnt + c].fw = 0; }
			st.seed_count[i] = 2 * cnt;
		}
	};
	parallel_for(nreads, nth, [&](size_t lo, size_t hi) { count_or_emit(lo, hi, nullptr); });
	for(size_t i = 0; i < nreads; i++) st.seed_first[i + 1] = st.seed_first[i] + st.seed_count[i];
	const size_t nseed = st.seed_first[nreads];
	if(st.sjobs.reserve(nseed * sizeof(bt2gpu_seed_job) + 64) || st.shits.reserve(nseed * sizeof(bt2gpu_seed_hit) + 64)) return 1;
	bt2gpu_seed_job *sj = st.sjobs.as<bt2gpu_seed_job>();
	bt2gpu_seed_hit *sh = st.shits.as<bt2gpu_seed_hit>();
	parallel_for(nreads, nth, [&](size_t lo, size_t hi) { count_or_emit(lo, hi, sj); });
	if(bt2gpu_seed_exact_batch(ctx, fw, bw, sj, nseed, sh)) return 1;
	if(nseed) { if(bt2gpu_ctx_kernel_ms(ctx, &ms)) return 1; S->ms_seed = ms; }
	S->seeds = nseed;
	S->h2d_bytes += nseed * sizeof(bt2gpu_seed_job);
	S->d2h_bytes += nseed * sizeof(bt2gpu_seed_hit);

	// ---- 3. extend every hit ----
	st.hit_idx.clear();
	for(size_t k = 0; k < nseed; k++) {
		S->seed_bwops += sh[k].bwops; S->seed_sides += sh[k].sides;
		if(sh[k].botf > sh[k].topf) st.hit_idx.push_back((uint32_t)k);
	}
	const size_t nhit = st.hit_idx.size();
	S->seed_hits = nhit;
	if(st.xjobs.reserve(nhit * sizeof(bt2gpu_extend_job) + 64) || st.xres.reserve(nhit * sizeof(bt2gpu_extend_result) + 64)) return 1;
	bt2gpu_extend_job *xj = st.xjobs.as<bt2gpu_extend_job>();
	bt2gpu_extend_result *xr = st.xres.as<bt2gpu_extend_result>();
	for(size_t h = 0; h < nhit; h++) {
		const uint32_t k = st.hit_idx[h];
		xj[h].read = sj[k].read; xj[h].topf = sh[k].topf; xj[h].botf = sh[k].botf; xj[h].topb = sh[k].topb; xj[h].botb = sh[k].botb;
		xj[h].off = sj[k].off; xj[h].len = sj[k].len; xj[h].fw = sj[k].fw;
	}
	if(bt2gpu_extend_batch(ctx, fw, bw, xj, nhit, xr)) return 1;
	if(nhit) { if(bt2gpu_ctx_kernel_ms(ctx, &ms)) return 1; S->ms_extend = ms; }
	S->h2d_bytes += nhit * sizeof(bt2gpu_extend_job);
	S->d2h_bytes += nhit * sizeof(bt2gpu_extend_result);

	// ---- 4. pick one hit per read: narrowest range, then longest exact extension, then first ----
	if(st.pickidx.reserve(nreads * 4 + 64) || st.rows.reserve(nreads * 4 + 64) || st.offs.reserve(nreads * 4 + 64) ||
	   st.steps.reserve(nreads * 4 + 64)) return 1;
	uint32_t *pick = st.pickidx.as<uint32_t>(), *rows = st.rows.as<uint32_t>();
	size_t npick = 0;
	for(size_t h = 0; h < nhit;) {
		size_t e = h, best = h;
		while(e < nhit && xj[e].read == xj[h].read) {
			S->extend_fmops += xr[e].fmops;
			const uint32_t sz = xj[e].botf - xj[e].topf, bsz = xj[best].botf - xj[best].topf;
			const uint32_t ex = (uint32_t)xr[e].nlex + xr[e].nrex, bex = (uint32_t)xr[best].nlex + xr[best].nrex;
			if(sz < bsz || (sz == bsz && ex > bex)) best = e;
			e++;
		}
		pick[npick] = (uint32_t)best; rows[npick] = xj[best].topf; npick++;
		h = e;
	}
	// ---- 5. locate ----
	uint32_t *offs = st.offs.as<uint32_t>(), *steps = st.steps.as<uint32_t>();
	if(bt2gpu_locate_batch(ctx, fw, rows, npick, offs, steps)) return 1;
	if(npick) { if(bt2gpu_ctx_kernel_ms(ctx, &ms)) return 1; S->ms_locate = ms; }
	S->locate_rows = npick;
	S->h2d_bytes += npick * 4; S->d2h_bytes += npick * 8;

	// ---- 6. frame the windows (frameSeedExtensionRect) and fetch them as masks (initRef) ----
	const int mg = P->maxgap;
	size_t tot_rd = 0, tot_win = 0;
	std::vector<uint32_t> &rdoff = st.seed_count;       // reuse: per-pick offsets into the DP pools
	rdoff.assign(npick + 1, 0);
	std::vector<uint32_t> winoff(npick + 1, 0);
	for(size_t k = 0; k < npick; k++) {
		const uint32_t rdlen = reads[xj[pick[k]].read].rd_len;
		S->locate_steps += steps[k];
		rdoff[k] = (uint32_t)tot_rd; winoff[k] = (uint32_t)tot_win;
		tot_rd += rdlen; tot_win += rdlen + 4 * mg + 1;
	}
	if(tot_rd > 0xfffffff0ull || tot_win > 0xfffffff0ull) BT2_FAIL("DP pools exceed 4 GiB: split the batch");
	if(st.dseq.reserve(tot_rd + 64) || st.dqual.reserve(tot_rd + 64) || st.dwin.reserve(tot_win + 64) ||
	   st.djobs.reserve(npick * sizeof(bt2gpu_dp_job) + 64)) return 1;
	uint8_t *dseq = st.dseq.as<uint8_t>(), *dqual = st.dqual.as<uint8_t>(), *dwin = st.dwin.as<uint8_t>();
	bt2gpu_dp_job *dj = st.djobs.as<bt2gpu_dp_job>();
	parallel_for(npick, nth, [&](size_t lo, size_t hi) {
		for(size_t k = lo; k < hi; k++) {
			const bt2gpu_extend_job &x = xj[pick[k]];
			const uint32_t ri = x.read, rdlen = reads[ri].rd_len;
			const uint8_t *rd = seq_pool + reads[ri].rd_off, *qu = qual_pool + reads[ri].rd_off;
			uint8_t *os = dseq + rdoff[k], *oq = dqual + rdoff[k], *ow = dwin + winoff[k];
			if(x.fw) { memcpy(os, rd, rdlen); memcpy(oq, qu, rdlen); }
			else for(uint32_t i = 0; i < rdlen; i++) { const uint8_t c = rd[rdlen - 1 - i]; os[i] = c > 3 ? 4 : (uint8_t)(3 - c); oq[i] = qu[rdlen - 1 - i]; }
			const int64_t seedoff = x.fw ? (int64_t)x.off : (int64_t)rdlen - x.off - x.len;   // in the oriented read
			const int64_t start = (int64_t)offs[k] - seedoff;
			const int64_t refl = start - 2 * mg;
			const uint32_t ncol = rdlen + 4 * (uint32_t)mg;
			for(uint32_t c = 0; c <= ncol; c++) {
				const int64_t p = refl + c;
				ow[c] = (p >= 0 && (uint64_t)p < text_len) ? (uint8_t)(1u << text[p]) : (uint8_t)16;
			}
			bt2gpu_dp_job j; memset(&j, 0, sizeof(j));
			j.rd_off = rdoff[k]; j.rd_len = rdlen; j.rf_off = winoff[k]; j.ncol = ncol;
			j.minsc = minsc[ri]; j.nceil = nceil[ri]; j.refl = refl; j.triml = 0;
			j.corel = (uint32_t)mg; j.corer = 3 * (uint32_t)mg; j.fw = x.fw; j.enable8 = 1;
			dj[k] = j;
			job_read[k] = ri; job_start[k] = start; job_fw[k] = x.fw;
		}
	});
	// ---- 7. fill + gather + nextAlignment loop ----
	if(npick) {
		if(bt2gpu_dp_stage(ctx, dseq, dqual, tot_rd, dwin, tot_win, dj, npick, max_aln, 256 * npick, edit_cap)) return 1;
		if(bt2gpu_dp_run(ctx, 7)) return 1;
		size_t nc = 0, ne = 0;
		if(bt2gpu_dp_fetch(ctx, results, nullptr, alns, edits, &nc, &ne)) return 1;
		if(bt2gpu_ctx_kernel_ms(ctx, &ms)) return 1;
		S->ms_dp = ms;
		float ph[3] = {0, 0, 0};
		if(bt2gpu_dp_phase_ms(ctx, ph) == 0) { S->ms_dp_fill = ph[0]; S->ms_dp_gather = ph[1]; S->ms_dp_backtrace = ph[2]; }
		*nedits_out = ne;
		S->dp_cells = bt2gpu_dp_staged_cells(ctx);
		for(size_t k = 0; k < npick; k++) if(results[k].naln > 0) S->aligned++;
		S->h2d_bytes += 2 * tot_rd + tot_win + npick * sizeof(bt2gpu_dp_job);
		S->d2h_bytes += npick * (sizeof(bt2gpu_dp_result) + max_aln * sizeof(bt2gpu_dp_aln)) + ne * sizeof(bt2gpu_edit);
	}
	S->dp_jobs = npick;
	*njobs_out = npick;
	return 0;
}
