/*
 * emul_abi.c — TEST INFRASTRUCTURE, NOT PRODUCT.  A test double of libbt2gpu.so for machines without a GPU.
 *
 * It exports the subset of the C ABI (include/bt2gpu.h) that the batched reference-side binding
 * (bowtie2-server_b200/shim/batch_shim.cpp) calls, answered by the CPU oracle (dp_oracle.c, fm_oracle.c: the pinned
 * restatement).  Its only use is `oracle/_ref/bowtie2-align-server-s-emul` (oracle/Makefile, target emul): the
 * unmodified reference + the shim + the fiber scheduler linked against THIS library, so that the scheduler, the batch
 * assembly and the scatter of results -- host logic with no arithmetic of its own -- can be tested for SAM identity in
 * the `-m "not gpu"` suite.  The product server (bowtie2-align-server-s-gpub) links libbt2gpu.so and has no CPU path;
 * nothing under bowtie2-server_b200/ references this file.
 */
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include "../include/bt2gpu.h"

int bt2oracle_dp_align_batch(const bt2gpu_scoring *sc, const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const uint8_t *win_pool, size_t win_len, const bt2gpu_dp_job *jobs, size_t njobs, uint32_t max_aln,
	bt2gpu_dp_result *results, bt2gpu_dp_cand *cands, size_t cand_cap, bt2gpu_dp_aln *alns, bt2gpu_edit *edits, size_t edit_cap,
	size_t *ncands, size_t *nedits);
int bt2oracle_ungapped_batch(const bt2gpu_scoring *sc, const uint8_t *seq_pool, const uint8_t *qual_pool,
	const uint8_t *win_pool, const bt2gpu_ungap_job *jobs, size_t njobs, bt2gpu_ungap_result *results, bt2gpu_edit *edits,
	size_t edit_cap, size_t *nedits);
int bt2oracle_seed_exact_batch(const bt2gpu_index_desc *fw, const bt2gpu_index_desc *bw, const uint8_t *seq_pool, const bt2gpu_read *reads,
	const bt2gpu_seed_job *jobs, size_t njobs, bt2gpu_seed_hit *out);
int bt2oracle_exact_sweep_batch(const bt2gpu_index_desc *fw, const uint8_t *seq_pool, const bt2gpu_read *reads, size_t nreads,
	int nofw, int norc, uint32_t mine_max, bt2gpu_sweep_result *out);
int bt2oracle_onemm_batch(const bt2gpu_index_desc *fw, const bt2gpu_index_desc *bw, const bt2gpu_scoring *scs, const uint8_t *seq_pool,
	const uint8_t *qual_pool, const bt2gpu_read *reads, size_t nreads, const bt2gpu_onemm_job *jobs, int local, int repex, int rep1mm,
	bt2gpu_onemm_result *out, bt2gpu_ee_hit *hits, size_t hit_cap, size_t *nhits);
int bt2oracle_extend_batch(const bt2gpu_index_desc *fw, const bt2gpu_index_desc *bw, const uint8_t *seq_pool, const bt2gpu_read *reads,
	const bt2gpu_extend_job *jobs, size_t njobs, bt2gpu_extend_result *out);
int bt2oracle_locate_batch(const bt2gpu_index_desc *ix, const uint32_t *rows, size_t n, uint32_t *offs_out, uint32_t *steps_out);

struct bt2gpu_ctx {
	bt2gpu_scoring sc; int have_sc;
	uint8_t *seq, *qual; bt2gpu_read *reads; size_t nreads;
	uint64_t calls;
};
struct bt2gpu_index { bt2gpu_index_desc d; };     /* host pointers: the Ebwt outlives the server's workers */

static __thread const char *g_err = "";
static int fail(const char *m) { g_err = m; return 1; }

int bt2gpu_abi_version(void) { return BT2GPU_ABI_VERSION; }
const char *bt2gpu_last_error(void) { return g_err; }
int bt2gpu_device_blocking_sync(int device, int on) { (void)device; (void)on; return 0; }
int bt2gpu_device_sync_mode(int device, int mode) { (void)device; (void)mode; return 0; }
int bt2gpu_ctx_set_wait(bt2gpu_ctx *ctx, uint32_t poll_ns) { (void)ctx; (void)poll_ns; return 0; }
int bt2gpu_ctx_set_timing(bt2gpu_ctx *ctx, int on) { (void)ctx; (void)on; return 0; }
int bt2gpu_ctx_create(int device, bt2gpu_ctx **out) {
	(void)device;
	if(getenv("BT2GPU_EMUL_NO_DEVICE")) return fail("no CUDA device (emulated)");
	*out = (bt2gpu_ctx*)calloc(1, sizeof(bt2gpu_ctx));
	return *out ? 0 : fail("out of memory");
}
void bt2gpu_ctx_destroy(bt2gpu_ctx *c) { if(c) { free(c->seq); free(c->qual); free(c->reads); free(c); } }
uint64_t bt2gpu_ctx_launches(bt2gpu_ctx *c) { return c ? c->calls : 0; }
int bt2gpu_scoring_set(bt2gpu_ctx *c, const bt2gpu_scoring *sc) { c->sc = *sc; c->have_sc = 1; return 0; }

int bt2gpu_index_create(int device, const bt2gpu_index_desc *d, bt2gpu_index **out) {
	(void)device;
	*out = (bt2gpu_index*)calloc(1, sizeof(bt2gpu_index));
	if(!*out) return fail("out of memory");
	(*out)->d = *d;
	return 0;
}
void bt2gpu_index_destroy(bt2gpu_index *ix) { free(ix); }

int bt2gpu_reads_stage(bt2gpu_ctx *c, const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
                       const bt2gpu_read *reads, size_t nreads) {
	free(c->seq); free(c->qual); free(c->reads);
	c->seq = (uint8_t*)malloc(pool_len + 1); c->qual = (uint8_t*)malloc(pool_len + 1);
	c->reads = (bt2gpu_read*)malloc((nreads + 1) * sizeof(bt2gpu_read));
	memcpy(c->seq, seq_pool, pool_len);
	if(qual_pool) memcpy(c->qual, qual_pool, pool_len); else memset(c->qual, 0xff, pool_len);   /* poison: a call that reads them shows */
	memcpy(c->reads, reads, nreads * sizeof(bt2gpu_read));
	c->nreads = nreads;
	return 0;
}
int bt2oracle_seed_mm_batch(const bt2gpu_index_desc *fw, const bt2gpu_index_desc *bw, const bt2gpu_scoring *scs, const uint8_t *seq_pool,
                            const uint8_t *qual_pool, const bt2gpu_read *reads, const bt2gpu_seed_job *jobs, size_t njobs, int mms,
                            uint32_t *hit_job, uint32_t *hit_ranges, uint64_t *hit_keys, size_t cap, size_t *nhits, uint32_t *out_n,
                            uint64_t *bwops);
int bt2gpu_seed_mm_batch(bt2gpu_ctx *c, const bt2gpu_index *fw, const bt2gpu_index *bw, const bt2gpu_seed_job *jobs, size_t njobs, int mms,
                         bt2gpu_mmseed_result *out, bt2gpu_mmseed_hit *hits, size_t hit_cap, size_t *nhits) {
	c->calls++;
	if(!c->have_sc) return fail("bt2gpu_scoring_set must be called first");
	*nhits = 0;
	if(njobs == 0) return 0;
	uint32_t *hj = (uint32_t*)malloc((hit_cap + 1) * 4), *hr = (uint32_t*)malloc((hit_cap + 1) * 16), *on = (uint32_t*)malloc(njobs * 4);
	uint64_t *hk = (uint64_t*)malloc((hit_cap + 1) * 8), ops = 0;
	size_t nh = 0;
	/* job by job: the double reports bwops per job as the library does */
	int rc = 0; size_t o = 0;
	for(size_t j = 0; j < njobs && !rc; j++) {
		size_t n1 = 0; uint64_t ops1 = 0;
		rc = bt2oracle_seed_mm_batch(&fw->d, &bw->d, &c->sc, c->seq, c->qual, c->reads, jobs + j, 1, mms, hj, hr + 4 * o, hk + o, hit_cap - o, &n1, on + j, &ops1);
		out[j].hit_off = (uint32_t)o; out[j].nhits = (uint32_t)n1; out[j].bwops = (uint32_t)ops1; out[j].bwedits = 0; out[j].ntypes = 0; out[j].pad = 0;
		o += n1; ops += ops1;
	}
	nh = o;
	for(size_t i = 0; i < nh && !rc; i++) { hits[i].topf = hr[4 * i]; hits[i].botf = hr[4 * i + 1]; hits[i].topb = hr[4 * i + 2]; hits[i].botb = hr[4 * i + 3]; hits[i].key = hk[i]; }
	free(hj); free(hr); free(on); free(hk);
	if(rc) return fail("seed hit pool overflow (raise hit_cap)");
	*nhits = nh;
	return 0;
}
int bt2gpu_exact_sweep_batch(bt2gpu_ctx *c, const bt2gpu_index *fw, int nofw, int norc, uint32_t mine_max, bt2gpu_sweep_result *out) {
	c->calls++;
	return bt2oracle_exact_sweep_batch(&fw->d, c->seq, c->reads, c->nreads, nofw, norc, mine_max, out) ? fail("sweep") : 0;
}
int bt2gpu_onemm_batch(bt2gpu_ctx *c, const bt2gpu_index *fw, const bt2gpu_index *bw, const bt2gpu_onemm_job *jobs, int local,
                       int repex, int rep1mm, bt2gpu_onemm_result *out, bt2gpu_ee_hit *hits, size_t hit_cap, size_t *nhits) {
	c->calls++;
	if(!c->have_sc) return fail("bt2gpu_scoring_set must be called first");
	return bt2oracle_onemm_batch(&fw->d, &bw->d, &c->sc, c->seq, c->qual, c->reads, c->nreads, jobs, local, repex, rep1mm, out, hits,
	                             hit_cap, nhits) ? fail("1-mismatch hit pool overflow") : 0;
}
int bt2gpu_seed_exact_batch(bt2gpu_ctx *c, const bt2gpu_index *fw, const bt2gpu_index *bw, const bt2gpu_seed_job *jobs, size_t njobs,
                            bt2gpu_seed_hit *out) {
	c->calls++;
	return bt2oracle_seed_exact_batch(&fw->d, &bw->d, c->seq, c->reads, jobs, njobs, out) ? fail("seed") : 0;
}
int bt2gpu_extend_batch(bt2gpu_ctx *c, const bt2gpu_index *fw, const bt2gpu_index *bw, const bt2gpu_extend_job *jobs, size_t njobs,
                        bt2gpu_extend_result *out) {
	c->calls++;
	return bt2oracle_extend_batch(&fw->d, &bw->d, c->seq, c->reads, jobs, njobs, out) ? fail("extend") : 0;
}
int bt2gpu_locate_batch(bt2gpu_ctx *c, const bt2gpu_index *fw, const uint32_t *rows, size_t n, uint32_t *offs_out, uint32_t *steps_out) {
	c->calls++;
	uint32_t *tmp = steps_out ? NULL : (uint32_t*)malloc((n + 1) * 4);
	int rc = bt2oracle_locate_batch(&fw->d, rows, n, offs_out, steps_out ? steps_out : tmp);
	free(tmp);
	return rc ? fail("locate") : 0;
}
int bt2gpu_dp_align_batch(bt2gpu_ctx *c, const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len, const uint8_t *win_pool,
                          size_t win_len, const bt2gpu_dp_job *jobs, size_t njobs, uint32_t max_aln, bt2gpu_dp_result *results,
                          bt2gpu_dp_cand *cands, size_t cand_cap, bt2gpu_dp_aln *alns, bt2gpu_edit *edits, size_t edit_cap) {
	size_t nc = 0, ne = 0;
	c->calls++;
	if(!c->have_sc) return fail("bt2gpu_scoring_set must be called first");
	const int rc = bt2oracle_dp_align_batch(&c->sc, seq_pool, qual_pool, pool_len, win_pool, win_len, jobs, njobs, max_aln, results, cands,
	                                        cand_cap, alns, edits, edit_cap, &nc, &ne);
	if(rc == 1) return fail("candidate pool overflow (raise cand_cap)");      /* the library's messages: the shim reads them */
	return rc ? fail("edit pool overflow (raise edit_cap)") : 0;
}
int bt2gpu_ungapped_batch(bt2gpu_ctx *c, const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len, const uint8_t *win_pool,
                          size_t win_len, const bt2gpu_ungap_job *jobs, size_t njobs, bt2gpu_ungap_result *results, bt2gpu_edit *edits,
                          size_t edit_cap, size_t *nedits) {
	(void)pool_len; (void)win_len;
	c->calls++;
	if(!c->have_sc) return fail("bt2gpu_scoring_set must be called first");
	return bt2oracle_ungapped_batch(&c->sc, seq_pool, qual_pool, win_pool, jobs, njobs, results, edits, edit_cap, nedits) ? fail("ungapped edit pool overflow") : 0;
}
