/*
 * bt2gpu.h — C ABI of the B200-native alignment hot path (libbt2gpu.so).
 *
 * This is the drop-in boundary for the two arithmetic cores of
 * sfiligoi/bowtie2-server (Bowtie 2 v2.5.4).  The reference has no FFI layer:
 * its "operator API" for this path is the C++ classes SeedAligner / SwAligner
 * plus the read-only Ebwt / BitPairReference / Scoring objects (SURVEY.md
 * §8b).  Each entry point below names the reference interface it replaces
 * (file:line relative to the reference tree).  Plain pointers and sizes only;
 * no C++ or torch types.  All functions return 0 on success, non-zero on
 * error (message via bt2gpu_last_error()).  There is NO CPU fallback: every
 * compute entry point fails if no CUDA device is usable.
 *
 * Threading: a bt2gpu_ctx owns one CUDA stream and its scratch arenas; use
 * one ctx per host thread (mirrors "one SwAligner/SeedAligner per worker
 * thread", bt2_search.cpp:3128-3130).  Index/reference handles are immutable
 * after creation and may be shared by all contexts of the same device.
 */
#ifndef BT2GPU_H_
#define BT2GPU_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BT2GPU_ABI_VERSION 1

typedef struct bt2gpu_ctx   bt2gpu_ctx;
typedef struct bt2gpu_index bt2gpu_index;   /* one Ebwt (forward or mirror)   */
typedef struct bt2gpu_ref   bt2gpu_ref;     /* BitPairReference (.3/.4.bt2)    */

/* ---- library / context ------------------------------------------------- */
int         bt2gpu_abi_version(void);
const char *bt2gpu_last_error(void);
int         bt2gpu_device_count(int *n);
int         bt2gpu_ctx_create(int device, bt2gpu_ctx **ctx);
void        bt2gpu_ctx_destroy(bt2gpu_ctx *ctx);
/* Block until everything enqueued on the context's stream has finished. */
int         bt2gpu_ctx_sync(bt2gpu_ctx *ctx);
/* Raw cudaStream_t of the context (for event timing by the harness). */
void       *bt2gpu_ctx_stream(bt2gpu_ctx *ctx);
/* Number of kernel launches issued through this context so far. */
uint64_t    bt2gpu_ctx_launches(bt2gpu_ctx *ctx);

/* ---- scoring (replaces class Scoring, scoring.h:96-486) ----------------- */
typedef struct {
	int32_t mmpens[256];   /* Scoring::mmpens  (scoring.h:116-120)            */
	int32_t npens[256];    /* Scoring::npens                                  */
	int32_t match_bonus;   /* Scoring::match(30)  (0 end-to-end, 2 --local)   */
	int32_t rd_gap_open;   /* Scoring::readGapOpen()   = rdGapConst+rdGapLinear (scoring.h:417) */
	int32_t rd_gap_ext;    /* Scoring::readGapExtend() = rdGapLinear                             */
	int32_t rf_gap_open;   /* Scoring::refGapOpen()                                              */
	int32_t rf_gap_ext;    /* Scoring::refGapExtend()                                            */
	int32_t gapbar;        /* Scoring::gapbar (--gbar)                                           */
	int32_t monotone;      /* Scoring::monotone: 1 = end-to-end, 0 = local                       */
} bt2gpu_scoring;
int bt2gpu_scoring_set(bt2gpu_ctx *ctx, const bt2gpu_scoring *sc);

/* ---- dynamic programming (replaces SwAligner, aligner_sw.h:241-374) ------ */

/* One DP problem = SwAligner::initRead + initRef (aligner_sw.cpp:34-144).
 * The read is given already oriented (rdfw_ or rdrc_ as initRef's `fw`
 * selects), as bytes 0..4 (A,C,G,T,N) with Phred+33 quality bytes.  The
 * window is given as reference masks 1,2,4,8 (16 = N), ncol+1 bytes: the
 * last byte is the extra right-hand column of aligner_sw.cpp:174-176. */
typedef struct {
	uint32_t rd_off;      /* offset of the read in seq_pool / qual_pool          */
	uint32_t rd_len;      /* rdf_ - rdi_  (= DP rows)                            */
	uint32_t rf_off;      /* offset of the window in win_pool                    */
	uint32_t ncol;        /* rff_ - rfi_  (= DP columns)                         */
	int32_t  minsc;       /* minsc_                                              */
	int32_t  nceil;       /* nceil_ = sc.nCeil.f(rdlen) (aligner_sw.cpp:44)      */
	int64_t  refl;        /* rect.refl                                           */
	uint32_t triml;       /* rect.triml                                          */
	uint32_t corel;       /* rect.corel                                          */
	uint32_t corer;       /* rect.corer                                          */
	uint8_t  fw;          /* read orientation (for invertEdits, aligner_sw.cpp:1137) */
	uint8_t  enable8;     /* enable8_ && !readSse16_  (aligner_sw.cpp:518,570)   */
	uint8_t  reserved[2];
} bt2gpu_dp_job;

/* SwAligner::align() outcome (aligner_sw.cpp:500-729). */
typedef struct {
	int64_t  best;        /* `best` out-parameter (INT64_MIN when the reference returns MIN_I64) */
	int32_t  aligned;     /* return value of align()                           */
	int32_t  width;       /* 8: sse8succ_, 16: sse16succ_, 0: neither           */
	uint32_t ncand;       /* btncand_.size()                                    */
	uint32_t cand_off;    /* first candidate of this job in the candidate pool  */
	uint32_t naln;        /* alignments produced by the nextAlignment() loop    */
	uint32_t aln_off;     /* first alignment of this job in the alignment pool  */
	uint32_t tail_tries;  /* backtraces attempted (RNG draws) after the last success */
	uint32_t cural;       /* cural_ after the loop                              */
} bt2gpu_dp_result;

/* DpBtCandidate (aligner_sw_nuc.h:40-142), sorted: score desc, row desc, col desc. */
typedef struct { uint32_t row, col; int32_t score; } bt2gpu_dp_cand;

/* One successful SwAligner::nextAlignment() (aligner_sw.cpp:737-1146). */
typedef struct {
	int32_t  score;       /* AlnScore::score_                                   */
	int32_t  ns;          /* AlnScore::ns_                                      */
	int32_t  gaps;        /* AlnScore::gaps_                                    */
	int32_t  refns;       /* AlnRes::refNs()                                    */
	int64_t  refoff;      /* AlnRes::refoff() = col + rfi_ + rect.refl          */
	uint32_t trim5;       /* soft trim at the read's 5' end                     */
	uint32_t trim3;       /* soft trim at the read's 3' end                     */
	uint32_t nedits;      /* ned_.size()                                        */
	uint32_t edit_off;    /* first edit in the edit pool                        */
	uint32_t cural;       /* cural_ after this call (candidate index + 1)       */
	uint32_t tries;       /* backtraces attempted for this call, incl. the successful one:
	                         each one is an rnd.nextU32() re-seed (aligner_sw.cpp:794,879,974,1062) */
} bt2gpu_dp_aln;

/* Edit (edit.h:58): pos w.r.t. the read's 5' end after invertEdits. */
typedef struct { uint32_t pos; uint8_t chr; uint8_t qchr; uint8_t type; uint8_t pad; } bt2gpu_edit;
#define BT2GPU_EDIT_READ_GAP 1   /* EDIT_TYPE_READ_GAP (edit.h:36) */
#define BT2GPU_EDIT_REF_GAP  2   /* EDIT_TYPE_REF_GAP             */
#define BT2GPU_EDIT_MM       3   /* EDIT_TYPE_MM                  */

/* Whole-batch initRef+align+nextAlignment* with HOST buffers (H2D and D2H
 * copies inside).  `max_aln` bounds the nextAlignment() loop per job.
 * Pools: cands[cand_cap], alns[njobs*max_aln], edits[edit_cap].  On overflow
 * of cand_cap / edit_cap the call fails (no silent truncation). */
int bt2gpu_dp_align_batch(
	bt2gpu_ctx *ctx,
	const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const uint8_t *win_pool, size_t win_len,
	const bt2gpu_dp_job *jobs, size_t njobs, uint32_t max_aln,
	bt2gpu_dp_result *results,
	bt2gpu_dp_cand *cands, size_t cand_cap,
	bt2gpu_dp_aln *alns,
	bt2gpu_edit *edits, size_t edit_cap);

/* The same pass split for device-resident timing: stage inputs once, run the
 * kernels any number of times, fetch the results. */
int bt2gpu_dp_stage(
	bt2gpu_ctx *ctx,
	const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const uint8_t *win_pool, size_t win_len,
	const bt2gpu_dp_job *jobs, size_t njobs, uint32_t max_aln,
	size_t cand_cap, size_t edit_cap);
/* phases: bit0 fill, bit1 gather, bit2 backtrace (7 = all). */
int bt2gpu_dp_run(bt2gpu_ctx *ctx, uint32_t phases);
int bt2gpu_dp_fetch(
	bt2gpu_ctx *ctx,
	bt2gpu_dp_result *results,
	bt2gpu_dp_cand *cands, bt2gpu_dp_aln *alns, bt2gpu_edit *edits,
	size_t *ncands, size_t *nedits);
/* Device time (ms, CUDA events on the context's stream) of the last bt2gpu_dp_run():
 * ms3[0] fill, ms3[1] gather+scan+sort, ms3[2] backtrace. Single-chunk batches only. */
int bt2gpu_dp_phase_ms(bt2gpu_ctx *ctx, float *ms3);
/* DP cells (rows x cols, the bowtie2-dp CUPS numerator bt2_dp.cpp:744) of the staged batch. */
uint64_t bt2gpu_dp_staged_cells(bt2gpu_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* BT2GPU_H_ */
