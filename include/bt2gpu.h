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
 * Scoring: the kernels read the scheme from device-wide constant memory, so a device holds ONE
 * scheme at a time: contexts that set the same scheme share it; bt2gpu_scoring_set fails for a
 * different scheme while another live context of the device holds the current one.
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

#define BT2GPU_ABI_VERSION 2

typedef struct bt2gpu_ctx   bt2gpu_ctx;
typedef struct bt2gpu_index bt2gpu_index;   /* one Ebwt (forward or mirror)   */
typedef struct bt2gpu_ref   bt2gpu_ref;     /* BitPairReference (.3/.4.bt2)    */

/* ---- library / context ------------------------------------------------- */
int         bt2gpu_abi_version(void);
const char *bt2gpu_last_error(void);
int         bt2gpu_device_count(int *n);
/* Host threads waiting for this device sleep (cudaDeviceScheduleBlockingSync) instead of spinning: for callers that
 * drive more contexts than they have cores to burn (the batched server binding: one helper thread per lane). */
int         bt2gpu_device_blocking_sync(int device, int on);
/* How host threads wait for this device (no counterpart in the reference: its hot path never waits for anything):
 * BT2GPU_SYNC_AUTO the driver's choice, BT2GPU_SYNC_SPIN poll, BT2GPU_SYNC_YIELD poll and give the core away between
 * polls (a few threads that wait for short calls next to busy worker threads: the batched server's lanes; a blocked
 * waiter costs 0.15-0.2 ms per wake-up, measured: profiles/r2_abi_latency.md), BT2GPU_SYNC_BLOCK sleep. */
#define BT2GPU_SYNC_AUTO  0
#define BT2GPU_SYNC_SPIN  1
#define BT2GPU_SYNC_YIELD 2
#define BT2GPU_SYNC_BLOCK 3
int         bt2gpu_device_sync_mode(int device, int mode);
int         bt2gpu_ctx_create(int device, bt2gpu_ctx **ctx);
/* How the batch entry points of THIS context wait for its stream (no counterpart in the reference): poll_ns = 0 (the
 * default) waits in the driver, in the device's sync mode; poll_ns > 0 polls the stream and sleeps that many nanoseconds
 * between polls.  For a thread that shares its cores with busy worker threads (the batched server's lanes): a sleeper
 * that wakes is scheduled at once, a spinning or yielding waiter queues behind the workers' time slices. */
int         bt2gpu_ctx_set_wait(bt2gpu_ctx *ctx, uint32_t poll_ns);
/* on = 0: the entry points of this context stop recording the CUDA events behind bt2gpu_ctx_kernel_ms /
 * bt2gpu_dp_phase_ms (two to six driver calls per entry point; a caller that makes thousands of small calls per second
 * and never asks for the times).  Default: on. */
int         bt2gpu_ctx_set_timing(bt2gpu_ctx *ctx, int on);
void        bt2gpu_ctx_destroy(bt2gpu_ctx *ctx);
/* Block until everything enqueued on the context's stream has finished. */
int         bt2gpu_ctx_sync(bt2gpu_ctx *ctx);
/* Raw cudaStream_t of the context (for event timing by the harness). */
void       *bt2gpu_ctx_stream(bt2gpu_ctx *ctx);
/* Number of kernel launches issued through this context so far. */
uint64_t    bt2gpu_ctx_launches(bt2gpu_ctx *ctx);
/* Device time (ms) of the kernels the last compute entry point launched: CUDA events on the
 * context's stream around the kernels only (host<->device copies excluded). */
int         bt2gpu_ctx_kernel_ms(bt2gpu_ctx *ctx, float *ms);

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
/* matrix cells visited by the nextAlignment walks of the last fetched run (SSEMetrics.btcell) */
uint64_t bt2gpu_dp_walked_cells(bt2gpu_ctx *ctx);

/* ---- SwAligner::ungappedAlign (aligner_sw.cpp:286-494) --------------------
 * The single-diagonal alignment SwDriver tries when the score budget affords no gap
 * (aligner_sw_driver.cpp:1031-1060): the read laid on the reference at `off`.
 * The window is given as reference codes 0..3 (4 = N), rd_len bytes, with 4 at
 * positions off the ends of the reference (the reference pads with Ns, :342-360).
 * ret: 1 = alignment, 0 = none, -1 = more than one local solution on the diagonal
 * (the caller falls through to the DP, :447-450). */
typedef struct {
	uint32_t rd_off, rd_len;   /* oriented read (rd/qu arguments) in seq_pool/qual_pool  */
	uint32_t rf_off;           /* window in win_pool                                      */
	int32_t  minsc;            /* minsc argument                                          */
	int32_t  nceil;            /* sc.nCeil.f(len)  (:297)                                 */
	uint8_t  fw;               /* coord.fw()                                              */
	uint8_t  ohang;            /* allow overhang? (:307-322)                              */
	uint8_t  pad[2];
	int64_t  off;              /* coord.off()                                             */
	int64_t  reflen;           /* length of the reference sequence                        */
} bt2gpu_ungap_job;
typedef struct {
	int32_t  ret;
	int32_t  score, ns, refns; /* AlnScore::score_/ns_, AlnRes::refNs()                   */
	int64_t  refoff;           /* coord.off() + rowi                                      */
	uint32_t trim5, trim3;     /* soft trims w.r.t. the read's 5'/3' ends                 */
	uint32_t nedits, edit_off; /* mismatch edits, pos w.r.t. the 5' end after invertEdits */
} bt2gpu_ungap_result;
int bt2gpu_ungapped_batch(
	bt2gpu_ctx *ctx,
	const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const uint8_t *win_pool, size_t win_len,
	const bt2gpu_ungap_job *jobs, size_t njobs,
	bt2gpu_ungap_result *results, bt2gpu_edit *edits, size_t edit_cap, size_t *nedits);

/* bt2gpu_ungapped_batch without any device allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, same results as bt2gpu_ungapped_batch; the device scratch that call
 * allocates per call (staged pools, job table, results, edit pool, counter block) lives in a static device array of
 * BT2GPU_NOALLOC_SCRATCH_BYTES instead, held by every GPU that loads the library until the process exits.
 * Limit: 512 bytes + 2*pool_len + win_len + njobs*80 + edit_cap*8 (each term rounded up to 256) must not exceed
 * BT2GPU_NOALLOC_SCRATCH_BYTES, else the call fails with an error naming the limit (split the batch, or use
 * bt2gpu_ungapped_batch).  The call makes no cudaMalloc / cudaFree / allocating library call.  All calls on a device share
 * the array: it keeps the previous call's contents (what must start at zero is cleared per call), and calls that overlap
 * on the GPU -- two contexts of one device running this entry point at the same time -- RACE: serialise them. */
#define BT2GPU_NOALLOC_SCRATCH_BYTES (4u << 20)
uint64_t bt2gpu_noalloc_scratch_bytes(void);
int bt2gpu_ungapped_batch_noalloc(
	bt2gpu_ctx *ctx,
	const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const uint8_t *win_pool, size_t win_len,
	const bt2gpu_ungap_job *jobs, size_t njobs,
	bt2gpu_ungap_result *results, bt2gpu_edit *edits, size_t edit_cap, size_t *nedits);

/* ======================================================================== */
/* ---- FM index (replaces Ebwt's search-side state, bt2_idx.h:2530-2570) --- */

/* One Ebwt as the reference holds it in memory after Ebwt::readIntoMemory
 * (bt2_io.cpp:131-614), small index (TIndexOffU = uint32_t, 64-byte sides).
 * The side layout is uploaded verbatim: bytes 0..47 = 192 bases (2 bit each,
 * base k of a byte at bits 2k..2k+1), bytes 48..63 = u32 occ[A,C,G,T]. */
typedef struct {
	const uint8_t  *ebwt;        /* Ebwt::ebwt()      numSides*64 bytes             */
	uint64_t        ebwt_bytes;
	const uint32_t *offs;        /* Ebwt::offs()      SA sample, may be NULL (mirror) */
	uint64_t        n_offs;
	const uint32_t *ftab;        /* Ebwt::ftab()      4^ftab_chars + 1 entries       */
	const uint32_t *eftab;       /* Ebwt::eftab()     2*ftab_chars entries           */
	uint32_t        fchr[5];     /* Ebwt::fchr()                                     */
	uint32_t        zoff;        /* Ebwt::_zOff                                      */
	uint32_t        len;         /* EbwtParams::_len                                 */
	int32_t         off_rate;    /* EbwtParams::_offRate                             */
	int32_t         ftab_chars;  /* EbwtParams::_ftabChars                           */
	int32_t         is_fw;       /* Ebwt::fw(): 1 forward index, 0 mirror (.rev)     */
} bt2gpu_index_desc;
int  bt2gpu_index_create(int device, const bt2gpu_index_desc *desc, bt2gpu_index **out);
void bt2gpu_index_destroy(bt2gpu_index *idx);
/* Bytes of HBM the index occupies. */
uint64_t bt2gpu_index_bytes(const bt2gpu_index *idx);
/* Roofline probe: bandwidth (GB/s) of `nloads` independent random 64-byte side gathers over this
 * index at full occupancy -- the HBM random-access ceiling of the seed kernels. */
int bt2gpu_index_gather_probe(bt2gpu_ctx *ctx, const bt2gpu_index *idx, uint64_t nloads, double *gbs);
/* Roofline probes on this device (probe.cu): independent random 64-byte gathers over a scratch table of
 * `table_bytes` (choose it far larger than the 126 MB L2), and the issue rate of the packed s16x2 DPX
 * instructions the DP fill is made of, in G thread-instructions/s. */
int bt2gpu_probe_random_gather(bt2gpu_ctx *ctx, uint64_t table_bytes, uint64_t nloads, double *gbs);
int bt2gpu_probe_dpx(bt2gpu_ctx *ctx, double *ginstr_per_s);

/* A read in the pools: bytes 0..4 (A,C,G,T,N), forward strand (Read::patFw);
 * quality bytes Phred+33 (Read::qual).  The kernels derive patRc/patFwRev/... */
typedef struct { uint32_t rd_off; uint32_t rd_len; } bt2gpu_read;

/* Stage a batch of reads on the device for the seed-side calls below (replaces the
 * per-read `const Read&` argument of the SeedAligner entry points).  qual_pool may be NULL when only
 * calls that never look at qualities follow (exact sweep, exact seeds, extend, locate); bt2gpu_onemm_batch
 * needs them (N penalties). */
int bt2gpu_reads_stage(bt2gpu_ctx *ctx, const uint8_t *seq_pool, const uint8_t *qual_pool,
                       size_t pool_len, const bt2gpu_read *reads, size_t nreads);

/* BwtTopBot (aligner_cache.h:68) + the bwops_ this seed cost (aligner_seed.cpp:1906,1999).
 * topf == botf means "no hit" (then topb/botb are 0).  sides = distinct 64-byte BWT sides the
 * search had to read (1 per mapLF1 step, 1 or 2 per mapBiLFEx step, bt2_idx.h:339-351): the
 * algorithmic HBM traffic of the seed in 64-byte units. */
typedef struct { uint32_t topf, botf, topb, botb, bwops, sides; } bt2gpu_seed_hit;

/* One exact seed = one InstantiatedSeed of type SEED_TYPE_EXACT
 * (aligner_seed.cpp:213-358): window [off, off+len) of the read, strand fw. */
typedef struct { uint32_t read; uint16_t off; uint8_t len; uint8_t fw; } bt2gpu_seed_job;

/* SeedAligner::searchAllSeeds for exact seeds (aligner_seed.cpp:597-718 ->
 * startSearchSeedBi :1633 -> searchSeedBi :1854): ftab jump + mapBiLFEx/mapLF1 steps. */
int bt2gpu_seed_exact_batch(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw,
                            const bt2gpu_seed_job *jobs, size_t njobs, bt2gpu_seed_hit *out);

/* ---- seeds with mismatches: -N 1 / -N 2 -----------------------------------------------------------------------
 * Seed::oneMmSeeds / twoMmSeeds (aligner_seed.cpp:380-448) searched by searchSeedBi with its mismatch branches
 * (:1854-2033).  One job = one seed offset of one read strand, as for bt2gpu_seed_exact_batch; the device instantiates
 * the offset's seed types itself (Seed::instantiate :213-358, Ns charged to their zones) and searches them in the
 * reference's order, so that the hits of an offset come back in the order of the reference's reportHit calls (:1575),
 * i.e. the order its SeedSearchCache and the RNG-driven choices downstream see.  mms = 0 runs the exact seed through the
 * same code (one hit at most).  Needs the reads staged WITH qualities and the scoring scheme set. */
#define BT2GPU_SEED_EXACT         1   /* SEED_TYPE_EXACT (aligner_seed.h:305-308) */
#define BT2GPU_SEED_LEFT_TO_RIGHT 2
#define BT2GPU_SEED_RIGHT_TO_LEFT 3
#define BT2GPU_SEED_INSIDE_OUT    4
typedef struct {
	uint32_t topf, botf, topb, botb;  /* BW range in the forward and in the mirror index                            */
	uint64_t key;                     /* the reference string the range stands for (the seed with its mismatches    */
	                                  /* applied), 2 bits per base, first base most significant: QKey / SAKey       */
} bt2gpu_mmseed_hit;
typedef struct {
	uint32_t hit_off, nhits;          /* slice of the hit pool                                                      */
	uint32_t bwops, bwedits;          /* SeedAligner::bwops_ / bwedits_ contributions                               */
	uint32_t ntypes;                  /* seed types that instantiated (0: the Ns of the window disqualify them all) */
	uint32_t pad;
} bt2gpu_mmseed_result;
/* Fails with "seed hit pool overflow (raise hit_cap)" when an offset has more hits than hit_cap / njobs. */
int bt2gpu_seed_mm_batch(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw,
                         const bt2gpu_seed_job *jobs, size_t njobs, int mms, bt2gpu_mmseed_result *out,
                         bt2gpu_mmseed_hit *hits, size_t hit_cap, size_t *nhits);

/* SeedAligner::exactSweep (aligner_seed.cpp:854-968), both strands of every staged read.
 * nedit = the reference's nedit[fwi] when the strand stopped; done = stopped because
 * nedit reached mine_max (then mineFw/mineRc = nedit); otherwise the sweep ran to the end
 * of the read: mineFw/mineRc = nedit and, if nedit == 0 and bot > top, [top,bot) is the
 * exact end-to-end hit.  Index 0 = forward read, 1 = reverse complement. */
typedef struct { uint32_t top[2], bot[2]; uint16_t nedit[2]; uint8_t done[2]; uint32_t bwops; } bt2gpu_sweep_result;
int bt2gpu_exact_sweep_batch(bt2gpu_ctx *ctx, const bt2gpu_index *fw, int nofw, int norc,
                             uint32_t mine_max, bt2gpu_sweep_result *out);

/* SeedAligner::oneMmSearch (aligner_seed.cpp:973-1323).  Per read: minsc,
 * nceil = sc.nCeil.f(len) and the nofw/norc arguments (the driver derives them per read
 * from exactSweep's bounds, bt2_search.cpp:3649-3661).  Hits are returned in the
 * reference's order of hits.add1mmEe / addExactEe* calls, grouped per read
 * (hit_off/nhits). */
typedef struct { int32_t minsc; int32_t nceil; uint8_t nofw, norc, pad[2]; } bt2gpu_onemm_job;
typedef struct {
	uint32_t top, bot;     /* BW range in the forward index (toprep/botrep)              */
	int32_t  score;        /* EEHit::score                                               */
	uint16_t pos;          /* Edit::pos (off5p); 0xffff for an exact hit                 */
	uint8_t  chr, qchr;    /* Edit::chr = "ACGT"[j], Edit::qchr = "ACGTN"[rdc]           */
	uint8_t  fw;           /* strand                                                     */
	uint8_t  pad[3];
} bt2gpu_ee_hit;
typedef struct { uint32_t hit_off, nhits, bwops; } bt2gpu_onemm_result;
int bt2gpu_onemm_batch(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw,
                       const bt2gpu_onemm_job *jobs, int local, int repex, int rep1mm,
                       bt2gpu_onemm_result *out, bt2gpu_ee_hit *hits, size_t hit_cap, size_t *nhits);

/* SwDriver::extend (aligner_sw_driver.cpp:299-484): exact extension of a seed hit. */
typedef struct { uint32_t read; uint32_t topf, botf, topb, botb; uint16_t off; uint8_t len; uint8_t fw; } bt2gpu_extend_job;
typedef struct { uint16_t nlex, nrex; uint32_t fmops; } bt2gpu_extend_result;
int bt2gpu_extend_batch(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw,
                        const bt2gpu_extend_job *jobs, size_t njobs, bt2gpu_extend_result *out);

/* Ebwt::getOffset(row) (bt2_idx.cpp:145-171) == what GroupWalk2S resolves per element
 * (group_walk.h:1105-1216): LF-walk to the next sampled row, then offs[] gather.
 * steps[i] (optional) = LF steps walked. */
int bt2gpu_locate_batch(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const uint32_t *rows, size_t n,
                        uint32_t *offs_out, uint32_t *steps_out);

/* Device-resident variants for timing: stage jobs once, run the kernel repeatedly. */
int bt2gpu_seed_exact_stage(bt2gpu_ctx *ctx, const bt2gpu_seed_job *jobs, size_t njobs);
int bt2gpu_seed_exact_run(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw);
int bt2gpu_seed_exact_fetch(bt2gpu_ctx *ctx, bt2gpu_seed_hit *out);
int bt2gpu_locate_stage(bt2gpu_ctx *ctx, const uint32_t *rows, size_t n);
int bt2gpu_locate_run(bt2gpu_ctx *ctx, const bt2gpu_index *fw);
int bt2gpu_locate_fetch(bt2gpu_ctx *ctx, uint32_t *offs_out, uint32_t *steps_out);

/* ---- window framing and paired-end geometry (between the FM and the DP kernels) -------- */

/* DPRect (dp_framer.h:34-90). */
typedef struct {
	int64_t  refl, refr;                 /* inclusive reference columns after trimming         */
	int64_t  refl_pretrim, refr_pretrim; /* ... before trimming                                */
	uint64_t triml, trimr;               /* columns trimmed off either side                    */
	uint64_t corel, corer;               /* core diagonals, w.r.t. the untrimmed rectangle     */
	uint64_t maxgap;
} bt2gpu_rect;

/* PairedEndPolicy (pe.h:169-321) + DynProgFramer::trimToRef_ (= !gReportOverhangs). */
#define BT2GPU_PE_FF 1   /* PE_POLICY_FF (pe.h:38-54) */
#define BT2GPU_PE_RR 2
#define BT2GPU_PE_FR 3
#define BT2GPU_PE_RF 4
#define BT2GPU_PE_ALS_NORMAL   1   /* PE_ALS_* (pe.h:62-95) */
#define BT2GPU_PE_ALS_OVERLAP  2
#define BT2GPU_PE_ALS_CONTAIN  3
#define BT2GPU_PE_ALS_DOVETAIL 4
#define BT2GPU_PE_ALS_DISCORD  5
typedef struct {
	int32_t  pol;            /* BT2GPU_PE_*                                   */
	int32_t  trim_to_ref;    /* DynProgFramer(trimToRef)                      */
	uint64_t maxfrag;        /* -X (gMaxInsert)                               */
	uint64_t minfrag;        /* -I (gMinInsert)                               */
	uint8_t  local, flip_ok, dovetail_ok, contain_ok, olap_ok, expand_to_fit;
	uint8_t  pad[2];
} bt2gpu_pe_policy;

/* DynProgFramer::frameSeedExtensionRect (dp_framer.cpp:81-129), one call per job. */
typedef struct {
	int64_t  off;            /* reference offset of the read implied by the seed hit */
	int64_t  reflen;
	int64_t  maxns;
	uint32_t rdlen, maxrdgap, maxrfgap, maxhalf;
} bt2gpu_frame_job;
int bt2gpu_frame_seed_ext_batch(bt2gpu_ctx *ctx, int trim_to_ref, const bt2gpu_frame_job *jobs, size_t n,
                                bt2gpu_rect *rects, uint8_t *ok);

/* PairedEndPolicy::otherMate (pe.cpp:161-353) followed by DynProgFramer::frameFindMateRect
 * (dp_framer.h:155-197) -- the call pair of SwDriver::extendSeedsPaired
 * (aligner_sw_driver.cpp:1994-2023).  len1/len2 = mate lengths; the opposite mate's length
 * (orows) is len2 when is1, else len1. */
typedef struct {
	int64_t  off;            /* anchor alignment's reference offset (AlnRes::refoff) */
	int64_t  maxalcols;      /* orows + oreadGaps, or -1                             */
	int64_t  reflen;
	int64_t  maxns;          /* opposite mate's N ceiling                            */
	uint32_t len1, len2, maxrdgap, maxrfgap, maxhalf;
	uint8_t  is1, fw, pad[2];
} bt2gpu_mate_job;
typedef struct {
	bt2gpu_rect rect;
	int64_t  oll, olr, orl, orr;
	uint8_t  found;          /* otherMate's return value                  */
	uint8_t  ok;             /* found && frameFindMateRect's return value */
	uint8_t  oleft, ofw;
	uint8_t  pad[4];
} bt2gpu_mate_frame;
int bt2gpu_frame_mate_batch(bt2gpu_ctx *ctx, const bt2gpu_pe_policy *pol, const bt2gpu_mate_job *jobs, size_t n,
                            bt2gpu_mate_frame *out);

/* DynProgFramer::frameFindMateRect alone (dp_framer.h:155-197 -> frameFindMateAnchorLeftRect /
 * frameFindMateAnchorRightRect, dp_framer.cpp:177-357), for callers that already hold otherMate's
 * window (ll, lr, rl, rr): the per-call seam of the reference-side shim. */
typedef struct {
	int64_t  ll, lr, rl, rr;
	int64_t  reflen, maxns;
	uint32_t rdlen, maxrdgap, maxrfgap, maxhalf;
	uint8_t  anchor_left, pad[7];
} bt2gpu_mate_rect_job;
int bt2gpu_frame_mate_rect_batch(bt2gpu_ctx *ctx, int trim_to_ref, const bt2gpu_mate_rect_job *jobs, size_t n,
                                 bt2gpu_rect *rects, uint8_t *ok);

/* PairedEndPolicy::peClassifyPair (pe.cpp:37-136): cls[i] = BT2GPU_PE_ALS_*. */
typedef struct { int64_t off1, off2; uint32_t len1, len2; uint8_t fw1, fw2, pad[6]; } bt2gpu_pair_job;
int bt2gpu_pe_classify_batch(bt2gpu_ctx *ctx, const bt2gpu_pe_policy *pol, const bt2gpu_pair_job *jobs, size_t n,
                             int32_t *cls);

/* Scoring::maxReadGaps / maxRefGaps (scoring.cpp:42-105) with the context's scoring scheme. */
int bt2gpu_max_gaps_batch(bt2gpu_ctx *ctx, const int64_t *minsc, const uint32_t *rdlen, size_t n,
                          int32_t *rdgaps, int32_t *rfgaps);

/* Ebwt::joinedToTextOff (bt2_idx.cpp:54-124).  rstarts = nfrag triples (Ebwt::rstarts()), plen = text
 * lengths, len = EbwtParams::_len, fw = Ebwt::fw().  tidx = 0xffffffff for rejected straddlers. */
int bt2gpu_joined_to_text_off_batch(bt2gpu_ctx *ctx, const uint32_t *rstarts, uint32_t nfrag, const uint32_t *plen,
                                    uint32_t npat, uint32_t len, int fw, const uint32_t *qlen, const uint32_t *off,
                                    size_t n, int reject_straddle,
                                    uint32_t *tidx, uint32_t *textoff, uint32_t *tlen, uint8_t *straddled);

/* ======================================================================== */
/* ---- batch runtime: one chained pass of the hot path --------------------- */

/* What multiseedSearchWorker + SwDriver::extendSeeds do for one unpaired read in seed round 0
 * (bt2_search.cpp:3440-4060, aligner_sw_driver.cpp:756-1297), for a whole batch with every
 * arithmetic step on the GPU and the glue (seed instantiation aligner_seed.cpp:498-587, hit
 * choice, DynProgFramer::frameSeedExtensionRect dp_framer.cpp:81-129, SwAligner::initRef
 * window extraction aligner_sw.cpp:155-271) in multi-threaded C++:
 *   exactSweep -> exact seeds -> extend -> pick one hit per read -> locate -> frame -> DP.
 * The pick policy is deterministic (narrowest BW range, then longest exact extension, then
 * first in seed order); the reference's RNG-driven policy (rankSeedHits / prioritizeSATups)
 * is NOT reproduced here -- this call is the batch runtime's building block and the
 * measurement path of bench.py, not a replacement of the driver's decisions. */
typedef struct {
	int32_t  seed_len;          /* -L (multiseedLen)                                       */
	double   ival_const, ival_coeff;   /* seed interval S,<const>,<coeff> (msIval, bt2_search.cpp:3391) */
	int32_t  maxgap;            /* DPRect::maxgap = min(max(rdgaps, rfgaps), maxhalf)      */
	uint32_t max_aln;           /* nextAlignment() calls per window                        */
	uint32_t mine_max;          /* exactSweep mineMax                                      */
	int32_t  pair_boost;        /* 1: both mates of a pair are searched: interval = (int)(interval * 1.2 + 0.5) (bt2_search.cpp:3392-3395) */
	uint32_t cand_per_job;      /* candidate pool = this many cells per DP window (0: 256 end-to-end, 4096 local) */
} bt2gpu_hotpath_params;

typedef struct {
	uint64_t seeds, seed_bwops, seed_sides, seed_hits, sweep_bwops, extend_fmops,
	         locate_rows, locate_steps, dp_jobs, dp_cells, aligned, h2d_bytes, d2h_bytes, bt_cells, dp_cands;
	float    ms_sweep, ms_seed, ms_extend, ms_locate, ms_dp, ms_dp_fill, ms_dp_gather, ms_dp_backtrace;
	/* host wall clock per stage (copies, syncs and glue included): reads+sweep, seed instantiation,
	 * seed search, extend, pick+locate, framing, DP staging, DP run+fetch */
	float    wall_sweep, wall_inst, wall_seed, wall_extend, wall_locate, wall_frame, wall_dp_stage, wall_dp;
	/* device span from the first kernel to the last (reads already in HBM), and the part of it that is
	 * not one of the phases above: glue kernels, scans and the bubbles of the three scalar read-backs */
	float    ms_total, ms_glue;
} bt2gpu_hotpath_stats;

/* The joined reference text on the device, 2 bits per base: what BitPairReference holds in
 * host memory (reference.h:180, the `.4.bt2` file) and SwAligner::initRef reads windows from
 * (aligner_sw.cpp:155-271).  text = bytes 0..3, host memory, the text the index was built over;
 * uploaded once per device and shared read-only by every context on it. */
typedef struct bt2gpu_ref bt2gpu_ref;
int      bt2gpu_ref_create(int device, const uint8_t *text, uint64_t text_len, bt2gpu_ref **out);   /* one gap-free sequence */
/* The same with the `.3.bt2` records: RefRecord{off,len,first} (reference.h:59-103) = `off` ambiguous characters, then
 * `len` unambiguous ones; `first` starts a new reference sequence.  text = the joined unambiguous stretches
 * (sum of len bytes, 0..3).  From the records the device derives what BitPairReference's constructor does
 * (reference.cpp:122-167: per-sequence record ranges, lengths, offsets into the joined text) -- so a window is
 * fetched per sequence, with N in the gaps and off the ends (BitPairReference::getStretch, reference.cpp:425-600) --
 * and the fragment list Ebwt::joinedToTextOff searches (rstarts / plen, bt2_idx.cpp:54-124), so a located seed hit
 * maps to (sequence, offset, sequence length).  _bitpairs takes the text as the `.4.bt2` file holds it
 * (BitPairReference::buf_: base i at bits 2*(i&3) of byte i>>2). */
typedef struct { uint32_t off, len; uint8_t first, pad[3]; } bt2gpu_ref_record;
int      bt2gpu_ref_create_records(int device, const uint8_t *text, uint64_t text_len,
                                   const bt2gpu_ref_record *recs, uint64_t nrec, bt2gpu_ref **out);
int      bt2gpu_ref_create_bitpairs(int device, const uint8_t *bitpairs, uint64_t text_len,
                                    const bt2gpu_ref_record *recs, uint64_t nrec, bt2gpu_ref **out);
void     bt2gpu_ref_destroy(bt2gpu_ref *ref);
uint64_t bt2gpu_ref_bytes(const bt2gpu_ref *ref);

/* Per read i: minsc[i], nceil[i].  Inputs are host buffers (pinned memory makes the copies
 * asynchronous); every intermediate stays in HBM.  Outputs (all host, caller-owned):
 * sweep[nreads]; job_read[nreads] = read id of DP job k (first *njobs entries valid);
 * job_start[k] = offset, inside its reference sequence, the oriented read was framed at; results[nreads],
 * alns[nreads*max_aln], edits[edit_cap] indexed by DP job. */
int bt2gpu_hotpath_round0_batch(
	bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw, const bt2gpu_ref *ref,
	const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const bt2gpu_read *reads, size_t nreads, const int32_t *minsc, const int32_t *nceil,
	const bt2gpu_hotpath_params *params,
	bt2gpu_sweep_result *sweep, uint32_t *job_read, int64_t *job_start, uint8_t *job_fw, size_t *njobs,
	bt2gpu_dp_result *results, bt2gpu_dp_aln *alns, bt2gpu_edit *edits, size_t edit_cap, size_t *nedits,
	bt2gpu_hotpath_stats *stats);

/* Reference sequence (tidx) of every DP job of the last bt2gpu_hotpath_*_batch call on this context (njobs entries):
 * job_start[k] and every alignment's refoff are offsets inside sequence job_tidx[k]. */
int bt2gpu_hotpath_job_tidx(bt2gpu_ctx *ctx, uint32_t *job_tidx);

/* ---- batch runtime, paired-end: anchor pass + mate search ------------------------------- */

/* What SwDriver::extendSeedsPaired adds to the unpaired flow (aligner_sw_driver.cpp:1925-2240):
 * for every anchor alignment, PairedEndPolicy::otherMate -> DynProgFramer::frameFindMateRect ->
 * SwAligner::initRef/align/nextAlignment of the OPPOSITE mate in that window -> peClassifyPair.
 * reads[2i] is mate 1 and reads[2i+1] mate 2 of pair i.  Pass 1 is exactly
 * bt2gpu_hotpath_round0_batch over all 2*npairs reads (every mate is tried as an anchor; its
 * outputs have the same meaning).  Pass 2 takes the FIRST alignment of every anchor DP job, frames
 * the opposite mate's window on the device and runs the DP kernels on those windows; mate job m
 * is described by mates[m] and answered by mresults[m], malns[m*max_aln..], medits.
 * The mate is searched in the anchor's reference sequence (reflen = that sequence's length). */
typedef struct {
	uint32_t anchor_job;     /* pass-1 DP job this window was framed from                     */
	uint32_t read;           /* read id of the opposite mate (the one aligned in this window)  */
	int32_t  cls;            /* BT2GPU_PE_ALS_* of (anchor, first mate alignment); 0 if the mate did not align */
	uint8_t  ofw, oleft, pad[2];
	int64_t  anchor_refoff;  /* AlnRes::refoff() of the anchor alignment                       */
	uint32_t anchor_extent;  /* AlnRes::refExtent() of the anchor alignment                    */
	uint32_t mate_extent;    /* ... of the first mate alignment (0 if none)                    */
	bt2gpu_rect rect;
} bt2gpu_mate_info;

typedef struct {
	uint64_t mate_jobs, mate_cells, mate_aligned, concordant, bt_cells, dp_cands, d2h_bytes;
	float    ms_frame, ms_dp, ms_dp_fill, ms_dp_gather, ms_dp_backtrace, ms_total;
} bt2gpu_pairs_stats;

int bt2gpu_hotpath_pairs_batch(
	bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw, const bt2gpu_ref *ref,
	const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const bt2gpu_read *reads, size_t nreads, const int32_t *minsc, const int32_t *nceil,
	const bt2gpu_hotpath_params *params, const bt2gpu_pe_policy *pol, uint32_t maxhalf,
	/* pass 1 (anchors), as bt2gpu_hotpath_round0_batch */
	bt2gpu_sweep_result *sweep, uint32_t *job_read, int64_t *job_start, uint8_t *job_fw, size_t *njobs,
	bt2gpu_dp_result *results, bt2gpu_dp_aln *alns, bt2gpu_edit *edits, size_t edit_cap, size_t *nedits,
	bt2gpu_hotpath_stats *stats,
	/* pass 2 (mates): arrays of nreads entries (at most one mate job per anchor job) */
	bt2gpu_mate_info *mates, size_t *nmates,
	bt2gpu_dp_result *mresults, bt2gpu_dp_aln *malns, bt2gpu_edit *medits, size_t medit_cap, size_t *nmedits,
	bt2gpu_pairs_stats *pstats);

#ifdef __cplusplus
}
#endif
#endif /* BT2GPU_H_ */
