// gpu_shim.h — the call layer between the reference-side shim bodies (sw_aligner_gpu.cpp, seed_aligner_gpu.cpp,
// sw_driver_gpu.cpp, group_walk_gpu.h: the reference's own member functions re-bodied over the C ABI) and libbt2gpu.so.
//
// Every function below takes ONE problem of ONE read, as the reference's per-read driver produces them, and returns
// its answer.  Two bindings implement it:
//   gpu_shim.cpp    one ABI call per problem (batch of one) on the calling thread's context -- the per-call seam;
//   batch_shim.cpp  the problem is queued, the calling worker (a fiber: fiber.h) is suspended, and the problems of
//                   all the reads in flight on that host thread travel together in one bt2gpu_*_batch call; the
//                   worker resumes with its slice of the results.  Same answers, thousands of problems per call.
// Neither binding has a CPU path: without a usable CUDA device the first call fails and the server stops.
#ifndef BT2GPU_SHIM_H_
#define BT2GPU_SHIM_H_

#include <stdint.h>
#include <vector>
#include "bt2gpu.h"

class Ebwt;
class Scoring;
struct Read;

namespace bt2shim {

enum Counter { C_ALIGN = 0, C_ALN, C_UNGAPPED, C_SWEEP, C_ONEMM, C_SEEDS, C_EXTEND, C_LOCATE, C_FRAME, C_OTHERMATE, C_MATERECT, C_PECLASS, C_JOINED,
               C_BATCHES, C_BATCH_JOBS, C_H2D, C_D2H,
               // batched binding only, microseconds summed over threads: host threads running workers (wall, CPU) / waiting;
               // lanes executing batches (wall, CPU); and of that, inside the ABI calls of each kind
               C_US_WORKERS, C_US_WORKERS_CPU, C_US_WAIT, C_US_LANES, C_US_LANES_CPU, C_US_ABI_STAGE, C_US_ABI_SWEEP, C_US_ABI_ONEMM, C_US_ABI_SEEDS, C_US_ABI_EXTEND,
               C_US_ABI_LOCATE, C_US_ABI_UNGAPPED, C_US_ABI_ALIGN,
               // times a host thread found nothing to run, and how many of its workers were then waiting for input (the rest: for the GPU)
               C_IDLE_WAITS, C_IDLE_WORKERS, C_IDLE_WORKERS_INPUT, C_NCOUNTERS };

// SwAligner::initRef + align + the nextAlignment loop of one window (bt2gpu_dp_align_batch)
struct DpCall {
	const Scoring *sc;
	const uint8_t *seq, *qual, *win;       // oriented read (rd_len), qualities (rd_len), window masks (ncol + 1)
	bt2gpu_dp_job job;                     // rd_off / rf_off are the binding's business
	uint32_t max_aln;                      // bound on the nextAlignment loop; res.naln == max_aln means "maybe more"
	bt2gpu_dp_result res;                  // out; cand_off = aln_off = 0
	std::vector<bt2gpu_dp_cand> cands;     // out; res.ncand entries
	std::vector<bt2gpu_dp_aln> alns;       // out; res.naln entries, edit_off relative to `edits`
	std::vector<bt2gpu_edit> edits;        // out
};
void dpAlign(DpCall& c);

// SwAligner::ungappedAlign of one diagonal (bt2gpu_ungapped_batch)
struct UngapCall {
	const Scoring *sc;
	const uint8_t *seq, *qual, *win;       // rd_len bytes each
	bt2gpu_ungap_job job;
	bt2gpu_ungap_result res;               // out; edit_off = 0
	std::vector<bt2gpu_edit> edits;        // out
};
void ungapped(UngapCall& c);

// SeedAligner::exactSweep of one read (bt2gpu_exact_sweep_batch)
struct SweepCall { const Ebwt *fw; const Scoring *sc; const Read *rd; bool nofw, norc; uint32_t mineMax; bt2gpu_sweep_result res; };
void sweep(SweepCall& c);

// SeedAligner::oneMmSearch of one read (bt2gpu_onemm_batch)
struct OneMmCall {
	const Ebwt *fw, *bw; const Scoring *sc; const Read *rd;
	bt2gpu_onemm_job job; bool local, repex, rep1mm;
	bt2gpu_onemm_result res;               // out; hit_off = 0
	std::vector<bt2gpu_ee_hit> hits;       // out
};
void oneMm(OneMmCall& c);

// SeedAligner::searchAllSeeds: all instantiated exact seeds of one read (bt2gpu_seed_exact_batch)
struct SeedsCall {
	const Ebwt *fw, *bw; const Scoring *sc; const Read *rd;
	std::vector<bt2gpu_seed_job> jobs;     // jobs[i].read is the binding's business
	std::vector<bt2gpu_seed_hit> out;      // out; one per job
};
void seeds(SeedsCall& c);

// SeedAligner::searchAllSeeds with multi-zone seeds (-N 1 / -N 2): all seed offsets of one read (bt2gpu_seed_mm_batch)
struct SeedsMmCall {
	const Ebwt *fw, *bw; const Scoring *sc; const Read *rd; int mms;
	std::vector<bt2gpu_seed_job> jobs;          // one per (strand, offset); jobs[i].read is the binding's business
	std::vector<bt2gpu_mmseed_result> res;      // out; one per job, hit_off relative to `hits`
	std::vector<bt2gpu_mmseed_hit> hits;        // out; in the order of the reference's reportHit calls
};
void seedsMm(SeedsMmCall& c);

// SwDriver::extend of one seed hit (bt2gpu_extend_batch)
struct ExtendCall { const Ebwt *fw, *bw; const Read *rd; bt2gpu_extend_job job; bt2gpu_extend_result res; };
void extend(ExtendCall& c);

// Ebwt::getOffset of one BW row == one GroupWalk2S element (bt2gpu_locate_batch)
struct LocateCall { const Ebwt *fw; uint32_t row, off, steps; };
void locate(LocateCall& c);

// ---- shared by the bindings and the geometry shims (frame_gpu.cpp) ----
bt2gpu_ctx   *ctx();                         // the calling thread's context (created on first use; throws 1 on failure)
bt2gpu_index *index(const Ebwt& e, int dev = -1);   // uploaded once per process and device (-1: BT2GPU_DEVICE), shared read-only by its contexts
void          toScoring(const Scoring& sc, bt2gpu_scoring& out);
void          fail(const char *what);        // prints bt2gpu_last_error() and throws 1 (bt2_search.cpp:5604-5611)
void          note(Counter c, uint64_t n);   // with BT2GPU_SHIM_STATS=<file>: totals rewritten in place (by a writer thread, every 20 ms while they change)
void          noteLaunches(bt2gpu_ctx *c, uint64_t& seen);

} // namespace bt2shim
#endif
