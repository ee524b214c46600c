// gpu_shim.h — shared plumbing of the reference-side shim (bowtie2-server_b200/shim/*.cpp): one bt2gpu_ctx per
// worker thread (the reference's "one SeedAligner/SwAligner per thread" rule, bt2_search.cpp:3128-3130), one
// uploaded bt2gpu_index per Ebwt for the process, the Scoring -> bt2gpu_scoring copy, and evidence counters.
#ifndef BT2GPU_SHIM_H_
#define BT2GPU_SHIM_H_

#include <stdint.h>
#include "bt2gpu.h"

class Ebwt;
class Scoring;
struct Read;

namespace bt2shim {

enum Counter { C_ALIGN = 0, C_ALN, C_UNGAPPED, C_SWEEP, C_ONEMM, C_SEEDS, C_EXTEND, C_LOCATE, C_FRAME, C_OTHERMATE, C_MATERECT, C_PECLASS, C_JOINED, C_NCOUNTERS };

bt2gpu_ctx   *ctx();                         // this thread's context (created on first use; throws 1 on failure)
void          bindScoring(const Scoring& sc);// bt2gpu_scoring_set once per (thread, Scoring object)
bt2gpu_index *index(const Ebwt& e);          // uploaded once per process, shared read-only by all contexts
void          stageRead(const Read& rd);     // Read::patFw + qual -> the context's read pool (a batch of one; cached per read)
void          fail(const char *what);        // prints bt2gpu_last_error() and throws 1 (bt2_search.cpp:5604-5611)
void          note(Counter c, uint64_t n);   // with BT2GPU_SHIM_STATS=<file>: totals rewritten in place per call

} // namespace bt2shim
#endif
