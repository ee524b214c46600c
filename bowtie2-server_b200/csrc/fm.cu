// fm.cu — FM-index kernels (placeholder until the seed path lands in this file).
#include "common.cuh"
namespace bt2 {
struct FmState { int unused; };
void fm_state_free(FmState *d) { delete d; }
}
