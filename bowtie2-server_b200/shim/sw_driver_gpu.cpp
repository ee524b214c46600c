// sw_driver_gpu.cpp — reference-side shim: SwDriver::extend (aligner_sw_driver.cpp:299-484), the exact left/right
// extension of a seed hit over the forward and mirror index, on the GPU through bt2gpu_extend_batch.
// prioritizeSATups (the caller, :490-738), RowSampler and everything RNG-driven stay the reference's.
#include <cstdio>
#include "aligner_sw_driver.h"
#include "gpu_shim.h"

void SwDriver::extend(
	const Read& rd, const Ebwt& ebwtFw, const Ebwt* ebwtBw, TIndexOffU topf, TIndexOffU botf, TIndexOffU topb,
	TIndexOffU botb, bool fw, size_t off, size_t len, PerReadMetrics& prm, size_t& nlex, size_t& nrex)
{
	if(ebwtBw == NULL) { fprintf(stderr, "bt2gpu shim: SwDriver::extend needs the mirror index\n"); throw 1; }
	bt2shim::ExtendCall x;
	x.fw = &ebwtFw; x.bw = ebwtBw; x.rd = &rd;
	bt2gpu_extend_job& j = x.job;
	j.read = 0; j.topf = topf; j.botf = botf; j.topb = topb; j.botb = botb;
	j.off = (uint16_t)off; j.len = (uint8_t)len; j.fw = fw ? 1 : 0;
	bt2shim::extend(x);
	const bt2gpu_extend_result& r = x.res;
	prm.nSdFmops += r.fmops;      // :362,386,427,451
	nlex = r.nlex; nrex = r.nrex;
}
