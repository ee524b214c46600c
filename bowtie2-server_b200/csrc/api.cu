// api.cu — library / context entry points of include/bt2gpu.h.
#include "common.cuh"

namespace bt2 {
thread_local std::string g_err;
void dp_state_free(DpState *d);
void fm_state_free(FmState *d);
void hp_state_free(HpState *d);
}

#include <mutex>
namespace bt2 {
// One scoring scheme per device at a time: the kernels read it from __constant__ memory (dp.cu c_sc/c_mmpens/c_npens,
// fm.cu c_fm_sc/c_fm_mmpens), which all contexts of a device share.  Contexts that set the same scheme share the
// registration; a different scheme is refused while any other context still holds the current one.
static std::mutex g_scmu;
struct ScReg { bt2gpu_scoring sc; int refs; bool uploaded; };
static ScReg g_screg[64];

int scoring_acquire(bt2gpu_ctx *ctx, const bt2gpu_scoring *sc, bool *upload) {
	if(ctx->device < 0 || ctx->device >= 64) BT2_FAIL("device ordinal out of range");
	std::lock_guard<std::mutex> lk(g_scmu);
	ScReg &r = g_screg[ctx->device];
	const bool same = r.uploaded && memcmp(&r.sc, sc, sizeof(*sc)) == 0;
	const int others = r.refs - (ctx->sc_registered ? 1 : 0);
	if(!same && others > 0)
		BT2_FAIL("another live context of this device uses a different scoring scheme (one scheme per device at a time: "
		         "the kernels read it from device-wide constant memory)");
	if(!ctx->sc_registered) { r.refs++; ctx->sc_registered = true; }
	*upload = !same;
	if(!same) { r.sc = *sc; r.uploaded = true; }
	return 0;
}

static void scoring_release(bt2gpu_ctx *ctx) {
	std::lock_guard<std::mutex> lk(g_scmu);
	if(ctx->sc_registered && ctx->device >= 0 && ctx->device < 64) { g_screg[ctx->device].refs--; ctx->sc_registered = false; }
}
}
using namespace bt2;

extern "C" int bt2gpu_abi_version(void) { return BT2GPU_ABI_VERSION; }
extern "C" const char *bt2gpu_last_error(void) { return g_err.c_str(); }

extern "C" int bt2gpu_device_count(int *n) {
	if(!n) BT2_FAIL("null argument");
	*n = 0;
	BT2_CUDA(cudaGetDeviceCount(n));
	return 0;
}

extern "C" int bt2gpu_device_blocking_sync(int device, int on) {
	int n = 0;
	BT2_CUDA(cudaGetDeviceCount(&n));
	if(n <= 0) BT2_FAIL("no CUDA device: libbt2gpu has no CPU fallback");
	if(device < 0 || device >= n) BT2_FAIL("bad device ordinal");
	BT2_CUDA(cudaSetDevice(device));
	BT2_CUDA(cudaSetDeviceFlags(on ? cudaDeviceScheduleBlockingSync : cudaDeviceScheduleAuto));
	return 0;
}

extern "C" int bt2gpu_ctx_set_wait(bt2gpu_ctx *ctx, uint32_t poll_ns) {
	if(!ctx) BT2_FAIL("null ctx");
	if(poll_ns > 100000000u) BT2_FAIL("poll interval above 100 ms");
	ctx->poll_ns = poll_ns;
	return 0;
}

extern "C" int bt2gpu_ctx_set_timing(bt2gpu_ctx *ctx, int on) {
	if(!ctx) BT2_FAIL("null ctx");
	ctx->timing = on != 0;
	return 0;
}

extern "C" int bt2gpu_device_sync_mode(int device, int mode) {
	int n = 0;
	BT2_CUDA(cudaGetDeviceCount(&n));
	if(n <= 0) BT2_FAIL("no CUDA device: libbt2gpu has no CPU fallback");
	if(device < 0 || device >= n) BT2_FAIL("bad device ordinal");
	unsigned int f = 0;
	switch(mode) {
		case BT2GPU_SYNC_AUTO:  f = cudaDeviceScheduleAuto; break;
		case BT2GPU_SYNC_SPIN:  f = cudaDeviceScheduleSpin; break;
		case BT2GPU_SYNC_YIELD: f = cudaDeviceScheduleYield; break;
		case BT2GPU_SYNC_BLOCK: f = cudaDeviceScheduleBlockingSync; break;
		default: BT2_FAIL("bad sync mode");
	}
	BT2_CUDA(cudaSetDevice(device));
	BT2_CUDA(cudaSetDeviceFlags(f));
	return 0;
}

extern "C" int bt2gpu_ctx_create(int device, bt2gpu_ctx **out) {
	if(!out) BT2_FAIL("null argument");
	*out = nullptr;
	int n = 0;
	BT2_CUDA(cudaGetDeviceCount(&n));
	if(n <= 0) BT2_FAIL("no CUDA device: libbt2gpu has no CPU fallback");
	if(device < 0 || device >= n) BT2_FAIL("bad device ordinal");
	BT2_CUDA(cudaSetDevice(device));
	cudaDeviceProp prop;
	BT2_CUDA(cudaGetDeviceProperties(&prop, device));
	if(prop.major < 10) BT2_FAIL("libbt2gpu is built for sm_100a (Blackwell) only");
	bt2gpu_ctx *c = new bt2gpu_ctx();
	c->device = device;
	c->sm_count = prop.multiProcessorCount;
	cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
	if(e != cudaSuccess) { delete c; return fail("cudaStreamCreate", __FILE__, __LINE__, e); }
	e = cudaStreamCreateWithFlags(&c->cstream, cudaStreamNonBlocking);
	if(e != cudaSuccess) { cudaStreamDestroy(c->stream); delete c; return fail("cudaStreamCreate", __FILE__, __LINE__, e); }
	for(int i = 0; i < 2; i++) cudaEventCreateWithFlags(&c->cev[i], cudaEventDisableTiming);
	*out = c;
	return 0;
}

extern "C" void bt2gpu_ctx_destroy(bt2gpu_ctx *ctx) {
	if(!ctx) return;
	cudaSetDevice(ctx->device);
	cudaStreamSynchronize(ctx->stream);
	scoring_release(ctx);
	dp_state_free(ctx->dp);
	fm_state_free(ctx->fm);
	hp_state_free(ctx->hp);
	for(int i = 0; i < 2; i++) if(ctx->kev[i]) cudaEventDestroy(ctx->kev[i]);
	for(int i = 0; i < 2; i++) if(ctx->cev[i]) cudaEventDestroy(ctx->cev[i]);
	if(ctx->cstream) { cudaStreamSynchronize(ctx->cstream); cudaStreamDestroy(ctx->cstream); }
	cudaStreamDestroy(ctx->stream);
	delete ctx;
}

extern "C" int bt2gpu_ctx_sync(bt2gpu_ctx *ctx) {
	if(!ctx) BT2_FAIL("null ctx");
	BT2_CUDA(cudaStreamSynchronize(ctx->stream));
	return 0;
}
extern "C" void *bt2gpu_ctx_stream(bt2gpu_ctx *ctx) { return ctx ? (void*)ctx->stream : nullptr; }
extern "C" uint64_t bt2gpu_ctx_launches(bt2gpu_ctx *ctx) { return ctx ? ctx->launches : 0; }

extern "C" int bt2gpu_ctx_kernel_ms(bt2gpu_ctx *ctx, float *ms) {
	if(!ctx || !ms) BT2_FAIL("null argument");
	*ms = 0.f;
	if(!ctx->kev_valid) BT2_FAIL("no kernel has been launched through this context yet");
	BT2_CUDA(cudaEventSynchronize(ctx->kev[1]));
	BT2_CUDA(cudaEventElapsedTime(ms, ctx->kev[0], ctx->kev[1]));
	return 0;
}
