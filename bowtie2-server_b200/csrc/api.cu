// api.cu — library / context entry points of include/bt2gpu.h.
#include "common.cuh"

namespace bt2 {
thread_local std::string g_err;
void dp_state_free(DpState *d);
void fm_state_free(FmState *d);
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
	dp_state_free(ctx->dp);
	fm_state_free(ctx->fm);
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
