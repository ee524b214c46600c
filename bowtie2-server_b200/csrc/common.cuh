// common.cuh — context, error plumbing and device buffers shared by the kernels of
// libbt2gpu.so (C ABI in include/bt2gpu.h).  sm_100a only; no CPU fallback anywhere.
#pragma once
#include <cuda_runtime.h>
#include <time.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/bt2gpu.h"

namespace bt2 {

extern thread_local std::string g_err;

inline int fail(const char *what, const char *file, int line, cudaError_t e = cudaSuccess) {
	char buf[512];
	if(e != cudaSuccess)
		snprintf(buf, sizeof(buf), "%s (%s) at %s:%d", what, cudaGetErrorString(e), file, line);
	else
		snprintf(buf, sizeof(buf), "%s at %s:%d", what, file, line);
	g_err = buf;
	return 1;
}

#define BT2_CUDA(x) do { cudaError_t e_ = (x); if(e_ != cudaSuccess) return bt2::fail(#x, __FILE__, __LINE__, e_); } while(0)
#define BT2_FAIL(msg) return bt2::fail(msg, __FILE__, __LINE__)

// Growable device buffer (never shrinks; freed with the context).
struct DevBuf {
	void *p = nullptr;
	size_t cap = 0;
	int reserve(size_t bytes) {
		if(bytes <= cap) return 0;
		if(p) cudaFree(p);
		p = nullptr; cap = 0;
		size_t want = bytes + (bytes >> 3) + 256;
		BT2_CUDA(cudaMalloc(&p, want));
		cap = want;
		return 0;
	}
	void release() { if(p) cudaFree(p); p = nullptr; cap = 0; }
	template<typename T> T *as() const { return reinterpret_cast<T*>(p); }
};

// Pinned host staging buffer.
struct HostBuf {
	void *p = nullptr;
	size_t cap = 0;
	int reserve(size_t bytes) {
		if(bytes <= cap) return 0;
		if(p) cudaFreeHost(p);
		p = nullptr; cap = 0;
		size_t want = bytes + (bytes >> 3) + 256;
		BT2_CUDA(cudaMallocHost(&p, want));
		cap = want;
		return 0;
	}
	void release() { if(p) cudaFreeHost(p); p = nullptr; cap = 0; }
	template<typename T> T *as() const { return reinterpret_cast<T*>(p); }
};

struct DpState;   // dp.cu
struct FmState;   // fm.cu
struct HpState;   // hotpath.cu

// ---- internal, device-resident plumbing of the chained batch runtime (hotpath.cu) ----
struct FmReadsDev { const uint8_t *seq, *qual; const bt2gpu_read *reads; };
struct DpDevPools { uint8_t *seq, *qual, *win; bt2gpu_dp_job *jobs; };

} // namespace bt2

struct bt2gpu_ctx {
	int device = 0;
	cudaStream_t stream = nullptr;
	cudaStream_t cstream = nullptr;       // copy stream of the batch runtime (transfers that overlap kernels)
	cudaEvent_t cev[2] = {nullptr, nullptr};
	uint64_t launches = 0;
	int sm_count = 0;
	bt2gpu_scoring sc;
	bool have_sc = false;
	bool sc_registered = false;           // holds a reference in the per-device scoring registry (api.cu)
	bt2::DpState *dp = nullptr;
	bt2::FmState *fm = nullptr;
	bt2::HpState *hp = nullptr;
	// how the batch entry points wait for the stream: 0 = cudaStreamSynchronize (the device's sync mode); > 0 = poll
	// cudaStreamQuery and sleep this many nanoseconds between polls (bt2gpu_ctx_set_wait)
	uint32_t poll_ns = 0;
	bool timing = true;                   // record the kernel-time events of every entry point (bt2gpu_ctx_set_timing)
	bool in_batch = false;                // inside a *_batch entry point: its stage step need not wait for its copies
	// device time of the kernels of the last entry point (events on `stream`)
	cudaEvent_t kev[2] = {nullptr, nullptr};
	bool kev_valid = false;
	int kstart() {
		if(!timing) { kev_valid = false; return 0; }
		for(int i = 0; i < 2; i++) if(!kev[i]) BT2_CUDA(cudaEventCreate(&kev[i]));
		kev_valid = false;
		BT2_CUDA(cudaEventRecord(kev[0], stream));
		return 0;
	}
	int kend() {
		if(!timing) return 0;
		BT2_CUDA(cudaEventRecord(kev[1], stream));
		kev_valid = true;
		return 0;
	}
};

namespace bt2 {
// Wait until everything enqueued on `s` has finished, the way the context was told to (bt2gpu_ctx_set_wait).
inline cudaError_t wait_stream(const bt2gpu_ctx *ctx, cudaStream_t s) {
	if(ctx->poll_ns == 0) return cudaStreamSynchronize(s);
	for(;;) {
		cudaError_t e = cudaStreamQuery(s);
		if(e != cudaErrorNotReady) return e;
		timespec ts = {0, (long)ctx->poll_ns};
		nanosleep(&ts, nullptr);
	}
}
// Before a device-to-host copy into pageable memory (which waits for the stream inside the driver, in the device's sync
// mode): with polling asked for, do the waiting here, so that the copy finds the stream idle.
inline cudaError_t settle(const bt2gpu_ctx *ctx, cudaStream_t s) { return ctx->poll_ns ? wait_stream(ctx, s) : cudaSuccess; }
// fm.cu: same kernels as the C-ABI entry points, device pointers in and out, no synchronisation
int fm_hp_reads_stage(bt2gpu_ctx *ctx, const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
                      const bt2gpu_read *reads, size_t nreads, FmReadsDev *view, cudaStream_t qual_stream);
int fm_hp_sweep(bt2gpu_ctx *ctx, const bt2gpu_index *fw, uint32_t mine_max, bt2gpu_sweep_result **dev_out);
int fm_hp_seed(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw, const bt2gpu_seed_job *djobs,
               size_t njobs, bt2gpu_seed_hit *dout);
int fm_hp_extend(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const bt2gpu_index *bw, const bt2gpu_extend_job *djobs,
                 size_t njobs, bt2gpu_extend_result *dout, const uint32_t *dlist, const uint32_t *dnlist);
int fm_hp_locate(bt2gpu_ctx *ctx, const bt2gpu_index *fw, const uint32_t *drows, size_t n, uint32_t *doffs, uint32_t *dsteps);
// dp.cu: DP problems whose reads, windows and job table are written by a kernel
int dp_dev_begin(bt2gpu_ctx *ctx, size_t njobs, size_t pool_len, size_t win_len, uint32_t max_aln, size_t cand_cap,
                 size_t edit_cap, DpDevPools *pools);
int dp_dev_finish(bt2gpu_ctx *ctx);
}
