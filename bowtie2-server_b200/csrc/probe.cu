// probe.cu — roofline probes measured on the device the kernels run on (bench.py reports the hot
// kernels against them): independent random 64-byte gathers over a table far larger than L2 (the
// HBM random-access ceiling of the FM kernels) and the issue rate of the packed
// 16-bit DPX instructions the DP fill is built from.
#include "common.cuh"
#include <algorithm>

namespace bt2 {

// quad-cooperative 64-byte gathers, the access shape of quad_lf (fm.cu): 4 lanes x LDG.128
__global__ void __launch_bounds__(256)
k_probe_gather(const uint4 *__restrict__ table, uint32_t nsides, uint32_t per_quad, uint32_t *__restrict__ sink)
{
	const uint32_t gq = (blockIdx.x * blockDim.x + threadIdx.x) >> 2;
	const int q = threadIdx.x & 3;
	uint32_t x = gq * 2654435761u + 12345u, acc = 0;
#pragma unroll 8
	for(uint32_t i = 0; i < per_quad; i++) {
		x = x * 1664525u + 1013904223u;
		const uint32_t side = (uint32_t)(((unsigned long long)(x ^ (x >> 15)) * nsides) >> 32);
		const uint4 v = __ldg(table + (size_t)side * 4 + q);
		acc ^= v.x ^ v.y ^ v.z ^ v.w;
	}
	if(acc == 0x9e3779b9u) sink[0] = acc;
}

// 8 independent chains per thread of the two instruction shapes of the fill's cell update
// (k_dp_fill: 4 x VIADDMNMX.S16x2 + 1 x VIMNMX3.S16x2 per packed pair of cells)
__global__ void __launch_bounds__(256)
k_probe_dpx(uint32_t iters, uint32_t seed, uint32_t *__restrict__ sink)
{
	uint32_t a[8];
#pragma unroll
	for(int k = 0; k < 8; k++) a[k] = seed * (k + 1) + threadIdx.x;
	const uint32_t p = seed | 0x00010001u, f = 0x80018001u;
	for(uint32_t i = 0; i < iters; i++) {
#pragma unroll
		for(int k = 0; k < 8; k++) {
			a[k] = __viaddmax_s16x2(a[k], p, f);
			a[k] = __viaddmax_s16x2(a[k], p, a[(k + 1) & 7]);
			a[k] = __viaddmax_s16x2(a[k], f, p);
			a[k] = __viaddmax_s16x2(a[k], p, f);
			a[k] = __vimax3_s16x2(a[k], a[(k + 3) & 7], p);
		}
	}
	uint32_t r = 0;
#pragma unroll
	for(int k = 0; k < 8; k++) r ^= a[k];
	if(r == 0x12345678u) sink[0] = r;
}

} // namespace bt2

using namespace bt2;

extern "C" int bt2gpu_probe_random_gather(bt2gpu_ctx *ctx, uint64_t table_bytes, uint64_t nloads, double *gbs) {
	if(!ctx || !gbs) BT2_FAIL("null argument");
	BT2_CUDA(cudaSetDevice(ctx->device));
	if(table_bytes < (1ull << 20) || table_bytes > (64ull << 30)) BT2_FAIL("table size must be 1 MiB .. 64 GiB");
	const uint64_t nsides64 = table_bytes / 64;
	if(nsides64 > 0xffffffffull) BT2_FAIL("table too large");
	DevBuf table, sink;
	if(table.reserve(nsides64 * 64) || sink.reserve(64)) { table.release(); sink.release(); return 1; }
	int rc = 0;
	cudaEvent_t e0 = nullptr, e1 = nullptr;
	if(cudaMemsetAsync(table.p, 0x5a, nsides64 * 64, ctx->stream) != cudaSuccess || cudaEventCreate(&e0) != cudaSuccess ||
	   cudaEventCreate(&e1) != cudaSuccess) rc = fail("probe setup", __FILE__, __LINE__);
	const uint32_t blocks = (uint32_t)ctx->sm_count * 8, quads = blocks * 64;
	const uint32_t per = (uint32_t)std::max<uint64_t>(1, nloads / quads);
	for(int rep = 0; !rc && rep < 2; rep++) {                      // first pass warms up
		cudaEventRecord(e0, ctx->stream);
		k_probe_gather<<<blocks, 256, 0, ctx->stream>>>(table.as<uint4>(), (uint32_t)nsides64, per, sink.as<uint32_t>());
		cudaEventRecord(e1, ctx->stream);
		ctx->launches++;
		if(cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = fail("k_probe_gather", __FILE__, __LINE__);
	}
	float ms = 0.f;
	if(!rc) { cudaEventElapsedTime(&ms, e0, e1); *gbs = (double)quads * per * 64.0 / (ms * 1e-3) / 1e9; }
	if(e0) cudaEventDestroy(e0);
	if(e1) cudaEventDestroy(e1);
	table.release(); sink.release();
	return rc;
}

/* G thread-instructions/s of packed s16x2 DPX (VIADDMNMX / VIMNMX3) at full occupancy. */
extern "C" int bt2gpu_probe_dpx(bt2gpu_ctx *ctx, double *ginstr_per_s) {
	if(!ctx || !ginstr_per_s) BT2_FAIL("null argument");
	BT2_CUDA(cudaSetDevice(ctx->device));
	DevBuf sink;
	if(sink.reserve(64)) return 1;
	cudaEvent_t e0 = nullptr, e1 = nullptr;
	int rc = 0;
	if(cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) rc = fail("probe setup", __FILE__, __LINE__);
	const uint32_t blocks = (uint32_t)ctx->sm_count * 8, iters = 4096;
	for(int rep = 0; !rc && rep < 2; rep++) {
		cudaEventRecord(e0, ctx->stream);
		k_probe_dpx<<<blocks, 256, 0, ctx->stream>>>(iters, 0x00030005u + rep, sink.as<uint32_t>());
		cudaEventRecord(e1, ctx->stream);
		ctx->launches++;
		if(cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = fail("k_probe_dpx", __FILE__, __LINE__);
	}
	float ms = 0.f;
	if(!rc) { cudaEventElapsedTime(&ms, e0, e1); *ginstr_per_s = (double)blocks * 256.0 * iters * 40.0 / (ms * 1e-3) / 1e9; }
	if(e0) cudaEventDestroy(e0);
	if(e1) cudaEventDestroy(e1);
	sink.release();
	return rc;
}
