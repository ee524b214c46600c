// scan.cuh — exclusive prefix sums over device arrays (three short launches), used by the glue
// kernels of the chained batch runtime (hotpath.cu) and the device-side DP staging (dp.cu).
// T needs operator+ and value-initialisation to zero.
#pragma once
#include "common.cuh"

namespace bt2 {

static constexpr int SCAN_THREADS = 256;
static constexpr int SCAN_ITEMS   = 8;
static constexpr int SCAN_TILE    = SCAN_THREADS * SCAN_ITEMS;

// inclusive scan of one value per thread across the block (Hillis-Steele in shared memory)
template<class T, int N> __device__ __forceinline__ T block_inclusive(T v, T *sm) {
	const int t = threadIdx.x;
	sm[t] = v;
	__syncthreads();
#pragma unroll
	for(int o = 1; o < N; o <<= 1) {
		T a = sm[t];
		if(t >= o) a = sm[t - o] + a;
		__syncthreads();
		sm[t] = a;
		__syncthreads();
	}
	return sm[t];
}

template<class T> __global__ void __launch_bounds__(SCAN_THREADS)
k_scan_part(const T *__restrict__ in, T *__restrict__ out, T *__restrict__ part, size_t n) {
	__shared__ T sm[SCAN_THREADS];
	const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
	T loc[SCAN_ITEMS]; T sum = T();
#pragma unroll
	for(int i = 0; i < SCAN_ITEMS; i++) {
		loc[i] = sum;
		if(base + i < n) sum = sum + in[base + i];
	}
	const T inc = block_inclusive<T, SCAN_THREADS>(sum, sm);
	const T off = threadIdx.x ? sm[threadIdx.x - 1] : T();
#pragma unroll
	for(int i = 0; i < SCAN_ITEMS; i++) if(base + i < n) out[base + i] = off + loc[i];
	if(threadIdx.x == SCAN_THREADS - 1) part[blockIdx.x] = inc;
}

// one block: exclusive scan of the per-tile totals in place, grand total to *total
template<class T> __global__ void __launch_bounds__(1024)
k_scan_top(T *__restrict__ part, size_t nparts, T *__restrict__ total) {
	__shared__ T sm[1024];
	__shared__ T carry;
	if(threadIdx.x == 0) carry = T();
	__syncthreads();
	for(size_t b = 0; b < nparts; b += 1024) {
		const size_t i = b + threadIdx.x;
		const T v = i < nparts ? part[i] : T();
		const T inc = block_inclusive<T, 1024>(v, sm);
		const T c = carry;
		if(i < nparts) part[i] = c + (threadIdx.x ? sm[threadIdx.x - 1] : T());
		__syncthreads();
		if(threadIdx.x == 1023) carry = c + inc;
		__syncthreads();
	}
	if(threadIdx.x == 0) *total = carry;
}

template<class T> __global__ void __launch_bounds__(SCAN_THREADS)
k_scan_add(T *__restrict__ out, const T *__restrict__ part, size_t n) {
	if(blockIdx.x == 0) return;                           // the first tile's offset is zero
	const T off = part[blockIdx.x];
	const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
#pragma unroll
	for(int i = 0; i < SCAN_ITEMS; i++) if(base + i < n) out[base + i] = off + out[base + i];
}

// scratch: at least scan_parts(n) elements of T; total_dev: one T.  `in` may equal `out`.
static inline size_t scan_parts(size_t n) { return (n + SCAN_TILE - 1) / SCAN_TILE; }

template<class T> static int dev_exclusive_scan(bt2gpu_ctx *ctx, const T *in, T *out, T *scratch, T *total_dev, size_t n) {
	if(n == 0) { BT2_CUDA(cudaMemsetAsync(total_dev, 0, sizeof(T), ctx->stream)); return 0; }
	const size_t np = scan_parts(n);
	k_scan_part<T><<<(unsigned)np, SCAN_THREADS, 0, ctx->stream>>>(in, out, scratch, n);
	k_scan_top<T><<<1, 1024, 0, ctx->stream>>>(scratch, np, total_dev);
	if(np > 1) k_scan_add<T><<<(unsigned)np, SCAN_THREADS, 0, ctx->stream>>>(out, scratch, n);
	ctx->launches += np > 1 ? 3 : 2;
	BT2_CUDA(cudaGetLastError());
	return 0;
}

} // namespace bt2
