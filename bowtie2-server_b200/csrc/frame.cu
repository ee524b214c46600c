// frame.cu — C-ABI batch entry points of the framing / paired-end geometry (frame.cuh): one thread
// per job.  These are tiny integer kernels; they exist so that the rectangles the DP kernels are
// given can be produced on the device by the batch runtime (hotpath.cu) and checked on their own
// against the reference's DynProgFramer / PairedEndPolicy.
#include "common.cuh"
#include "frame.cuh"

using namespace bt2;

namespace {

__global__ void __launch_bounds__(256)
k_frame_seed_ext(const bt2gpu_frame_job *__restrict__ jobs, uint32_t n, int trim, bt2gpu_rect *__restrict__ rects,
                 uint8_t *__restrict__ ok)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n) return;
	const bt2gpu_frame_job j = jobs[i];
	bt2gpu_rect r;
	ok[i] = frame_seed_ext(j.off, j.rdlen, j.reflen, j.maxrdgap, j.maxrfgap, j.maxns, j.maxhalf, trim != 0, r) ? 1 : 0;
	rects[i] = r;
}

__global__ void __launch_bounds__(256)
k_frame_mate(const bt2gpu_mate_job *__restrict__ jobs, uint32_t n, bt2gpu_pe_policy pol, bt2gpu_mate_frame *__restrict__ out)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n) return;
	const bt2gpu_mate_job j = jobs[i];
	bt2gpu_mate_frame f; memset(&f, 0, sizeof(f));
	MateWin w;
	const bool found = other_mate(pol, j.is1 != 0, j.fw != 0, j.off, j.maxalcols, j.len1, j.len2, w);
	f.found = found; f.oleft = w.oleft; f.ofw = w.ofw;
	if(found) {
		f.oll = w.oll; f.olr = w.olr; f.orl = w.orl; f.orr = w.orr;
		const uint32_t orows = j.is1 ? j.len2 : j.len1;
		f.ok = frame_mate(!w.oleft, w.oll, w.olr, w.orl, w.orr, orows, j.reflen, j.maxrdgap, j.maxrfgap, j.maxns, j.maxhalf,
		                  pol.trim_to_ref != 0, f.rect) ? 1 : 0;
	}
	out[i] = f;
}

__global__ void __launch_bounds__(256)
k_frame_mate_rect(const bt2gpu_mate_rect_job *__restrict__ jobs, uint32_t n, int trim, bt2gpu_rect *__restrict__ rects,
                  uint8_t *__restrict__ ok)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n) return;
	const bt2gpu_mate_rect_job j = jobs[i];
	bt2gpu_rect r; memset(&r, 0, sizeof(r));
	ok[i] = frame_mate(j.anchor_left != 0, j.ll, j.lr, j.rl, j.rr, j.rdlen, j.reflen, j.maxrdgap, j.maxrfgap, j.maxns,
	                   j.maxhalf, trim != 0, r) ? 1 : 0;
	rects[i] = r;
}

__global__ void __launch_bounds__(256)
k_pe_classify(const bt2gpu_pair_job *__restrict__ jobs, uint32_t n, bt2gpu_pe_policy pol, int32_t *__restrict__ cls)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n) return;
	const bt2gpu_pair_job j = jobs[i];
	cls[i] = pe_classify(pol, j.off1, j.len1, j.fw1 != 0, j.off2, j.len2, j.fw2 != 0);
}

__global__ void __launch_bounds__(256)
k_max_gaps(const long long *__restrict__ minsc, const uint32_t *__restrict__ rdlen, uint32_t n, int match,
           int rdo, int rde, int rfo, int rfe, int32_t *__restrict__ rdg, int32_t *__restrict__ rfg)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n) return;
	rdg[i] = max_gaps(minsc[i], rdlen[i], match, rdo, rde, false);
	rfg[i] = max_gaps(minsc[i], rdlen[i], match, rfo, rfe, true);
}

__global__ void __launch_bounds__(256)
k_joined_to_text(const uint32_t *__restrict__ rstarts, uint32_t nfrag, const uint32_t *__restrict__ plen, uint32_t len, int fw,
                 const uint32_t *__restrict__ qlen, const uint32_t *__restrict__ off, uint32_t n, int reject,
                 uint32_t *__restrict__ tidx, uint32_t *__restrict__ textoff, uint32_t *__restrict__ tlen,
                 uint8_t *__restrict__ straddled)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n) return;
	uint32_t t, o, l; bool s;
	joined_to_text(rstarts, nfrag, plen, len, fw != 0, qlen[i], off[i], reject != 0, t, o, l, s);
	tidx[i] = t; textoff[i] = o; tlen[i] = l; straddled[i] = s ? 1 : 0;
}

// scratch: one device buffer, carved
struct Carve {
	uint8_t *base; size_t at = 0;
	template<typename T> T *take(size_t n) { at = (at + 255) & ~(size_t)255; T *p = reinterpret_cast<T*>(base + at); at += n * sizeof(T); return p; }
};
static size_t carve_size(std::initializer_list<size_t> parts) { size_t t = 0; for(size_t p : parts) t = ((t + 255) & ~(size_t)255) + p; return t + 256; }

int check_policy(const bt2gpu_pe_policy *pol) {
	if(!pol) BT2_FAIL("null paired-end policy");
	if(pol->pol < BT2GPU_PE_FF || pol->pol > BT2GPU_PE_RF) BT2_FAIL("bad paired-end policy");
	if(pol->maxfrag == 0 || pol->maxfrag < pol->minfrag) BT2_FAIL("bad fragment length range");
	return 0;
}

} // namespace

namespace bt2 { int check_pe_policy(const bt2gpu_pe_policy *pol) { return check_policy(pol); } }

#define FRAME_PROLOGUE(n_) \
	if(!ctx) BT2_FAIL("null context"); \
	if((n_) > 0x7fffffffull) BT2_FAIL("too many jobs"); \
	BT2_CUDA(cudaSetDevice(ctx->device))

extern "C" int bt2gpu_frame_seed_ext_batch(bt2gpu_ctx *ctx, int trim_to_ref, const bt2gpu_frame_job *jobs, size_t n,
                                           bt2gpu_rect *rects, uint8_t *ok)
{
	FRAME_PROLOGUE(n);
	if(n == 0) return 0;
	DevBuf buf;
	if(buf.reserve(carve_size({n * sizeof(*jobs), n * sizeof(*rects), n}))) return 1;
	Carve c{buf.as<uint8_t>()};
	auto *dj = c.take<bt2gpu_frame_job>(n); auto *dr = c.take<bt2gpu_rect>(n); auto *dk = c.take<uint8_t>(n);
	int rc = 0;
	do {
		if(cudaMemcpyAsync(dj, jobs, n * sizeof(*jobs), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) { rc = fail("H2D", __FILE__, __LINE__); break; }
		ctx->kstart();
		k_frame_seed_ext<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(dj, (uint32_t)n, trim_to_ref, dr, dk);
		ctx->launches++;
		ctx->kend();
		cudaMemcpyAsync(rects, dr, n * sizeof(*rects), cudaMemcpyDeviceToHost, ctx->stream);
		cudaMemcpyAsync(ok, dk, n, cudaMemcpyDeviceToHost, ctx->stream);
		cudaError_t e = cudaStreamSynchronize(ctx->stream);
		if(e != cudaSuccess) rc = fail("frame_seed_ext", __FILE__, __LINE__, e);
	} while(0);
	buf.release();
	return rc;
}

extern "C" int bt2gpu_frame_mate_batch(bt2gpu_ctx *ctx, const bt2gpu_pe_policy *pol, const bt2gpu_mate_job *jobs, size_t n,
                                       bt2gpu_mate_frame *out)
{
	FRAME_PROLOGUE(n);
	if(check_policy(pol)) return 1;
	if(n == 0) return 0;
	DevBuf buf;
	if(buf.reserve(carve_size({n * sizeof(*jobs), n * sizeof(*out)}))) return 1;
	Carve c{buf.as<uint8_t>()};
	auto *dj = c.take<bt2gpu_mate_job>(n); auto *dout = c.take<bt2gpu_mate_frame>(n);
	int rc = 0;
	do {
		if(cudaMemcpyAsync(dj, jobs, n * sizeof(*jobs), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) { rc = fail("H2D", __FILE__, __LINE__); break; }
		ctx->kstart();
		k_frame_mate<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(dj, (uint32_t)n, *pol, dout);
		ctx->launches++;
		ctx->kend();
		cudaMemcpyAsync(out, dout, n * sizeof(*out), cudaMemcpyDeviceToHost, ctx->stream);
		cudaError_t e = cudaStreamSynchronize(ctx->stream);
		if(e != cudaSuccess) rc = fail("frame_mate", __FILE__, __LINE__, e);
	} while(0);
	buf.release();
	return rc;
}

extern "C" int bt2gpu_frame_mate_rect_batch(bt2gpu_ctx *ctx, int trim_to_ref, const bt2gpu_mate_rect_job *jobs, size_t n,
                                            bt2gpu_rect *rects, uint8_t *ok)
{
	FRAME_PROLOGUE(n);
	if(n == 0) return 0;
	DevBuf buf;
	if(buf.reserve(carve_size({n * sizeof(*jobs), n * sizeof(*rects), n}))) return 1;
	Carve c{buf.as<uint8_t>()};
	auto *dj = c.take<bt2gpu_mate_rect_job>(n); auto *dr = c.take<bt2gpu_rect>(n); auto *dk = c.take<uint8_t>(n);
	int rc = 0;
	do {
		if(cudaMemcpyAsync(dj, jobs, n * sizeof(*jobs), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) { rc = fail("H2D", __FILE__, __LINE__); break; }
		ctx->kstart();
		k_frame_mate_rect<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(dj, (uint32_t)n, trim_to_ref, dr, dk);
		ctx->launches++;
		ctx->kend();
		cudaMemcpyAsync(rects, dr, n * sizeof(*rects), cudaMemcpyDeviceToHost, ctx->stream);
		cudaMemcpyAsync(ok, dk, n, cudaMemcpyDeviceToHost, ctx->stream);
		cudaError_t e = cudaStreamSynchronize(ctx->stream);
		if(e != cudaSuccess) rc = fail("frame_mate_rect", __FILE__, __LINE__, e);
	} while(0);
	buf.release();
	return rc;
}

extern "C" int bt2gpu_pe_classify_batch(bt2gpu_ctx *ctx, const bt2gpu_pe_policy *pol, const bt2gpu_pair_job *jobs, size_t n,
                                        int32_t *cls)
{
	FRAME_PROLOGUE(n);
	if(check_policy(pol)) return 1;
	if(n == 0) return 0;
	DevBuf buf;
	if(buf.reserve(carve_size({n * sizeof(*jobs), n * sizeof(*cls)}))) return 1;
	Carve c{buf.as<uint8_t>()};
	auto *dj = c.take<bt2gpu_pair_job>(n); auto *dc = c.take<int32_t>(n);
	int rc = 0;
	do {
		if(cudaMemcpyAsync(dj, jobs, n * sizeof(*jobs), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) { rc = fail("H2D", __FILE__, __LINE__); break; }
		ctx->kstart();
		k_pe_classify<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(dj, (uint32_t)n, *pol, dc);
		ctx->launches++;
		ctx->kend();
		cudaMemcpyAsync(cls, dc, n * sizeof(*cls), cudaMemcpyDeviceToHost, ctx->stream);
		cudaError_t e = cudaStreamSynchronize(ctx->stream);
		if(e != cudaSuccess) rc = fail("pe_classify", __FILE__, __LINE__, e);
	} while(0);
	buf.release();
	return rc;
}

extern "C" int bt2gpu_max_gaps_batch(bt2gpu_ctx *ctx, const int64_t *minsc, const uint32_t *rdlen, size_t n,
                                     int32_t *rdgaps, int32_t *rfgaps)
{
	FRAME_PROLOGUE(n);
	if(!ctx->have_sc) BT2_FAIL("scoring not set");
	if(n == 0) return 0;
	const bt2gpu_scoring &s = ctx->sc;
	if(s.rd_gap_ext <= 0 || s.rf_gap_ext + s.match_bonus <= 0) BT2_FAIL("gap extension penalties must be positive");
	DevBuf buf;
	if(buf.reserve(carve_size({n * 8, n * 4, n * 4, n * 4}))) return 1;
	Carve c{buf.as<uint8_t>()};
	auto *dm = c.take<long long>(n); auto *dl = c.take<uint32_t>(n); auto *d1 = c.take<int32_t>(n); auto *d2 = c.take<int32_t>(n);
	int rc = 0;
	do {
		if(cudaMemcpyAsync(dm, minsc, n * 8, cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) { rc = fail("H2D", __FILE__, __LINE__); break; }
		cudaMemcpyAsync(dl, rdlen, n * 4, cudaMemcpyHostToDevice, ctx->stream);
		ctx->kstart();
		k_max_gaps<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(dm, dl, (uint32_t)n, s.match_bonus, s.rd_gap_open,
		                                                                   s.rd_gap_ext, s.rf_gap_open, s.rf_gap_ext, d1, d2);
		ctx->launches++;
		ctx->kend();
		cudaMemcpyAsync(rdgaps, d1, n * 4, cudaMemcpyDeviceToHost, ctx->stream);
		cudaMemcpyAsync(rfgaps, d2, n * 4, cudaMemcpyDeviceToHost, ctx->stream);
		cudaError_t e = cudaStreamSynchronize(ctx->stream);
		if(e != cudaSuccess) rc = fail("max_gaps", __FILE__, __LINE__, e);
	} while(0);
	buf.release();
	return rc;
}

extern "C" int bt2gpu_joined_to_text_off_batch(bt2gpu_ctx *ctx, const uint32_t *rstarts, uint32_t nfrag, const uint32_t *plen,
                                               uint32_t npat, uint32_t len, int fw, const uint32_t *qlen, const uint32_t *off,
                                               size_t n, int reject_straddle,
                                               uint32_t *tidx, uint32_t *textoff, uint32_t *tlen, uint8_t *straddled)
{
	FRAME_PROLOGUE(n);
	if(nfrag == 0 || npat == 0) BT2_FAIL("empty fragment table");
	for(uint32_t f = 0; f < nfrag; f++) if(rstarts[f * 3 + 1] >= npat) BT2_FAIL("fragment names a text that does not exist");
	for(size_t i = 0; i < n; i++) if(off[i] >= len) BT2_FAIL("joined offset beyond the text");
	if(n == 0) return 0;
	DevBuf buf;
	if(buf.reserve(carve_size({(size_t)nfrag * 12, (size_t)npat * 4, n * 4, n * 4, n * 4, n * 4, n * 4, n}))) return 1;
	Carve c{buf.as<uint8_t>()};
	auto *drs = c.take<uint32_t>((size_t)nfrag * 3); auto *dpl = c.take<uint32_t>(npat);
	auto *dq = c.take<uint32_t>(n); auto *dof = c.take<uint32_t>(n);
	auto *dt = c.take<uint32_t>(n); auto *dx = c.take<uint32_t>(n); auto *dl = c.take<uint32_t>(n); auto *ds = c.take<uint8_t>(n);
	int rc = 0;
	do {
		if(cudaMemcpyAsync(drs, rstarts, (size_t)nfrag * 12, cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) { rc = fail("H2D", __FILE__, __LINE__); break; }
		cudaMemcpyAsync(dpl, plen, (size_t)npat * 4, cudaMemcpyHostToDevice, ctx->stream);
		cudaMemcpyAsync(dq, qlen, n * 4, cudaMemcpyHostToDevice, ctx->stream);
		cudaMemcpyAsync(dof, off, n * 4, cudaMemcpyHostToDevice, ctx->stream);
		ctx->kstart();
		k_joined_to_text<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(drs, nfrag, dpl, len, fw, dq, dof, (uint32_t)n,
		                                                                         reject_straddle, dt, dx, dl, ds);
		ctx->launches++;
		ctx->kend();
		cudaMemcpyAsync(tidx, dt, n * 4, cudaMemcpyDeviceToHost, ctx->stream);
		cudaMemcpyAsync(textoff, dx, n * 4, cudaMemcpyDeviceToHost, ctx->stream);
		cudaMemcpyAsync(tlen, dl, n * 4, cudaMemcpyDeviceToHost, ctx->stream);
		cudaMemcpyAsync(straddled, ds, n, cudaMemcpyDeviceToHost, ctx->stream);
		cudaError_t e = cudaStreamSynchronize(ctx->stream);
		if(e != cudaSuccess) rc = fail("joined_to_text_off", __FILE__, __LINE__, e);
	} while(0);
	buf.release();
	return rc;
}
