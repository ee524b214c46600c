// gpu_shim.cpp — the per-call binding of gpu_shim.h: every problem is one ABI call (a batch of one) on the calling
// worker thread's own context, as the reference calls its SeedAligner / SwAligner once per read / seed hit / window
// ("one SeedAligner/SwAligner per thread", bt2_search.cpp:3128-3130).  The batched binding is batch_shim.cpp.
#include <cstdio>
#include <cstring>
#include <vector>
#include "bt2_idx.h"
#include "scoring.h"
#include "read.h"
#include "gpu_shim.h"

namespace bt2shim {
int device();
namespace {

struct Thread {
	bt2gpu_ctx *ctx = NULL;
	bt2gpu_scoring sc; bool haveSc = false;
	uint64_t launches = 0;
	std::vector<uint8_t> seq, qual;           // the staged read
	const Read *rd = NULL; uint64_t rdid = 0; int mate = 0;
	~Thread() { if(ctx != NULL) bt2gpu_ctx_destroy(ctx); }
};
thread_local Thread g_t;

// bt2gpu_scoring_set once per (thread, scheme): compared by content, so a Scoring rebuilt at the same address with
// other penalties is uploaded again
void bindScoring(const Scoring& sc) {
	bt2gpu_ctx *c = ctx();
	bt2gpu_scoring s;
	toScoring(sc, s);
	if(g_t.haveSc && memcmp(&s, &g_t.sc, sizeof(s)) == 0) return;
	if(bt2gpu_scoring_set(c, &s)) fail("bt2gpu_scoring_set");
	g_t.sc = s; g_t.haveSc = true;
}

// Read::patFw (codes 0..4) + Read::qual (Phred+33) -> the context's read pool.  The reference hands the same Read to
// exactSweep, oneMmSearch, searchAllSeeds and every SwDriver::extend call; it is staged once.
void stageRead(const Read& rd) {
	Thread& t = g_t;
	bt2gpu_ctx *c = ctx();
	const size_t len = rd.length();
	if(t.rd == &rd && t.rdid == (uint64_t)rd.rdid && t.mate == rd.mate && t.seq.size() == len) {
		bool same = true;
		for(size_t i = 0; i < len && same; i++) same = t.seq[i] == (uint8_t)rd.patFw[i] && t.qual[i] == (uint8_t)rd.qual[i];
		if(same) return;
	}
	t.seq.resize(len); t.qual.resize(len);
	for(size_t i = 0; i < len; i++) { t.seq[i] = (uint8_t)rd.patFw[i]; t.qual[i] = (uint8_t)rd.qual[i]; }
	bt2gpu_read r; r.rd_off = 0; r.rd_len = (uint32_t)len;
	if(bt2gpu_reads_stage(c, t.seq.data(), t.qual.data(), len, &r, 1)) fail("bt2gpu_reads_stage");
	t.rd = &rd; t.rdid = (uint64_t)rd.rdid; t.mate = rd.mate;
}

void noted(Counter c, uint64_t n) { noteLaunches(g_t.ctx, g_t.launches); note(c, n); }

} // namespace

bt2gpu_ctx *ctx() {
	if(g_t.ctx == NULL && bt2gpu_ctx_create(device(), &g_t.ctx)) fail("bt2gpu_ctx_create");
	return g_t.ctx;
}

void dpAlign(DpCall& c) {
	bindScoring(*c.sc);
	const size_t cc = c.sc->monotone ? (size_t)c.job.ncol + 64 : (size_t)c.job.rd_len * c.job.ncol + 64;
	const size_t ec = (size_t)1 << 18;
	c.job.rd_off = 0; c.job.rf_off = 0;
	c.cands.resize(cc); c.alns.resize(c.max_aln); c.edits.resize(ec);
	if(bt2gpu_dp_align_batch(ctx(), c.seq, c.qual, c.job.rd_len, c.win, (size_t)c.job.ncol + 1, &c.job, 1, c.max_aln,
	                         &c.res, c.cands.data(), cc, c.alns.data(), c.edits.data(), ec))
		fail("bt2gpu_dp_align_batch");
	noted(C_ALIGN, 1);
}

void ungapped(UngapCall& c) {
	bindScoring(*c.sc);
	const size_t len = c.job.rd_len;
	c.job.rd_off = 0; c.job.rf_off = 0;
	c.edits.resize(len + 1);
	size_t ne = 0;
	if(bt2gpu_ungapped_batch(ctx(), c.seq, c.qual, len, c.win, len, &c.job, 1, &c.res, c.edits.data(), len + 1, &ne))
		fail("bt2gpu_ungapped_batch");
	noted(C_UNGAPPED, 1);
}

void sweep(SweepCall& c) {
	bindScoring(*c.sc);
	stageRead(*c.rd);
	if(bt2gpu_exact_sweep_batch(ctx(), index(*c.fw), c.nofw ? 1 : 0, c.norc ? 1 : 0, c.mineMax, &c.res))
		fail("bt2gpu_exact_sweep_batch");
	noted(C_SWEEP, 1);
}

void oneMm(OneMmCall& c) {
	bindScoring(*c.sc);
	stageRead(*c.rd);
	const size_t cap = 8 * c.rd->length() + 64;
	size_t n = 0;
	c.hits.resize(cap);
	if(bt2gpu_onemm_batch(ctx(), index(*c.fw), index(*c.bw), &c.job, c.local ? 1 : 0, c.repex ? 1 : 0, c.rep1mm ? 1 : 0,
	                      &c.res, c.hits.data(), cap, &n))
		fail("bt2gpu_onemm_batch");
	noted(C_ONEMM, 1);
}

void seeds(SeedsCall& c) {
	c.out.resize(c.jobs.size());
	if(c.jobs.empty()) return;
	bindScoring(*c.sc);
	stageRead(*c.rd);
	for(size_t i = 0; i < c.jobs.size(); i++) c.jobs[i].read = 0;
	if(bt2gpu_seed_exact_batch(ctx(), index(*c.fw), index(*c.bw), c.jobs.data(), c.jobs.size(), c.out.data()))
		fail("bt2gpu_seed_exact_batch");
	noted(C_SEEDS, c.jobs.size());
}

void seedsMm(SeedsMmCall& c) {
	c.res.assign(c.jobs.size(), bt2gpu_mmseed_result());
	c.hits.clear();
	if(c.jobs.empty()) return;
	bindScoring(*c.sc);
	stageRead(*c.rd);
	for(size_t i = 0; i < c.jobs.size(); i++) c.jobs[i].read = 0;
	size_t cap = 64 * c.jobs.size() + 256, n = 0;
	for(int attempt = 0; ; attempt++) {
		c.hits.resize(cap);
		if(bt2gpu_seed_mm_batch(ctx(), index(*c.fw), index(*c.bw), c.jobs.data(), c.jobs.size(), c.mms, c.res.data(), c.hits.data(), cap, &n) == 0) break;
		if(attempt < 6 && strstr(bt2gpu_last_error(), "seed hit pool overflow")) { cap *= 4; continue; }
		fail("bt2gpu_seed_mm_batch");
	}
	c.hits.resize(n);
	noted(C_SEEDS, c.jobs.size());
}

void extend(ExtendCall& c) {
	stageRead(*c.rd);
	c.job.read = 0;
	if(bt2gpu_extend_batch(ctx(), index(*c.fw), index(*c.bw), &c.job, 1, &c.res)) fail("bt2gpu_extend_batch");
	noted(C_EXTEND, 1);
}

void locate(LocateCall& c) {
	if(bt2gpu_locate_batch(ctx(), index(*c.fw), &c.row, 1, &c.off, &c.steps)) fail("bt2gpu_locate_batch");
	noted(C_LOCATE, 1);
}

} // namespace bt2shim
