// shim_common.cpp — what both bindings of gpu_shim.h share: the uploaded index per Ebwt, the Scoring -> bt2gpu_scoring
// copy, error plumbing and the evidence counters.  Our code, compiled against the reference's headers.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <mutex>
#include <thread>
#include <time.h>
#include <map>
#include <tuple>
#include <fcntl.h>
#include <unistd.h>
#include "bt2_idx.h"
#include "scoring.h"
#include "gpu_shim.h"

namespace bt2shim {
namespace {

std::mutex g_ixmu;
// keyed on the object AND what it holds, so that an Ebwt rebuilt at the same address (another index loaded by a
// long-lived process) is uploaded again instead of answering with the previous one's arrays
typedef std::tuple<const Ebwt*, const uint8_t*, uint32_t, uint32_t, int> IxKey;
std::map<IxKey, bt2gpu_index*> g_ix;
std::atomic<uint64_t> g_cnt[C_NCOUNTERS], g_launches(0);
std::atomic<bool> g_dirty(false);

} // namespace

int device() { const char *d = getenv("BT2GPU_DEVICE"); return d ? atoi(d) : 0; }

void fail(const char *what) {
	fprintf(stderr, "bt2gpu shim: %s: %s\n", what, bt2gpu_last_error());
	throw 1;
}

void toScoring(const Scoring& sc, bt2gpu_scoring& s) {
	memset(&s, 0, sizeof(s));
	for(int q = 0; q < 256; q++) { s.mmpens[q] = sc.mmpens[q]; s.npens[q] = sc.npens[q]; }
	s.match_bonus = (int32_t)sc.match(30);
	s.rd_gap_open = sc.readGapOpen();  s.rd_gap_ext = sc.readGapExtend();
	s.rf_gap_open = sc.refGapOpen();   s.rf_gap_ext = sc.refGapExtend();
	s.gapbar = sc.gapbar;              s.monotone = sc.monotone ? 1 : 0;
}

// Ebwt as loaded by Ebwt::readIntoMemory (bt2_io.cpp:131-614) -> bt2gpu_index_desc, field for field
bt2gpu_index *index(const Ebwt& e, int dev) {
	if(dev < 0) dev = device();
	const IxKey key(&e, e.ebwt(), (uint32_t)e.eh().len(), (uint32_t)e.zOff(), dev);
	std::lock_guard<std::mutex> lk(g_ixmu);
	auto it = g_ix.find(key);
	if(it != g_ix.end()) return it->second;
	bt2gpu_index_desc d;
	d.ebwt = e.ebwt();             d.ebwt_bytes = e.eh().ebwtTotLen();
	d.offs = e.offs();             d.n_offs = e.offs() ? e.eh().offsLen() : 0;
	d.ftab = e.ftab();             d.eftab = e.eftab();
	for(int i = 0; i < 5; i++) d.fchr[i] = e.fchr()[i];
	d.zoff = e.zOff();             d.len = e.eh().len();
	d.off_rate = e.eh().offRate(); d.ftab_chars = e.eh().ftabChars();
	d.is_fw = e.fw() ? 1 : 0;
	bt2gpu_index *h = NULL;
	if(bt2gpu_index_create(dev, &d, &h)) fail("bt2gpu_index_create");
	g_ix[key] = h;
	return h;
}

void noteLaunches(bt2gpu_ctx *c, uint64_t& seen) {
	if(c == NULL) return;
	const uint64_t l = bt2gpu_ctx_launches(c);
	if(l != seen) { g_launches += l - seen; seen = l; g_dirty.store(true, std::memory_order_relaxed); }
}

namespace {
const char *g_names[C_NCOUNTERS] = { "align_calls_on_gpu", "alignments", "ungapped_calls_on_gpu", "sweep_calls_on_gpu", "onemm_calls_on_gpu",
                                     "seeds_on_gpu", "extend_calls_on_gpu", "locate_rows_on_gpu", "frame_calls_on_gpu", "other_mate_on_gpu", "mate_rect_on_gpu", "pe_classify_on_gpu",
                                     "joined_to_text_off_on_gpu", "abi_batches", "abi_batch_jobs", "h2d_bytes", "d2h_bytes",
                                     "us_workers", "us_workers_cpu", "us_wait", "us_lanes", "us_lanes_cpu", "us_abi_stage", "us_abi_sweep", "us_abi_onemm", "us_abi_seeds", "us_abi_extend",
                                     "us_abi_locate", "us_abi_ungapped", "us_abi_align", "idle_waits", "idle_workers", "idle_workers_wait_input" };
// The totals are rewritten in place by this thread, not by the callers: note() is on the per-alignment path of every
// worker (a formatted pwrite there cost more than the alignment, and serialised the host threads on the file's lock).
void statsWriter(int fd) {
	for(;;) {
		timespec ts = {0, 20 * 1000000L};
		nanosleep(&ts, NULL);
		if(!g_dirty.exchange(false)) continue;
		char buf[4096]; int n2 = 0;
		for(int i = 0; i < C_NCOUNTERS; i++)
			n2 += snprintf(buf + n2, sizeof buf - n2, "%s %12llu ", g_names[i], (unsigned long long)g_cnt[i].load());
		n2 += snprintf(buf + n2, sizeof buf - n2, "kernel_launches %12llu\n", (unsigned long long)g_launches.load());
		if(pwrite(fd, buf, (size_t)n2, 0) < 0) { /* evidence only */ }
	}
}
int statsStart() {
	const char *path = getenv("BT2GPU_SHIM_STATS");
	const int fd = path ? open(path, O_CREAT | O_WRONLY | O_TRUNC, 0644) : -1;
	if(fd >= 0) std::thread(statsWriter, fd).detach();
	return fd;
}
} // namespace

void note(Counter c, uint64_t n) {
	static const int fd = statsStart();
	g_cnt[c].fetch_add(n, std::memory_order_relaxed);
	if(fd >= 0 && !g_dirty.load(std::memory_order_relaxed)) g_dirty.store(true, std::memory_order_relaxed);
}

} // namespace bt2shim
