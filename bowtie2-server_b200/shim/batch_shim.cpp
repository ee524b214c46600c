// batch_shim.cpp — the batched binding of gpu_shim.h and the worker scheduler of fiber.h.
//
// A call (dpAlign, sweep, seeds, seedsMm, extend, locate, ...) made on a worker's stack appends the problem to what the
// worker's host thread has queued and suspends the worker.  The host thread runs its runnable workers round robin and
// hands the queued problems -- at the latest every 16 workers -- to the LANE of their kind: one lane per kind of problem
// and device (exactSweep + oneMmSearch + ungapped | seeds + extend | locate | DP windows), shared by all host threads.
// A lane's thread takes whatever is queued for it, makes ONE bt2gpu_*_batch call per kind -- host buffers in, host
// buffers out -- writes every answer into its caller's own call struct and gives the workers back to their host
// threads, which make them runnable again.  A worker therefore finds its answer where the per-call binding would have
// left it: the shim bodies (sw_aligner_gpu.cpp, seed_aligner_gpu.cpp, ...) are the same code in both builds, and so is
// everything the reference's driver sees.  Batches grow with the load; the kinds run side by side on their own streams.
//
// No CPU path: a lane that cannot create its bt2gpu_ctx, or an ABI call that fails, ends the process with the error.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include <malloc.h>
#include <pthread.h>
#include <sys/mman.h>
#include <sys/prctl.h>
#include <sys/resource.h>
#include <sys/syscall.h>
#include <sched.h>
#include <time.h>
#include <unistd.h>
#include "bt2_idx.h"
#include "scoring.h"
#include "read.h"
#include "gpu_shim.h"
#include "fiber.h"

// ---- context switch (x86-64 System V): callee-saved registers on the old stack, stack pointers swapped --------------
extern "C" void bt2_fiber_switch(void **save_sp, void *load_sp);
asm(R"(
	.text
	.globl bt2_fiber_switch
	.type bt2_fiber_switch,@function
bt2_fiber_switch:
	pushq %rbp
	pushq %rbx
	pushq %r12
	pushq %r13
	pushq %r14
	pushq %r15
	movq %rsp, (%rdi)
	movq %rsi, %rsp
	popq %r15
	popq %r14
	popq %r13
	popq %r12
	popq %rbx
	popq %rbp
	ret
	.size bt2_fiber_switch,.-bt2_fiber_switch
)");

namespace bt2shim {
int device();
namespace {

struct Worker;
struct Lane;
struct Fiber {
	enum State { RUNNABLE, WAIT_BATCH, WAIT_INPUT, DONE };
	void *sp = NULL;
	void *stack = NULL; size_t stackBytes = 0;
	void (*fn)(void*) = NULL; void *arg = NULL;
	State state = RUNNABLE;
	Worker *w = NULL;
	bool queued = false;             // on its host thread's run queue
};

enum Kind { K_DP = 0, K_UNGAP, K_SWEEP, K_ONEMM, K_SEEDS, K_SEEDMM, K_EXTEND, K_LOCATE, K_NKINDS };

// Problems waiting to travel, by kind, and the workers that wait for them.
struct Batch {
	std::vector<DpCall*> dps;
	std::vector<UngapCall*> ungaps;
	std::vector<SweepCall*> sweeps;
	std::vector<OneMmCall*> onemms;
	std::vector<SeedsCall*> seeds;
	std::vector<SeedsMmCall*> seedmms;
	std::vector<ExtendCall*> extends;
	std::vector<LocateCall*> locates;
	std::vector<Fiber*> waiters;
	bool empty() const { return waiters.empty(); }
	void clear() { dps.clear(); ungaps.clear(); sweeps.clear(); onemms.clear(); seeds.clear(); seedmms.clear(); extends.clear(); locates.clear(); waiters.clear(); }
	template<typename T> static void cat(std::vector<T>& d, std::vector<T>& s) { d.insert(d.end(), s.begin(), s.end()); s.clear(); }
};

// A submission lane: one bt2gpu_ctx (stream + arenas) of one device and the thread that blocks in its ABI calls.  A
// lane serves a fixed set of problem kinds for ALL host threads that drive its device: whatever they have queued when
// the lane is free travels in one bt2gpu_*_batch call per kind, so the batch grows with the load (thousands of problems
// per call when the GPU is the slower side, a handful when the workers are), and the kinds run side by side on their
// own streams.
struct Lane {
	int dev = 0, id = 0;
	bt2gpu_ctx *ctx = NULL;
	uint64_t launches = 0;
	bt2gpu_scoring sc; bool haveSc = false;
	std::thread th;
	std::mutex mu; std::condition_variable cv;
	Batch inbox;                     // filled by the host threads (under mu)
	std::atomic<int> load{0};        // problems queued or being executed: the host threads pick the lighter lane of a kind
	Batch work;                      // the lane's own
	// host staging, reused
	std::vector<uint8_t> seq, qual, win;
	std::vector<bt2gpu_read> reads;
	std::vector<bt2gpu_dp_job> djobs; std::vector<bt2gpu_dp_result> dres; std::vector<bt2gpu_dp_cand> dcands;
	std::vector<bt2gpu_dp_aln> dalns; std::vector<bt2gpu_edit> dedits;
	std::vector<bt2gpu_ungap_job> ujobs; std::vector<bt2gpu_ungap_result> ures;
	std::vector<bt2gpu_sweep_result> swres;
	std::vector<bt2gpu_onemm_job> ojobs; std::vector<bt2gpu_onemm_result> ores; std::vector<bt2gpu_ee_hit> ohits;
	std::vector<bt2gpu_seed_job> sjobs; std::vector<bt2gpu_seed_hit> shits;
	std::vector<bt2gpu_mmseed_result> mres; std::vector<bt2gpu_mmseed_hit> mhits;
	std::vector<bt2gpu_extend_job> xjobs; std::vector<bt2gpu_extend_result> xres;
	std::vector<uint32_t> rows, offs, steps;
	void run();
	void execute(Batch& b);
};

// The lanes of one device.
struct Hub {
	int dev = 0;
	std::vector<Lane*> lanes;
	std::vector<Lane*> of[K_NKINDS];   // the lanes that serve a kind
};

struct Worker {
	int id = 0;
	Hub *hub = NULL;
	std::thread th;
	std::mutex inboxMu; std::vector<Fiber*> inbox; std::atomic<int> inboxN{0};
	std::vector<Fiber*> fibers;
	std::vector<Fiber*> runq, nextq;
	Batch pending;                   // queued by this thread's workers since the last hand-over
	std::vector<Fiber*> pw[K_NKINDS];// ... and who waits for each kind
	size_t npending = 0;
	void *schedSp = NULL;
	Fiber *cur = NULL;
	std::mutex mu; std::condition_variable cv;     // answered workers, handed back by the lanes
	std::vector<Fiber*> ready; std::atomic<int> readyN{0};
	bool inputOk = true;             // the last look for input found some
	bool inputMiss = false;          // a worker looked and found nothing in this pass
	void loop();
	void runFiber(Fiber *f);
	void publish();
};

thread_local Worker *tl_worker = NULL;
std::mutex g_mu;
std::vector<Worker*> g_workers;
std::vector<Hub*> g_hubs;
size_t g_next = 0;
std::atomic<bool> g_spawned{false};
int g_nthreads = 0, g_nlanes = 1;
std::vector<int> g_devices;        // BT2GPU_DEVICES=0,1,...: host thread i drives device g_devices[i % n] (one process, several GPUs)
size_t g_stackBytes = 1u << 20;
size_t g_batchMin = 64, g_batchWaitUs = 100;
size_t g_localCands = 8192;        // BT2GPU_LOCAL_CANDS: candidate slots first asked for per local-mode window
size_t g_publishAt = 32;           // BT2GPU_PUBLISH: a host thread hands its queued problems over at the latest after this many

// glibc's allocator, as the reference uses it, gives memory back to the kernel whenever the top of a thread's arena
// is free (madvise / mprotect) and asks for it again page by page -- with thousands of workers' buffers per host thread,
// and connection threads that build 4 * (-p) + 1 read buffers each, a profile of the batched server showed more time in
// those two system calls than in the reference's own code.  Keep freed memory, grow in large steps.  Process set-up of
// the batched binding only; BT2GPU_MALLOC_TUNING=0 leaves the allocator alone.
struct MallocTuning {
	MallocTuning() {
		const char *v = getenv("BT2GPU_MALLOC_TUNING");
		if(v != NULL && atoi(v) == 0) return;
		mallopt(M_TRIM_THRESHOLD, 1 << 30);
		mallopt(M_TOP_PAD, 64 << 20);      // a thread arena's heap (64 MB) is made accessible at once, not page by page (one mprotect, under the process's mmap lock, per page otherwise)
		mallopt(M_MMAP_THRESHOLD, 32 << 20);
	}
} g_mallocTuning;

int envInt(const char *name, int dflt, int lo, int hi) {
	const char *v = getenv(name);
	if(v == NULL || *v == 0) return dflt;
	int x = atoi(v);
	return x < lo ? lo : (x > hi ? hi : x);
}

[[noreturn]] void die(const char *what, const char *detail) {
	fprintf(stderr, "bt2gpu shim: %s: %s\n", what, detail ? detail : "");
	fflush(stderr);
	_exit(1);
}

// ---- worker entry ----------------------------------------------------------------------------------------------------
void fiberMain() {
	Worker *w = tl_worker;
	Fiber *f = w->cur;
	try {
		f->fn(f->arg);
	} catch(int e) {
		die("worker stopped by a fatal error", "see above");       // what bowtie() reports for `throw 1` (bt2_search.cpp:5604-5611)
	} catch(std::exception& e) {
		die("worker stopped by an exception", e.what());
	} catch(...) {
		die("worker stopped by an exception", "unknown");
	}
	f->state = Fiber::DONE;
	bt2_fiber_switch(&f->sp, w->schedSp);
	die("finished worker resumed", "");
}

void prepare(Fiber *f) {
	const size_t guard = 4096;
	void *m = mmap(NULL, f->stackBytes + guard, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE | MAP_STACK, -1, 0);
	if(m == MAP_FAILED) die("mmap of a worker stack failed", "");
	mprotect(m, guard, PROT_NONE);
	f->stack = m;
	uintptr_t top = ((uintptr_t)m + guard + f->stackBytes) & ~(uintptr_t)15;
	void **s = (void**)top;
	*--s = NULL;                       // fake return address of fiberMain
	*--s = (void*)&fiberMain;          // `ret` of the first switch lands here (16-byte aligned slot: ABI-conformant entry)
	for(int i = 0; i < 6; i++) *--s = NULL;   // rbp rbx r12 r13 r14 r15
	f->sp = (void*)s;
}

// suspends the calling worker; returns when the scheduler resumes it
void yieldTo(Fiber::State st) {
	Worker *w = tl_worker;
	Fiber *f = w->cur;
	f->state = st;
	bt2_fiber_switch(&f->sp, w->schedSp);
}

void Worker::runFiber(Fiber *f) {
	cur = f;
	bt2_fiber_switch(&schedSp, f->sp);
	cur = NULL;
}

void napMicros(long us) { timespec ts = {0, us * 1000}; nanosleep(&ts, NULL); }

inline uint64_t cpuUs() { timespec ts; clock_gettime(CLOCK_THREAD_CPUTIME_ID, &ts); return (uint64_t)ts.tv_sec * 1000000u + (uint64_t)ts.tv_nsec / 1000u; }
inline uint64_t nowUs() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return (uint64_t)ts.tv_sec * 1000000u + (uint64_t)ts.tv_nsec / 1000u; }
// adds the time a scope took to one of the us_* counters
struct Timed {
	Counter c; uint64_t t0;
	explicit Timed(Counter c_) : c(c_), t0(nowUs()) {}
	~Timed() { note(c, nowUs() - t0); }
};

void nameThread(const char *kind, int a, int b) {      // readable in top -H / a sampling profiler
	char nm[16];
	snprintf(nm, sizeof nm, b < 0 ? "bt2-%s-%d" : "bt2-%s-%d.%d", kind, a, b);
	pthread_setname_np(pthread_self(), nm);
}

// hands what this thread's workers have queued to the lanes of their kinds (the lighter one where a kind has several)
void Worker::publish() {
	if(npending == 0) return;
	auto to = [&](Kind k, auto Batch::*list) {
		std::vector<Fiber*>& who = pw[k];
		if(who.empty()) return;
		const std::vector<Lane*>& ls = hub->of[k];
		Lane *l = ls[0];
		for(size_t i = 1; i < ls.size(); i++) if(ls[i]->load.load(std::memory_order_relaxed) < l->load.load(std::memory_order_relaxed)) l = ls[i];
		l->load.fetch_add((int)who.size(), std::memory_order_relaxed);
		bool wasEmpty;
		{
			std::lock_guard<std::mutex> lk(l->mu);
			wasEmpty = l->inbox.empty();
			Batch::cat(l->inbox.*list, pending.*list);
			Batch::cat(l->inbox.waiters, who);
		}
		if(wasEmpty) l->cv.notify_one();
	};
	to(K_SWEEP, &Batch::sweeps); to(K_ONEMM, &Batch::onemms); to(K_SEEDS, &Batch::seeds); to(K_SEEDMM, &Batch::seedmms); to(K_EXTEND, &Batch::extends);
	to(K_LOCATE, &Batch::locates); to(K_UNGAP, &Batch::ungaps); to(K_DP, &Batch::dps);
	npending = 0;
}

void Worker::loop() {
	tl_worker = this;
	nameThread("host", id, -1);
	for(;;) {
		if(inboxN.load(std::memory_order_acquire) > 0) {
			std::lock_guard<std::mutex> lk(inboxMu);
			for(Fiber *f : inbox) { prepare(f); fibers.push_back(f); f->queued = true; runq.push_back(f); }
			inbox.clear(); inboxN.store(0);
		}
		// 1. answered workers may go on
		if(readyN.load(std::memory_order_acquire) > 0) {
			std::lock_guard<std::mutex> lk(mu);
			for(Fiber *f : ready) { f->state = Fiber::RUNNABLE; f->queued = true; runq.push_back(f); }
			ready.clear(); readyN.store(0, std::memory_order_relaxed);
		}
		// of the workers that wait for input only one looks while the last look found nothing
		{
			int looks = inputOk ? (int)fibers.size() : 1;
			for(size_t i = 0; i < fibers.size() && looks > 0; i++) {
				Fiber *f = fibers[i];
				if(f->state == Fiber::WAIT_INPUT && !f->queued) { f->queued = true; runq.push_back(f); looks--; }
			}
		}
		// 2. every runnable worker runs until its next hot-path call
		bool progress = false;
		inputMiss = false;
		const uint64_t tRun = nowUs(), cRun = cpuUs();
		for(size_t i = 0; i < runq.size(); i++) {
			Fiber *f = runq[i];
			f->queued = false;
			if(f->state == Fiber::WAIT_INPUT && inputMiss) continue;     // an earlier look of this pass found nothing
			runFiber(f);
			if(f->state != Fiber::WAIT_INPUT) progress = true;
			if(f->state == Fiber::RUNNABLE) { f->queued = true; nextq.push_back(f); }    // yieldNow()
			// 3. what they queue travels while the rest of the pass runs (at the latest every 16 workers), and workers
			//    whose answers arrive meanwhile join this pass (a bounded one: input is looked for between passes)
			if(npending >= g_publishAt || (npending > 0 && (i & 15) == 15)) publish();
			if(i < 4096 && readyN.load(std::memory_order_acquire) > 0) {
				std::lock_guard<std::mutex> lk(mu);
				for(Fiber *r : ready) { r->state = Fiber::RUNNABLE; r->queued = true; runq.push_back(r); }
				ready.clear(); readyN.store(0, std::memory_order_relaxed);
			}
		}
		runq.clear(); runq.swap(nextq);
		inputOk = !inputMiss;
		publish();
		if(progress) { note(C_US_WORKERS, nowUs() - tRun); note(C_US_WORKERS_CPU, cpuUs() - cRun); }
		// 4. nothing to do: wait for an answer, or for input
		if(!progress && runq.empty()) {
			if(!fibers.empty()) {
				size_t ni = 0;
				for(Fiber *f : fibers) if(f->state == Fiber::WAIT_INPUT) ni++;
				note(C_IDLE_WAITS, 1); note(C_IDLE_WORKERS, fibers.size()); note(C_IDLE_WORKERS_INPUT, ni);
			}
			Timed tw(C_US_WAIT);
			std::unique_lock<std::mutex> lk(mu);
			if(ready.empty()) cv.wait_for(lk, std::chrono::microseconds(fibers.empty() ? 2000 : 300));
		}
	}
}

// ---- lane: executes what the host threads have queued for its kinds ----------------------------------------------------
void Lane::run() {
	nameThread("lane", dev, id);
	// A lane runs for microseconds at a time (assemble, a few driver calls, sleep until the next poll) and every worker
	// waits for it: it should get a core the moment it wakes, ahead of the host threads' time slices.  Real-time
	// round-robin at the lowest priority where the process may ask for it (CAP_SYS_NICE), a negative nice value where it
	// may ask for that; BT2GPU_LANE_PRIORITY=0 leaves the lanes as they are.
	if(envInt("BT2GPU_LANE_PRIORITY", 1, 0, 1)) {
		sched_param sp; memset(&sp, 0, sizeof sp); sp.sched_priority = 1;
		if(pthread_setschedparam(pthread_self(), SCHED_RR, &sp) != 0) {
			if(setpriority(PRIO_PROCESS, (id_t)syscall(SYS_gettid), -10) != 0) { /* unprivileged: stay as we are */ }
		}
	}
	std::vector<std::pair<Worker*, Fiber*> > back;
	for(;;) {
		{
			std::unique_lock<std::mutex> lk(mu);
			cv.wait(lk, [this] { return !inbox.empty(); });
			// BT2GPU_BATCH_MIN (64) / BT2GPU_BATCH_WAIT_US (100): a lane that finds fewer problems than that waits once, that
			// long, for more -- fewer, fuller calls: less lane CPU per problem on a machine whose cores are all taken
			// (346 k -> 363 k reads/s measured); BT2GPU_BATCH_MIN=0 turns the wait off
			if(g_batchMin > 0 && inbox.waiters.size() < g_batchMin) {
				lk.unlock();
				napMicros((long)g_batchWaitUs);
				lk.lock();
			}
			std::swap(inbox, work);
		}
		const int n = (int)work.waiters.size();
		try {
			// The lane threads wait in the ABI calls.  They are few (one per kind), their calls are short and every
			// worker waits for them: they poll and give the core away between polls (a blocked waiter wakes 0.15-0.2 ms
			// late, several times per call).  BT2GPU_SYNC=block|spin|auto for other machines.
			static std::once_flag once;
			std::call_once(once, [] {
				const char *m = getenv("BT2GPU_SYNC");
				int mode = BT2GPU_SYNC_YIELD;
				if(m != NULL) mode = !strcmp(m, "block") ? BT2GPU_SYNC_BLOCK : !strcmp(m, "spin") ? BT2GPU_SYNC_SPIN : !strcmp(m, "auto") ? BT2GPU_SYNC_AUTO : BT2GPU_SYNC_YIELD;
				for(int d : g_devices) if(bt2gpu_device_sync_mode(d, mode)) fail("bt2gpu_device_sync_mode");
			});
			if(ctx == NULL) {
				if(bt2gpu_ctx_create(dev, &ctx)) fail("bt2gpu_ctx_create");
				// BT2GPU_POLL_US (default 20): this lane sleeps between polls of its stream instead of waiting in the
				// driver -- on a machine whose cores are all busy with workers a waiter that spins or yields gets the
				// core back a time slice later, a sleeper that wakes gets it at once; 0 = wait in the driver
				const int us = envInt("BT2GPU_POLL_US", 20, 0, 10000);
				if(us > 0) {
					prctl(PR_SET_TIMERSLACK, 1000UL, 0, 0, 0);        // this thread's sleeps end within 1 us of their time
					if(bt2gpu_ctx_set_wait(ctx, (uint32_t)us * 1000u)) fail("bt2gpu_ctx_set_wait");
				}
				if(bt2gpu_ctx_set_timing(ctx, 0)) fail("bt2gpu_ctx_set_timing");   // nobody asks a lane for kernel times
			}
			Timed tl(C_US_LANES);
			const uint64_t c0 = cpuUs();
			execute(work);
			note(C_US_LANES_CPU, cpuUs() - c0);
		} catch(int) {
			die("batched ABI call failed", "the server has no CPU path and stops");
		}
		noteLaunches(ctx, launches);
		// the answers are in the callers' own call structs: hand the workers back to their host threads
		back.clear();
		for(Fiber *f : work.waiters) back.push_back(std::make_pair(f->w, f));
		std::sort(back.begin(), back.end());
		for(size_t i = 0; i < back.size(); ) {
			Worker *w = back[i].first;
			size_t j = i;
			{
				std::lock_guard<std::mutex> lk(w->mu);
				for(; j < back.size() && back[j].first == w; j++) w->ready.push_back(back[j].second);
				w->readyN.store((int)w->ready.size(), std::memory_order_release);
			}
			w->cv.notify_one();
			i = j;
		}
		work.clear();
		load.fetch_sub(n, std::memory_order_relaxed);
	}
}

void appendRead(Lane& l, const Read& rd) {
	const size_t len = rd.length();
	bt2gpu_read r; r.rd_off = (uint32_t)l.seq.size(); r.rd_len = (uint32_t)len;
	const size_t o = l.seq.size();
	l.seq.resize(o + len); l.qual.resize(o + len);
	if(len > 0) { memcpy(&l.seq[o], rd.patFw.buf(), len); memcpy(&l.qual[o], rd.qual.buf(), len); }   // codes 0..4, Phred+33
	l.reads.push_back(r);
}

// `quals`: only oneMmSearch reads the qualities of the staged reads (N penalties); the other FM calls skip their copy
void stage(Lane& l, bool quals) {
	Timed t(C_US_ABI_STAGE);
	if(bt2gpu_reads_stage(l.ctx, l.seq.data(), quals ? l.qual.data() : NULL, l.seq.size(), l.reads.data(), l.reads.size())) fail("bt2gpu_reads_stage");
}

void bind(Lane& l, const Scoring *sc) {
	bt2gpu_scoring s;
	toScoring(*sc, s);
	if(l.haveSc && memcmp(&s, &l.sc, sizeof(s)) == 0) return;
	if(bt2gpu_scoring_set(l.ctx, &s)) fail("bt2gpu_scoring_set");
	l.sc = s; l.haveSc = true;
}

// evidence: problems per ABI call, and the bytes of the host buffers that went in and came back
void noteBatch(Counter c, size_t n, size_t in_bytes, size_t out_bytes) { note(c, n); note(C_BATCHES, 1); note(C_BATCH_JOBS, n); note(C_H2D, in_bytes); note(C_D2H, out_bytes); }

void Lane::execute(Batch& b) {
	// ---- exactSweep: reads that share (nofw, norc, mineMax) travel together ----
	{
		std::vector<char> taken(b.sweeps.size(), 0);
		for(size_t i0 = 0; i0 < b.sweeps.size(); i0++) {
			if(taken[i0]) continue;
			const SweepCall& k = *b.sweeps[i0];
			std::vector<size_t> grp;
			for(size_t i = i0; i < b.sweeps.size(); i++) {
				const SweepCall& c = *b.sweeps[i];
				if(!taken[i] && c.fw == k.fw && c.nofw == k.nofw && c.norc == k.norc && c.mineMax == k.mineMax) { taken[i] = 1; grp.push_back(i); }
			}
			bind(*this, k.sc);
			seq.clear(); qual.clear(); reads.clear();
			for(size_t i : grp) appendRead(*this, *b.sweeps[i]->rd);
			stage(*this, false);
			swres.resize(grp.size());
			const uint64_t t_sweep = nowUs();
			if(bt2gpu_exact_sweep_batch(ctx, index(*k.fw, dev), k.nofw ? 1 : 0, k.norc ? 1 : 0, k.mineMax, swres.data())) fail("bt2gpu_exact_sweep_batch");
			note(C_US_ABI_SWEEP, nowUs() - t_sweep);
			for(size_t j = 0; j < grp.size(); j++) b.sweeps[grp[j]]->res = swres[j];
			noteBatch(C_SWEEP, grp.size(), 2 * seq.size() + reads.size() * sizeof(bt2gpu_read), grp.size() * sizeof(bt2gpu_sweep_result));
		}
	}
	// ---- oneMmSearch: by (local, repex, rep1mm) ----
	{
		std::vector<char> taken(b.onemms.size(), 0);
		for(size_t i0 = 0; i0 < b.onemms.size(); i0++) {
			if(taken[i0]) continue;
			const OneMmCall& k = *b.onemms[i0];
			std::vector<size_t> grp;
			for(size_t i = i0; i < b.onemms.size(); i++) {
				const OneMmCall& c = *b.onemms[i];
				if(!taken[i] && c.fw == k.fw && c.bw == k.bw && c.local == k.local && c.repex == k.repex && c.rep1mm == k.rep1mm) { taken[i] = 1; grp.push_back(i); }
			}
			bind(*this, k.sc);
			seq.clear(); qual.clear(); reads.clear(); ojobs.clear();
			size_t cap = 0;
			for(size_t i : grp) { appendRead(*this, *b.onemms[i]->rd); ojobs.push_back(b.onemms[i]->job); cap += 8 * b.onemms[i]->rd->length() + 64; }
			stage(*this, true);
			ores.resize(grp.size()); ohits.resize(cap);
			size_t n = 0;
			const uint64_t t_onemm = nowUs();
			if(bt2gpu_onemm_batch(ctx, index(*k.fw, dev), index(*k.bw, dev), ojobs.data(), k.local ? 1 : 0, k.repex ? 1 : 0, k.rep1mm ? 1 : 0,
			                      ores.data(), ohits.data(), cap, &n)) fail("bt2gpu_onemm_batch");
			note(C_US_ABI_ONEMM, nowUs() - t_onemm);
			for(size_t j = 0; j < grp.size(); j++) {
				OneMmCall& c = *b.onemms[grp[j]];
				c.res = ores[j];
				c.hits.assign(ohits.begin() + ores[j].hit_off, ohits.begin() + ores[j].hit_off + ores[j].nhits);
				c.res.hit_off = 0;
			}
			noteBatch(C_ONEMM, grp.size(), 2 * seq.size() + reads.size() * sizeof(bt2gpu_read) + ojobs.size() * sizeof(bt2gpu_onemm_job), grp.size() * sizeof(bt2gpu_onemm_result) + n * sizeof(bt2gpu_ee_hit));
		}
	}
	// ---- exact seeds and seed-hit extensions share one staged read pool ----
	if(!b.seeds.empty() || !b.extends.empty()) {
		const Ebwt *fw = !b.seeds.empty() ? b.seeds[0]->fw : b.extends[0]->fw, *bw = !b.seeds.empty() ? b.seeds[0]->bw : b.extends[0]->bw;
		if(!b.seeds.empty()) bind(*this, b.seeds[0]->sc);
		seq.clear(); qual.clear(); reads.clear(); sjobs.clear(); xjobs.clear();
		for(SeedsCall *c : b.seeds) {
			if(c->fw != fw || c->bw != bw) { fprintf(stderr, "bt2gpu shim: workers of one server use different indexes\n"); throw 1; }
			const uint32_t ri = (uint32_t)reads.size();
			appendRead(*this, *c->rd);
			for(bt2gpu_seed_job j : c->jobs) { j.read = ri; sjobs.push_back(j); }
		}
		for(ExtendCall *c : b.extends) {
			if(c->fw != fw || c->bw != bw) { fprintf(stderr, "bt2gpu shim: workers of one server use different indexes\n"); throw 1; }
			bt2gpu_extend_job j = c->job;
			j.read = (uint32_t)reads.size();
			appendRead(*this, *c->rd);
			xjobs.push_back(j);
		}
		stage(*this, false);
		if(!sjobs.empty()) {
			shits.resize(sjobs.size());
			const uint64_t t_seeds = nowUs();
			if(bt2gpu_seed_exact_batch(ctx, index(*fw, dev), index(*bw, dev), sjobs.data(), sjobs.size(), shits.data())) fail("bt2gpu_seed_exact_batch");
			note(C_US_ABI_SEEDS, nowUs() - t_seeds);
			size_t o = 0;
			for(SeedsCall *c : b.seeds) { c->out.assign(shits.begin() + o, shits.begin() + o + c->jobs.size()); o += c->jobs.size(); }
			noteBatch(C_SEEDS, sjobs.size(), 2 * seq.size() + reads.size() * sizeof(bt2gpu_read) + sjobs.size() * sizeof(bt2gpu_seed_job), sjobs.size() * sizeof(bt2gpu_seed_hit));
		} else for(SeedsCall *c : b.seeds) c->out.clear();
		if(!xjobs.empty()) {
			xres.resize(xjobs.size());
			const uint64_t t_extend = nowUs();
			if(bt2gpu_extend_batch(ctx, index(*fw, dev), index(*bw, dev), xjobs.data(), xjobs.size(), xres.data())) fail("bt2gpu_extend_batch");
			note(C_US_ABI_EXTEND, nowUs() - t_extend);
			for(size_t j = 0; j < xjobs.size(); j++) b.extends[j]->res = xres[j];
			noteBatch(C_EXTEND, xjobs.size(), xjobs.size() * sizeof(bt2gpu_extend_job), xjobs.size() * sizeof(bt2gpu_extend_result));
		}
	}
	// ---- multi-zone seeds (-N 1 / -N 2): by mismatch count; reads staged with their qualities (N penalties) ----
	{
		std::vector<char> taken(b.seedmms.size(), 0);
		for(size_t i0 = 0; i0 < b.seedmms.size(); i0++) {
			if(taken[i0]) continue;
			const SeedsMmCall& k = *b.seedmms[i0];
			std::vector<size_t> grp;
			for(size_t i = i0; i < b.seedmms.size(); i++) {
				const SeedsMmCall& c = *b.seedmms[i];
				if(!taken[i] && c.fw == k.fw && c.bw == k.bw && c.mms == k.mms) { taken[i] = 1; grp.push_back(i); }
			}
			bind(*this, k.sc);
			seq.clear(); qual.clear(); reads.clear(); sjobs.clear();
			for(size_t i : grp) {
				SeedsMmCall& c = *b.seedmms[i];
				const uint32_t ri = (uint32_t)reads.size();
				appendRead(*this, *c.rd);
				for(bt2gpu_seed_job j : c.jobs) { j.read = ri; sjobs.push_back(j); }
			}
			stage(*this, true);
			mres.resize(sjobs.size());
			size_t cap = 64 * sjobs.size() + 256, n = 0;
			const uint64_t t_mm = nowUs();
			for(int attempt = 0; ; attempt++) {
				mhits.resize(cap);
				if(bt2gpu_seed_mm_batch(ctx, index(*k.fw, dev), index(*k.bw, dev), sjobs.data(), sjobs.size(), k.mms, mres.data(), mhits.data(), cap, &n) == 0) break;
				if(attempt < 6 && strstr(bt2gpu_last_error(), "seed hit pool overflow")) { cap *= 4; continue; }
				fail("bt2gpu_seed_mm_batch");
			}
			note(C_US_ABI_SEEDS, nowUs() - t_mm);
			size_t o = 0;
			for(size_t i : grp) {
				SeedsMmCall& c = *b.seedmms[i];
				c.res.assign(mres.begin() + o, mres.begin() + o + c.jobs.size());
				c.hits.clear();
				for(bt2gpu_mmseed_result& r : c.res) {
					const uint32_t ho = (uint32_t)c.hits.size();
					c.hits.insert(c.hits.end(), mhits.begin() + r.hit_off, mhits.begin() + r.hit_off + r.nhits);
					r.hit_off = ho;
				}
				o += c.jobs.size();
			}
			noteBatch(C_SEEDS, sjobs.size(), 2 * seq.size() + reads.size() * sizeof(bt2gpu_read) + sjobs.size() * sizeof(bt2gpu_seed_job),
			          sjobs.size() * sizeof(bt2gpu_mmseed_result) + n * sizeof(bt2gpu_mmseed_hit));
		}
	}
	// ---- BW rows to text offsets ----
	if(!b.locates.empty()) {
		const size_t n = b.locates.size();
		rows.resize(n); offs.resize(n); steps.resize(n);
		for(size_t j = 0; j < n; j++) rows[j] = b.locates[j]->row;
		const uint64_t t_locate = nowUs();
		if(bt2gpu_locate_batch(ctx, index(*b.locates[0]->fw, dev), rows.data(), n, offs.data(), steps.data())) fail("bt2gpu_locate_batch");
		note(C_US_ABI_LOCATE, nowUs() - t_locate);
		for(size_t j = 0; j < n; j++) { b.locates[j]->off = offs[j]; b.locates[j]->steps = steps[j]; }
		noteBatch(C_LOCATE, n, n * 4, n * 8);
	}
	// ---- ungapped diagonals ----
	if(!b.ungaps.empty()) {
		const size_t n = b.ungaps.size();
		bind(*this, b.ungaps[0]->sc);
		seq.clear(); qual.clear(); win.clear(); ujobs.clear();
		size_t cap = 0;
		for(UngapCall *c : b.ungaps) {
			bt2gpu_ungap_job j = c->job;
			j.rd_off = (uint32_t)seq.size(); j.rf_off = (uint32_t)win.size();
			seq.insert(seq.end(), c->seq, c->seq + j.rd_len); qual.insert(qual.end(), c->qual, c->qual + j.rd_len);
			win.insert(win.end(), c->win, c->win + j.rd_len);
			ujobs.push_back(j);
			cap += j.rd_len + 1;
		}
		ures.resize(n); dedits.resize(cap);
		size_t ne = 0;
		const uint64_t t_ungapped = nowUs();
		if(bt2gpu_ungapped_batch(ctx, seq.data(), qual.data(), seq.size(), win.data(), win.size(), ujobs.data(), n, ures.data(),
		                         dedits.data(), cap, &ne)) fail("bt2gpu_ungapped_batch");
		note(C_US_ABI_UNGAPPED, nowUs() - t_ungapped);
		for(size_t j = 0; j < n; j++) {
			UngapCall& c = *b.ungaps[j];
			c.res = ures[j];
			c.edits.assign(dedits.begin() + ures[j].edit_off, dedits.begin() + ures[j].edit_off + (ures[j].ret == 1 ? ures[j].nedits : 0));
			c.res.edit_off = 0;
		}
		noteBatch(C_UNGAPPED, n, 2 * seq.size() + win.size() + n * sizeof(bt2gpu_ungap_job), n * sizeof(bt2gpu_ungap_result) + ne * sizeof(bt2gpu_edit));
	}
	// ---- DP windows: by the bound on the nextAlignment loop (the first ask is short, the rare re-ask is complete) ----
	{
		std::vector<char> taken(b.dps.size(), 0);
		for(size_t i0 = 0; i0 < b.dps.size(); i0++) {
			if(taken[i0]) continue;
			const DpCall& k = *b.dps[i0];
			std::vector<size_t> grp;
			for(size_t i = i0; i < b.dps.size(); i++) if(!taken[i] && b.dps[i]->max_aln == k.max_aln && b.dps[i]->sc == k.sc) { taken[i] = 1; grp.push_back(i); }
			bind(*this, k.sc);
			const size_t n = grp.size();
			seq.clear(); qual.clear(); win.clear(); djobs.clear();
			size_t ccap = 0, cfull = 0, ecap = 1024;
			for(size_t i : grp) {
				DpCall& c = *b.dps[i];
				c.job.rd_off = (uint32_t)seq.size(); c.job.rf_off = (uint32_t)win.size();
				seq.insert(seq.end(), c.seq, c.seq + c.job.rd_len); qual.insert(qual.end(), c.qual, c.qual + c.job.rd_len);
				win.insert(win.end(), c.win, c.win + c.job.ncol + 1);
				djobs.push_back(c.job);
				// candidates: end-to-end, at most one per column; local, any cell can be one (rd_len x ncol) but a window has
				// a few thousand -- ask for BT2GPU_LOCAL_CANDS (8192) per window and come back with more if the pool overflows
				const size_t cells = (size_t)c.job.rd_len * c.job.ncol;
				ccap += k.sc->monotone ? (size_t)c.job.ncol + 64 : (cells < g_localCands ? cells : g_localCands) + 64;
				cfull += k.sc->monotone ? (size_t)c.job.ncol + 64 : cells + 64;
				ecap += (size_t)(c.job.rd_len + 64) * (k.max_aln < 4 ? k.max_aln : 4);
			}
			dres.resize(n); dalns.resize(n * (size_t)k.max_aln);
			for(int attempt = 0; ; attempt++) {
				dcands.resize(ccap); dedits.resize(ecap);
				const uint64_t t_align = nowUs();
				const int rc = bt2gpu_dp_align_batch(ctx, seq.data(), qual.data(), seq.size(), win.data(), win.size(), djobs.data(), n, k.max_aln,
				                                     dres.data(), dcands.data(), ccap, dalns.data(), dedits.data(), ecap);
				note(C_US_ABI_ALIGN, nowUs() - t_align);
				if(rc == 0) break;
				const char *e = bt2gpu_last_error();
				if(attempt < 6 && strstr(e, "edit pool overflow")) { ecap *= 4; continue; }
				if(ccap < cfull && strstr(e, "candidate pool overflow")) { ccap = ccap * 4 < cfull ? ccap * 4 : cfull; continue; }
				fail("bt2gpu_dp_align_batch");
			}
			for(size_t j = 0; j < n; j++) {
				DpCall& c = *b.dps[grp[j]];
				const bt2gpu_dp_result& r = dres[j];
				c.res = r;
				c.cands.assign(dcands.begin() + r.cand_off, dcands.begin() + r.cand_off + r.ncand);
				c.alns.assign(dalns.begin() + r.aln_off, dalns.begin() + r.aln_off + r.naln);
				c.edits.clear();
				for(bt2gpu_dp_aln& a : c.alns) {
					const uint32_t o = (uint32_t)c.edits.size();
					c.edits.insert(c.edits.end(), dedits.begin() + a.edit_off, dedits.begin() + a.edit_off + a.nedits);
					a.edit_off = o;
				}
				c.res.cand_off = 0; c.res.aln_off = 0;
			}
			size_t nc = 0, ne = 0;
			for(size_t j = 0; j < n; j++) { nc += dres[j].ncand; ne += b.dps[grp[j]]->edits.size(); }
			noteBatch(C_ALIGN, n, 2 * seq.size() + win.size() + n * sizeof(bt2gpu_dp_job),
			          n * (sizeof(bt2gpu_dp_result) + k.max_aln * sizeof(bt2gpu_dp_aln)) + nc * sizeof(bt2gpu_dp_cand) + ne * sizeof(bt2gpu_edit));
		}
	}
}

// ---- the binding: queue, suspend, resume with the answer ---------------------------------------------------------------
template<typename T> void travel(Kind k, std::vector<T*> Batch::*list, T& c) {
	Worker *w = tl_worker;
	if(w == NULL || w->cur == NULL) { fprintf(stderr, "bt2gpu shim: hot-path call outside a worker (batched binding)\n"); throw 1; }
	(w->pending.*list).push_back(&c);
	w->pw[k].push_back(w->cur);
	w->npending++;
	yieldTo(Fiber::WAIT_BATCH);
}

} // namespace

bt2gpu_ctx *ctx() { fprintf(stderr, "bt2gpu shim: per-thread context requested in the batched binding\n"); throw 1; }

void dpAlign(DpCall& c)   { travel(K_DP, &Batch::dps, c); }
void ungapped(UngapCall& c) { travel(K_UNGAP, &Batch::ungaps, c); }
void sweep(SweepCall& c)  { travel(K_SWEEP, &Batch::sweeps, c); }
void oneMm(OneMmCall& c)  { travel(K_ONEMM, &Batch::onemms, c); }
void seeds(SeedsCall& c)  { c.out.clear(); if(c.jobs.empty()) return; travel(K_SEEDS, &Batch::seeds, c); }
void seedsMm(SeedsMmCall& c) { c.res.clear(); c.hits.clear(); if(c.jobs.empty()) return; travel(K_SEEDMM, &Batch::seedmms, c); }
void extend(ExtendCall& c) { travel(K_EXTEND, &Batch::extends, c); }
void locate(LocateCall& c) { travel(K_LOCATE, &Batch::locates, c); }

// ---- fiber.h ------------------------------------------------------------------------------------------------------
bool inWorker() { return tl_worker != NULL && tl_worker->cur != NULL; }
void *currentWorkerArg() { return inWorker() ? tl_worker->cur->arg : NULL; }

void waitInput() {
	Worker *w = tl_worker;
	if(w == NULL || w->cur == NULL) { napMicros(200); return; }
	w->inputMiss = true;
	yieldTo(Fiber::WAIT_INPUT);
}

void yieldNow() {
	if(tl_worker == NULL || tl_worker->cur == NULL) { napMicros(50); return; }
	yieldTo(Fiber::RUNNABLE);
}

void sleepMs(int ms) {
	if(g_spawned.exchange(false)) return;
	timespec ts = {ms / 1000, (long)(ms % 1000) * 1000000L};
	nanosleep(&ts, NULL);
}

void spawnWorker(void (*fn)(void*), void *arg) {
	std::lock_guard<std::mutex> lk(g_mu);
	if(g_nthreads == 0) {
		int hw = (int)std::thread::hardware_concurrency();
		if(hw < 1) hw = 1;
		g_nthreads = envInt("BT2GPU_THREADS", hw > 16 ? 16 : hw, 1, 1024);
		g_nlanes = envInt("BT2GPU_LANES", 1, 1, 8);
		g_stackBytes = (size_t)envInt("BT2GPU_STACK_KB", 1024, 64, 1 << 20) * 1024;
		if(const char *dv = getenv("BT2GPU_DEVICES")) {
			for(const char *p = dv; *p; ) {
				char *e = NULL;
				long d = strtol(p, &e, 10);
				if(e == p) break;
				g_devices.push_back((int)d);
				p = (*e == ',') ? e + 1 : e;
			}
		}
		if(g_devices.empty()) g_devices.push_back(device());
		g_publishAt = (size_t)envInt("BT2GPU_PUBLISH", 32, 1, 1 << 20);
		g_localCands = (size_t)envInt("BT2GPU_LOCAL_CANDS", 8192, 1, 1 << 30);
		g_batchMin = (size_t)envInt("BT2GPU_BATCH_MIN", 64, 0, 1 << 20);
		g_batchWaitUs = (size_t)envInt("BT2GPU_BATCH_WAIT_US", 100, 1, 100000);
		// per device: BT2GPU_FM_LANES x { exactSweep + oneMmSearch + ungapped | seeds + extend | locate } and
		// BT2GPU_LANES x { DP windows }
		const int fmLanes = envInt("BT2GPU_FM_LANES", 1, 1, 8);
		for(int d : g_devices) {
			Hub *h = new Hub();
			h->dev = d;
			auto add = [&](std::initializer_list<Kind> kinds) {
				Lane *l = new Lane();
				l->dev = d; l->id = (int)h->lanes.size();
				h->lanes.push_back(l);
				for(Kind k : kinds) h->of[k].push_back(l);
			};
			for(int i = 0; i < fmLanes; i++) { add({K_SWEEP, K_ONEMM, K_UNGAP}); add({K_SEEDS, K_SEEDMM, K_EXTEND}); add({K_LOCATE}); }
			for(int i = 0; i < g_nlanes; i++) add({K_DP});
			for(Lane *l : h->lanes) { l->th = std::thread([l] { l->run(); }); l->th.detach(); }
			g_hubs.push_back(h);
		}
	}
	Worker *w = NULL;
	if((int)g_workers.size() < g_nthreads) {
		w = new Worker();
		w->id = (int)g_workers.size();
		w->hub = g_hubs[(size_t)w->id % g_hubs.size()];
		g_workers.push_back(w);
		w->th = std::thread([w] { w->loop(); });
		w->th.detach();
	} else w = g_workers[g_next++ % g_workers.size()];
	Fiber *f = new Fiber();
	f->fn = fn; f->arg = arg; f->w = w; f->stackBytes = g_stackBytes;
	{
		std::lock_guard<std::mutex> lk2(w->inboxMu);
		w->inbox.push_back(f);
		w->inboxN.fetch_add(1, std::memory_order_release);
	}
	g_spawned.store(true);
}

} // namespace bt2shim
