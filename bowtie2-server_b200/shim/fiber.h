// fiber.h — user-space workers for the batched binding (batch_shim.cpp).
//
// The reference aligns one read per worker thread with blocking calls (multiseedSearchWorker, bt2_search.cpp:3050-4197:
// exactSweep, searchAllSeeds, SwDriver::extend, GroupWalk2S::advanceElement, SwAligner::align ... one after the other).
// To hand the GPU thousands of problems per ABI call WITHOUT touching that driver -- it decides, with the per-read RNG,
// which problems exist at all -- every worker the reference starts (`-p N`) becomes a fiber: a worker with its own
// stack that runs the reference's unmodified per-read code until it reaches a hot-path call, queues the problem, and
// switches (in user space, ~20 ns) to the next worker of the same host thread.  When every worker of the thread is
// waiting (and at the latest every 16 workers), the queued problems are handed to the lanes of their kinds, which make
// one bt2gpu_*_batch call per kind; each worker resumes with its own answer.
// N workers are spread over BT2GPU_THREADS host threads (default: the machine's cores, at most 16).  Each device has one
// submission lane per kind of problem (a bt2gpu_ctx with its own stream and arenas, driven by its own thread), shared by
// all host threads: whatever is queued for a kind when its lane is free travels together, and the kinds overlap.
//
// Nothing here computes an alignment: it only decides when problems travel.  x86-64 System V only.
#ifndef BT2GPU_FIBER_H_
#define BT2GPU_FIBER_H_

namespace bt2shim {

// starts `fn(arg)` as a worker on one of the host threads (round robin); returns at once
void spawnWorker(void (*fn)(void*), void *arg);

// what the reference's `new std::thread(fn, arg)` / `->detach()` become in the worker start-up loop
// (bt2_search.cpp:4913-4925) of the batched build
class FiberThread {
public:
	template<typename F> FiberThread(F fn, void *arg) { spawnWorker((void (*)(void*))fn, arg); }
	void detach() {}
	void join() {}
};

bool inWorker();        // true on a worker's stack
void *currentWorkerArg(); // the `arg` the running worker was started with (NULL outside a worker)
void waitInput();       // a worker found nothing to read: suspend until the scheduler lets it look again
void yieldNow();        // let the other workers of this host thread run; the caller stays runnable
void sleepMs(int ms);   // SLEEP(): skipped right after a spawn (10 ms per worker x thousands of workers otherwise)

} // namespace bt2shim
#endif
