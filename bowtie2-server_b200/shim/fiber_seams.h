// fiber_seams.h — FORCE-INCLUDED (-include, with -DBT2SHIM_SEAMS_SEARCH) when the unmodified bt2_search.cpp is compiled
// for the batched GPU-bound server (oracle/Makefile, target gpub).  It stands for the three `#ifdef BT2_GPU` edits a
// maintainer would make in that file; nothing else of the reference changes:
//
//   1. worker start-up (bt2_search.cpp:4913-4925): `new std::thread(multiseedSearchWorker, &tps[i])` starts a worker
//      FIBER (fiber.h) instead of an OS thread -- `-p N` is then the number of reads in flight, spread over
//      BT2GPU_THREADS host threads -- and the 10 ms pause after each start is skipped;
//   2. input (bt2_search.cpp:3201): `PSFactory::ReadAhead psrah(factory)` blocks the calling thread on the server's
//      ready queue (pat.h:1885-1911, :1987-1994); a fiber must not block its host thread, so the batched build takes
//      the next filled read buffer if there is one and otherwise lets the other workers of the thread run;
//      When a worker hands a read buffer back it also writes out the results it has buffered for that connection.
//      The reference's OutputQueue keeps the last readsPerBatch (16) results of every worker per connection until that
//      worker's 17th arrives (outq.cpp:84-96) -- 256 read pairs with 16 threads, but 32768 with 2048 workers, more than
//      the 20000 pairs a client may have in flight (PatternSourceWebClient::LockedOrigBufMap, pat.h:2464-2550: two
//      halves of 10000, a half is reused only when ALL its results are back).  Held-back results then starve the
//      input and, with few clients, deadlock client and server; for the same reason the connection's OutFileBuf is
//      handed to its writer then, instead of waiting for 16 KB of further output.  Same records, same stream; only
//      when a worker's records are written changes (their order was thread-dependent already).
//   3. GroupWalk2S -> the locate call of the binding (group_walk_gpu.h, as in the per-call build).
//   4. read buffers per connection (bt2_search.cpp:4859 asks for 4 * nthreads + 1, pat.cpp:2045-2050 builds them for every
//      connection): with `-p` in the thousands that is 8193+ PatternSourcePerThread of 32 Reads each, built and torn
//      down per connection, of which the reference's client can never have more than 20000 / 16 = 1250 in flight (its
//      window, above).  The batched build caps the count at BT2GPU_READ_BUFFERS (default 1280).
//
// All reference headers bt2_search.cpp uses are included FIRST, so the renames below touch that one file only.
#ifndef BT2GPU_FIBER_SEAMS_H_
#define BT2GPU_FIBER_SEAMS_H_

#include "group_walk_gpu.h"

#ifdef BT2SHIM_SEAMS_SEARCH
#include <thread>
#include <mutex>
#include <condition_variable>
#include <queue>
#include "alphabet.h"
#include "assert_helpers.h"
#include "endian_swap.h"
#include "bt2_idx.h"
#include "formats.h"
#include "sequence_io.h"
#include "tokenize.h"
#include "aln_sink.h"
#include "pat.h"
#include "threading.h"
#include "ds.h"
#include "aligner_metrics.h"
#include "sam.h"
#include "aligner_seed.h"
#include "aligner_seed_policy.h"
#include "aligner_driver.h"
#include "aligner_sw.h"
#include "aligner_sw_driver.h"
#include "aligner_cache.h"
#include "util.h"
#include "pe.h"
#include "simple_func.h"
#include "presets.h"
#include "opts.h"
#include "outq.h"
#include "aligner_seed2.h"
#include "fiber.h"

namespace bt2shim {

// PatternSourceServiceFactory::ReadAhead (pat.h:1885-1911) with a non-blocking first step: same members, same
// protocol (the element goes back to the connection's idle queue when the holder is destroyed).
// Compiled with -fno-access-control: the ready queue and its lock are private members of the factory.
class FiberReadAhead {
public:
	typedef PatternSourceServiceFactory::ReadElement ReadElement;
	FiberReadAhead(PatternSourceServiceFactory& fact) : fact_(fact), re_(take(fact)), first_(true) {}
	~FiberReadAhead() {
		if(re_.ps != NULL) flushWorker(re_.ps->msink().outq());      // before returnUnready: it may delete ps
		fact_.returnUnready(re_);
	}
	const std::pair<bool, bool>& nextReadResult() {
		if(first_) first_ = false;       // nextReadPair was already called by the connection's thread (pat.cpp:2064-2079)
		else re_.nextReadPair();
		return re_.readResult;
	}
	PatternSourcePerThread* ptr() { return re_.ps; }
private:
	// OutputQueue::flushImpl (outq.cpp:101-115) for the calling worker's buffer only
	static void flushWorker(OutputQueue& oq) {
		thread_tracking_pair *tp = (thread_tracking_pair*)currentWorkerArg();
		if(tp == NULL || oq.reorder_ || oq.perThreadCounter == NULL) return;
		const size_t tid = (size_t)tp->tid;
		if(tid >= oq.nthreads_ || oq.perThreadCounter[tid] == 0) return;
		{
			ThreadSafe ts(oq.mutex_m);
			for(int j = 0; j < oq.perThreadCounter[tid]; j++) {
				oq.obuf_.writeString(oq.perThreadBuf[tid][j]);
				oq.nfinished_++;
				oq.nflushed_++;
			}
			oq.perThreadCounter[tid] = 0;
		}
		// ... and on to the socket: OutFileBuf keeps up to 16 KB (16 MB while its writer is busy, filebuf.h:999-1000,
		// :757-773) until more output pushes it out.  With a thousand workers finishing in any order the records left
		// there can be ones the client's window is waiting for, and then nothing more is ever written.  flush() waits
		// for the connection's writer thread: while that one is busy this worker lets the others run instead.
		for(;;) {
			{
				ThreadSafe ts(oq.mutex_m);
				if(oq.obuf_.cur_ == 0) break;
				if(!oq.obuf_.flushBlocking()) { oq.obuf_.flush(); break; }
			}
			yieldNow();
		}
	}
	static ReadElement take(PatternSourceServiceFactory& fact) {
		for(;;) {
			{
				std::unique_lock<std::mutex> lk(fact.psq_ready_.m_);
				if(!fact.psq_ready_.q_.empty()) {
					ReadElement r(fact.psq_ready_.q_.front());
					fact.psq_ready_.q_.pop();
					return r;
				}
			}
			waitInput();
		}
	}
	PatternSourceServiceFactory& fact_;
	ReadElement re_;
	bool first_;
};

// 4. the factory with the number of read buffers per connection capped
class BatchedServiceFactory : public PatternSourceServiceFactory {
public:
	BatchedServiceFactory(int server_port, PatternComposer& composer, const PatternParams& pp, size_t n_readahead,
	                      AlnSinkSam& msink, const Config& config)
		: PatternSourceServiceFactory(server_port, composer, pp, readBuffers(n_readahead), msink, config) {}
private:
	static size_t readBuffers(size_t asked) {
		const char *v = getenv("BT2GPU_READ_BUFFERS");
		size_t cap = (v != NULL && atol(v) > 0) ? (size_t)atol(v) : 1280;
		return asked < cap ? asked : cap;
	}
};

} // namespace bt2shim

namespace std { typedef ::bt2shim::FiberThread bt2shim_fiber_thread; }

// 1. std::thread -> worker fibers (the only uses of the token in bt2_search.cpp are `std::thread`)
#define thread bt2shim_fiber_thread
// 2. `PSFactory::ReadAhead psrah(f);` -> `PSFactory::ReadElement *bt2shim_unused_ = 0; bt2shim::FiberReadAhead psrah(f);`
#define ReadAhead ReadElement *bt2shim_unused_ = 0; ::bt2shim::FiberReadAhead
// 4. `PatternSourceServiceFactory readahead_factory(port, patsrc, pp, 4*nthreads+1, ...)` -> the capped factory
#define PatternSourceServiceFactory ::bt2shim::BatchedServiceFactory
// the pause after each worker start (threading.h:44-51)
#undef SLEEP
#define SLEEP(x) ::bt2shim::sleepMs(x)
#endif  // BT2SHIM_SEAMS_SEARCH

#endif
