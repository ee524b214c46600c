#!/usr/bin/env python3
"""bench.py -- aligned reads/s of the B200 build next to the reference's CPU server, on BASELINE.json's own shape.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--batch READS] [--genome-mbp 3100] [--nseq 24] [--mode pairs|e2e|local]

Workload (config C5 of BASELINE.json at N=1): an i.i.d. random reference of 3.1 Gbp in 24 sequences (small index,
~3.6 GB with both BWTs, the SA sample and the 2-bit text; built at set-up on the GPU by bt2gpu.build, byte-identical to
bowtie2-build-s output, and written to /tmp so that every arm loads the same files), 2x150 bp --fr pairs from the
simulate.pl model, `-I 0 -X 500`, defaults otherwise.  A step = one batch of reads (262144 = 131072 pairs per GPU).

Two measurements per run, both in the one JSON line:
  e2e    THE HEADLINE.  The batch goes through the reference's own front end: FASTQ files -> `bowtie2-align-l` clients ->
         the reference server whose SeedAligner / SwAligner / SwDriver::extend / GroupWalk2S are bound to libbt2gpu.so by
         the batched binding (oracle/_ref/bowtie2-align-server-s-gpub: every worker of the unmodified server is a fiber,
         the hot-path calls of all reads in flight travel in one C-ABI call per kind, HOST buffers in and out) -> SAM
         files.  The SAM is byte-identical to the stock server's (checked in the same run on the CPU sample).  reads /
         wall time from the first client's start to the last client's exit.
  value  the hot path alone with the reads already in HBM (bt2gpu_hotpath_pairs_batch: exactSweep -> exact seeds ->
         extend -> locate -> seed-extension DP -> mate-window DP, every intermediate resident on the device; CUDA events
         from the first kernel to the last).  It explains the kernels (rooflines, ncu) and is NOT a SAM-producing path:
         it extends one deterministically chosen hit per read (include/bt2gpu.h).
`--impl reference` = the unmodified reference server (oracle/_ref/bowtie2-align-server-s, `-p <all host cores>`) with the
same clients on the same reads: reads / wall time.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "bowtie2-server_b200"))

import numpy as np  # noqa: E402

RDLEN = 150
MAXGAP = 15
SEED_LEN = 22
IVAL = (1.0, 1.15)
LOCAL = False
PAIRED = False
MAXFRAG, MINFRAG, MAXHALF = 500, 0, 15
PE_FLAGS = (0, 0, 0, 1, 1, 1)   # local, flip, dovetail, contain, overlap, expand-to-fit: the reference defaults (bt2_search.cpp:381-385)
SERVER_ARGS = []
METRIC = "aligned reads/sec (2x150 end-to-end)"
UNIT = "reads/s"


def workload_name(batch, gmbp, nseq):
    ref = "FM index of an i.i.d. random %d Mbp reference in %d sequences" % (gmbp, nseq)
    if PAIRED:
        return ("synthetic 2x%d bp --fr pairs (simulate.pl model, fragments N(250,45)) vs %s, -I %d -X %d, defaults "
                "otherwise (BASELINE.json configs[4] shape); %d reads (%d pairs) per step per GPU"
                % (RDLEN, ref, MINFRAG, MAXFRAG, batch, batch // 2))
    return ("synthetic %d bp reads (simulate.pl model) vs %s, %s; %d reads per step per GPU"
            % (RDLEN, ref, " ".join(SERVER_ARGS) or "defaults", batch))


# ---------------------------------------------------------------------------------------------
# set-up shared by both arms: genome, index files, reads
# ---------------------------------------------------------------------------------------------
def seq_lengths(gmbp, nseq):
    """nseq sequence lengths adding up to gmbp * 10^6, longest first (a chromosome-like spread)."""
    total = int(gmbp * 1000000)
    w = np.array([1.0 / (1.0 + 0.12 * i) for i in range(nseq)])
    ln = np.floor(total * w / w.sum()).astype(np.int64)
    ln[0] += total - int(ln.sum())
    return [int(x) for x in ln]


def make_genome(seed, gmbp, nseq):
    rng = np.random.default_rng(seed)
    return [rng.integers(0, 4, size=n, dtype=np.uint8) for n in seq_lengths(gmbp, nseq)]


def get_index(args, build_ok=True):
    """(index base path, genome as list of code arrays).  Built once per box (GPU, bt2gpu.build) and kept in /tmp."""
    from bt2gpu import build, index
    base = os.path.join(tempfile.gettempdir(), "bt2gpu_bench_idx_%d_%d_%d" % (args.genome_mbp, args.nseq, args.seed))
    gpath = base + ".genome.npy"
    if index.exists(base) and os.path.exists(base + ".ok"):
        flat = np.load(gpath, mmap_mode="r")
        cuts = np.concatenate([[0], np.cumsum(seq_lengths(args.genome_mbp, args.nseq))])
        return base, [flat[cuts[i]:cuts[i + 1]] for i in range(args.nseq)], None
    if not build_ok:
        raise RuntimeError("index files missing")
    t0 = time.time()
    seqs = make_genome(args.seed, args.genome_mbp, args.nseq)
    t1 = time.time()
    ix, extras = build.build_index(seqs, ["chr%d" % (i + 1) for i in range(args.nseq)])
    t2 = time.time()
    build.write_index(ix, extras, base)
    np.save(gpath, extras["joined"])
    open(base + ".ok", "w").write("ok")
    sys.stderr.write("bench: genome %.1f s, index build %.1f s, files %.1f s\n" % (t1 - t0, t2 - t1, time.time() - t2))
    return base, seqs, ix


def make_reads(seed, seqs, n):
    """n reads sampled over the sequences in proportion to their lengths (pairs never span two sequences)."""
    from bt2gpu import synth
    rng = np.random.default_rng(seed)
    lens = np.array([len(s) for s in seqs], dtype=np.float64)
    units = n // 2 if PAIRED else n
    per = rng.multinomial(units, lens / lens.sum())
    rs, qs = [], []
    for s, k in zip(seqs, per):
        if k == 0:
            continue
        if PAIRED:
            r, q, _, _ = synth.simulate_pairs(rng, s, int(k), RDLEN)
        else:
            r, q, _, _ = synth.simulate_reads(rng, s, int(k), RDLEN)
            rc = rng.random(int(k)) < 0.5          # strand coin-flip as simulate.pl does
            r[rc] = np.array([3, 2, 1, 0, 4], dtype=np.uint8)[r[rc][:, ::-1]]
            q[rc] = q[rc][:, ::-1]
        rs.append(r); qs.append(q)
    reads, quals = np.concatenate(rs), np.concatenate(qs)
    # shuffle units so that every client shard and every batch sees all sequences
    o = rng.permutation(units)
    if PAIRED:
        o = np.stack([2 * o, 2 * o + 1], axis=1).reshape(-1)
    return np.ascontiguousarray(reads[o]), np.ascontiguousarray(quals[o])


def write_shards(td, tag, reads, quals, nclients):
    """FASTQ shards, one per client; returns the clients' argument lists."""
    from bt2gpu import serverbench as SB
    n = len(reads)
    units = n // 2 if PAIRED else n
    cuts = [units * i // nclients for i in range(nclients + 1)]
    shards = []
    for k in range(nclients):
        a, b = cuts[k], cuts[k + 1]
        if b <= a:
            continue
        if PAIRED:
            f1, f2 = os.path.join(td, "%s_%d_1.fq" % (tag, k)), os.path.join(td, "%s_%d_2.fq" % (tag, k))
            SB.write_fastq(f1, reads[2 * a:2 * b:2], quals[2 * a:2 * b:2], a, "/1")
            SB.write_fastq(f2, reads[2 * a + 1:2 * b:2], quals[2 * a + 1:2 * b:2], a, "/2")
            shards.append(["-1", f1, "-2", f2])
        else:
            f = os.path.join(td, "%s_%d.fq" % (tag, k))
            SB.write_fastq(f, reads[a:b], quals[a:b], a)
            shards.append(["-U", f])
    return shards


class ClockSampler(threading.Thread):
    def __init__(self, device):
        super().__init__(daemon=True)
        self.device, self.stop_flag, self.samples, self.reasons = device, False, [], set()
        self.max_mhz = None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                o = subprocess.check_output(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + q,
                                             "--format=csv,noheader,nounits"], timeout=5).decode().strip()
                t = [x.strip() for x in o.split(",")]
                self.samples.append(float(t[0]))
                self.max_mhz = float(t[1])
                for n, v in zip(names, t[2:]):
                    if v.lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


_T0 = time.time()


def progress(msg):
    """One time-stamped line on stderr per stage: a run that is cut off shows where it was."""
    sys.stderr.write("bench [%6.1f s] %s\n" % (time.time() - _T0, msg))
    sys.stderr.flush()


def default_clients(nreads):
    """Client processes per server: enough that no client sends more than 16384 read pairs (or reads) per pass -- the
    reference client keeps at most 20000 in flight (pat.h:2464-2550) and its flow control can stall beyond that, with the
    stock server too -- and at least 2, at most 16."""
    units = nreads // 2 if PAIRED else nreads
    return max(2, min(16, -(-units // 16384)))


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ---------------------------------------------------------------------------------------------
# server runs (both arms)
# ---------------------------------------------------------------------------------------------
def server_steps(binary, base, workers, env, shards, td, tag, warmup, steps, keep_sam=False, pass_timeout=600, retries=1):
    """Start one server, run warmup + steps passes of the shard set through it.  Returns (seconds of the timed steps,
    server start seconds, RSS MB, sam paths of the last step, shim counters or None).
    A pass that does not end within pass_timeout abandons that server (its clients are killed) and the whole series is
    run again on a fresh one, `retries` times: the reference's client / server pair can stall on its own flow control
    (PatternSourceWebClient's window of 20000 pairs against results the server has not flushed yet), the stock server
    included, and one stalled pass must not cost the run its JSON line."""
    from bt2gpu import serverbench as SB
    stats = os.path.join(td, tag + "_shim_stats.txt")
    e = dict(env or {})
    if binary == SB.GPUB_SERVER:
        e["BT2GPU_SHIM_STATS"] = stats
    for attempt in range(retries + 1):
        srv = SB.Server(binary, base, workers, SERVER_ARGS, env=e, log=os.path.join(td, "%s_server%d.err" % (tag, attempt)), start_timeout=300)
        progress("%s: server up in %.1f s (-p %d)" % (tag, srv.start_s, workers))
        try:
            for w in range(warmup):
                dt, _ = srv.align(shards, os.path.join(td, "%s_w%d" % (tag, w)), timeout=pass_timeout)
                progress("%s: warm-up pass %d %.1f s" % (tag, w, dt))
            t, sams = 0.0, []
            for s in range(steps):
                dt, sams = srv.align(shards, os.path.join(td, "%s_s%d" % (tag, s)), timeout=pass_timeout)
                progress("%s: timed pass %d %.1f s" % (tag, s, dt))
                t += dt
            rss = srv.rss_mb()
            break
        except Exception as ex:
            progress("%s: attempt %d failed: %s" % (tag, attempt, str(ex).splitlines()[0][:200]))
            if attempt == retries:
                raise
        finally:
            srv.close()
    cnt = None
    if os.path.exists(stats):
        import re
        cnt = dict((k, int(v)) for k, v in re.findall(r"(\w+)\s+(\d+)", open(stats).read()))
    return t, srv.start_s, rss, sams, cnt


def stock_server_steps(base, ncores, shards, td, tag, warmup, steps):
    """The unmodified reference server in both of its builds (SSE2 default, SSE_AVX2=1): the faster one counts
    (SURVEY 8d).  Returns (seconds, start_s, rss, sams, build name, {build: seconds})."""
    from bt2gpu import serverbench as SB
    best, times = None, {}
    for name, binary in (("sse2", SB.CPU_SERVER), ("avx2", SB.CPU_SERVER_AVX2)):
        if not os.path.exists(binary):
            continue
        try:
            t, start_s, rss, sams, _ = server_steps(binary, base, ncores, None, shards, td, "%s_%s" % (tag, name), warmup, steps, pass_timeout=120)
        except Exception as e:      # e.g. a host CPU without AVX2
            sys.stderr.write("bench: stock server build %s failed: %s\n" % (name, str(e).splitlines()[0][:200]))
            continue
        times[name] = t
        if best is None or t < best[0]:
            best = (t, start_s, rss, sams, name)
    if best is None:
        raise RuntimeError("no stock reference server could be run")
    return best + (times,)


def run_reference(args, rank, world):
    if rank != 0:
        return
    from bt2gpu import serverbench as SB
    if not SB.have_binaries():
        emit({"impl": "reference", "unavailable": "oracle/_ref binaries missing (built by __graft_entry__.build() where /root/reference is mounted)"})
        return
    ncores = host_cores()
    base, seqs, _ = get_index(args)
    n = min(args.batch, args.cpu_sample) if args.cpu_sample else args.batch
    reads, quals = make_reads(args.seed + 7, seqs, n)
    nclients = args.clients or default_clients(len(reads))
    with tempfile.TemporaryDirectory() as td:
        shards = write_shards(td, "ref", reads, quals, nclients)
        t, start_s, rss, _, build, times = stock_server_steps(base, ncores, shards, td, "ref", args.warmup, args.steps)
    val = len(reads) * args.steps / t
    sample = ("%d reads per step through the unmodified reference server (bowtie2-align-server-s -p %d, the faster of its "
              "SSE2 and SSE_AVX2=1 builds: %s) and %d bowtie2-align-l clients; wall time from the first client's start to the "
              "last client's exit" % (len(reads), ncores, build, nclients))
    emit({"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
          "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
          "vs_baseline": None, "dtype": "u8/i16/u32", "data": "synthetic",
          "config": {"workload": workload_name(args.batch, args.genome_mbp, args.nseq), "reads_per_step": len(reads),
                     "server_threads": ncores, "clients": nclients, "server_args": SERVER_ARGS, "server_start_s": start_s,
                     "server_rss_mb": rss, "server_build": build,
                     "seconds_by_build": times},
          "cpu_baseline": {"value": val, "unit": UNIT, "cores": ncores, "kind": "reference", "sample": sample},
          "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def load_traffic(args):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch, from the committed ncu capture of THIS command
    (profiles/r2_traffic.json, written by scripts/ncu_traffic.py from the ncu CSV; it records the command and the
    workload it was taken on).  Other workloads: None."""
    p = os.path.join(ROOT, "profiles", "r2_traffic.json")
    if not os.path.exists(p):
        return {}, None
    t = json.load(open(p))
    w = t.get("workload", {})
    if (w.get("batch"), w.get("genome_mbp"), w.get("nseq"), w.get("mode")) != (args.batch, args.genome_mbp, args.nseq, args.mode):
        return {}, None
    return t.get("kernels", {}), t.get("source")


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from bt2gpu import index, lib, scoring, structs
    from bt2gpu import serverbench as SB
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device and the product has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1 and rank != 0:
        dist.barrier()                                        # rank 0 builds and writes the files first
    base, seqs, ix = get_index(args)
    if world > 1 and rank == 0:
        dist.barrier()
    progress("index files ready")
    if ix is None:
        ix = index.Index.load(base)
    progress("index in host memory")
    torch.cuda.empty_cache()
    sc = scoring.Scoring(local=LOCAL)
    dix = lib.DeviceIndex(ix, local_rank)
    recs = index.read_ref_records(base)
    joined = np.load(base + ".genome.npy", mmap_mode="r")
    dref = lib.DeviceRef(joined, local_rank, recs=recs)       # 2-bit text + .3.bt2 records in HBM (windows are cut on the device)
    ctx = lib.Context(local_rank)
    ctx.set_scoring(sc)
    ctx.bind_index(dix)
    progress("index and reference in HBM")
    reads, quals = make_reads(args.seed + 7 + 1000 * rank, seqs, args.batch)   # each rank its own shard of the read stream
    n = len(reads)
    progress("%d reads made" % n)

    keep = []

    def pin(a):
        t_ = torch.empty(a.nbytes, dtype=torch.uint8, pin_memory=True)
        v = t_.numpy().view(a.dtype).reshape(a.shape)
        v[...] = a
        keep.append(t_)
        return v
    hseq, hqual = pin(reads.reshape(-1)), pin(quals.reshape(-1))
    max_aln = 2
    rd = np.zeros(n, dtype=structs.read)
    rd["rd_off"] = np.arange(n, dtype=np.int64) * RDLEN
    rd["rd_len"] = RDLEN
    io = dict(seq=hseq, qual=hqual, rd=rd,
              minsc=np.full(n, max(1, sc.score_min.f_int(RDLEN)) if LOCAL else sc.score_min.f_int(RDLEN), dtype=np.int32),
              nceil=np.full(n, sc.nceil(RDLEN), dtype=np.int32),
              out=dict(sweep=pin(np.zeros(n, dtype=structs.sweep_result)), job_read=np.zeros(n, dtype=np.uint32),
                       job_start=np.zeros(n, dtype=np.int64), job_fw=np.zeros(n, dtype=np.uint8),
                       results=pin(np.zeros(n, dtype=structs.dp_result)),
                       alns=pin(np.zeros(n * max_aln, dtype=structs.dp_aln)),
                       edits=pin(np.zeros(64 * n + 1024, dtype=structs.edit)),
                       **(dict(mates=pin(np.zeros(n, dtype=structs.mate_info)), mresults=pin(np.zeros(n, dtype=structs.dp_result)),
                               malns=pin(np.zeros(n * max_aln, dtype=structs.dp_aln)),
                               medits=pin(np.zeros(64 * n + 1024, dtype=structs.edit))) if PAIRED else {})))
    pe_pol = np.zeros(1, dtype=structs.pe_policy)
    pe_pol["pol"], pe_pol["trim_to_ref"], pe_pol["maxfrag"], pe_pol["minfrag"] = 3, 1, MAXFRAG, MINFRAG
    for k_, v_ in zip(("local", "flip_ok", "dovetail_ok", "contain_ok", "olap_ok", "expand_to_fit"), PE_FLAGS):
        pe_pol[k_] = v_

    def chain():
        if PAIRED:
            o_, st_, ps_ = ctx.hotpath_pairs(dref, io["seq"], io["qual"], io["rd"], io["minsc"], io["nceil"], pe_pol, maxhalf=MAXHALF,
                                             seed_len=SEED_LEN, ival=IVAL, maxgap=MAXGAP, max_aln=max_aln, mine_max=2, out=io["out"])
            st_ = dict(st_)
            st_.update({("mate_" + k_ if not k_.startswith("mate_") else k_): v_ for k_, v_ in ps_.items()})
            st_["ms_total"] = st_["ms_total"] + ps_["ms_total"]          # device span of both passes
            st_["d2h_bytes"] += ps_["d2h_bytes"]
            return o_, st_
        return ctx.hotpath_round0(dref, io["seq"], io["qual"], io["rd"], io["minsc"], io["nceil"], seed_len=SEED_LEN,
                                  ival=IVAL, maxgap=MAXGAP, max_aln=max_aln, mine_max=2, nthreads=0, out=io["out"])

    def barrier(with_ctx=True):
        if world > 1:
            dist.barrier()
        if with_ctx:
            ctx.sync()
        torch.cuda.synchronize()

    # ---- (1) `value`: the device-resident stage chain ----
    for _ in range(args.warmup):
        chain()
    barrier()
    progress("chain: %d warm-up steps done" % args.warmup)
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = ctx.launches
    dev_ms = 0.0
    kms, cnt = {}, None
    for _ in range(args.steps):
        o, cnt = chain()
        dev_ms += cnt["ms_total"]
        for k in ("sweep", "seed", "extend", "locate", "dp", "dp_fill", "dp_gather", "dp_backtrace", "glue", "total"):
            kms[k] = kms.get(k, 0.0) + cnt["ms_" + k]
        if PAIRED:
            for k in ("frame", "dp", "dp_fill", "dp_gather", "dp_backtrace", "total"):
                kms["mate_" + k] = kms.get("mate_" + k, 0.0) + cnt["mate_ms_" + k]
    barrier()
    chain_launches = ctx.launches - l0
    progress("chain: %d timed steps done, %.1f ms on the device" % (args.steps, dev_ms))
    cnt = {k: (int(v) if not k.startswith(("ms_", "wall_", "mate_ms_")) else float(v)) for k, v in cnt.items()}
    probe_ix = ctx.gather_probe("fw", 1 << 27)                # random sides of the index itself (L2 helps when it fits)
    probe = ctx.probe_random_gather(8 << 30, 1 << 27)         # random 64-B gathers over 8 GiB: the HBM random-access ceiling
    dpx_peak = ctx.probe_dpx()                                # G thread-instr/s of packed s16x2 DPX
    index_bytes = dix.nbytes + dref.nbytes
    progress("probes done")
    ctx.close(); dref.close(); dix.close()
    del keep[:]
    torch.cuda.empty_cache()

    # ---- (2) `e2e`: the SAM-identical path, reads as FASTQ through the reference's own front end ----
    ncores = host_cores()
    threads = args.host_threads or max(1, min(16, ncores // world))
    workers = args.workers or threads * args.workers_per_thread
    nclients = args.clients or default_clients(n)
    e2e_s, srv_start, rss, shim, sams, parity = None, None, None, None, [], None
    e2e_error = None
    td_obj = tempfile.TemporaryDirectory()
    td = td_obj.name
    have = SB.have_binaries() and not args.chain_only
    if have:
        shards = write_shards(td, "r%d" % rank, reads, quals, nclients)
        progress("FASTQ shards for %d clients written" % nclients)
        env = {"BT2GPU_DEVICE": str(local_rank), "BT2GPU_THREADS": str(threads), "BT2GPU_LANES": str(args.lanes)}
        if ncores // world < 8:      # several servers on few cores (N ranks on one box): their lanes poll less often
            env["BT2GPU_POLL_US"] = "50"
        barrier(False)
        try:
            e2e_s, srv_start, rss, sams, shim = server_steps(SB.GPUB_SERVER, base, workers, env, shards, td, "gpub%d" % rank,
                                                              args.warmup, args.steps, pass_timeout=args.pass_timeout)
        except Exception as ex:      # reported in the line (e2e null, e2e_error), not a crash without a line
            e2e_error = str(ex).splitlines()[0][:300]
            progress("e2e failed: " + e2e_error)
    sampler.stop_flag = True
    sampler.join()
    t = torch.tensor([dev_ms, e2e_s or 0.0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_s = [float(x) for x in t.tolist()]
    if rank != 0:
        td_obj.cleanup()
        return
    total_reads = n * world * args.steps
    value = total_reads / (dev_ms / 1e3)
    e2e = total_reads / e2e_s if e2e_s else None

    # ---- CPU baseline + parity on rank 0 at N=1: the stock server on the same reads (bounded sample) ----
    cpu_line = None
    if world == 1 and have and not args.no_cpu_baseline and e2e_error is None:
        m = min(n, args.cpu_sample) if args.cpu_sample else n
        m -= m % 2
        cshards = shards if m == n else write_shards(td, "cpu", reads[:m], quals[:m], nclients)
        t_cpu, cstart, crss, csams, cbuild, ctimes = stock_server_steps(base, ncores, cshards, td, "cpu", 1, 1)
        cpu_line = {"value": m / t_cpu, "unit": UNIT, "cores": ncores, "kind": "reference", "build": cbuild,
                    "reads_per_s_by_build": {k: m / v for k, v in ctimes.items()},
                    "sample": "%d reads through the unmodified reference server (bowtie2-align-server-s -p %d, the faster of its SSE2 "
                              "and SSE_AVX2=1 builds: %s) and %d bowtie2-align-l clients, one pass after one warm-up pass"
                              % (m, ncores, cbuild, nclients)}
        progress("stock server done")
        if m == n:
            a, b = SB.sam_records(csams), SB.sam_records(sams)
        else:
            t_g, _, _, gsams, _ = server_steps(SB.GPUB_SERVER, base, workers, env, cshards, td, "par", 0, 1)
            a, b = SB.sam_records(csams), SB.sam_records(gsams)
        parity = {"sam_records": len(a), "differing_records": SB.count_differing(a, b),
                  "aligned_records": sum(1 for ln in a if ln.split("\t")[2] != "*")}
    progress("parity compared")
    td_obj.cleanup()

    peaks = {}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    K = args.steps
    traffic, traffic_src = load_traffic(args)

    def rl(kernel, alg_bytes, ms, extra=None):
        a = alg_bytes / (ms / 1e3) / 1e9 if ms > 0 else 0.0
        d = {"bound": "hbm", "kernel": kernel, "achieved": a, "peak": hbm, "unit": "GB/s", "frac": a / hbm,
             "traffic": traffic.get(kernel), "ms": ms, "algorithmic_bytes": alg_bytes}
        d.update(extra or {})
        return d

    # FM kernels: dependent random 64-byte gathers.  achieved = algorithmic bytes (64 B per distinct BWT side the
    # reference would touch, + ftab / offs words) / kernel time; the bound is the measured random-gather ceiling of this
    # device (bt2gpu_probe_random_gather), the stream peak is listed beside it.
    def fm(kernel, sides, extra_bytes, ms, steps=None):
        b = 64.0 * sides + extra_bytes
        a = b / (ms / 1e3) / 1e9 if ms > 0 else 0.0
        d = {"bound": "hbm-random-gather", "kernel": kernel, "achieved": a, "peak": probe, "unit": "GB/s",
             "frac": (a / probe) if probe else 0.0, "traffic": traffic.get(kernel), "ms": ms, "algorithmic_bytes": b,
             "hbm_stream_peak_gbs": hbm, "frac_of_stream_peak": a / hbm, "index_gather_gbs": probe_ix}
        if steps is not None and ms > 0:
            d["lf_steps_per_s"] = steps / (ms / 1e3)
        return d

    # DP fill: bound = DPX issue rate.  peak GCUPS = measured packed-s16x2 DPX thread-instructions/s x 2 cells
    # / 4 instructions (the survey's algorithmic count); the kernel itself spends DPX_PER_PAIR per pair of cells.
    def fill(kernel, cells, ms, rec_bytes):
        g = cells / (ms / 1e3) / 1e9 if ms > 0 else 0.0
        peak = dpx_peak * 2.0 / 4.0
        return {"bound": "dpx", "kernel": kernel, "achieved": g, "peak": peak, "unit": "GCUPS", "frac": g / peak if peak else 0.0,
                "traffic": traffic.get(kernel), "ms": ms, "cells": cells, "dpx_peak_ginstr_per_s": dpx_peak,
                "dpx_instr_per_cell_pair_in_kernel": lib.dpx_per_pair(), "dpx_pipe_frac": (g * lib.dpx_per_pair() / 2.0) / dpx_peak if dpx_peak else 0.0,
                "design_stream": {"what": "bytes the kernel writes per cell by design (not algorithmic traffic)", "bytes_per_cell": rec_bytes,
                                  "gbs": rec_bytes * g, "frac_of_hbm_peak": rec_bytes * g / hbm}}

    fill_ms = kms["dp_fill"] / K
    rls = [
        fm("k_seed_exact", cnt["seed_sides"], 16.0 * cnt["seeds"], kms["seed"] / K, cnt["seed_bwops"]),
        fm("k_exact_sweep", cnt["sweep_bwops"], 0.0, kms["sweep"] / K, cnt["sweep_bwops"]),
        fm("k_extend", cnt["extend_fmops"], 0.0, kms["extend"] / K, cnt["extend_fmops"]),
        fm("k_locate", cnt["locate_steps"], 4.0 * cnt["locate_steps"] + 4.0 * cnt["locate_rows"], kms["locate"] / K, cnt["locate_steps"]),
        fill("k_dp_fill", cnt["dp_cells"], fill_ms, lib.fill_bytes_per_cell(LOCAL)),
        rl("k_dp_backtrace", lib.backtrace_bytes_per_cell(LOCAL) * cnt["bt_cells"] + 12.0 * cnt["dp_cands"], kms["dp_backtrace"] / K,
           {"note": "dependent walk: latency/issue-bound, the byte figure is the cells it must read, not a bandwidth claim",
            "walked_cells_per_job": cnt["bt_cells"] / max(cnt["dp_jobs"], 1), "candidates_per_job": cnt["dp_cands"] / max(cnt["dp_jobs"], 1)}),
        rl("k_dp_gather", 4.0 * cnt["dp_jobs"] * (RDLEN + 4 * MAXGAP), kms["dp_gather"] / K),
    ]
    if PAIRED:
        mfill = kms["mate_dp_fill"] / K
        rls += [
            fill("k_dp_fill[mate windows]", cnt["mate_cells"], mfill, lib.fill_bytes_per_cell(LOCAL)),
            rl("k_dp_backtrace[mate windows]", lib.backtrace_bytes_per_cell(LOCAL) * cnt["mate_bt_cells"] + 12.0 * cnt["mate_dp_cands"],
               kms["mate_dp_backtrace"] / K,
               {"walked_cells_per_job": cnt["mate_bt_cells"] / max(cnt["mate_jobs"], 1),
                "candidates_per_job": cnt["mate_dp_cands"] / max(cnt["mate_jobs"], 1)}),
            rl("k_dp_gather[mate windows]", 4.0 * cnt["mate_cells"] / RDLEN, kms["mate_dp_gather"] / K),
        ]
    dominant = max(rls, key=lambda r: r["ms"])
    h2d = d2h = None
    launches = int(chain_launches)
    if shim:
        launches += int(shim.get("kernel_launches", 0))
        h2d, d2h = shim.get("h2d_bytes"), shim.get("d2h_bytes")
    per_step = lambda v: int(v / (args.warmup + args.steps)) if v else 0  # noqa: E731  (the server ran warm-up + timed passes)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8/i16/u32", "data": "synthetic",
        "config": {"workload": workload_name(args.batch, args.genome_mbp, args.nseq),
                   "sam_identical": {"e2e": True, "value": False,
                                     "note": "e2e is the reference server bound to libbt2gpu.so (batched binding, worker fibers): its SAM is "
                                             "byte-identical to the stock server's; value is the device-resident stage chain (one "
                                             "deterministically chosen hit per read), which explains the kernels and prints no SAM",
                                     "checked_in_this_run": parity},
                   "l2": "index + reference %.0f MB and the per-step DP working set both exceed the 126 MB L2; inputs change every step "
                         "only in e2e (the stage chain re-reads the same %d reads: %.0f MB of reads + > 1 GB of DP arenas per step)"
                         % (index_bytes / 1e6, n, 2.0 * n * RDLEN / 1e6),
                   "index_bytes": index_bytes, "counters": cnt,
                   "e2e_path": {"server": "oracle/_ref/bowtie2-align-server-s-gpub", "workers": workers, "host_threads": threads,
                                "dp_lanes": args.lanes, "clients": nclients, "server_args": SERVER_ARGS,
                                "server_start_s": srv_start, "server_rss_mb": rss, "seconds_per_step": (e2e_s / args.steps) if e2e_s else None,
                                "shim_counters": shim, "host_cores": ncores, "error": e2e_error},
                   "peak_source": "measured" if peaks else "fallback", "traffic_source": traffic_src},
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": per_step(h2d), "d2h_bytes_per_step": per_step(d2h)},
        "gpu_launches": launches,
        "roofline": dominant, "rooflines": rls,
        "phase_ms": {k: v / K for k, v in kms.items()},
        "clocks": sampler.summary(),
    }
    if cpu_line is not None:
        line["cpu_baseline"] = cpu_line
    elif world == 1:
        line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "unavailable: " + (e2e_error or "oracle/_ref binaries missing, or --no-cpu-baseline / --chain-only")}
    emit(line)


_REAL_STDOUT = None


def quiet_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner on the first
    communicator whatever NCCL_DEBUG level is set above VERSION): from here on file descriptor 1 goes to stderr and the
    JSON line is written to the original stdout by emit()."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--batch", type=int, default=262144, help="reads per step per GPU")
    ap.add_argument("--genome-mbp", type=int, default=3100)
    ap.add_argument("--nseq", type=int, default=24)
    ap.add_argument("--mode", default="pairs", choices=["pairs", "e2e", "local"],
                    help="pairs: 2x150 --fr pairs, -I 0 -X 500 (C5, the metric's shape); e2e: unpaired 150 bp reads, --very-sensitive (C3); "
                         "local: 250 bp reads, --local (C4)")
    ap.add_argument("--workers-per-thread", type=int, default=64, help="worker fibers (reads in flight) per host thread of the GPU-bound server")
    ap.add_argument("--workers", type=int, default=0)
    ap.add_argument("--host-threads", type=int, default=0)
    ap.add_argument("--lanes", type=int, default=1, help="DP lanes per device of the GPU-bound server (BT2GPU_LANES)")
    ap.add_argument("--clients", type=int, default=0)
    ap.add_argument("--cpu-sample", type=int, default=0, help="reads of the batch the CPU server sees (0 = the whole batch)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--pass-timeout", type=int, default=120, help="seconds one pass of a batch through a server may take")
    ap.add_argument("--chain-only", action="store_true", help="device stage chain only (for ncu: no server processes)")
    ap.add_argument("--seed", type=int, default=20261104)
    args = ap.parse_args()
    global RDLEN, SEED_LEN, IVAL, LOCAL, PAIRED, SERVER_ARGS
    if args.mode == "pairs":
        PAIRED = True
        args.batch -= args.batch % 2
        SERVER_ARGS = ["-I", str(MINFRAG), "-X", str(MAXFRAG)]
    elif args.mode == "e2e":
        SEED_LEN, IVAL = 20, (1.0, 0.5)
        SERVER_ARGS = ["--very-sensitive"]
    else:
        RDLEN, SEED_LEN, IVAL, LOCAL = 250, 20, (1.0, 0.75), True
        SERVER_ARGS = ["--local"]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"        # keep NCCL's version banner off stdout: one JSON line only
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    run_ours(args, rank, world, local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
