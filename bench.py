#!/usr/bin/env python3
"""bench.py -- hot-path throughput of the B200 build vs the reference's CPU path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--batch B] [--genome-mbp G]

A "step" is one pass of the alignment hot path over one batch of synthetic 150 bp reads
(simulate.pl model; 262144 reads by default) against an FM index of a random genome (1024 Mbp
by default: index and E/F/H stream both far larger than L2):
exactSweep -> exact seeds (-L 22, interval S,1,1.15) -> extend -> locate -> one
seed-extension DP window per read (fill + gather + nextAlignment loop), every intermediate
resident in HBM (bt2gpu_hotpath_round0_batch).  The index is built at set-up by bt2gpu.build
(byte-identical to bowtie2-build-s output, tests/test_index_build.py) and written to /tmp so
the reference arm loads the same files.

Prints ONE JSON line (contract in the task statement): `value` = reads/s over the device span
of a step (CUDA events on the launching stream from the first kernel to the last, reads already
in HBM, glue kernels and the three scalar read-backs included), `e2e` = the same step through
the C ABI with HOST buffers (H2D and D2H inside the timed region), `roofline` for the dominant
kernel (`rooflines` lists every kernel, FM kernels also against the measured HBM random-gather
ceiling, the fill also against the measured DPX issue rate), `cpu_baseline` = the unmodified
reference (SeedAligner, SwDriver::extend, Ebwt::getOffset, SwAligner via oracle/_ref harnesses)
on a bounded sample.
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
PAIRED = False            # --mode pairs: reads are 2 x RDLEN pairs (mate 1, mate 2 interleaved), -I 0 -X 500 --fr
MAXFRAG, MINFRAG, MAXHALF = 500, 0, 15
PE_FLAGS = (0, 0, 0, 1, 1, 1)   # local, flip, dovetail, contain, overlap, expand-to-fit: the reference defaults (bt2_search.cpp:381-385)
METRIC = "aligned reads/sec (2x150 end-to-end)"
UNIT = "reads/s"


def workload_name(batch, gmbp):
    if PAIRED:
        return ("synthetic 2x%d bp --fr pairs (simulate.pl model, fragments N(250,45) >= 2x%d) vs FM index of a random %d Mbp "
                "genome, -I %d -X %d: every mate through exactSweep + exact seeds (-L %d, S,%g,%g) + extend + locate + one "
                "seed-extension DP window (%d x %d, end-to-end u8: fill + gather + nextAlignment loop); then, per anchor "
                "alignment, otherMate + frameFindMateRect + the opposite mate's DP in the mate window (%d x ~%d) + "
                "peClassifyPair; %d reads (%d pairs) per step per GPU"
                % (RDLEN, RDLEN, gmbp, MINFRAG, MAXFRAG, SEED_LEN, IVAL[0], IVAL[1], RDLEN, RDLEN + 4 * MAXGAP, RDLEN,
                   MAXFRAG + RDLEN + 56, batch, batch // 2))
    return ("synthetic %d bp reads (simulate.pl model) vs FM index of a random %d Mbp genome: exactSweep + "
            "exact seeds (-L %d, S,%g,%g) + extend + locate + one seed-extension DP window (%d x %d, "
            "%s: fill + gather + nextAlignment loop) per read, %d reads per step per GPU"
            % (RDLEN, gmbp, SEED_LEN, IVAL[0], IVAL[1], RDLEN, RDLEN + 4 * MAXGAP,
               "local u8 with i16 re-runs" if LOCAL else "end-to-end u8", batch))


def make_genome(seed, gmbp):
    rng = np.random.default_rng(seed)
    return rng.integers(0, 4, size=gmbp << 20, dtype=np.uint8)


def get_index(genome, seed, gmbp, need_files):
    """Build (or reload from /tmp) the index of `genome`; returns (Index, base path or None)."""
    from bt2gpu import build, index
    base = os.path.join(tempfile.gettempdir(), "bt2gpu_bench_idx_%d_%d" % (gmbp, seed))
    if index.exists(base) and os.path.exists(base + ".ok"):
        return index.Index.load(base), base
    ix, extras = build.build_index([genome], ["chr1"])
    if need_files:
        build.write_index(ix, extras, base)
        open(base + ".ok", "w").write("ok")
        return ix, base
    return ix, None


def make_reads(seed, genome, n):
    from bt2gpu import synth
    rng = np.random.default_rng(seed)
    if PAIRED:
        reads, quals, pos, fw = synth.simulate_pairs(rng, genome, n // 2, RDLEN)
        return np.ascontiguousarray(reads), np.ascontiguousarray(quals), pos, fw
    reads, quals, pos, _ = synth.simulate_reads(rng, genome, n, RDLEN)
    # strand coin-flip as simulate.pl does: half of the reads are given reverse-complemented
    rc = rng.random(n) < 0.5
    comp = np.array([3, 2, 1, 0, 4], dtype=np.uint8)
    reads[rc] = comp[reads[rc][:, ::-1]]
    quals[rc] = quals[rc][:, ::-1]
    return np.ascontiguousarray(reads), np.ascontiguousarray(quals), pos, ~rc


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


# ---------------------------------------------------------------------------------------------
# reference arm: the unmodified reference on the host cores
# ---------------------------------------------------------------------------------------------
def write_fastq(path, reads, quals):
    c2a = np.frombuffer(b"ACGTN", dtype=np.uint8)
    with open(path, "wb") as fh:
        for i in range(len(reads)):
            fh.write(b"@r%d\n" % i + c2a[reads[i]].tobytes() + b"\n+\n" + quals[i].tobytes() + b"\n")


def _dp_problem_file(path, genome, sc, items):
    """items: (oriented read, oriented quals, rect tuple (refl, refr, ..., triml, trimr, corel, corer, maxgap)) per problem
    -> "P" lines for ref_dp_dump (window fetched like SwAligner::initRef: N outside the text, one pad column)."""
    from bt2gpu import synth, structs
    G = len(genome)
    n = len(items)
    jobs = np.zeros(n, dtype=structs.dp_job)
    seqs, quals, wins = [], [], []
    ro = wo = 0
    for k, (rd, qu, rect) in enumerate(items):
        refl, ncol = rect[0], rect[1] - rect[0] + 1
        cols = refl + np.arange(ncol + 1)
        codes = np.where((cols >= 0) & (cols < G), genome[np.clip(cols, 0, G - 1)], 4).astype(np.uint8)
        seqs.append(rd); quals.append(qu); wins.append((1 << codes).astype(np.uint8))
        j = jobs[k]
        j["rd_off"], j["rd_len"], j["rf_off"], j["ncol"] = ro, len(rd), wo, ncol
        j["minsc"], j["nceil"] = sc.score_min.f_int(len(rd)), sc.nceil(len(rd))
        j["refl"], j["triml"], j["corel"], j["corer"] = refl, rect[4], rect[6], rect[7]
        j["fw"], j["enable8"] = 1, 1        # the window is framed for the oriented read; '+' keeps it as given
        ro += len(rd); wo += ncol + 1
    synth.problems_to_plines(np.concatenate(seqs), np.concatenate(quals), np.concatenate(wins), jobs, path, reflen=G)
    return n


def reference_step(base, genome, reads, quals, pos, fw, threads, td):
    """One bounded step through the reference's own code with `threads` threads:
      ref_seed_dump --bench   exactSweep, searchAllSeeds, SwDriver::extend, getOffset for every read; it also reports
                              which reads had an exact seed hit (--hits-out) -- the reads the chain takes to DP;
      ref_dp_dump --bench     SwAligner (initRead, initRef, align, nextAlignment x2) on the seed-extension window of
                              each of those reads, framed on the read's origin;
      --mode pairs            ... and on the mate window (otherMate + frameFindMateRect from that anchor) of the
                              opposite mate, again through ref_dp_dump --bench.
    The windows are framed with oracle/frame_oracle.c (pinned on the reference's DynProgFramer / PairedEndPolicy);
    framing is not timed on either arm's DP kernels.  Returns (seconds, detail)."""
    from bt2gpu import scoring
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers as H
    bin_ = os.path.join(ROOT, "oracle", "_ref")
    fq = os.path.join(td, "reads.fq")
    hits = os.path.join(td, "hits.bin")
    if not os.path.exists(fq):
        write_fastq(fq, reads, quals)
    o1 = json.loads(subprocess.check_output([os.path.join(bin_, "ref_seed_dump"), "--index", base, "--reads", fq,
                                             "--bench", "--threads", str(threads), "--hits-out", hits]).decode())
    pf, mf = os.path.join(td, "p.txt"), os.path.join(td, "m.txt")
    sc = scoring.Scoring(local=False)
    if not os.path.exists(pf):
        hit = np.fromfile(hits, dtype=np.uint8).astype(bool)
        fo = H.FrameOracle()
        G = len(genome)
        comp = np.array([3, 2, 1, 0, 4], dtype=np.uint8)
        scs = sc.to_struct().reshape(-1)[0]
        minsc, nceil = sc.score_min.f_int(RDLEN), sc.nceil(RDLEN)
        rdg = fo.lib.bt2o_max_read_gaps(minsc, RDLEN, int(scs["match_bonus"]), int(scs["rd_gap_open"]), int(scs["rd_gap_ext"]))
        rfg = fo.lib.bt2o_max_ref_gaps(minsc, RDLEN, int(scs["match_bonus"]), int(scs["rf_gap_open"]), int(scs["rf_gap_ext"]))
        pol = fo.policy(3, MAXFRAG, MINFRAG, PE_FLAGS)
        anchors, mates = [], []
        for i in np.nonzero(hit)[0]:
            rd, qu = (reads[i], quals[i]) if fw[i] else (comp[reads[i][::-1]], quals[i][::-1])
            ok, rect = fo.frame_seed_ext(int(pos[i]), RDLEN, G, MAXGAP, MAXGAP, nceil, MAXGAP, 1)
            if ok:
                anchors.append((rd, qu, rect))
            if PAIRED and ok:
                m = i ^ 1
                is1 = int(i % 2 == 0)
                found, oleft, o4, ofw = fo.other_mate(pol, is1, int(fw[i]), int(pos[i]), RDLEN + rdg, G, RDLEN, RDLEN)
                if found:
                    ok2, mrect = fo.frame_mate(int(not oleft), *o4, RDLEN, G, rdg, rfg, nceil, MAXHALF, 1)
                    if ok2:
                        mrd, mqu = (reads[m], quals[m]) if ofw else (comp[reads[m][::-1]], quals[m][::-1])
                        mates.append((mrd, mqu, mrect))
        _dp_problem_file(pf, genome, sc, anchors)
        if PAIRED:
            _dp_problem_file(mf, genome, sc, mates)
    o2 = json.loads(subprocess.check_output([os.path.join(bin_, "ref_dp_dump"), "--problems", pf, "--bench",
                                             "--threads", str(threads)]).decode())
    det = {"seed_s": o1["seconds"], "dp_s": o2["seconds"], "lf_steps_per_s": o1["lf_steps_per_s"],
           "dp_problems": o2.get("problems"), "gcups": o2["cells"] / o2["seconds"] / 1e9}
    tot = o1["seconds"] + o2["seconds"]
    if PAIRED:
        o3 = json.loads(subprocess.check_output([os.path.join(bin_, "ref_dp_dump"), "--problems", mf, "--bench",
                                                 "--threads", str(threads)]).decode())
        det.update({"mate_dp_s": o3["seconds"], "mate_dp_problems": o3.get("problems"), "mate_gcups": o3["cells"] / o3["seconds"] / 1e9})
        tot += o3["seconds"]
    return tot, det


def run_reference(args, rank, world):
    if rank != 0:
        return
    ncores = os.cpu_count() or 1
    sample = min(args.batch, args.cpu_sample)
    genome = make_genome(args.seed, args.genome_mbp)
    ix, base = get_index(genome, args.seed, args.genome_mbp, True)
    reads, quals, pos, fw = make_reads(args.seed + 7, genome, sample)
    with tempfile.TemporaryDirectory() as td:
        for _ in range(args.warmup):
            reference_step(base, genome, reads, quals, pos, fw, ncores, td)
        t, det = 0.0, None
        for _ in range(args.steps):
            s, det = reference_step(base, genome, reads, quals, pos, fw, ncores, td)
            t += s
    val = sample * args.steps / t
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/i16/u32",
            "data": "synthetic", "config": {"workload": workload_name(args.batch, args.genome_mbp),
                                            "sample_per_step": sample, "detail": det},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": ncores, "kind": "reference",
                             "sample": "%d reads per step through the unmodified reference (oracle/_ref/ref_seed_dump "
                                       "--bench + ref_dp_dump --bench, %d threads)" % (sample, ncores)},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from bt2gpu import lib, scoring
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device and the product has no CPU path")
    torch.cuda.set_device(local_rank)
    genome = make_genome(args.seed, args.genome_mbp)          # same genome / index on every rank (replicated)
    if world > 1 and rank != 0:
        dist.barrier()                                        # rank 0 builds and writes the files first
    ix, base = get_index(genome, args.seed, args.genome_mbp, world > 1)
    if world > 1 and rank == 0:
        dist.barrier()
    torch.cuda.empty_cache()
    sc = scoring.Scoring(local=LOCAL)
    dix = lib.DeviceIndex(ix, local_rank)
    dref = lib.DeviceRef(genome, local_rank)                  # 2-bit text in HBM (windows are cut on the device)
    # C worker contexts per GPU, each with its own stream and arenas and a contiguous shard of the batch -- the
    # reference's own model (one SeedAligner / SwAligner per worker thread, bt2_search.cpp:3128) -- so that one
    # shard's latency-bound FM and backtrace kernels and its PCIe copies overlap the other's E/F/H stream
    C = max(1, args.contexts)
    ctxs = [lib.Context(local_rank) for _ in range(C)]
    for c_ in ctxs:
        c_.set_scoring(sc)
        c_.bind_index(dix)
    ctx = ctxs[0]
    reads, quals, pos, fw = make_reads(args.seed + 7 + 1000 * rank, genome, args.batch)   # each rank its own shard
    n = len(reads)

    def pinned(a):
        t_ = torch.empty(a.nbytes, dtype=torch.uint8, pin_memory=True)
        v = t_.numpy().view(a.dtype).reshape(a.shape)
        v[...] = a
        return t_, v
    keep = []

    def pin(a):
        t_, v = pinned(a)
        keep.append(t_)
        return v
    from bt2gpu import structs
    from concurrent.futures import ThreadPoolExecutor
    hseq, hqual = pin(reads.reshape(-1)), pin(quals.reshape(-1))
    max_aln = 2

    def make_io(lo, hi):
        m = hi - lo
        rd = np.zeros(m, dtype=structs.read)
        rd["rd_off"] = np.arange(m, dtype=np.int64) * RDLEN
        rd["rd_len"] = RDLEN
        return dict(seq=hseq[lo * RDLEN:hi * RDLEN], qual=hqual[lo * RDLEN:hi * RDLEN], rd=rd, lo=lo,
                    minsc=np.full(m, max(1, sc.score_min.f_int(RDLEN)) if LOCAL else sc.score_min.f_int(RDLEN), dtype=np.int32),
                    nceil=np.full(m, sc.nceil(RDLEN), dtype=np.int32),
                    out=dict(sweep=pin(np.zeros(m, dtype=structs.sweep_result)), job_read=np.zeros(m, dtype=np.uint32),
                             job_start=np.zeros(m, dtype=np.int64), job_fw=np.zeros(m, dtype=np.uint8),
                             results=pin(np.zeros(m, dtype=structs.dp_result)),
                             alns=pin(np.zeros(m * max_aln, dtype=structs.dp_aln)),
                             edits=pin(np.zeros(64 * m + 1024, dtype=structs.edit)),
                             **(dict(mates=pin(np.zeros(m, dtype=structs.mate_info)), mresults=pin(np.zeros(m, dtype=structs.dp_result)),
                                     malns=pin(np.zeros(m * max_aln, dtype=structs.dp_aln)),
                                     medits=pin(np.zeros(64 * m + 1024, dtype=structs.edit))) if PAIRED else {})))
    bounds = [n * i // C for i in range(C + 1)]
    shards = [make_io(bounds[i], bounds[i + 1]) for i in range(C)]
    pool = ThreadPoolExecutor(max_workers=C)

    pe_pol = np.zeros(1, dtype=structs.pe_policy)
    pe_pol["pol"], pe_pol["trim_to_ref"], pe_pol["maxfrag"], pe_pol["minfrag"] = 3, 1, MAXFRAG, MINFRAG
    for k_, v_ in zip(("local", "flip_ok", "dovetail_ok", "contain_ok", "olap_ok", "expand_to_fit"), PE_FLAGS):
        pe_pol[k_] = v_

    def run(c_, io):
        if PAIRED:
            o_, st_, ps_ = c_.hotpath_pairs(dref, io["seq"], io["qual"], io["rd"], io["minsc"], io["nceil"], pe_pol, maxhalf=MAXHALF,
                                            seed_len=SEED_LEN, ival=IVAL, maxgap=MAXGAP, max_aln=max_aln, mine_max=2, out=io["out"])
            st_ = dict(st_)
            st_.update({("mate_" + k_ if not k_.startswith("mate_") else k_): v_ for k_, v_ in ps_.items()})
            st_["ms_anchor_total"] = st_["ms_total"]
            st_["ms_total"] = st_["ms_total"] + ps_["ms_total"]          # device span of both passes
            st_["d2h_bytes"] += ps_["d2h_bytes"]
            return o_, st_
        return c_.hotpath_round0(dref, io["seq"], io["qual"], io["rd"], io["minsc"], io["nceil"], seed_len=SEED_LEN,
                                 ival=IVAL, maxgap=MAXGAP, max_aln=max_aln, mine_max=2, nthreads=0, out=io["out"])

    def step():
        if C == 1:
            return [run(ctxs[0], shards[0])]
        return list(pool.map(run, ctxs, shards))       # ctypes drops the GIL: the C contexts run concurrently

    def barrier():
        if world > 1:
            dist.barrier()
        for c_ in ctxs:
            c_.sync()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = sum(c_.launches for c_ in ctxs)
    dev_ms = 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res = step()
        # device span of the step: the longest of the contexts' spans (first kernel to last, reads already in HBM,
        # glue kernels and scalar read-backs included; the contexts run side by side)
        dev_ms += max(r[1]["ms_total"] for r in res)
    barrier()
    wall_e2e = time.perf_counter() - t0
    launches = sum(c_.launches for c_ in ctxs) - l0
    sampler.stop_flag = True
    sampler.join()
    h2d = sum(r[1]["h2d_bytes"] for r in res)
    d2h = sum(r[1]["d2h_bytes"] for r in res)
    # per-kernel times for the rooflines: the same batch through ONE context, every kernel alone on the device
    # (outside the timed region)
    solo = make_io(0, n) if C > 1 else shards[0]
    for _ in range(2):
        run(ctx, solo)
    kms = {}
    KS = 3
    for _ in range(KS):
        o, cnt = run(ctx, solo)
        for k in ("sweep", "seed", "extend", "locate", "dp", "dp_fill", "dp_gather", "dp_backtrace", "glue", "total"):
            kms[k] = kms.get(k, 0.0) + cnt["ms_" + k]
        if PAIRED:
            for k in ("frame", "dp", "dp_fill", "dp_gather", "dp_backtrace", "total"):
                kms["mate_" + k] = kms.get("mate_" + k, 0.0) + cnt["mate_ms_" + k]
    # sanity: the located diagonal of a read is its simulated origin (indel shifts aside)
    nj = o["njobs"]
    ok = np.abs(o["job_start"][:nj] - pos[o["job_read"][:nj]]) <= 64
    frac_true = float(ok.mean()) if nj else 0.0
    cnt = {k: (int(v) if not k.startswith(("ms_", "wall_", "mate_ms_")) else float(v)) for k, v in cnt.items()}
    probe_ix = ctx.gather_probe("fw", 1 << 27)                # random sides of the index itself (L2 helps when it fits)
    probe = ctx.probe_random_gather(8 << 30, 1 << 27)         # random 64-B gathers over 8 GiB: the HBM random-access ceiling
    dpx_peak = ctx.probe_dpx()                                # G thread-instr/s of packed s16x2 DPX

    t = torch.tensor([dev_ms, wall_e2e], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, wall_e2e = [float(x) for x in t.tolist()]
    if rank != 0:
        return
    total_reads = n * world * args.steps
    value = total_reads / (dev_ms / 1e3)
    e2e = total_reads / wall_e2e
    peaks = {}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    K = KS            # the per-kernel times come from the KS solo steps

    def rl(kernel, alg_bytes, ms, extra=None):
        a = alg_bytes / (ms / 1e3) / 1e9 if ms > 0 else 0.0
        d = {"bound": "hbm", "kernel": kernel, "achieved": a, "peak": hbm, "unit": "GB/s", "frac": a / hbm,
             "traffic": None, "ms": ms}
        d.update(extra or {})
        return d
    # algorithmic bytes per launch: 64 B per distinct BWT side + 16 B of ftab per seed;
    # 64 B per LF step (+4 B offs) in locate; 64 B per side in sweep/extend (bwops counts sides);
    # E, F, H of every DP cell (3 B/cell in the u8 path) for the fill
    def fm(kernel, sides, extra_bytes, ms, steps=None):
        d = rl(kernel, 64.0 * sides + extra_bytes, ms)
        d.update({"random_gather_peak_gbs": probe, "index_gather_gbs": probe_ix,
                  "frac_of_random_gather": (d["achieved"] / probe) if probe else 0.0})
        if steps is not None and ms > 0:
            d["lf_steps_per_s"] = steps / (ms / 1e3)
        return d
    fill_ms = kms["dp_fill"] / K
    dpx = 2.5 * cnt["dp_cells"] / (fill_ms / 1e3) / 1e9 if fill_ms > 0 else 0.0   # 5 DPX instr per packed pair of cells
    rls = [
        fm("k_seed_exact", cnt["seed_sides"], 16.0 * cnt["seeds"], kms["seed"] / K, cnt["seed_bwops"]),
        fm("k_exact_sweep", cnt["sweep_bwops"], 0.0, kms["sweep"] / K, cnt["sweep_bwops"]),
        fm("k_extend", cnt["extend_fmops"], 0.0, kms["extend"] / K, cnt["extend_fmops"]),
        fm("k_locate", cnt["locate_steps"], 4.0 * cnt["locate_steps"] + 4.0 * cnt["locate_rows"], kms["locate"] / K, cnt["locate_steps"]),
        rl("k_dp_fill", (6.0 if LOCAL else 3.0) * cnt["dp_cells"], fill_ms,
           {"gcups": cnt["dp_cells"] / (fill_ms / 1e3) / 1e9 if fill_ms > 0 else 0.0,
            "dpx_ginstr_per_s": dpx, "dpx_peak_ginstr_per_s": dpx_peak, "dpx_frac": dpx / dpx_peak if dpx_peak else 0.0}),
        rl("k_dp_backtrace", 3.0 * cnt["bt_cells"] + 12.0 * cnt["dp_cands"], kms["dp_backtrace"] / K,
           {"note": "dependent walk over the stored matrices (E,F,H of each visited cell + the candidate list): "
                    "latency/issue-bound, a byte roofline does not describe it",
            "walked_cells_per_job": cnt["bt_cells"] / max(cnt["dp_jobs"], 1), "candidates_per_job": cnt["dp_cands"] / max(cnt["dp_jobs"], 1)}),
        rl("k_dp_gather_e2e", 1.0 * cnt["dp_jobs"] * (RDLEN + 4 * MAXGAP), kms["dp_gather"] / K),
    ]
    if PAIRED:          # the same DP kernels on the mate windows (second launch of each per step)
        mfill = kms["mate_dp_fill"] / K
        mdpx = 2.5 * cnt["mate_cells"] / (mfill / 1e3) / 1e9 if mfill > 0 else 0.0
        rls += [
            rl("k_dp_fill[mate windows]", 3.0 * cnt["mate_cells"], mfill,
               {"gcups": cnt["mate_cells"] / (mfill / 1e3) / 1e9 if mfill > 0 else 0.0, "dpx_ginstr_per_s": mdpx,
                "dpx_peak_ginstr_per_s": dpx_peak, "dpx_frac": mdpx / dpx_peak if dpx_peak else 0.0}),
            rl("k_dp_backtrace[mate windows]", 3.0 * cnt["mate_bt_cells"] + 12.0 * cnt["mate_dp_cands"], kms["mate_dp_backtrace"] / K,
               {"walked_cells_per_job": cnt["mate_bt_cells"] / max(cnt["mate_jobs"], 1),
                "candidates_per_job": cnt["mate_dp_cands"] / max(cnt["mate_jobs"], 1)}),
            rl("k_dp_gather_e2e[mate windows]", 4.0 * cnt["mate_cells"] / RDLEN, kms["mate_dp_gather"] / K),
        ]
    # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` captures of this
    # command at its default batch and genome (profiles/r1r_summary.md for --mode pairs, profiles/r1z_summary.md for
    # --mode e2e, whose backtrace capture predates the diagonal-first move test: not restated); other workloads: not captured
    if args.batch == 262144 and args.genome_mbp == 1024 and not LOCAL:
        if PAIRED:
            traffic = {"k_exact_sweep": 0.453e9, "k_seed_exact": 0.713e9, "k_extend": 0.532e9, "k_locate": 0.162e9,
                       "k_dp_fill": 10.624e9, "k_dp_gather_e2e": 0.112e9, "k_dp_backtrace": 4.885e9,
                       "k_dp_fill[mate windows]": 33.921e9, "k_dp_gather_e2e[mate windows]": 0.304e9,
                       "k_dp_backtrace[mate windows]": 5.493e9}
        else:
            traffic = {"k_exact_sweep": 0.409e9, "k_seed_exact": 0.710e9, "k_extend": 0.526e9, "k_locate": 0.159e9,
                       "k_dp_fill": 10.582e9, "k_dp_gather_e2e": 0.112e9, "k_dp_backtrace": None}
        for r_ in rls:
            r_["traffic"] = traffic.get(r_["kernel"])
    dominant = max(rls, key=lambda r: r["ms"])
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8/i16/u32", "data": "synthetic",
        "config": {"workload": workload_name(args.batch, args.genome_mbp),
                   "l2": "index %.0f MB and the per-step E/F/H stream (%.1f GB) both exceed the 126 MB L2" % (dix.nbytes / 1e6, 3.3 * cnt["dp_cells"] / 1e9),
                   "contexts_per_gpu": C,
                   "timing": "value: per step the longest device span among the %d worker contexts of a GPU, which run "
                             "side by side on their own streams; phase_ms / rooflines: the same batch through one "
                             "context, every kernel alone on the device (outside the timed region)" % C,
                   "index_bytes": dix.nbytes, "counters": cnt, "located_at_origin_frac": frac_true,
                   "peak_source": "measured" if peaks else "fallback"},
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
        "gpu_launches": int(launches),
        "roofline": dominant, "rooflines": rls,
        "phase_ms": {k: v / K for k, v in kms.items()},
        "clocks": sampler.summary(),
    }
    # CPU baseline on rank 0 at N=1: the unmodified reference on the host cores, bounded sample
    if world == 1 and not args.no_cpu_baseline:
        ncores = os.cpu_count() or 1
        sample = min(n, args.cpu_sample)
        if base is None:
            from bt2gpu import build
            base = os.path.join(tempfile.gettempdir(), "bt2gpu_bench_idx_%d_%d" % (args.genome_mbp, args.seed))
            ix2, extras = ix, {"recs": [(0, len(genome), 1)], "joined": genome}
            build.write_index(ix2, extras, base)
            open(base + ".ok", "w").write("ok")
        try:
            with tempfile.TemporaryDirectory() as td:
                s, det = reference_step(base, genome, reads[:sample], quals[:sample], pos[:sample], fw[:sample], ncores, td)
            line["cpu_baseline"] = {"value": sample / s, "unit": UNIT, "cores": ncores, "kind": "reference", "detail": det,
                                    "sample": "%d reads through the unmodified reference (oracle/_ref/ref_seed_dump --bench + "
                                              "ref_dp_dump --bench, %d threads)" % (sample, ncores)}
        except Exception as e:   # oracle/_ref missing on this box
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "unavailable: %r" % (e,)}
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
    ap.add_argument("--batch", type=int, default=262144)
    ap.add_argument("--contexts", type=int, default=1, help="worker contexts (streams) per GPU; more than one did not pay (DESIGN.md 5)")
    ap.add_argument("--genome-mbp", type=int, default=1024)
    ap.add_argument("--mode", default="pairs", choices=["pairs", "e2e", "local"],
                    help="pairs: 2x150 --fr pairs, anchor pass + mate-window pass (the metric's own shape); e2e: unpaired 150 bp reads, anchor pass only; local: 250 bp reads, --sensitive-local seeds (-L 20, S,1,0.75), u8 local DP with i16 re-runs (config C4); our arm only")
    ap.add_argument("--cpu-sample", type=int, default=40000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--seed", type=int, default=20261104)
    args = ap.parse_args()
    global RDLEN, SEED_LEN, IVAL, LOCAL, PAIRED
    if args.mode == "pairs":
        PAIRED = True
        args.batch -= args.batch % 2
    if args.mode == "local":
        RDLEN, SEED_LEN, IVAL, LOCAL = 250, 20, (1.0, 0.75), True
        args.no_cpu_baseline = True
        if args.impl == "reference":
            emit({"impl": "reference", "unavailable": "the reference arm times the end-to-end workload only"})
            return
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
