#!/usr/bin/env python
"""Latency of the per-kind C-ABI batch calls as the batched server makes them (host buffers in, host buffers out), by
batch size and by the number of contexts calling at once.  What one round trip of a worker costs, and how many
problems a call must carry before the GPU rather than the call is the limit.

    python scripts/abi_latency.py [--genome-mbp 64] [--threads 1,8,32] [--blocking 0,1]

One JSON line per (kind, problems per call, concurrent contexts, blocking sync): ms per call (median over the
contexts), problems/s over all contexts."""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bowtie2-server_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genome-mbp", type=int, default=64)
    ap.add_argument("--sizes", default="1,8,64,512,4096")
    ap.add_argument("--threads", default="1,8,32")
    ap.add_argument("--blocking", default="1")
    ap.add_argument("--kinds", default="dp,dpmate,sweep,seeds,locate")
    ap.add_argument("--reps", type=int, default=12)
    args = ap.parse_args()
    import ctypes as C
    from bt2gpu import build, fmio, lib, scoring, structs, synth
    rng = np.random.default_rng(5)
    genome = rng.integers(0, 4, size=args.genome_mbp * 1000000, dtype=np.uint8)
    ix, extras = build.build_index([genome], ["chr1"])
    dix = lib.DeviceIndex(ix, 0)
    sc = scoring.Scoring(local=False)
    sizes = [int(x) for x in args.sizes.split(",")]
    nmax = max(sizes)
    rdlen = 150
    seq, qual, win, jobs = synth.dp_extend_batch(rng, genome, nmax, rdlen, sc)[:4]
    # mate-finding windows: 150 x ~700
    ncol = 700
    reads_m, quals_m, pos_m, _ = synth.simulate_reads(rng, genome, nmax, rdlen)
    refl = pos_m - 300
    cols = refl[:, None] + np.arange(ncol + 1)[None, :]
    codes = np.where((cols >= 0) & (cols < len(genome)), genome[np.clip(cols, 0, len(genome) - 1)], 4).astype(np.uint8)
    win_m = (1 << codes).astype(np.uint8).reshape(-1)
    jobs_m = np.zeros(nmax, dtype=structs.dp_job)
    jobs_m["rd_off"] = np.arange(nmax, dtype=np.int64) * rdlen
    jobs_m["rd_len"] = rdlen
    jobs_m["rf_off"] = np.arange(nmax, dtype=np.int64) * (ncol + 1)
    jobs_m["ncol"] = ncol
    jobs_m["minsc"] = sc.score_min.f_int(rdlen)
    jobs_m["nceil"] = sc.nceil(rdlen)
    jobs_m["refl"] = refl
    jobs_m["corel"] = 0
    jobs_m["corer"] = ncol - 1
    jobs_m["fw"] = 1
    jobs_m["enable8"] = 1
    seq_m, qual_m = reads_m.reshape(-1), quals_m.reshape(-1)
    pools = {n: fmio.ReadPool(list(reads_m[:n]), list(quals_m[:n])) for n in sizes}
    rows = rng.integers(0, len(genome), size=nmax).astype(np.uint32)

    def seed_jobs(n):
        per = 16
        j = np.zeros(n * per, dtype=structs.seed_job)
        j["read"] = np.repeat(np.arange(n), per)
        j["off"] = np.tile(np.arange(per) * 8, n)
        j["len"] = 22
        j["fw"] = np.tile(np.arange(per) % 2, n)
        return j

    def one(ctx, kind, n):
        if kind == "dp":
            e = int(jobs["rd_off"][n - 1] + rdlen)
            w = int(jobs["rf_off"][n - 1] + jobs["ncol"][n - 1] + 1)
            ctx.dp_align_batch(seq[:e], qual[:e], win[:w], jobs[:n], max_aln=16, cand_cap=n * 300 + 64, edit_cap=n * 856 + 1024)
        elif kind == "dpmate":
            ctx.dp_align_batch(seq_m[:n * rdlen], qual_m[:n * rdlen], win_m[:n * (ncol + 1)], jobs_m[:n], max_aln=16,
                               cand_cap=n * 800 + 64, edit_cap=n * 856 + 1024)
        elif kind == "sweep":
            ctx.stage_reads(pools[n])
            ctx.exact_sweep()
        elif kind == "seeds":
            ctx.stage_reads(pools[n])
            ctx.seed_exact(sjobs[n])
        elif kind == "locate":
            ctx.locate(rows[:n])

    sjobs = {n: seed_jobs(n) for n in sizes}
    nthreads_list = [int(x) for x in args.threads.split(",")]
    ctxs = []
    for _ in range(max(nthreads_list)):
        c = lib.Context(0)
        c.set_scoring(sc)
        c.bind_index(dix)
        ctxs.append(c)
    for blocking in [int(x) for x in args.blocking.split(",")]:
        if ctxs[0].lib.bt2gpu_device_blocking_sync(C.c_int(0), C.c_int(blocking)):
            raise SystemExit("bt2gpu_device_blocking_sync failed")
        for kind in args.kinds.split(","):
            for n in sizes:
                for T in nthreads_list:
                    if T > 1 and n > 512:
                        continue
                    per = [None] * T

                    def work(t):
                        c = ctxs[t]
                        for _ in range(3):
                            one(c, kind, n)
                        t0 = time.perf_counter()
                        for _ in range(args.reps):
                            one(c, kind, n)
                        per[t] = (time.perf_counter() - t0) / args.reps
                    ths = [threading.Thread(target=work, args=(t,)) for t in range(T)]
                    t0 = time.perf_counter()
                    for th in ths:
                        th.start()
                    for th in ths:
                        th.join()
                    ms = 1e3 * float(np.median(per))
                    print(json.dumps({"kind": kind, "problems_per_call": n, "contexts": T, "blocking_sync": blocking,
                                      "ms_per_call": round(ms, 3), "problems_per_s": round(T * n / (ms / 1e3))}), flush=True)
    for c in ctxs:
        c.close()


if __name__ == "__main__":
    main()
