"""Sweep of the batched server's scheduling knobs on the index bench.py left in /tmp (measurement helper, GPU box):
host threads x worker fibers per thread x lanes x clients -> reads/s, one JSON line per point.
  python scripts/server_sweep.py --genome-mbp 3100 --nseq 24 --batch 262144 --points 8:128:2:8,16:128:2:8"""
import argparse
import json
import os
import re
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "bowtie2-server_b200"))
import bench  # noqa: E402
from bt2gpu import serverbench as SB  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genome-mbp", type=int, default=3100)
    ap.add_argument("--nseq", type=int, default=24)
    ap.add_argument("--batch", type=int, default=262144)
    ap.add_argument("--seed", type=int, default=20261104)
    ap.add_argument("--mode", default="pairs")
    ap.add_argument("--points", default="16:128:2:8")
    ap.add_argument("--stock", action="store_true", help="also time the stock server (-p all cores) per client count")
    ap.add_argument("--passes", type=int, default=2)
    ap.add_argument("--pass-timeout", type=int, default=90, help="seconds a single pass may take before the point is abandoned")
    ap.add_argument("--devices", default="", help="BT2GPU_DEVICES of every point (one server process driving several GPUs), e.g. 0,1")
    a = ap.parse_args()
    if a.mode == "pairs":
        bench.PAIRED = True
        bench.SERVER_ARGS = ["-I", "0", "-X", "500"]
    elif a.mode == "local":
        bench.RDLEN, bench.LOCAL, bench.SERVER_ARGS = 250, True, ["--local"]
    else:
        bench.SERVER_ARGS = ["--very-sensitive"]
    base, seqs, _ = bench.get_index(a)
    reads, quals = bench.make_reads(a.seed + 7, seqs, a.batch)
    cores = bench.host_cores()
    done_stock = set()
    with tempfile.TemporaryDirectory() as td:
        for pt in a.points.split(","):
            f = pt.split(":")                      # th:wpt:lanes:clients[:KEY=VALUE[:KEY=VALUE...]] (extra environment of the server)
            th, wpt, lanes, ncl = (int(x) for x in f[:4])
            extra = dict(x.split("=", 1) for x in f[4:])
            shards = bench.write_shards(td, "c%d" % ncl, reads, quals, ncl)
            if a.stock and ncl not in done_stock:
                done_stock.add(ncl)
                t, st, rss, _, _ = bench.server_steps(SB.CPU_SERVER, base, cores, None, shards, td, "stock%d" % ncl, 1, a.passes, pass_timeout=a.pass_timeout)
                print(json.dumps({"server": "stock", "threads": cores, "clients": ncl, "reads_per_s": len(reads) * a.passes / t,
                                  "start_s": st, "rss_mb": rss}), flush=True)
            env = {"BT2GPU_THREADS": str(th), "BT2GPU_LANES": str(lanes)}
            if a.devices:
                env["BT2GPU_DEVICES"] = a.devices
            env.update(extra)
            try:
                t, st, rss, _, cnt = bench.server_steps(SB.GPUB_SERVER, base, th * wpt, env, shards, td, "g_" + re.sub(r"[^A-Za-z0-9]+", "_", pt), 1, a.passes,
                                                        pass_timeout=a.pass_timeout)
            except Exception as e:      # a point that hangs or fails must not take the rest of the sweep (and the GPU lease) with it
                print(json.dumps({"server": "gpub", "point": pt, "failed": str(e).splitlines()[0][:200]}), flush=True)
                continue
            jobs = (cnt or {}).get("abi_batch_jobs", 0) / max((cnt or {}).get("abi_batches", 1), 1)
            print(json.dumps({"server": "gpub", "host_threads": th, "workers": th * wpt, "lanes": lanes, "clients": ncl, "env": extra, "devices": a.devices or "0",
                              "reads_per_s": len(reads) * a.passes / t, "start_s": st, "rss_mb": rss, "jobs_per_abi_call": jobs,
                              "counters": cnt}), flush=True)


if __name__ == "__main__":
    main()
