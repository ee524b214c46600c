"""Scratch: backtrace work counters from a -DBT_STATS build of the library (scripts/libbt2gpu_stats.so)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "bowtie2-server_b200"))
from bt2gpu import lib
if os.environ.get("BT_STATS_LIB", "1") == "1":
    lib.LIB_PATH = os.path.join(ROOT, "scripts", "libbt2gpu_stats.so")
import numpy as np, bench
from bt2gpu import scoring, structs
g = bench.make_genome(1, 64)
ix, _ = bench.get_index(g, 1, 64, False)
ctx = lib.Context(0); dix = lib.DeviceIndex(ix, 0); sc = scoring.Scoring(local=False)
ctx.set_scoring(sc); ctx.bind_index(dix); dref = lib.DeviceRef(g, 0)
reads, quals, pos, fw = bench.make_reads(3, g, 65536)
n = len(reads)
rd = np.zeros(n, dtype=structs.read); rd["rd_off"] = np.arange(n) * 150; rd["rd_len"] = 150
minsc = np.full(n, sc.score_min.f_int(150), dtype=np.int32); nceil = np.full(n, sc.nceil(150), dtype=np.int32)
L = lib.load()
out = (C.c_ulonglong * 16)()
ctx.hotpath_round0(dref, np.ascontiguousarray(reads.reshape(-1)), np.ascontiguousarray(quals.reshape(-1)), rd, minsc, nceil)
L.bt2gpu_dbg_btstats(out, 1) if hasattr(L, "bt2gpu_dbg_btstats") and os.environ.get("BT_STATS_LIB", "1") == "1" else None
o, st = ctx.hotpath_round0(dref, np.ascontiguousarray(reads.reshape(-1)), np.ascontiguousarray(quals.reshape(-1)), rd, minsc, nceil)
L.bt2gpu_dbg_btstats(out, 1) if hasattr(L, "bt2gpu_dbg_btstats") and os.environ.get("BT_STATS_LIB", "1") == "1" else None
v = list(out)
names = ["jobs", "cands", "trial groups", "trial iterations", "trial accepted", "trial failed", "H batches", "E/F scalar steps", "full walks", "trial active lanes"]
j = max(v[0], 1)
for nme, x in zip(names, v):
    print("%-20s %10d  per job %.2f" % (nme, x, x / j))
print("backtrace ms", st["ms_dp_backtrace"])
