import sys, os, time, cProfile, pstats
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "bowtie2-server_b200"))
import numpy as np
import bench
from bt2gpu import lib, scoring, pipeline
g = bench.make_genome(1, 64)
ix, _ = bench.get_index(g, 1, 64, False)
ctx = lib.Context(0); dix = lib.DeviceIndex(ix, 0)
hp = pipeline.HotPath(ctx, dix, g, scoring.Scoring(local=False))
reads, quals, pos, fw = bench.make_reads(3, g, 65536)
for _ in range(2): hp.step(reads, quals)
t=time.perf_counter(); hp.step(reads, quals); print("step s", time.perf_counter()-t)
pr = cProfile.Profile(); pr.enable(); hp.step(reads, quals); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
