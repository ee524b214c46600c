"""SAM-level parity run at a larger scale than tests/test_gpu_dropin.py (test infrastructure; needs oracle/_ref):
the stock reference server and the GPU-bound one (oracle/_ref/bowtie2-align-server-s-gpu) on the C3 / C4 / C5 shapes
of BASELINE.json over an i.i.d. random genome.  Prints one JSON line per config.
  python scripts/dropin_parity.py --genome-mbp 32 --reads 20000 --threads 8"""
import argparse
import json
import os
import re
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import dropin as D  # noqa: E402
from bt2gpu import synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genome-mbp", type=float, default=32.0)
    ap.add_argument("--reads", type=int, default=20000)
    ap.add_argument("--threads", type=int, default=8)
    a = ap.parse_args()
    work = tempfile.mkdtemp(prefix="dropin_parity_")
    rng = np.random.default_rng(20261104)
    n = int(a.genome_mbp * 1e6) // 10
    chroms = [synth.random_genome(rng, n) for _ in range(10)]
    fa = os.path.join(work, "rand.fa")
    with open(fa, "w") as fh:
        for i, c in enumerate(chroms):
            fh.write(">chr%d\n" % i)
            s = np.frombuffer(b"ACGT", np.uint8)[c].tobytes().decode()
            fh.write("\n".join(s[k:k + 100000] for k in range(0, len(s), 100000)) + "\n")
    t0 = time.time()
    base = os.path.join(work, "rand")
    D.build_index(fa, base, threads=os.cpu_count() or 1)
    print(json.dumps({"index_build_s": round(time.time() - t0, 1), "genome_bp": 10 * n}), flush=True)
    g = np.concatenate(chroms)
    configs = []
    r, q, _, _ = synth.simulate_reads(rng, g, a.reads, 150)
    f = os.path.join(work, "c3.fq"); D.write_fastq(f, r, q, "c3r")
    configs.append(("C3 150bp --very-sensitive", ["--very-sensitive"], ["-U", f]))
    r, q, _, _ = synth.simulate_reads(rng, g, a.reads // 2, 250)
    f = os.path.join(work, "c4.fq"); D.write_fastq(f, r, q, "c4r")
    configs.append(("C4 250bp --local", ["--local"], ["-U", f]))
    r, q, _, _ = synth.simulate_pairs(rng, g, a.reads // 2, 150)
    f1, f2 = os.path.join(work, "c5_1.fq"), os.path.join(work, "c5_2.fq")
    D.write_fastq(f1, r[0::2], q[0::2], "c5p"); D.write_fastq(f2, r[1::2], q[1::2], "c5p")
    configs.append(("C5 2x150 -I 0 -X 500", ["-I", "0", "-X", "500"], ["-1", f1, "-2", f2]))
    rc = 0
    for k, (name, sargs, cargs) in enumerate(configs):
        t0 = time.time()
        cpu, _ = D.run(D.CPU_SERVER, base, sargs, cargs, os.path.join(work, "c%d_cpu.sam" % k), threads=a.threads, timeout=3000)
        t1 = time.time()
        stats = os.path.join(work, "c%d_stats.txt" % k)
        gpu, _ = D.run(D.GPU_SERVER, base, sargs, cargs, os.path.join(work, "c%d_gpu.sam" % k), threads=a.threads,
                       timeout=3000, env={"BT2GPU_SHIM_STATS": stats})
        t2 = time.time()
        st = dict((x, int(v)) for x, v in re.findall(r"(\w+)\s+(\d+)", open(stats).read()))
        diff = sum(1 for x, y in zip(cpu, gpu) if x != y) + abs(len(cpu) - len(gpu))
        aligned = sum(1 for ln in cpu if ln.split("\t")[2] != "*")
        print(json.dumps({"config": name, "sam_records": len(cpu), "aligned_records": aligned, "differing_records": diff,
                          "stock_reference_s": round(t1 - t0, 1), "gpu_bound_reference_s": round(t2 - t1, 1),
                          "threads": a.threads, "shim_counters": st}), flush=True)
        rc |= diff != 0
    return rc


if __name__ == "__main__":
    sys.exit(main())
