"""TEST INFRASTRUCTURE.  Golden vectors for SwAligner::ungappedAlign (aligner_sw.cpp:286-494): problems simulated on
tests/golden/rep.fa (three sequences, an N gap) are run through the UNMODIFIED reference (oracle/_ref/ref_dp_dump
--ungapped, built by oracle/Makefile from /root/reference) in end-to-end and local mode; the dumps are committed as
tests/golden/ungapped_{ee,loc}.txt.  Run where /root/reference is mounted."""
import math
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref")
GOLD = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, os.path.join(ROOT, "tests"))
import dropin as D  # noqa: E402


def problems(rng, chroms, local, n):
    comp = np.array([3, 2, 1, 0, 4], dtype=np.uint8)
    out = []
    for k in range(n):
        ci = int(rng.integers(0, len(chroms)))
        c = chroms[ci]
        L = int(rng.choice([20, 33, 50, 76, 100, 150, 251]))
        kind = k % 8
        if kind == 0:      # hanging off the left end
            off = -int(rng.integers(1, 12))
        elif kind == 1:    # hanging off the right end
            off = len(c) - L + int(rng.integers(1, 12))
        elif kind == 2 and ci == 1:   # across the N gap of rep1
            npos = np.nonzero(c == 4)[0]
            off = int(npos[0]) - int(rng.integers(0, L))
        else:
            off = int(rng.integers(0, len(c) - L))
        idx = np.clip(np.arange(off, off + L), 0, len(c) - 1)
        rd = c[idx].copy()
        rd[rd > 3] = rng.integers(0, 4, size=int((rd > 3).sum()))
        nmm = int(rng.choice([0, 0, 1, 2, 4, 8, 20]))
        for p in rng.choice(L, size=min(nmm, L), replace=False):
            rd[p] = (rd[p] + rng.integers(1, 4)) % 4
        if rng.random() < 0.25:
            rd[rng.integers(0, L)] = 4
        if local and rng.random() < 0.4:      # two good stretches split by junk: >1 solution / trimming
            a = int(rng.integers(L // 6, L // 3 + 1)); b = min(L, 2 * a)
            rd[a:b] = (rd[a:b] + 1 + rng.integers(0, 3, size=b - a)) % 4
        q = rng.integers(2, 41, size=L) + 33
        fw = bool(rng.random() < 0.5)
        shift = int(rng.choice([0, 0, 0, 0, 1, -1]))
        if local:
            minsc = int(max(20 + 8.0 * math.log(L), 0)) + int(rng.choice([0, 0, 10, -10]))
        else:
            minsc = int(min(-0.6 - 0.6 * L, 0)) + int(rng.choice([0, 0, 20, -30]))
        out.append("U\t%s\t%s\t%s\t%d\t%d\t%d\t%d" % ("".join("ACGTN"[x] for x in rd), bytes(q.astype(np.uint8)).decode(),
                                                    "+" if fw else "-", ci, off + shift, int(rng.random() < 0.7), minsc))
    return out


def main():
    work = tempfile.mkdtemp()
    base = os.path.join(work, "rep")
    D.build_index(os.path.join(GOLD, "rep.fa"), base)
    chroms = D.genome_codes("rep.fa")
    for tag, local, seed in (("ee", False, 21), ("loc", True, 22)):
        rng = np.random.default_rng(seed)
        inp = os.path.join(work, tag + ".in")
        with open(inp, "w") as fh:
            fh.write("\n".join(problems(rng, chroms, local, 600)) + "\n")
        with open(os.path.join(GOLD, "ungapped_%s.txt" % tag), "w") as out:
            subprocess.check_call([os.path.join(BIN, "ref_dp_dump"), "--ungapped", inp, "--index", base]
                                  + (["--local"] if local else []), stdout=out)


if __name__ == "__main__":
    main()
