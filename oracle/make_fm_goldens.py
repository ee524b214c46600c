#!/usr/bin/env python3
"""oracle/make_fm_goldens.py -- TEST INFRASTRUCTURE.  Regenerates the seed-side golden vectors
(tests/golden/fm_*.txt, *_index.npz) by running the UNMODIFIED reference (oracle/_ref/
bowtie2-build-s and oracle/_ref/ref_seed_dump, built by oracle/Makefile) on
  * the reference's own lambda example (example/reference/lambda_virus.fa + example/reads), and
  * a small synthetic genome with tandem / dispersed repeats (so BW ranges wider than one row,
    the mapBiLFEx path, ftab/eftab corner cases and multi-row locate are exercised).
Needs /root/reference: runs only in the build container; its outputs are committed.

  python oracle/make_fm_goldens.py
"""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("BT2_REFERENCE", "/root/reference")
BIN = os.path.join(ROOT, "oracle", "_ref")
GOLD = os.path.join(ROOT, "tests", "golden")
IDX = os.path.join(BIN, "idx")
sys.path.insert(0, os.path.join(ROOT, "bowtie2-server_b200"))
from bt2gpu import index  # noqa: E402


def build_index(fa, base):
    subprocess.check_call([os.path.join(BIN, "bowtie2-build-s"), "-q", fa, base],
                          stdout=subprocess.DEVNULL)


def dump(base, reads, out, extra=()):
    with open(out, "w") as fh:
        subprocess.check_call([os.path.join(BIN, "ref_seed_dump"), "--index", base, "--reads", reads] + list(extra),
                              stdout=fh)
    print(out, os.path.getsize(out))


def make_repeat_genome(rng):
    """~90 kbp over 3 sequences: random backbone, a 400-bp block dispersed 12x (with and without
    point mutations), tandem repeats of periods 1..7, and an N gap (-> 2 fragments in one sequence)."""
    def rnd(n):
        return "".join("ACGT"[i] for i in rng.integers(0, 4, size=n))
    block = rnd(400)
    seqs = []
    for s in range(3):
        parts = []
        for k in range(8):
            parts.append(rnd(int(rng.integers(1500, 4000))))
            b = list(block)
            if k % 3 == 1:
                for p in rng.integers(0, 400, size=3):
                    b[p] = "ACGT"[int(rng.integers(0, 4))]
            if k % 2 == 0:
                parts.append("".join(b))
            unit = rnd(int(rng.integers(1, 8)))
            parts.append(unit * int(rng.integers(10, 60)))
        if s == 1:
            parts.insert(5, "N" * 37)
        seqs.append("".join(parts))
    return seqs


def simulate_reads(rng, seqs, n, rdlen_range):
    out = []
    comp = str.maketrans("ACGTN", "TGCAN")
    for i in range(n):
        s = seqs[int(rng.integers(0, len(seqs)))]
        ln = int(rng.integers(*rdlen_range))
        p = int(rng.integers(0, len(s) - ln))
        r = list(s[p:p + ln])
        mode = i % 5
        nmm = [0, 1, 1, 2, 4][mode]
        for q in rng.integers(0, ln, size=nmm):
            r[q] = "ACGT"[int(rng.integers(0, 4))]
        if i % 11 == 0:
            r[int(rng.integers(0, ln))] = "N"
        r = "".join(r)
        if rng.random() < 0.5:
            r = r.translate(comp)[::-1]
        q = "".join(chr(33 + int(x)) for x in rng.integers(2, 41, size=ln))
        out.append("@s%d\n%s\n+\n%s\n" % (i, r, q))
    return "".join(out)


def main():
    os.makedirs(IDX, exist_ok=True)
    lam = os.path.join(IDX, "lambda_virus")
    if not index.exists(lam):
        build_index(os.path.join(REF, "example/reference/lambda_virus.fa"), lam)
    index.Index.load(lam).save_npz(os.path.join(GOLD, "lambda_index.npz"))
    import shutil
    shutil.copyfile(os.path.join(REF, "example/reference/lambda_virus.fa"), os.path.join(GOLD, "lambda.fa"))
    ex = os.path.join(REF, "example/reads")
    dump(lam, os.path.join(ex, "reads_1.fq"), os.path.join(GOLD, "fm_lambda_r1.txt"), ["--limit", "200"])
    dump(lam, os.path.join(ex, "longreads.fq"), os.path.join(GOLD, "fm_lambda_long.txt"), ["--limit", "40"])
    dump(lam, os.path.join(ex, "reads_2.fq"), os.path.join(GOLD, "fm_lambda_r2_local.txt"),
         ["--limit", "100", "--local", "-L", "20", "--interval", "1", "0.75"])

    rng = np.random.default_rng(20261104)
    seqs = make_repeat_genome(rng)
    fa = os.path.join(IDX, "rep.fa")
    with open(fa, "w") as fh:
        for i, s in enumerate(seqs):
            fh.write(">rep%d\n" % i)
            for k in range(0, len(s), 70):
                fh.write(s[k:k + 70] + "\n")
    rep = os.path.join(IDX, "rep")
    build_index(fa, rep)
    index.Index.load(rep).save_npz(os.path.join(GOLD, "rep_index.npz"))
    shutil.copyfile(fa, os.path.join(GOLD, "rep.fa"))     # inputs of tests/test_index_build.py
    fq = os.path.join(IDX, "rep_reads.fq")
    with open(fq, "w") as fh:
        fh.write(simulate_reads(rng, seqs, 240, (30, 160)))
    dump(rep, fq, os.path.join(GOLD, "fm_rep.txt"), ["--maxloc", "24"])
    dump(rep, fq, os.path.join(GOLD, "fm_rep_local.txt"), ["--limit", "120", "--local", "-L", "20", "--interval", "1", "0.75",
                                                          "--maxloc", "4"])
    # multi-zone seeds (-N 1, -N 2): ranges and their reference strings in the order of the reportHit calls
    dump(lam, os.path.join(ex, "reads_1.fq"), os.path.join(GOLD, "fm_mm1_lambda_r1.txt"), ["--limit", "80", "--seedmms", "1", "--maxloc", "1"])
    dump(rep, fq, os.path.join(GOLD, "fm_mm1_rep.txt"), ["--limit", "100", "--seedmms", "1", "--maxloc", "1"])
    dump(rep, fq, os.path.join(GOLD, "fm_mm2_rep.txt"), ["--limit", "60", "--seedmms", "2", "--maxloc", "1"])
    dump(lam, os.path.join(ex, "reads_2.fq"), os.path.join(GOLD, "fm_mm2_lambda_r2_L20.txt"),
         ["--limit", "40", "--seedmms", "2", "-L", "20", "--interval", "1", "0.75", "--maxloc", "1"])


if __name__ == "__main__":
    main()
