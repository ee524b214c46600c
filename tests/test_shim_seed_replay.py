"""CPU: the shim's SeedAligner / SwDriver bodies (shim/seed_aligner_gpu.cpp: exactSweep, oneMmSearch, instantiateSeeds,
searchAllSeeds for -N 0 / 1 / 2; shim/sw_driver_gpu.cpp: extend) against the reference's own, read by read.

oracle/_ref/ref_seed_dump drives the unmodified reference; oracle/_ref/ref_seed_dump_emul is the same harness with those
members replaced by the shim's (per-call binding) over the CPU oracle behind the ABI's names (TEST INFRASTRUCTURE:
oracle/Makefile).  Everything both print must be identical: sweep bounds and exact hits, 1-mismatch hit lists in order,
the BW ranges of every seed offset in the order the alignment cache holds them (with their reference strings for
-N 1 / 2), extension lengths, SA offsets, bwops.  One field is exempt: oneMmSearch's return value, which its only caller
ignores (bt2_search.cpp:3654) and which the reference sets even when the local-mode legality check drops the hit."""
import gzip
import os
import shutil
import subprocess

import pytest

import dropin as D
import helpers as H

BIN = os.path.join(H.ROOT, "oracle", "_ref")
STOCK, SHIM = os.path.join(BIN, "ref_seed_dump"), os.path.join(BIN, "ref_seed_dump_emul")


@pytest.fixture(scope="module")
def work(tmp_path_factory):
    if not (D.have_binaries() and os.path.exists(STOCK) and os.path.exists(SHIM)):
        pytest.skip("oracle/_ref harnesses missing: run __graft_entry__.build() where /root/reference is mounted")
    w = str(tmp_path_factory.mktemp("seedreplay"))
    D.build_index(os.path.join(D.GOLDEN, "lambda.fa"), os.path.join(w, "lambda"))
    D.build_index(os.path.join(D.GOLDEN, "rep.fa"), os.path.join(w, "rep"))
    for n in ("example_reads_1.fq.gz", "example_longreads.fq.gz"):
        with gzip.open(os.path.join(D.GOLDEN, n), "rb") as s, open(os.path.join(w, n[:-3]), "wb") as o:
            shutil.copyfileobj(s, o)
    return w


CASES = [("lambda", "example_reads_1.fq", ["--limit", "150"]),
         ("lambda", "example_reads_1.fq", ["--limit", "100", "--local", "-L", "20", "--interval", "1", "0.75"]),
         ("lambda", "example_longreads.fq", ["--limit", "30"]),
         ("lambda", "example_reads_1.fq", ["--limit", "80", "--seedmms", "1"]),
         ("lambda", "example_reads_1.fq", ["--limit", "50", "--seedmms", "2", "-L", "20"]),
         ("rep", "example_reads_1.fq", ["--limit", "60", "--seedmms", "1"])]


@pytest.mark.parametrize("idx,fq,extra", CASES, ids=[" ".join([c[0]] + c[2][2:]) or c[0] for c in CASES])
def test_shim_seed_side_replays_the_reference(work, idx, fq, extra):
    def run(binary):
        out = subprocess.check_output([binary, "--index", os.path.join(work, idx), "--reads", os.path.join(work, fq)] + extra)
        lines = out.decode().splitlines()
        return [("M1 _ " + ln.split(" ", 2)[2]) if ln.startswith("M1 ") else ln for ln in lines]
    a, b = run(STOCK), run(SHIM)
    assert len(a) > 1000
    bad = [(x, y) for x, y in zip(a, b) if x != y]
    assert len(a) == len(b) and not bad, bad[:3]
