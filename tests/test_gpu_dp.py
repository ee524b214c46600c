"""GPU parity (through the C ABI): CUDA fill/gather/backtrace vs the reference dumps
(tests/golden) and vs the CPU oracle on seeded random problems.  Bit-exact."""
import os

import numpy as np
import pytest

import helpers as H
from bt2gpu import dpio, scoring, lib

pytestmark = pytest.mark.gpu

CASES = [
    ("dp_c2_ext.txt", False), ("dp_c2_mate.txt", False), ("dp_c1_ext.txt", False),
    ("dp_c2loc_ext.txt", True), ("dp_c2loc_mate.txt", True), ("dp_c1loc_ext.txt", True),
]


@pytest.fixture(scope="module")
def ctx():
    c = lib.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("name,local", CASES)
def test_cuda_matches_reference_dump(ctx, name, local):
    sc = scoring.Scoring(local=local)
    ctx.set_scoring(sc)
    probs = list(dpio.parse_dump(os.path.join(H.GOLDEN, name)))
    seq, qual, win, jobs = dpio.batch_from_problems(probs, sc)
    res = ctx.dp_align_batch(seq, qual, win, jobs, max_aln=8, cand_cap=40_000_000, edit_cap=2_000_000)
    assert H.compare_with_dump(probs, *res) == len(probs)


def _random_batch(rng, sc, n, local, rdlen_range, mutate=0.06, tandem=False):
    b = dpio.DpBatch()
    for _ in range(n):
        rdlen = int(rng.integers(*rdlen_range))
        maxgap = int(rng.integers(0, 16))
        width = rdlen + 2 * maxgap
        ref = rng.integers(0, 4, size=width + 1).astype(np.uint8)
        if tandem:   # tandem repeat: many equally good placements, long runs of tied moves, many candidates
            unit = rng.integers(0, 4, size=int(rng.integers(1, 9))).astype(np.uint8)
            ref = np.resize(unit, width + 1).copy()
            drift = rng.random(width + 1) < 0.02
            ref[drift] = rng.integers(0, 4, size=int(drift.sum()))
        start = int(rng.integers(0, 2 * maxgap + 1))
        read = ref[start:start + rdlen].copy()
        # substitutions, a deletion and an insertion now and then, a few Ns
        mm = rng.random(rdlen) < mutate
        read[mm] = rng.integers(0, 4, size=int(mm.sum()))
        if rdlen > 30 and rng.random() < 0.4:
            p = int(rng.integers(10, rdlen - 10)); g = int(rng.integers(1, 4))
            read = np.concatenate([read[:p], read[p + g:], rng.integers(0, 4, size=g).astype(np.uint8)])
        if rdlen > 30 and rng.random() < 0.4:
            p = int(rng.integers(10, rdlen - 10)); g = int(rng.integers(1, 4))
            read = np.concatenate([read[:p], rng.integers(0, 4, size=g).astype(np.uint8), read[p:]])[:rdlen]
        if rng.random() < 0.2:
            read[int(rng.integers(0, rdlen))] = 4
        if rng.random() < 0.1:
            ref[int(rng.integers(0, width))] = 4
        qual = (rng.integers(2, 41, size=rdlen) + 33).astype(np.uint8)
        masks = (1 << ref).astype(np.uint8)
        fw = bool(rng.integers(0, 2))
        if not fw:  # DpBatch.add reverse-complements for '-' problems; undo so the planted hit stays
            read = dpio.revcomp(read); qual = qual[::-1]
        minsc = int(sc.score_min.f_int(rdlen)) if not local else max(1, int(sc.score_min.f_int(rdlen)))
        if rng.random() < 0.15:
            minsc = minsc // 4 if not local else minsc * 2
        b.add(read, qual, fw, masks, minsc, sc.nceil(rdlen), int(rng.integers(-5, 1000)), 0,
              maxgap, 3 * maxgap if maxgap else 0, enable8=bool(rng.random() < 0.8))
    return b.finalize()


@pytest.mark.parametrize("local,rdlen_range,n,seed", [
    (False, (20, 160), 600, 1), (False, (150, 151), 400, 2), (False, (300, 900), 60, 3),
    (True, (20, 160), 400, 4), (True, (250, 251), 200, 5), (True, (300, 700), 40, 6),
])
def test_cuda_matches_oracle_random(ctx, local, rdlen_range, n, seed):
    rng = np.random.default_rng(seed)
    sc = scoring.Scoring(local=local)
    ctx.set_scoring(sc)
    seq, qual, win, jobs = _random_batch(rng, sc, n, local, rdlen_range)
    ref = H.oracle_dp(sc.to_struct(), seq, qual, win, jobs, cand_cap=30_000_000, edit_cap=2_000_000)
    got = ctx.dp_align_batch(seq, qual, win, jobs, max_aln=8, cand_cap=30_000_000, edit_cap=2_000_000)
    for i in range(n):
        a = dpio.results_to_py(*ref, i)
        b = dpio.results_to_py(*got, i)
        assert a == b, (i, a[0]["best"], b[0]["best"], a[2], b[2])


@pytest.mark.parametrize("seed,gapbar", [(21, 4), (22, 1), (23, 8)])
def test_cuda_matches_oracle_tandem_repeats(ctx, seed, gapbar):
    # end-to-end windows over tandem repeats: dozens of candidates per window whose walks cross each
    # other, so the order in which nextAlignment tries them (and what each one marks) decides the answer
    rng = np.random.default_rng(seed)
    sc = scoring.Scoring(local=False, gapbar=gapbar)
    ctx.set_scoring(sc)
    seq, qual, win, jobs = _random_batch(rng, sc, 300, False, (30, 180), mutate=0.03, tandem=True)
    ref = H.oracle_dp(sc.to_struct(), seq, qual, win, jobs, max_aln=16, cand_cap=30_000_000, edit_cap=2_000_000)
    got = ctx.dp_align_batch(seq, qual, win, jobs, max_aln=16, cand_cap=30_000_000, edit_cap=2_000_000)
    ntries = 0
    for i in range(len(jobs)):
        a = dpio.results_to_py(*ref, i)
        b = dpio.results_to_py(*got, i)
        assert a == b, (i, a[0], b[0])
        ntries += a[0]["ncand"]
    assert ntries > 15 * len(jobs)


def test_non_default_scoring(ctx):
    rng = np.random.default_rng(11)
    sc = scoring.Scoring(local=False, mmp_max=5, mmp_min=3, np_=2, rdg=(4, 2), rfg=(6, 4), gapbar=2)
    ctx.set_scoring(sc)
    seq, qual, win, jobs = _random_batch(rng, sc, 300, False, (30, 200))
    ref = H.oracle_dp(sc.to_struct(), seq, qual, win, jobs)
    got = ctx.dp_align_batch(seq, qual, win, jobs, max_aln=8)
    for i in range(len(jobs)):
        assert dpio.results_to_py(*ref, i) == dpio.results_to_py(*got, i), i


def test_one_scoring_scheme_per_device(ctx):
    """The kernels read the scheme from device-wide constant memory: a second, different scheme is refused while another
    live context holds the first (instead of silently overwriting it under that context's kernels)."""
    ctx.set_scoring(scoring.Scoring(local=False))      # the module's context is alive too: all three share one scheme
    a, b = lib.Context(0), lib.Context(0)
    try:
        a.set_scoring(scoring.Scoring(local=False))
        b.set_scoring(scoring.Scoring(local=False))            # the same scheme is shared
        with pytest.raises(lib.Bt2GpuError, match="different scoring scheme"):
            b.set_scoring(scoring.Scoring(local=True))
        a.close()
        with pytest.raises(lib.Bt2GpuError, match="different scoring scheme"):
            b.set_scoring(scoring.Scoring(local=True))         # the module's context still holds the end-to-end scheme
        b.close()
        b = lib.Context(0)
        ctx.set_scoring(scoring.Scoring(local=True))           # a scheme nobody else holds may be replaced
        b.set_scoring(scoring.Scoring(local=True))
        probs = list(dpio.parse_dump(os.path.join(H.GOLDEN, "dp_c2loc_ext.txt"), limit=40))
        sc = scoring.Scoring(local=True)
        seq, qual, win, jobs = dpio.batch_from_problems(probs, sc)
        res = b.dp_align_batch(seq, qual, win, jobs, max_aln=8, cand_cap=4_000_000, edit_cap=200_000)
        assert H.compare_with_dump(probs, *res) == len(probs)
    finally:
        a.close(); b.close()


def test_candidate_pool_overflow_is_an_error_not_a_scribble(ctx):
    """Local mode with a candidate pool far too small: the call fails cleanly (no out-of-bounds emit / sort / backtrace
    on the jobs that did not fit) and the context is usable afterwards."""
    sc = scoring.Scoring(local=True)
    ctx.set_scoring(sc)
    probs = list(dpio.parse_dump(os.path.join(H.GOLDEN, "dp_c2loc_ext.txt"), limit=200))
    seq, qual, win, jobs = dpio.batch_from_problems(probs, sc)
    for cap in (16, 700, 5000):
        with pytest.raises(lib.Bt2GpuError, match="candidate pool overflow"):
            ctx.dp_align_batch(seq, qual, win, jobs, max_aln=8, cand_cap=cap, edit_cap=200_000)
    res = ctx.dp_align_batch(seq, qual, win, jobs, max_aln=8, cand_cap=8_000_000, edit_cap=400_000)
    assert H.compare_with_dump(probs, *res) == len(probs)
