"""GPU: bt2gpu_ungapped_batch (SwAligner::ungappedAlign, aligner_sw.cpp:286-494) against the reference dumps and,
on a larger random batch, against the oracle."""
import os

import numpy as np
import pytest

import helpers as H
from bt2gpu import lib, scoring, structs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("tag,local", [("ee", False), ("loc", True)])
def test_gpu_ungapped_matches_reference(tag, local):
    sc = scoring.Scoring(local=local)
    seq, qual, win, jobs, exp = H.parse_ungapped_golden(os.path.join(H.GOLDEN, "ungapped_%s.txt" % tag), sc)
    ctx = lib.Context(0)
    ctx.set_scoring(sc)
    res, edits = ctx.ungapped_batch(seq, qual, win, jobs)
    for i in range(len(jobs)):
        assert H.ungapped_to_py(res, edits, i) == exp[i], (i, jobs[i])
    ctx.close()


@pytest.mark.parametrize("local", [False, True])
def test_gpu_ungapped_random_vs_oracle(local):
    rng = np.random.default_rng(31 + local)
    sc = scoring.Scoring(local=local)
    n = 20000
    lens = rng.integers(1, 400, size=n)
    offs = np.concatenate([[0], np.cumsum(lens)[:-1]])
    tot = int(lens.sum())
    win = rng.integers(0, 4, size=tot).astype(np.uint8)
    seq = win.copy()
    mm = rng.random(tot) < rng.choice([0.0, 0.02, 0.1, 0.5], size=n).repeat(lens)
    seq[mm] = rng.integers(0, 5, size=int(mm.sum()))
    win[rng.random(tot) < 0.01] = 4
    qual = (rng.integers(0, 42, size=tot) + 33).astype(np.uint8)
    jobs = np.zeros(n, dtype=structs.ungap_job)
    jobs["rd_off"] = offs; jobs["rf_off"] = offs; jobs["rd_len"] = lens
    jobs["minsc"] = (20 + 8 * np.log(lens)).astype(np.int64) if local else (-0.6 - 0.6 * lens).astype(np.int64)
    jobs["nceil"] = [sc.nceil(int(x)) for x in lens]
    jobs["fw"] = rng.integers(0, 2, size=n); jobs["ohang"] = rng.integers(0, 2, size=n)
    jobs["reflen"] = 100000
    jobs["off"] = np.where(rng.random(n) < 0.1, rng.integers(-5, 0, size=n), rng.integers(0, 99000, size=n))
    ctx = lib.Context(0)
    ctx.set_scoring(sc)
    got, ge = ctx.ungapped_batch(seq, qual, win, jobs)
    ref, re_ = H.oracle_ungapped(sc.to_struct(), seq, qual, win, jobs)
    for i in range(n):
        assert H.ungapped_to_py(got, ge, i) == H.ungapped_to_py(ref, re_, i), (i, jobs[i])
    assert (ref["ret"] == 1).sum() > n // 10
    ctx.close()
