"""GPU: bt2gpu_ungapped_batch_noalloc -- the no-device-allocation variant of bt2gpu_ungapped_batch (its scratch arena is
a static device array of BT2GPU_NOALLOC_SCRATCH_BYTES; include/bt2gpu.h) -- against the existing entry point, through the
C ABI, bit for bit (integer work: the tolerance of every test here is zero), on two calls in a row that share the
array.  The entry point has no counterpart in the reference.

This file is named to be collected LAST: a failure here must not hide the parity
tests.  It passes alone in a fresh process and after the rest of the suite (it asks for nothing another test left behind:
its own context, its own scoring scheme, and the shared array is cleared where it must be by the call itself)."""
import gc
import os

import numpy as np
import pytest

import helpers as H
from bt2gpu import lib, scoring, structs

pytestmark = pytest.mark.gpu


def _random_batch(rng, sc, n, maxlen):
    lens = rng.integers(1, maxlen, size=n)
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
    jobs["minsc"] = (-0.6 - 0.6 * lens).astype(np.int64)
    jobs["nceil"] = [sc.nceil(int(x)) for x in lens]
    jobs["fw"] = rng.integers(0, 2, size=n); jobs["ohang"] = rng.integers(0, 2, size=n)
    jobs["reflen"] = 100000
    jobs["off"] = np.where(rng.random(n) < 0.1, rng.integers(-5, 0, size=n), rng.integers(0, 99000, size=n))
    return seq, qual, win, jobs


def _same(a, b, n):
    (ra, ea), (rb, eb) = a, b
    assert len(ea) == len(eb)
    for i in range(n):
        assert H.ungapped_to_py(ra, ea, i) == H.ungapped_to_py(rb, eb, i), i


def test_noalloc_ungapped_matches_existing_entry_point_twice():
    gc.collect()                    # contexts other tests dropped without closing release their scoring registration
    sc = scoring.Scoring(local=False)
    ctx = lib.Context(0)
    try:
        ctx.set_scoring(sc)
        cap = int(ctx.lib.bt2gpu_noalloc_scratch_bytes())
        assert cap == 4 << 20
        # call 1: the reference's own dump (also pins both entry points on the reference), a prefix that fits the array
        seq, qual, win, jobs, exp = H.parse_ungapped_golden(os.path.join(H.GOLDEN, "ungapped_ee.txt"), sc)
        n1 = 600
        j1 = jobs[:n1]
        end = int(j1["rd_off"][-1] + j1["rd_len"][-1])
        a = ctx.ungapped_batch_noalloc(seq[:end], qual[:end], win[:end], j1)
        b = ctx.ungapped_batch(seq[:end], qual[:end], win[:end], j1)
        _same(a, b, n1)
        for i in range(n1):
            assert H.ungapped_to_py(a[0], a[1], i) == exp[i], i
        assert (a[0]["ret"] == 1).sum() > 50
        # call 2, straight after, other and smaller inputs: the array still holds call 1's pools, results, edits and
        # counters; nothing of them may show
        rng = np.random.default_rng(77)
        s2, q2, w2, j2 = _random_batch(rng, sc, 1500, 300)
        a2 = ctx.ungapped_batch_noalloc(s2, q2, w2, j2)
        b2 = ctx.ungapped_batch(s2, q2, w2, j2)
        _same(a2, b2, len(j2))
        assert (a2[0]["ret"] == 1).sum() > 100 and len(a2[1]) > 0
        # and the first batch again gives the first answer again
        a3 = ctx.ungapped_batch_noalloc(seq[:end], qual[:end], win[:end], j1)
        _same(a3, a, n1)
        # a batch beyond the array is refused with an error that names the limit; the context stays usable
        s4, q4, w4, j4 = _random_batch(rng, sc, 20000, 400)
        with pytest.raises(lib.Bt2GpuError, match="BT2GPU_NOALLOC_SCRATCH_BYTES"):
            ctx.ungapped_batch_noalloc(s4, q4, w4, j4)
        a5 = ctx.ungapped_batch_noalloc(s2, q2, w2, j2)
        _same(a5, b2, len(j2))
    finally:
        ctx.close()
