"""GPU parity (through the C ABI): the CUDA FM-index kernels vs the dumps of the unmodified
reference (tests/golden/fm_*.txt) and vs the CPU oracle on seeded random reads.  Bit-exact."""
import os

import numpy as np
import pytest

import helpers as H
from bt2gpu import fmio, scoring, lib, structs
from test_oracle_fm import CASES, fm_case_params

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = lib.Context(0)
    yield c
    c.close()


_IDX = {}


def dev_index(name):
    if name not in _IDX:
        ix = H.load_index_fixture(name)
        _IDX[name] = (ix, lib.DeviceIndex(ix, 0))
    return _IDX[name]


@pytest.mark.parametrize("dump,idx,local", CASES)
def test_cuda_fm_matches_reference_dump(ctx, dump, idx, local):
    sc = scoring.Scoring(local=local)
    ctx.set_scoring(sc)
    ix, dix = dev_index(idx)
    ctx.bind_index(dix)
    recs = fmio.parse_seed_dump(os.path.join(H.GOLDEN, dump))
    seed_len, ival = fm_case_params(local)
    assert H.compare_fm_with_dump(ctx, recs, sc, local, seed_len, ival) > 500


def _random_reads(rng, ix_name, n, rdlen_range):
    """Reads sampled from the indexed text itself is not possible from the fixture (no text),
    so mix: random reads (mostly no hit) + reads rebuilt from golden dumps' sequences."""
    recs = []
    for dump, idx, _ in CASES:
        if idx == ix_name:
            recs += fmio.parse_seed_dump(os.path.join(H.GOLDEN, dump))
    seqs, quals = [], []
    for i in range(n):
        ln = int(rng.integers(*rdlen_range))
        if recs and i % 3:
            r = recs[int(rng.integers(0, len(recs)))]
            s = r["seq"]
            if len(s) > ln:
                p = int(rng.integers(0, len(s) - ln + 1))
                s = s[p:p + ln]
            s = list(s)
            for _ in range(int(rng.integers(0, 3))):
                s[int(rng.integers(0, len(s)))] = "ACGTN"[int(rng.integers(0, 5))]
            s = "".join(s)
        else:
            s = "".join("ACGT"[k] for k in rng.integers(0, 4, size=ln))
        seqs.append(s)
        quals.append("".join(chr(33 + int(x)) for x in rng.integers(2, 41, size=len(s))))
    return fmio.ReadPool(seqs, quals)


@pytest.mark.parametrize("idx,local,seed", [("lambda_index.npz", False, 1), ("rep_index.npz", False, 2),
                                           ("rep_index.npz", True, 3)])
def test_cuda_fm_matches_oracle_random(ctx, idx, local, seed):
    rng = np.random.default_rng(seed)
    sc = scoring.Scoring(local=local)
    ctx.set_scoring(sc)
    ix, dix = dev_index(idx)
    ctx.bind_index(dix)
    orc = H.FmOracle(ix, sc.to_struct())
    pool = _random_reads(rng, idx, 1500, (12, 260))
    ctx.stage_reads(pool)
    orc.stage_reads(pool)
    # exactSweep with different mineMax / strand switches
    for nofw, norc, mm in ((False, False, 2), (True, False, 1), (False, True, 3)):
        a, b = orc.exact_sweep(nofw, norc, mm), ctx.exact_sweep(nofw, norc, mm)
        assert a.tobytes() == b.tobytes(), ("sweep", nofw, norc, mm)
    # oneMmSearch
    jobs = np.zeros(len(pool), dtype=structs.onemm_job)
    lens = pool.reads["rd_len"].astype(np.int64)
    jobs["minsc"] = [int(sc.score_min.f_int(int(x))) if not local else max(1, int(sc.score_min.f_int(int(x)))) for x in lens]
    jobs["nceil"] = [sc.nceil(int(x)) for x in lens]
    jobs["nofw"] = rng.integers(0, 4, size=len(pool)) == 0
    jobs["norc"] = rng.integers(0, 4, size=len(pool)) == 0
    for repex, rep1mm in ((True, True), (False, True), (True, False)):
        (ra, ha), (rb, hb) = orc.one_mm(jobs, local, repex, rep1mm), ctx.one_mm(jobs, local, repex, rep1mm)
        assert ra.tobytes() == rb.tobytes() and ha.tobytes() == hb.tobytes(), ("onemm", repex, rep1mm)
    # exact seeds of several lengths (incl. shorter than ftabChars) and intervals
    for sl, per in ((22, 7), (20, 3), (32, 11), (8, 5), (10, 9)):
        sj, _ = fmio.instantiate_exact_seeds(pool, sl, lambda n: per)
        a, b = orc.seed_exact(sj), ctx.seed_exact(sj)
        assert a.tobytes() == b.tobytes(), ("seed", sl, per)
        hit = np.nonzero(a["botf"] > a["topf"])[0]
        xj = np.zeros(len(hit), dtype=structs.extend_job)
        for k in ("topf", "botf", "topb", "botb"):
            xj[k] = a[k][hit]
        for k in ("read", "off", "len", "fw"):
            xj[k] = sj[k][hit]
        assert orc.extend(xj).tobytes() == ctx.extend(xj).tobytes(), ("extend", sl)
        rows = np.concatenate([np.arange(t, min(bt, t + 40), dtype=np.uint32) for t, bt in zip(a["topf"][hit], a["botf"][hit])]
                              + [rng.integers(0, ix.fw.len + 1, size=500).astype(np.uint32),
                                 np.array([ix.fw.zoff, 0, ix.fw.len], dtype=np.uint32)])
        (oa, sa), (ob, sb) = orc.locate(rows), ctx.locate(rows)
        assert np.array_equal(oa, ob) and np.array_equal(sa, sb), ("locate", sl)


def test_fm_errors_are_loud(ctx):
    ix, dix = dev_index("lambda_index.npz")
    ctx.bind_index(dix)
    pool = fmio.ReadPool(["ACGTACGTACGTACGTACGTACGTACGT"], ["I" * 28])
    ctx.stage_reads(pool)
    bad = np.zeros(1, dtype=structs.seed_job)
    bad["read"], bad["off"], bad["len"], bad["fw"] = 0, 20, 22, 1
    with pytest.raises(lib.Bt2GpuError):
        ctx.seed_exact(bad)
    bad["read"] = 5
    with pytest.raises(lib.Bt2GpuError):
        ctx.seed_exact(bad)
