"""GPU: the device framer / paired-end geometry (csrc/frame.cu, through the C ABI) against the
vectors dumped from the unmodified reference (tests/golden/frame_vectors.txt) and, on a larger
seeded random batch, against the C oracle."""
import numpy as np
import pytest

import helpers as H
from bt2gpu import lib, scoring, structs

pytestmark = pytest.mark.gpu

RECT = ("refl", "refr", "refl_pretrim", "refr_pretrim", "triml", "trimr", "corel", "corer", "maxgap")


def _rect(r):
    return tuple(int(r[k]) for k in RECT)


def test_seed_extension_rects_match_reference_vectors():
    vec = H.parse_frame_vectors()["S"]
    ctx = lib.Context(0)
    for trim in (0, 1):
        sel = [(i, o) for (i,), o in vec if i[7] == trim]
        jobs = np.zeros(len(sel), dtype=structs.frame_job)
        for k, (i, _) in enumerate(sel):
            jobs[k] = (i[0], i[2], i[5], i[1], i[3], i[4], i[6])
        rects, ok = ctx.frame_seed_ext(jobs, trim_to_ref=bool(trim))
        for k, (_, o) in enumerate(sel):
            assert (int(ok[k]),) + _rect(rects[k]) == tuple(o), sel[k][0]
    ctx.close()


def test_mate_frames_match_reference_vectors():
    vec = H.parse_frame_vectors()["M"]
    ctx = lib.Context(0)
    fo = H.FrameOracle()
    nchk = 0
    for (a, b, c), out in vec:            # one policy per vector: one call each (n = 1 batches)
        pol, maxfrag, minfrag, flags, is1, fw, off, maxalcols, reflen, len1, len2 = a
        orows, rdgap, rfgap, maxns, maxhalf, trim = c
        p = fo.policy(pol, maxfrag, minfrag, flags)
        p["trim_to_ref"] = trim
        if maxfrag == 0 or maxfrag < minfrag:
            continue
        j = np.zeros(1, dtype=structs.mate_job)
        j[0] = (off, maxalcols, reflen, maxns, len1, len2, rdgap, rfgap, maxhalf, is1, fw, (0, 0))
        f = ctx.frame_mate(p, j)[0]
        assert int(f["found"]) == b[0]
        assert (int(f["oleft"]), int(f["ofw"])) == (b[1], b[6])
        if b[0]:
            assert tuple(int(f[k]) for k in ("oll", "olr", "orl", "orr")) == tuple(b[2:6])
            assert (int(f["ok"]),) + _rect(f["rect"]) == tuple(out), (a, c)
            nchk += 1
    assert nchk > 1300
    ctx.close()


def test_pair_classes_and_gap_budgets_match_reference_vectors():
    vec = H.parse_frame_vectors()
    ctx = lib.Context(0)
    fo = H.FrameOracle()
    for (a,), out in vec["C"]:
        pol, maxfrag, minfrag, flags, off1, len1, fw1, off2, len2, fw2 = a
        if maxfrag == 0:
            continue
        j = np.zeros(1, dtype=structs.pair_job)
        j[0] = (off1, off2, len1, len2, fw1, fw2, (0,) * 6)
        assert int(ctx.pe_classify(fo.policy(pol, maxfrag, minfrag, flags), j)[0]) == out[0], a
    for local in (0, 1):
        ctx.set_scoring(scoring.Scoring(local=bool(local)))
        g = [(a, o) for (a,), o in vec["G"] if a[0] == local]
        rd, rf = ctx.max_gaps([a[1] for a, _ in g], [a[2] for a, _ in g])
        assert [(int(x), int(y)) for x, y in zip(rd, rf)] == [tuple(o) for _, o in g]
    ctx.close()


def test_joined_to_text_off_matches_reference_vectors():
    vec = H.parse_frame_vectors()["J"]
    e = H.load_index_fixture("rep_index.npz").fw
    ctx = lib.Context(0)
    for reject in (0, 1):
        sel = [(a, o) for (a,), o in vec if a[3] == reject]
        t, x, l, s = ctx.joined_to_text_off(e, [a[1] for a, _ in sel], [a[2] for a, _ in sel], bool(reject))
        for k, (a, o) in enumerate(sel):
            assert (int(t[k]), int(x[k]), int(l[k]), int(s[k])) == tuple(o), a
    with pytest.raises(lib.Bt2GpuError):
        ctx.joined_to_text_off(e, [5], [e.len], True)
    ctx.close()


def test_large_random_batch_matches_oracle():
    rng = np.random.default_rng(7)
    n = 200000
    ctx = lib.Context(0)
    fo = H.FrameOracle()
    pol = fo.policy(3, 500, 0, (0, 0, 0, 1, 1, 1))
    pol["trim_to_ref"] = 1
    j = np.zeros(n, dtype=structs.mate_job)
    reflen = 50_000_000
    j["off"] = rng.integers(-100, reflen + 100, n)
    j["len1"] = rng.integers(30, 260, n)
    j["len2"] = rng.integers(30, 260, n)
    j["is1"] = rng.integers(0, 2, n)
    j["fw"] = rng.integers(0, 2, n)
    j["maxrdgap"] = rng.integers(0, 40, n)
    j["maxrfgap"] = rng.integers(0, 40, n)
    j["maxhalf"] = 15
    j["maxns"] = rng.integers(0, 40, n)
    j["reflen"] = reflen
    j["maxalcols"] = np.where(j["is1"] == 1, j["len2"], j["len1"]) + j["maxrdgap"]
    out = ctx.frame_mate(pol, j)
    assert out["found"].all()
    for k in rng.choice(n, 3000, replace=False):
        jk = j[k]
        found, oleft, o4, ofw = fo.other_mate(pol, int(jk["is1"]), int(jk["fw"]), int(jk["off"]), int(jk["maxalcols"]), reflen,
                                              int(jk["len1"]), int(jk["len2"]))
        orows = int(jk["len2"] if jk["is1"] else jk["len1"])
        ok, r = fo.frame_mate(int(not oleft), *o4, orows, reflen, int(jk["maxrdgap"]), int(jk["maxrfgap"]), int(jk["maxns"]), 15, 1)
        f = out[k]
        assert (int(f["oleft"]), int(f["ofw"]), int(f["ok"])) == (oleft, ofw, ok)
        assert _rect(f["rect"]) == r
    ctx.close()


def test_mate_rect_alone_matches_oracle():
    """bt2gpu_frame_mate_rect_batch (frameFindMateRect for a given otherMate window) against the framing oracle."""
    rng = np.random.default_rng(77)
    n = 20000
    ctx = lib.Context(0)
    fo = H.FrameOracle()
    reflen = 100000
    j = np.zeros(n, dtype=structs.mate_rect_job)
    ll = rng.integers(-300, reflen + 100, n)
    j["ll"] = ll
    j["lr"] = ll + rng.integers(0, 400, n)
    j["rl"] = j["lr"] + rng.integers(0, 200, n)
    j["rr"] = j["rl"] + rng.integers(0, 400, n)
    j["reflen"] = reflen
    j["maxns"] = rng.integers(0, 40, n)
    j["rdlen"] = rng.integers(20, 260, n)
    j["maxrdgap"] = rng.integers(0, 30, n)
    j["maxrfgap"] = rng.integers(0, 30, n)
    j["maxhalf"] = 15
    j["anchor_left"] = rng.integers(0, 2, n)
    for trim in (0, 1):
        rects, ok = ctx.frame_mate_rect(trim, j)
        for k in rng.choice(n, 3000, replace=False):
            jk = j[k]
            eok, er = fo.frame_mate(int(jk["anchor_left"]), int(jk["ll"]), int(jk["lr"]), int(jk["rl"]), int(jk["rr"]),
                                    int(jk["rdlen"]), reflen, int(jk["maxrdgap"]), int(jk["maxrfgap"]), int(jk["maxns"]), 15, trim)
            assert int(ok[k]) == eok, jk
            if eok:
                assert _rect(rects[k]) == er, jk
    ctx.close()
