"""CPU suite: oracle/frame_oracle.c (window framing, paired-end geometry, gap budgets, joined ->
text offsets) against vectors dumped from the unmodified reference by oracle/ref_frame_dump.cpp
(tests/golden/frame_vectors.txt: DynProgFramer, PairedEndPolicy, Scoring::max*Gaps,
Ebwt::joinedToTextOff over the 3-sequence repeat genome, which has an N gap inside a sequence)."""
import helpers as H


def test_seed_extension_rects_match_reference():
    fo, vec = H.FrameOracle(), H.parse_frame_vectors()
    assert len(vec["S"]) == 1500
    ntrim = 0
    for (inp,), out in vec["S"]:
        ok, r = fo.frame_seed_ext(*inp)
        assert (ok,) + r == tuple(out), inp
        ntrim += (out[5] > 0) + (out[6] > 0)
    assert ntrim > 200          # rectangles trimmed against a reference end are well represented


def test_other_mate_and_mate_rects_match_reference():
    fo, vec = H.FrameOracle(), H.parse_frame_vectors()
    nfound = nleft = 0
    for (a, b, c), out in vec["M"]:
        pol, maxfrag, minfrag, flags, is1, fw, off, maxalcols, reflen, len1, len2 = a
        found, oleft, oll, olr, orl, orr, ofw = b
        orows, rdgap, rfgap, maxns, maxhalf, trim = c
        p = fo.policy(pol, maxfrag, minfrag, flags)
        got = fo.other_mate(p, is1, fw, off, maxalcols, reflen, len1, len2)
        if not found:
            assert got[0] == 0
            continue
        assert got == (1, oleft, [oll, olr, orl, orr], ofw), a
        ok, r = fo.frame_mate(int(not oleft), oll, olr, orl, orr, orows, reflen, rdgap, rfgap, maxns, maxhalf, trim)
        assert (ok,) + r == tuple(out), (a, b, c)
        nfound += 1
        nleft += oleft
    assert nfound > 1300 and 300 < nleft < nfound - 300


def test_pair_classes_match_reference():
    fo, vec = H.FrameOracle(), H.parse_frame_vectors()
    seen = set()
    for (a,), out in vec["C"]:
        pol, maxfrag, minfrag, flags, off1, len1, fw1, off2, len2, fw2 = a
        cl = fo.classify(fo.policy(pol, maxfrag, minfrag, flags), off1, len1, fw1, off2, len2, fw2)
        assert cl == out[0], a
        seen.add(cl)
    assert seen == {1, 2, 3, 4, 5}


def test_gap_budgets_match_reference():
    fo, vec = H.FrameOracle(), H.parse_frame_vectors()
    for (a,), out in vec["G"]:
        local, minsc, rdlen, match, rdo, rde, rfo, rfe = a
        assert fo.lib.bt2o_max_read_gaps(minsc, rdlen, match, rdo, rde) == out[0]
        assert fo.lib.bt2o_max_ref_gaps(minsc, rdlen, match, rfo, rfe) == out[1]


def test_joined_to_text_off_matches_reference():
    fo, vec = H.FrameOracle(), H.parse_frame_vectors()
    e = H.load_index_fixture("rep_index.npz").fw
    assert e.nfrag == 4 and e.npat == 3
    nstr = nrej = 0
    for (a,), out in vec["J"]:
        fw, qlen, off, reject = a
        tidx, textoff, tlen, straddled = fo.joined_to_text_off(e, qlen, off, reject)
        if out[0] == 0xffffffff:
            assert tidx == 0xffffffff and straddled == 1
            nrej += 1
        else:
            assert (tidx, textoff, tlen, straddled) == tuple(out), a
        nstr += out[3]
    assert nrej > 50 and nstr > nrej
