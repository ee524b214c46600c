"""CPU suite: bt2gpu.build (torch-sort index construction) against the reference's own
bowtie2-build-s output, array by array, on the committed index fixtures (lambda phage and the
repeat genome with an N gap), and file by file when oracle/_ref holds the reference build."""
import os

import numpy as np
import pytest

import helpers as H
from bt2gpu import build, dpio, index


def _read_fasta(path):
    names, seqs, cur = [], [], []
    for line in open(path):
        if line.startswith(">"):
            if names:
                seqs.append("".join(cur))
            names.append(line[1:].strip())
            cur = []
        else:
            cur.append(line.strip())
    seqs.append("".join(cur))
    return names, [dpio.encode_seq(s) for s in seqs]


def _same(a, b):
    for k in ("len", "off_rate", "ftab_chars", "zoff", "npat", "nfrag"):
        assert getattr(a, k) == getattr(b, k), k
    for k in ("fchr", "plen", "rstarts", "ebwt", "ftab", "eftab"):
        assert np.array_equal(np.asarray(getattr(a, k)).reshape(-1), np.asarray(getattr(b, k)).reshape(-1)), k
    if b.offs is not None:
        assert np.array_equal(a.offs, b.offs), "offs"


@pytest.mark.parametrize("fa,fixture", [("rep.fa", "rep_index.npz"), ("lambda.fa", "lambda_index.npz")])
def test_builder_matches_bowtie2_build(fa, fixture, tmp_path):
    path = os.path.join(H.GOLDEN, fa)
    names, seqs = _read_fasta(path)
    ix, extras = build.build_index(seqs, [n.split()[0] for n in names], device="cpu")
    ref = H.load_index_fixture(fixture)
    _same(ix.fw, ref.fw)
    _same(ix.bw, ref.bw)
    # written files load back identically, and equal the reference's files when those exist here
    base = str(tmp_path / "idx")
    build.write_index(ix, extras, base)
    back = index.Index.load(base)
    _same(back.fw, ref.fw)
    _same(back.bw, ref.bw)
    refbase = os.path.join(H.REFBIN, "idx", "rep" if fa == "rep.fa" else "lambda_virus")
    if fa == "rep.fa" and index.exists(refbase):
        for suf in (".1.bt2", ".2.bt2", ".rev.1.bt2", ".rev.2.bt2", ".3.bt2", ".4.bt2"):
            assert open(base + suf, "rb").read() == open(refbase + suf, "rb").read(), suf


def test_random_genome_index_is_self_consistent():
    """LF-walking the built index from the '$' row must spell the text backwards."""
    rng = np.random.default_rng(5)
    s = rng.integers(0, 4, size=5000).astype(np.uint8)
    ix, _ = build.build_index([s], device="cpu")
    e = ix.fw
    orc = H.FmOracle(ix, None)
    offs, _ = orc.locate(np.arange(0, e.len + 1, dtype=np.uint32))
    assert sorted(int(x) for x in offs[:-1]) == list(range(e.len)) or sorted(int(x) for x in offs) == list(range(e.len + 1))


def test_bucketed_suffix_sort_matches_single_bucket():
    """The memory-lean path (key-prefix buckets, used for human-scale genomes) gives the same index as the
    single-sort path: forced with a tiny bucket size on a repetitive text with an N gap and on a random one."""
    path = os.path.join(H.GOLDEN, "rep.fa")
    names, seqs = _read_fasta(path)
    ref = H.load_index_fixture("rep_index.npz")
    for rows in (64, 1000, 7919):
        ix, _ = build.build_index(seqs, [n.split()[0] for n in names], device="cpu", bucket_rows=rows)
        _same(ix.fw, ref.fw)
        _same(ix.bw, ref.bw)
    rng = np.random.default_rng(11)
    s = rng.integers(0, 4, size=30011).astype(np.uint8)
    a, _ = build.build_index([s[:17000], s[17000:]], device="cpu")
    b, _ = build.build_index([s[:17000], s[17000:]], device="cpu", bucket_rows=2048)
    _same(a.fw, b.fw)
    _same(a.bw, b.bw)
