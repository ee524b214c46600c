"""GPU: the drop-in statement.  The unmodified reference (oracle/_ref/bowtie2-align-server-s + client) and the same
reference with SeedAligner::exactSweep / oneMmSearch / searchAllSeeds and SwAligner::align / nextAlignment bound
to libbt2gpu.so through bowtie2-server_b200/shim/ must print byte-identical SAM records for the same reads, --seed 0:
  bowtie2-align-server-s-gpu    per-call binding (one problem per ABI call; also binds the framer / paired-end geometry),
  bowtie2-align-server-s-gpub   batched binding (worker fibers: the problems of all reads in flight per ABI call) -- the
                                path bench.py times as `e2e`.
SwDriver, AlnSinkWrap and the SAM writer are the reference's in all runs (SURVEY §8b, §8d parity gate)."""
import os
import re

import numpy as np
import pytest

import dropin as D
from bt2gpu import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def work(tmp_path_factory):
    if not D.have_binaries():
        # test infrastructure, not product: the stock and the GPU-bound reference servers are built by
        # __graft_entry__.build() where /root/reference is mounted and travel to the GPU box as files
        pytest.skip("oracle/_ref binaries missing: run __graft_entry__.build() where /root/reference is mounted")
    w = str(tmp_path_factory.mktemp("dropin"))
    D.build_index(os.path.join(D.GOLDEN, "lambda.fa"), os.path.join(w, "lambda"))
    return w


def _both(work, tag, server_args, client_args, index="lambda", threads=2):
    base = os.path.join(work, index)
    cpu, _ = D.run(D.CPU_SERVER, base, server_args, client_args, os.path.join(work, tag + "_cpu.sam"), threads=threads)
    stats = os.path.join(work, tag + "_shim_stats.txt")
    gpu, err = D.run(D.GPU_SERVER, base, server_args, client_args, os.path.join(work, tag + "_gpu.sam"),
                     env={"BT2GPU_SHIM_STATS": stats}, threads=threads)
    st = dict((k, int(v)) for k, v in re.findall(r"(\w+)\s+(\d+)", open(stats).read()))
    print(tag, st)
    for k in ("align_calls_on_gpu", "sweep_calls_on_gpu", "seeds_on_gpu", "extend_calls_on_gpu", "locate_rows_on_gpu", "frame_calls_on_gpu",
              "joined_to_text_off_on_gpu", "kernel_launches"):
        assert st.get(k, 0) > 0, "%s: nothing ran on the GPU:\n%s\n%s" % (k, st, err[-2000:])
    _batched(work, tag, cpu, base, server_args, client_args)
    return cpu, gpu, st["align_calls_on_gpu"]


def _batched(work, tag, cpu, base, server_args, client_args, workers=96, host_threads=3):
    """The same reads through the batched binding: identical records, and the problems did travel together."""
    stats = os.path.join(work, tag + "_batched_stats.txt")
    got, err = D.run(D.GPUB_SERVER, base, server_args, client_args, os.path.join(work, tag + "_gpub.sam"), threads=workers,
                     env={"BT2GPU_SHIM_STATS": stats, "BT2GPU_THREADS": str(host_threads)})
    st = dict((k, int(v)) for k, v in re.findall(r"(\w+)\s+(\d+)", open(stats).read()))
    print(tag, "batched", st)
    assert st.get("kernel_launches", 0) > 0 and st.get("abi_batches", 0) > 0, (st, err[-2000:])
    _check(cpu, got, need_aligned=False)
    return st


def _check(cpu, gpu, need_aligned=True):
    assert len(cpu) == len(gpu)
    bad = [(a, b) for a, b in zip(cpu, gpu) if a != b]
    assert not bad, "%d of %d SAM records differ, first:\n%s%s" % (len(bad), len(cpu), bad[0][0], bad[0][1])
    aligned = sum(1 for ln in cpu if ln.split("\t")[2] != "*")
    assert aligned > len(cpu) // 4 or not need_aligned


def test_paired_end_to_end_sam_identical(work):
    """C2 shape: paired-end, end-to-end, default options (u8 path, mate-window DP through osw)."""
    rng = np.random.default_rng(11)
    r, q, _, _ = synth.simulate_pairs(rng, D.lambda_genome(), 1500, 100)
    f1, f2 = os.path.join(work, "pe_1.fq"), os.path.join(work, "pe_2.fq")
    D.write_fastq(f1, r[0::2], q[0::2], "pair")
    D.write_fastq(f2, r[1::2], q[1::2], "pair")
    cpu, gpu, calls = _both(work, "pe", [], ["-1", f1, "-2", f2])
    assert calls > 500
    st = dict((k, int(v)) for k, v in re.findall(r"(\w+)\s+(\d+)", open(os.path.join(work, "pe_shim_stats.txt")).read()))
    assert st["pe_classify_on_gpu"] > 0 and st["other_mate_on_gpu"] > 0 and st["mate_rect_on_gpu"] > 0, st
    _check(cpu, gpu)


def test_unpaired_ragged_sensitive_sam_identical(work):
    """C1 shape: unpaired reads of 30..700 bp, --sensitive (u8 and, for minsc < -254, the i16 end-to-end path)."""
    rng = np.random.default_rng(12)
    L = 700
    r, q, _, _ = synth.simulate_reads(rng, D.lambda_genome(), 600, L)
    lens = rng.integers(30, L + 1, size=600)
    f = os.path.join(work, "long.fq")
    D.write_fastq(f, r, q, "long", lens)
    cpu, gpu, _ = _both(work, "long", ["--sensitive"], ["-U", f])
    _check(cpu, gpu)


def test_unpaired_local_sam_identical(work):
    """C4 shape: --local (u8 local DP, i16 re-run on saturation, local candidate gather and SQ filter)."""
    rng = np.random.default_rng(13)
    r, q, _, _ = synth.simulate_reads(rng, D.lambda_genome(), 800, 250)
    f = os.path.join(work, "loc.fq")
    D.write_fastq(f, r, q, "loc")
    cpu, gpu, _ = _both(work, "loc", ["--local"], ["-U", f])
    _check(cpu, gpu)


def test_paired_local_sam_identical(work):
    rng = np.random.default_rng(14)
    r, q, _, _ = synth.simulate_pairs(rng, D.lambda_genome(), 600, 150)
    f1, f2 = os.path.join(work, "pl_1.fq"), os.path.join(work, "pl_2.fq")
    D.write_fastq(f1, r[0::2], q[0::2], "pl")
    D.write_fastq(f2, r[1::2], q[1::2], "pl")
    cpu, gpu, _ = _both(work, "pl", ["--local"], ["-1", f1, "-2", f2])
    _check(cpu, gpu)


def test_near_exact_reads_one_mismatch_path(work):
    """High-quality reads with zero or one substitution, both strands: exactSweep's exact end-to-end hits and the
    up-front oneMmSearch (bt2_search.cpp:3634-3668) decide these reads; a third carry several substitutions so the
    seed / extend / DP path runs too."""
    rng = np.random.default_rng(15)
    g = D.lambda_genome()
    n, L = 900, 80
    pos = rng.integers(0, len(g) - L, size=n)
    r = g[pos[:, None] + np.arange(L)[None, :]].copy()
    nsub = rng.choice([0, 1, 1, 3], size=n)
    for i in range(n):
        for p in rng.choice(L, size=int(nsub[i]), replace=False):
            r[i, p] = (r[i, p] + rng.integers(1, 4)) % 4
    rc = rng.random(n) < 0.5
    comp = np.array([3, 2, 1, 0, 4], dtype=np.uint8)
    r[rc] = comp[r[rc, ::-1]]
    q = np.full((n, L), ord("I"), dtype=np.uint8)
    f = os.path.join(work, "near.fq")
    D.write_fastq(f, r, q, "near")
    # the last two leave no room for a gap in the score budget: SwDriver goes through ungappedAlign
    # (aligner_sw_driver.cpp:1031-1060) before / instead of the DP
    for tag, args in (("near", []), ("nearloc", ["--local"]), ("nearung", ["--score-min", "L,0,-0.2", "--rdg", "20,3", "--rfg", "20,3"]),
                      ("nearungloc", ["--local", "--score-min", "L,0,1.95"])):
        base = os.path.join(work, "lambda")
        cpu, _ = D.run(D.CPU_SERVER, base, args, ["-U", f], os.path.join(work, tag + "_cpu.sam"))
        stats = os.path.join(work, tag + "_shim_stats.txt")
        gpu, err = D.run(D.GPU_SERVER, base, args, ["-U", f], os.path.join(work, tag + "_gpu.sam"),
                         env={"BT2GPU_SHIM_STATS": stats})
        st = dict((k, int(v)) for k, v in re.findall(r"(\w+)\s+(\d+)", open(stats).read()))
        print(tag, st)
        assert st.get("onemm_calls_on_gpu", 0) > 0 and st.get("sweep_calls_on_gpu", 0) > 0, st
        if "ung" in tag:
            assert st.get("ungapped_calls_on_gpu", 0) > 0, st
        _check(cpu, gpu)
        bst = _batched(work, tag, cpu, base, args, ["-U", f])
        assert bst.get("onemm_calls_on_gpu", 0) > 0 and ("ung" not in tag or bst.get("ungapped_calls_on_gpu", 0) > 0), bst


def test_repeat_genome_three_chromosomes(work):
    """tests/golden/rep.fa: three sequences with tandem and dispersed repeats and an N gap -- wide BW ranges (RowSampler,
    redundant-alignment filters, many nextAlignment candidates), reads hanging over sequence ends and across the gap
    (joinedToTextOff straddle rejection, window trimming), -k 5 so several alignments per read are printed."""
    D.build_index(os.path.join(D.GOLDEN, "rep.fa"), os.path.join(work, "rep"))
    rng = np.random.default_rng(16)
    chroms = D.genome_codes("rep.fa")
    rs, qs = [], []
    for c in chroms:
        r, q, _, _ = synth.simulate_pairs(rng, c, 400, 120)
        rs.append(r); qs.append(q)
        # reads hanging over both ends of the sequence (soft ends / trimmed windows)
        edge = np.concatenate([c[:90], c[-90:]])
        for k in range(12):
            o = 6 * k
            x = np.concatenate([rng.integers(0, 4, size=30 - 2 * k).astype(np.uint8), edge[o:o + 90 + 2 * k]])
            rs.append(np.stack([x, x[::-1].copy()])); qs.append(np.full((2, 120), ord("F"), dtype=np.uint8))
    r, q = np.concatenate(rs), np.concatenate(qs)
    f1, f2 = os.path.join(work, "rep_1.fq"), os.path.join(work, "rep_2.fq")
    D.write_fastq(f1, r[0::2], q[0::2], "rep")
    D.write_fastq(f2, r[1::2], q[1::2], "rep")
    cpu, gpu, _ = _both(work, "rep", ["-k", "5"], ["-1", f1, "-2", f2], index="rep")
    _check(cpu, gpu)
    cpu, gpu, _ = _both(work, "repl", ["--local", "--very-sensitive-local"], ["-U", f1 + "," + f2], index="rep")
    _check(cpu, gpu)


def _gunzip(work, name):
    import gzip
    import shutil
    dst = os.path.join(work, name[:-3])
    with gzip.open(os.path.join(D.GOLDEN, name), "rb") as src, open(dst, "wb") as out:
        shutil.copyfileobj(src, out)
    return dst


def test_config_c1_example_longreads(work):
    """BASELINE.json configs[0]: lambda_virus + example/reads/longreads.fq, unpaired end-to-end --sensitive (the first
    500 reads shorter than the checkpointing threshold, scripts/make_example_fixtures.py)."""
    f = _gunzip(work, "example_longreads.fq.gz")
    cpu, gpu, _ = _both(work, "c1", ["--sensitive"], ["-U", f])
    _check(cpu, gpu)


def test_config_c2_example_pairs(work):
    """BASELINE.json configs[1]: lambda_virus paired-end example/reads/reads_{1,2}.fq, end-to-end u8 path (first 1200
    pairs)."""
    f1, f2 = _gunzip(work, "example_reads_1.fq.gz"), _gunzip(work, "example_reads_2.fq.gz")
    cpu, gpu, _ = _both(work, "c2", [], ["-1", f1, "-2", f2])
    st = dict((k, int(v)) for k, v in re.findall(r"(\w+)\s+(\d+)", open(os.path.join(work, "c2_shim_stats.txt")).read()))
    assert st["pe_classify_on_gpu"] > 0 and st["other_mate_on_gpu"] > 0 and st["mate_rect_on_gpu"] > 0, st
    _check(cpu, gpu)
    # the same pairs in --local mode (the C4 policy on the C2 reads)
    cpu, gpu, _ = _both(work, "c2l", ["--local"], ["-1", f1, "-2", f2])
    _check(cpu, gpu)


def test_configs_c1_c2_whole_files_through_the_batched_server(work):
    """BASELINE.json configs[0] and [1] in full through the product path (the batched binding): all 10000 pairs of
    example/reads/reads_{1,2}.fq end-to-end and --local, and every read of longreads.fq except its 7 reads of >= 2000 rows
    (the checkpointed DP is not on the CUDA path; the server refuses those loudly), --sensitive."""
    base = os.path.join(work, "lambda")
    f1, f2 = _gunzip(work, "example_reads_1_all.fq.gz"), _gunzip(work, "example_reads_2_all.fq.gz")
    fl = _gunzip(work, "example_longreads_all.fq.gz")
    for tag, sargs, cargs, nrec in (("c2all", [], ["-1", f1, "-2", f2], 20000), ("c2lall", ["--local"], ["-1", f1, "-2", f2], 20000),
                                    ("c1all", ["--sensitive"], ["-U", fl], 5993)):
        cpu, _ = D.run(D.CPU_SERVER, base, sargs, cargs, os.path.join(work, tag + "_cpu.sam"), threads=8)
        assert len(cpu) == nrec
        st = _batched(work, tag, cpu, base, sargs, cargs, workers=512, host_threads=8)
        assert st["align_calls_on_gpu"] > 3000 and st["abi_batch_jobs"] > 2 * st["abi_batches"], st


def test_config_c3_c4_c5_shapes_on_random_genome(work):
    """BASELINE.json configs[2..4] at test scale: a 3 Mbp i.i.d. random reference in 10 sequences, 150 bp
    reads --very-sensitive (C3), 250 bp reads --local (C4), 2x150 pairs -I 0 -X 500 (C5); 4 worker threads, i.e. 4
    GPU contexts driven concurrently."""
    rng = np.random.default_rng(20261104)
    chroms = [synth.random_genome(rng, 300000) for _ in range(10)]
    fa = os.path.join(work, "rand.fa")
    with open(fa, "w") as fh:
        for i, c in enumerate(chroms):
            fh.write(">chr%d\n" % i)
            s = "".join("ACGT"[x] for x in c)
            fh.write("\n".join(s[k:k + 60000] for k in range(0, len(s), 60000)) + "\n")
    D.build_index(fa, os.path.join(work, "rand"))
    g = np.concatenate(chroms)          # reads that straddle two sequences exercise the straddle rejection
    r, q, _, _ = synth.simulate_reads(rng, g, 1500, 150)
    f = os.path.join(work, "c3.fq")
    D.write_fastq(f, r, q, "c3r")
    cpu, gpu, _ = _both(work, "c3", ["--very-sensitive"], ["-U", f], index="rand", threads=4)
    _check(cpu, gpu)
    r, q, _, _ = synth.simulate_reads(rng, g, 800, 250)
    f = os.path.join(work, "c4.fq")
    D.write_fastq(f, r, q, "c4r")
    cpu, gpu, _ = _both(work, "c4", ["--local"], ["-U", f], index="rand", threads=4)
    _check(cpu, gpu)
    r, q, _, _ = synth.simulate_pairs(rng, g, 800, 150)
    f1, f2 = os.path.join(work, "c5_1.fq"), os.path.join(work, "c5_2.fq")
    D.write_fastq(f1, r[0::2], q[0::2], "c5p")
    D.write_fastq(f2, r[1::2], q[1::2], "c5p")
    cpu, gpu, _ = _both(work, "c5", ["-I", "0", "-X", "500"], ["-1", f1, "-2", f2], index="rand", threads=4)
    _check(cpu, gpu)


# Cases restated from the reference's own regression table (scripts/test/simple_tests.pl: "Gap penalties 1-7" :3760-3875,
# "N filtering 1-3" :3877-...): tiny references (16-31 bp, far shorter than the ftab), seeds of length 2, gap budgets
# that just allow / just forbid one gap.  (reference, read, --policy string, expected offset or None, expected CIGAR)
_GP = "MMP=C30;SEED=0;SEEDLEN=2;IVAL=C,1,0;%s;MIN=L,-3,-3"
SIMPLE_CASES = [
    ("gap1", "TTGTTCGTTTGTTCGT", "TTGTTCTTTGTT", _GP % "RDG=29,10;RFG=25,15", 0, "6M1D6M"),
    ("gap2", "TTGTTCGTTTGTTCGT", "TTGTTCTTTGTT", _GP % "RDG=30,10;RFG=25,15", None, "*"),
    ("gap3", "TTGTTCGTTTGTTCGT", "TTGTTCGATTTGTT", _GP % "RDG=25,15;RFG=30,15", 0, "7M1I6M"),
    ("gap4", "TTGTTCGTTTGTTCGT", "TTGTTCGATTTGTT", _GP % "RDG=25,15;RFG=30,16", None, "*"),
    ("gap5", "TTGTTCGTTTGTTCGT", "TTGTTCGATTTGTT", _GP % "RDG=25,15;RFG=31,15", None, "*"),
    ("gap6", "ATTGTTCGTTTGTTCGTA", "ATTGTTGTTTGATTCGTA", _GP % "RDG=19,10;RFG=18,10", 0, "6M1D5M1I6M"),
    ("gap7", "TTGTTCGTTTGTTCGT", "TTGTTGTTTGATTCGT", _GP % "RDG=16,10;RFG=16,10", None, "*"),
    ("nfilt1", "GAGACTTTATACGCATCGAACTATCGCTCTA", "ATACGCATCGAAC", "NCEIL=L,0,0", 8, "13M"),
    ("nfilt2", "GAGACTTTATNCGCATCGAACTATCGCTCTA", "ATACGCATCGAAC", "NCEIL=L,0,0", None, "*"),
    ("nfilt3", "GAGACTTTATACGCATCGAANTATCGCTCTA", "ATACGCATCGAAC", "NCEIL=L,0,0", None, "*"),
]


@pytest.mark.parametrize("case", SIMPLE_CASES, ids=[c[0] for c in SIMPLE_CASES])
def test_reference_regression_cases(work, case):
    name, ref, read, policy, off, cigar = case
    fa = os.path.join(work, name + ".fa")
    with open(fa, "w") as fh:
        fh.write(">0\n%s\n" % ref)
    D.build_index(fa, os.path.join(work, name))
    fq = os.path.join(work, name + ".fq")
    with open(fq, "w") as fh:
        fh.write("@r0\n%s\n+\n%s\n" % (read, "I" * len(read)))
    base = os.path.join(work, name)
    args = ["-a", "--policy", policy]
    cpu, _ = D.run(D.CPU_SERVER, base, args, ["-U", fq], os.path.join(work, name + "_cpu.sam"), threads=1)
    gpu, _ = D.run(D.GPU_SERVER, base, args, ["-U", fq], os.path.join(work, name + "_gpu.sam"), threads=1,
                   env={"BT2GPU_SHIM_STATS": os.path.join(work, name + "_stats.txt")})
    assert cpu == gpu, (cpu, gpu)
    gpub, _ = D.run(D.GPUB_SERVER, base, args, ["-U", fq], os.path.join(work, name + "_gpub.sam"), threads=4, env={"BT2GPU_THREADS": "2"})
    assert cpu == gpub, (cpu, gpub)
    f = cpu[0].split("\t")
    assert f[5] == cigar, cpu[0]
    if off is None:
        assert f[2] == "*"
    else:
        assert int(f[3]) - 1 == off, cpu[0]


# The reference's regression table, restated (scripts/restate_simple_tests.py -> tests/golden/simple_cases.json: every
# case of scripts/test/simple_tests.pl that the stock server accepts, seed-mismatch cases (-N 1, SEED=1) included).  EVERY case runs through the batched binding (the path bench.py times); the per-call binding takes
# every fourth case (BT2GPU_ALL_SIMPLE_CASES=1: all of them).
import json

_ALL = json.load(open(os.path.join(D.GOLDEN, "simple_cases.json")))


@pytest.mark.parametrize("ci", range(len(_ALL)), ids=["%03d" % i for i in range(len(_ALL))])
def test_restated_regression_table(work, ci):
    case = _ALL[ci]
    base, sargs, cargs = D.write_case_inputs(case, work, "sc%03d" % ci)
    cpu, _ = D.run(D.CPU_SERVER, base, sargs, cargs, base + "_cpu.sam", threads=1)
    gpub, _ = D.run(D.GPUB_SERVER, base, sargs, cargs, base + "_gpub.sam", threads=4, env={"BT2GPU_THREADS": "2"})
    assert cpu == gpub, (case["name"], case["args"], cpu, gpub)
    if ci % 4 == 0 or os.environ.get("BT2GPU_ALL_SIMPLE_CASES"):
        gpu, _ = D.run(D.GPU_SERVER, base, sargs, cargs, base + "_gpu.sam", threads=1)
        assert cpu == gpu, (case["name"], case["args"], cpu, gpu)
    if "cigar" in case and "reads" in case and len(case["cigar"]) == len(case["reads"]) and not case.get("nofw"):
        got = {ln.split("\t")[0].split("/")[0]: ln.split("\t")[5] for ln in cpu}
        for i, cg in enumerate(case["cigar"]):
            if "r%d" % i in got and "-k" not in case["args"] and "-a" not in case["args"]:
                assert got["r%d" % i] == cg, (case["name"], got, case["cigar"])
