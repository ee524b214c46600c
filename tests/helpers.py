"""Shared test helpers: loads the CPU oracle (oracle/libbt2oracle.so, test infrastructure only)
and the product library through ctypes, and compares result sets with reference dumps."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bowtie2-server_b200"))
from bt2gpu import structs, dpio  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
REFBIN = os.path.join(ROOT, "oracle", "_ref")


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


_oracle = None


def oracle_lib():
    """Build (if needed) and load the CPU oracle."""
    global _oracle
    if _oracle is None:
        so = os.path.join(ROOT, "oracle", "libbt2oracle.so")
        srcs = [os.path.join(ROOT, "oracle", f) for f in ("dp_oracle.c", "fm_oracle.c")]
        srcs = [s for s in srcs if os.path.exists(s)]
        if (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
            subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-o", so] + srcs)
        _oracle = C.CDLL(so)
        _oracle.bt2oracle_dp_align_batch.restype = C.c_int
    return _oracle


def oracle_dp(sc_struct, seq, qual, win, jobs, max_aln=8, cand_cap=None, edit_cap=None):
    lib = oracle_lib()
    n = len(jobs)
    cand_cap = cand_cap or max(1024, int(sum(int(j["rd_len"]) * int(j["ncol"]) for j in jobs[:0])) + 4096 * max(n, 1))
    edit_cap = edit_cap or 4096 * max(n, 1)
    results = np.zeros(n, dtype=structs.dp_result)
    cands = np.zeros(cand_cap, dtype=structs.dp_cand)
    alns = np.zeros(n * max_aln, dtype=structs.dp_aln)
    edits = np.zeros(edit_cap, dtype=structs.edit)
    nc, ne = C.c_size_t(0), C.c_size_t(0)
    rc = lib.bt2oracle_dp_align_batch(
        _ptr(sc_struct), _ptr(seq), _ptr(qual), C.c_size_t(len(seq)), _ptr(win), C.c_size_t(len(win)),
        _ptr(jobs), C.c_size_t(n), C.c_uint32(max_aln), _ptr(results), _ptr(cands), C.c_size_t(cand_cap),
        _ptr(alns), _ptr(edits), C.c_size_t(edit_cap), C.byref(nc), C.byref(ne))
    assert rc == 0, "oracle overflow"
    return results, cands[: nc.value], alns, edits[: ne.value]


class RandomSource:
    """random_source.h:33-63 (LCG); used to replay SwAligner::nextAlignment's re-seed protocol."""
    A, Cc = 1664525, 1013904223

    def __init__(self, last):
        self.last = last & 0xFFFFFFFF

    def init(self, seed):
        self.last = seed & 0xFFFFFFFF

    def next_u32(self):
        self.last = (self.A * self.last + self.Cc) & 0xFFFFFFFF
        ret = self.last >> 16
        self.last = (self.A * self.last + self.Cc) & 0xFFFFFFFF
        return ret ^ self.last

    def copy(self):
        return RandomSource(self.last)


def replay_reseed(rnd, tries, width):
    """aligner_sw.cpp:794-795,877 (u8: init(reseed) before, init(reseed+1) after) and
    :879,932 (i16: init(reseed) after) -- once per attempted backtrace."""
    for _ in range(tries):
        reseed = (rnd.next_u32() + 1) & 0xFFFFFFFF
        rnd.init(reseed + 1 if width == 8 else reseed)


def compare_with_dump(problems, results, cands, alns, edits, max_aln=8, check_cands=True):
    """Assert bit-exact equality with the reference dump; returns number of problems checked."""
    for i, p in enumerate(problems):
        R, A, E = dpio.results_to_py(results, cands, alns, edits, i)
        ref = p["R"]
        ctx = "problem %d" % p["pid"]
        assert R["aligned"] == ref["aligned"], (ctx, R, ref)
        assert R["best"] == ref["best"], (ctx, R["best"], ref["best"])
        assert R["sse8"] == ref["sse8"] and R["sse16"] == ref["sse16"], (ctx, R, ref)
        assert R["ncand"] == ref["ncand"], (ctx, R["ncand"], ref["ncand"])
        if ref["aligned"] and check_cands:
            k = len(ref["cands"])  # goldens keep only the first 64 sorted candidates
            assert R["cands"][:k] == ref["cands"], (ctx, R["cands"][:5], ref["cands"][:5])
        assert len(A) == len(p["A"]), (ctx, len(A), len(p["A"]))
        rnd = RandomSource(p["pid"] * 7919 + 17)
        width = 8 if ref["sse8"] else 16
        for a, ra in zip(A, p["A"]):
            for k in ("score", "refoff", "trim5", "trim3", "ns", "gaps", "refns", "nedits", "edits", "cural"):
                assert a[k] == ra[k], (ctx, k, a[k], ra[k])
            replay_reseed(rnd, a["tries"], width)
            assert rnd.copy().next_u32() == ra["rnd"], (ctx, "rng after alignment")
        if len(A) < max_aln:
            replay_reseed(rnd, E["tail_tries"], width)
            assert E["cural"] == p["E"]["cural"], (ctx, E, p["E"])
            assert rnd.copy().next_u32() == p["E"]["rnd"], (ctx, "rng at end")
    return len(problems)
