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
        srcs = [os.path.join(ROOT, "oracle", f) for f in ("dp_oracle.c", "fm_oracle.c", "frame_oracle.c")]
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


def oracle_ungapped(sc_struct, seq, qual, win, jobs):
    lib = oracle_lib()
    n = len(jobs)
    cap = int(jobs["rd_len"].sum()) + 16
    res = np.zeros(n, dtype=structs.ungap_result)
    edits = np.zeros(cap, dtype=structs.edit)
    ne = C.c_size_t(0)
    rc = lib.bt2oracle_ungapped_batch(_ptr(sc_struct), _ptr(seq), _ptr(qual), _ptr(win), _ptr(jobs), C.c_size_t(n),
                                      _ptr(res), _ptr(edits), C.c_size_t(cap), C.byref(ne))
    assert rc == 0
    return res, edits[: ne.value]


def parse_ungapped_golden(path, sc):
    """tests/golden/ungapped_*.txt (oracle/ref_dp_dump --ungapped) -> (seq, qual, win, jobs, expected)."""
    code = {c: i for i, c in enumerate("ACGTN")}
    seqs, quals, wins, jobs, exp = [], [], [], [], []
    off = 0
    for ln in open(path):
        t = ln.rstrip("\n").split("\t")
        if t[0] == "U":
            rd = np.array([code[c] for c in t[1]], dtype=np.uint8)
            seqs.append(rd); quals.append(np.frombuffer(t[2].encode(), dtype=np.uint8))
            wins.append(np.array([code[c] for c in t[9]], dtype=np.uint8))
            jobs.append((off, len(rd), off, int(t[7]), sc.nceil(len(rd)), 1 if t[3] == "+" else 0, int(t[6]),
                         int(t[5]), int(t[8])))
            off += len(rd)
        elif t[0] == "UR":
            if int(t[1]) == 1:
                ed = [] if not t[9] else [(int(a), b, c, int(d)) for a, b, c, d in (x.split(":") for x in t[9].split(";"))]
                exp.append((1, int(t[2]), int(t[3]), int(t[4]), int(t[5]), int(t[6]), int(t[7]), ed))
            else:
                exp.append((int(t[1]),))
    j = np.zeros(len(jobs), dtype=structs.ungap_job)
    a = np.array(jobs, dtype=np.int64)
    for c, name in enumerate(("rd_off", "rd_len", "rf_off", "minsc", "nceil", "fw", "ohang", "off", "reflen")):
        j[name] = a[:, c]
    return np.concatenate(seqs), np.concatenate(quals), np.concatenate(wins), j, exp


def ungapped_to_py(res, edits, i):
    r = res[i]
    if int(r["ret"]) != 1:
        return (int(r["ret"]),)
    e = edits[int(r["edit_off"]): int(r["edit_off"]) + int(r["nedits"])]
    return (1, int(r["score"]), int(r["ns"]), int(r["refns"]), int(r["refoff"]), int(r["trim5"]), int(r["trim3"]),
            [(int(x["pos"]), chr(x["chr"]), chr(x["qchr"]), int(x["type"])) for x in e])


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


# ---------------------------------------------------------------------------------------
# FM-index side: CPU oracle bindings (oracle/fm_oracle.c) and golden comparison
# ---------------------------------------------------------------------------------------
from bt2gpu import fmio, index as bt2index  # noqa: E402


def load_index_fixture(name):
    return bt2index.Index.load_npz(os.path.join(GOLDEN, name))


class FmOracle:
    """Same call surface as bt2gpu.lib.FmContext, answered by the C restatement on the CPU."""

    def __init__(self, ix, sc_struct):
        self.lib = oracle_lib()
        self.ix = ix
        self.dfw, self.dbw = structs.index_desc(ix.fw), structs.index_desc(ix.bw)
        self.sc = sc_struct

    def stage_reads(self, pool):
        self.pool = pool

    def seed_exact(self, jobs):
        out = np.zeros(len(jobs), dtype=structs.seed_hit)
        rc = self.lib.bt2oracle_seed_exact_batch(C.byref(self.dfw), C.byref(self.dbw), _ptr(self.pool.seq),
                                                 _ptr(self.pool.reads), _ptr(jobs), C.c_size_t(len(jobs)), _ptr(out))
        assert rc == 0
        return out

    def exact_sweep(self, nofw=False, norc=False, mine_max=2):
        out = np.zeros(len(self.pool), dtype=structs.sweep_result)
        rc = self.lib.bt2oracle_exact_sweep_batch(C.byref(self.dfw), _ptr(self.pool.seq), _ptr(self.pool.reads),
                                                  C.c_size_t(len(self.pool)), C.c_int(nofw), C.c_int(norc),
                                                  C.c_uint32(mine_max), _ptr(out))
        assert rc == 0
        return out

    def one_mm(self, jobs, local=False, repex=True, rep1mm=True, hit_cap=None):
        n = len(self.pool)
        hit_cap = hit_cap or 64 * n + 1024
        out = np.zeros(n, dtype=structs.onemm_result)
        hits = np.zeros(hit_cap, dtype=structs.ee_hit)
        nh = C.c_size_t(0)
        rc = self.lib.bt2oracle_onemm_batch(
            C.byref(self.dfw), C.byref(self.dbw), _ptr(self.sc), _ptr(self.pool.seq), _ptr(self.pool.qual),
            _ptr(self.pool.reads), C.c_size_t(n), _ptr(jobs), C.c_int(local),
            C.c_int(repex), C.c_int(rep1mm), _ptr(out), _ptr(hits), C.c_size_t(hit_cap), C.byref(nh))
        assert rc == 0, "oracle hit pool overflow"
        return out, hits[: nh.value]

    def seed_mm(self, jobs, mms, cap=None):
        """Multi-zone seeds (-N mms): per job (= seed offset of a read strand) the hits in the order of the reference's
        reportHit calls -> (list of [(topf, botf, topb, botb, key), ...] per job, total bwops)."""
        n = len(jobs)
        cap = cap or 256 * n + 1024
        hit_job = np.zeros(cap, dtype=np.uint32); rng_ = np.zeros(4 * cap, dtype=np.uint32); keys = np.zeros(cap, dtype=np.uint64)
        out_n = np.zeros(n, dtype=np.uint32)
        nh, ops = C.c_size_t(0), C.c_uint64(0)
        rc = self.lib.bt2oracle_seed_mm_batch(
            C.byref(self.dfw), C.byref(self.dbw), _ptr(self.sc), _ptr(self.pool.seq), _ptr(self.pool.qual), _ptr(self.pool.reads),
            _ptr(jobs), C.c_size_t(n), C.c_int(mms), _ptr(hit_job), _ptr(rng_), _ptr(keys), C.c_size_t(cap), C.byref(nh),
            _ptr(out_n), C.byref(ops))
        assert rc == 0, "oracle multi-zone seed search failed (%d)" % rc
        out, k = [], 0
        for j in range(n):
            out.append([(int(rng_[4 * i]), int(rng_[4 * i + 1]), int(rng_[4 * i + 2]), int(rng_[4 * i + 3]), int(keys[i]))
                        for i in range(k, k + int(out_n[j]))])
            k += int(out_n[j])
        assert k == nh.value
        return out, int(ops.value)

    def extend(self, jobs):
        out = np.zeros(len(jobs), dtype=structs.extend_result)
        rc = self.lib.bt2oracle_extend_batch(C.byref(self.dfw), C.byref(self.dbw), _ptr(self.pool.seq),
                                             _ptr(self.pool.reads), _ptr(jobs), C.c_size_t(len(jobs)), _ptr(out))
        assert rc == 0
        return out

    def locate(self, rows):
        rows = np.ascontiguousarray(rows, dtype=np.uint32)
        offs = np.zeros(len(rows), dtype=np.uint32)
        steps = np.zeros(len(rows), dtype=np.uint32)
        rc = self.lib.bt2oracle_locate_batch(C.byref(self.dfw), _ptr(rows), C.c_size_t(len(rows)), _ptr(offs), _ptr(steps))
        assert rc == 0
        return offs, steps


def compare_fm_with_dump(eng, recs, sc, local, seed_len, interval_fn):
    """Drive `eng` (FmOracle or the CUDA FmContext) with the reads of a ref_seed_dump golden and
    assert bit-exact equality with everything the reference printed.  Returns #checks."""
    pool = fmio.ReadPool([r["seq"] for r in recs], [r["qual"] for r in recs])
    eng.stage_reads(pool)
    n = len(recs)
    checks = 0
    # ---- exactSweep ----
    sw = eng.exact_sweep(False, False, 2)
    for i, r in enumerate(recs):
        nelt, mine_fw, mine_rc, ft, fb, rt, rb, bwops = r["SW"]
        s = sw[i]
        ctx = ("sweep", i)
        assert int(s["nedit"][0]) == mine_fw and int(s["nedit"][1]) == mine_rc, (ctx, s, r["SW"])
        got_nelt = 0
        for k, (t, b) in enumerate(((ft, fb), (rt, rb))):
            hit = (not s["done"][k]) and s["nedit"][k] == 0 and s["bot"][k] > s["top"][k]
            if hit:
                assert (int(s["top"][k]), int(s["bot"][k])) == (t, b), (ctx, k, s, r["SW"])
                got_nelt += int(s["bot"][k]) - int(s["top"][k])
            else:
                assert (t, b) == (0, 0), (ctx, k, s, r["SW"])
        assert got_nelt == nelt and int(s["bwops"]) == bwops, (ctx, s, r["SW"])
        checks += 1
    # ---- oneMmSearch ----
    jobs = np.zeros(n, dtype=structs.onemm_job)
    for i, r in enumerate(recs):
        jobs["minsc"][i] = r["M1"][7]
        jobs["nceil"][i] = sc.nceil(len(r["seq"]))
    res, hits = eng.one_mm(jobs, local, True, True)
    for i, r in enumerate(recs):
        ret, nmm, ft, fb, rt, rb, bwops, _ = r["M1"]
        h = hits[int(res["hit_off"][i]): int(res["hit_off"][i]) + int(res["nhits"][i])]
        mm = [(int(x["fw"]), int(x["top"]), int(x["bot"]), int(x["score"]), int(x["pos"]), chr(x["chr"]), chr(x["qchr"]))
              for x in h if x["pos"] != 0xFFFF]
        ex = {int(x["fw"]): (int(x["top"]), int(x["bot"])) for x in h if x["pos"] == 0xFFFF}
        ctx = ("onemm", i)
        assert mm == r["H"], (ctx, mm, r["H"])
        assert ex.get(1, (0, 0)) == (ft, fb) and ex.get(0, (0, 0)) == (rt, rb), (ctx, ex, r["M1"])
        assert int(res["bwops"][i]) == bwops, (ctx, int(res["bwops"][i]), bwops)
        checks += 1
    # ---- exact seeds ----
    sjobs, per_read = fmio.instantiate_exact_seeds(pool, seed_len, interval_fn)
    sh = eng.seed_exact(sjobs)
    by_key = {(int(j["read"]), int(j["fw"]), int(j["off"])): k for k, j in enumerate(sjobs)}
    bw_tot = np.zeros(n, dtype=np.int64)
    np.add.at(bw_tot, sjobs["read"].astype(np.int64), sh["bwops"].astype(np.int64))
    xjobs, xexp, lrows, lexp = [], [], [], []
    for i, r in enumerate(recs):
        sd = r["SD"]
        assert sd[0] == seed_len and sd[1] == per_read[i][0], (i, sd, per_read[i])
        if sd[2] == 0:
            assert not any(int(j["read"]) == i for j in sjobs[:0]), i
            continue
        assert sd[2] == per_read[i][1] and sd[3] == int(bw_tot[i]), ("seeds", i, sd, per_read[i], int(bw_tot[i]))
        for fw, offidx, off, ranges in r["S"]:
            k = by_key.get((i, fw, off))
            if k is None:
                assert ranges == [], ("seed with N", i, fw, off)
                continue
            g = sh[k]
            got = [] if g["topf"] == g["botf"] else [(int(g["topf"]), int(g["botf"]), int(g["topb"]), int(g["botb"]))]
            assert got == ranges, ("seed", i, fw, off, got, ranges)
            checks += 1
        for x in r["X"]:
            xjobs.append((i,) + x[3:7] + (x[1], x[2], x[0]))
            xexp.append(x[7:10])
        for row, off in r["L"]:
            lrows.append(row)
            lexp.append(off)
    xj = np.zeros(len(xjobs), dtype=structs.extend_job)
    if xjobs:
        a = np.array(xjobs, dtype=np.int64)
        for c, name in enumerate(("read", "topf", "botf", "topb", "botb", "off", "len", "fw")):
            xj[name] = a[:, c]
    xr = eng.extend(xj)
    for k in range(len(xjobs)):
        got = (int(xr["nlex"][k]), int(xr["nrex"][k]), int(xr["fmops"][k]))
        assert got == tuple(xexp[k]), ("extend", xjobs[k], got, xexp[k])
        checks += 1
    offs, _ = eng.locate(np.array(lrows, dtype=np.uint32))
    assert [int(x) for x in offs] == lexp, "locate"
    checks += len(lrows)
    return checks


# ---------------------------------------------------------------------------------------
# framing / paired-end geometry: golden vectors of the reference (oracle/ref_frame_dump.cpp)
# ---------------------------------------------------------------------------------------
def parse_frame_vectors(path=None):
    """-> dict kind -> list of (inputs, outputs) integer tuples; flag strings become tuples of 0/1."""
    path = path or os.path.join(GOLDEN, "frame_vectors.txt")
    out = {k: [] for k in "SMCGJ"}

    def ints(tok):
        r = []
        for t in tok:
            if len(t) == 6 and set(t) <= {"0", "1"}:
                r.append(tuple(int(c) for c in t))
            else:
                r.append(int(t))
        return r
    with open(path) as fh:
        for ln in fh:
            kind, rest = ln[0], ln[2:]
            lhs, rhs = rest.split(" -> ")
            parts = [ints(p.split()) for p in lhs.split(" | ")]
            out[kind].append((parts, ints(rhs.split())))
    return out


class FrameOracle:
    """ctypes face of oracle/frame_oracle.c."""

    def __init__(self):
        self.lib = oracle_lib()
        L = self.lib
        i64, u64, ci = C.c_int64, C.c_uint64, C.c_int
        L.bt2o_frame_seed_ext.argtypes = [i64, u64, i64, u64, u64, i64, u64, ci, C.c_void_p]
        L.bt2o_frame_mate.argtypes = [ci, i64, i64, i64, i64, u64, i64, u64, u64, i64, u64, ci, C.c_void_p]
        L.bt2o_pe_other_mate.argtypes = [C.c_void_p, ci, ci, i64, i64, u64, u64, u64] + [C.c_void_p] * 6
        L.bt2o_pe_classify.argtypes = [C.c_void_p, i64, u64, ci, i64, u64, ci]
        L.bt2o_max_read_gaps.argtypes = [i64, u64, i64, i64, i64]
        L.bt2o_max_ref_gaps.argtypes = [i64, u64, i64, i64, i64]
        L.bt2o_joined_to_text_off.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, ci, C.c_uint32, C.c_uint32, ci] + [C.c_void_p] * 4

    @staticmethod
    def policy(pol, maxfrag, minfrag, flags):
        p = np.zeros(1, dtype=structs.o_pe_policy)
        p["pol"], p["maxfrag"], p["minfrag"] = pol, maxfrag, minfrag
        for k, v in zip(("local", "flip_ok", "dovetail_ok", "contain_ok", "olap_ok", "expand_to_fit"), flags):
            p[k] = v
        return p

    @staticmethod
    def rect_tuple(r):
        r = r.reshape(-1)[0]
        return tuple(int(r[k]) for k in ("refl", "refr", "refl_pretrim", "refr_pretrim", "triml", "trimr", "corel", "corer", "maxgap"))

    def frame_seed_ext(self, off, rdlen, reflen, rdgap, rfgap, maxns, maxhalf, trim):
        r = np.zeros(1, dtype=structs.rect)
        ok = self.lib.bt2o_frame_seed_ext(off, rdlen, reflen, rdgap, rfgap, maxns, maxhalf, trim, _ptr(r))
        return ok, self.rect_tuple(r)

    def frame_mate(self, anchor_left, ll, lr, rl, rr, rdlen, reflen, rdgap, rfgap, maxns, maxhalf, trim):
        r = np.zeros(1, dtype=structs.rect)
        ok = self.lib.bt2o_frame_mate(anchor_left, ll, lr, rl, rr, rdlen, reflen, rdgap, rfgap, maxns, maxhalf, trim, _ptr(r))
        return ok, self.rect_tuple(r)

    def other_mate(self, pol, is1, fw, off, maxalcols, reflen, len1, len2):
        o = np.zeros(4, dtype=np.int64)
        fl = np.zeros(2, dtype=np.int32)
        found = self.lib.bt2o_pe_other_mate(_ptr(pol), is1, fw, off, maxalcols, reflen, len1, len2,
                                            fl[0:].ctypes.data_as(C.c_void_p), o[0:].ctypes.data_as(C.c_void_p),
                                            o[1:].ctypes.data_as(C.c_void_p), o[2:].ctypes.data_as(C.c_void_p),
                                            o[3:].ctypes.data_as(C.c_void_p), fl[1:].ctypes.data_as(C.c_void_p))
        return found, int(fl[0]), [int(x) for x in o], int(fl[1])

    def classify(self, pol, off1, len1, fw1, off2, len2, fw2):
        return self.lib.bt2o_pe_classify(_ptr(pol), off1, len1, fw1, off2, len2, fw2)

    def joined_to_text_off(self, e, qlen, off, reject):
        o = np.zeros(3, dtype=np.uint32)
        st = np.zeros(1, dtype=np.int32)
        rs = np.ascontiguousarray(e.rstarts, dtype=np.uint32)
        pl = np.ascontiguousarray(e.plen, dtype=np.uint32)
        self.lib.bt2o_joined_to_text_off(_ptr(rs), len(rs), _ptr(pl), e.len, 1 if e.fw else 0, qlen, off, reject,
                                         o[0:].ctypes.data_as(C.c_void_p), o[1:].ctypes.data_as(C.c_void_p),
                                         o[2:].ctypes.data_as(C.c_void_p), _ptr(st))
        return int(o[0]), int(o[1]), int(o[2]), int(st[0])
