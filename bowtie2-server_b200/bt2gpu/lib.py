"""ctypes binding of libbt2gpu.so (C ABI: include/bt2gpu.h).  Fails loudly when the CUDA
library is missing or no device is usable -- there is no CPU path in the product."""
import ctypes as C
import os

import numpy as np

from . import structs

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "libbt2gpu.so")

# every symbol include/bt2gpu.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "bt2gpu_abi_version", "bt2gpu_last_error", "bt2gpu_device_count", "bt2gpu_ctx_create",
    "bt2gpu_ctx_destroy", "bt2gpu_ctx_sync", "bt2gpu_ctx_stream", "bt2gpu_ctx_launches", "bt2gpu_ctx_kernel_ms",
    "bt2gpu_scoring_set", "bt2gpu_dp_align_batch", "bt2gpu_dp_stage", "bt2gpu_dp_run",
    "bt2gpu_dp_fetch", "bt2gpu_dp_staged_cells", "bt2gpu_dp_phase_ms",
    "bt2gpu_index_create", "bt2gpu_index_destroy", "bt2gpu_index_bytes", "bt2gpu_index_gather_probe",
    "bt2gpu_reads_stage",
    "bt2gpu_seed_exact_batch", "bt2gpu_exact_sweep_batch", "bt2gpu_onemm_batch", "bt2gpu_extend_batch",
    "bt2gpu_locate_batch", "bt2gpu_seed_exact_stage", "bt2gpu_seed_exact_run", "bt2gpu_seed_exact_fetch",
    "bt2gpu_locate_stage", "bt2gpu_locate_run", "bt2gpu_locate_fetch", "bt2gpu_hotpath_round0_batch",
    "bt2gpu_ref_create", "bt2gpu_ref_destroy", "bt2gpu_ref_bytes", "bt2gpu_probe_random_gather", "bt2gpu_probe_dpx", "bt2gpu_dp_walked_cells",
    "bt2gpu_frame_seed_ext_batch", "bt2gpu_frame_mate_batch", "bt2gpu_pe_classify_batch", "bt2gpu_max_gaps_batch",
    "bt2gpu_joined_to_text_off_batch", "bt2gpu_hotpath_pairs_batch", "bt2gpu_ungapped_batch", "bt2gpu_frame_mate_rect_batch",
    "bt2gpu_device_blocking_sync", "bt2gpu_device_sync_mode", "bt2gpu_ctx_set_wait", "bt2gpu_seed_mm_batch", "bt2gpu_ctx_set_timing", "bt2gpu_noalloc_scratch_bytes", "bt2gpu_ungapped_batch_noalloc", "bt2gpu_ref_create_records", "bt2gpu_ref_create_bitpairs", "bt2gpu_hotpath_job_tidx",
]

_lib = None


class Bt2GpuError(RuntimeError):
    pass


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise Bt2GpuError("libbt2gpu.so not built (run __graft_entry__.build()); no CPU fallback exists")
        lib = C.CDLL(LIB_PATH)
        lib.bt2gpu_last_error.restype = C.c_char_p
        lib.bt2gpu_ctx_stream.restype = C.c_void_p
        lib.bt2gpu_ctx_launches.restype = C.c_uint64
        lib.bt2gpu_dp_staged_cells.restype = C.c_uint64
        lib.bt2gpu_dp_walked_cells.restype = C.c_uint64
        lib.bt2gpu_index_bytes.restype = C.c_uint64
        lib.bt2gpu_ref_bytes.restype = C.c_uint64
        lib.bt2gpu_noalloc_scratch_bytes.restype = C.c_uint64
        lib.bt2gpu_ref_destroy.restype = None
        lib.bt2gpu_index_bytes.argtypes = [C.c_void_p]
        lib.bt2gpu_index_destroy.argtypes = [C.c_void_p]
        lib.bt2gpu_index_destroy.restype = None
        for name in SYMBOLS:
            getattr(lib, name)
        _lib = lib
    return _lib


def _ptr(a):
    if a is None:
        return C.c_void_p(0)
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data_as(C.c_void_p)
    return C.c_void_p(int(a))  # raw address (e.g. a pinned torch tensor's data_ptr())


def _check(rc):
    if rc != 0:
        raise Bt2GpuError(load().bt2gpu_last_error().decode())


class Context:
    """One bt2gpu_ctx (one CUDA stream + scratch arenas); one per host thread."""

    def __init__(self, device=0):
        self.lib = load()
        self._h = C.c_void_p(0)
        _check(self.lib.bt2gpu_ctx_create(C.c_int(device), C.byref(self._h)))
        self.device = device

    def close(self):
        if self._h:
            self.lib.bt2gpu_ctx_destroy(self._h)
            self._h = C.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        return self.lib.bt2gpu_ctx_stream(self._h)

    @property
    def launches(self):
        return int(self.lib.bt2gpu_ctx_launches(self._h))

    def sync(self):
        _check(self.lib.bt2gpu_ctx_sync(self._h))

    def kernel_ms(self):
        """Device time of the kernels of the last compute call (CUDA events on the ctx stream)."""
        ms = C.c_float(0)
        _check(self.lib.bt2gpu_ctx_kernel_ms(self._h, C.byref(ms)))
        return float(ms.value)

    def set_wait(self, poll_ns):
        """bt2gpu_ctx_set_wait: 0 = wait in the driver, > 0 = poll the stream and sleep poll_ns between polls."""
        _check(self.lib.bt2gpu_ctx_set_wait(self._h, C.c_uint32(poll_ns)))

    def set_timing(self, on):
        """bt2gpu_ctx_set_timing: whether the entry points record the events behind kernel_ms / dp_phase_ms."""
        _check(self.lib.bt2gpu_ctx_set_timing(self._h, C.c_int(1 if on else 0)))

    def set_scoring(self, sc):
        s = sc.to_struct() if hasattr(sc, "to_struct") else sc
        _check(self.lib.bt2gpu_scoring_set(self._h, _ptr(s)))

    # ---- DP -------------------------------------------------------------------------------
    def dp_stage(self, seq, qual, win, jobs, max_aln=8, cand_cap=None, edit_cap=None,
                 pool_len=None, win_len=None, njobs=None):
        njobs = len(jobs) if njobs is None else njobs
        pool_len = len(seq) if pool_len is None else pool_len
        win_len = len(win) if win_len is None else win_len
        self._dp_n, self._dp_k = njobs, max_aln
        self._cand_cap = cand_cap if cand_cap is not None else 512 * max(njobs, 1)
        self._edit_cap = edit_cap if edit_cap is not None else 256 * max(njobs, 1)
        _check(self.lib.bt2gpu_dp_stage(
            self._h, _ptr(seq), _ptr(qual), C.c_size_t(pool_len), _ptr(win), C.c_size_t(win_len),
            _ptr(jobs), C.c_size_t(njobs), C.c_uint32(max_aln),
            C.c_size_t(self._cand_cap), C.c_size_t(self._edit_cap)))

    def dp_run(self, phases=7):
        _check(self.lib.bt2gpu_dp_run(self._h, C.c_uint32(phases)))

    def dp_fetch(self, want_cands=True, out=None):
        n, k = self._dp_n, self._dp_k
        if out is None:
            results = np.zeros(n, dtype=structs.dp_result)
            cands = np.zeros(self._cand_cap, dtype=structs.dp_cand) if want_cands else None
            alns = np.zeros(n * k, dtype=structs.dp_aln)
            edits = np.zeros(self._edit_cap, dtype=structs.edit)
        else:
            results, cands, alns, edits = out
        nc, ne = C.c_size_t(0), C.c_size_t(0)
        _check(self.lib.bt2gpu_dp_fetch(self._h, _ptr(results), _ptr(cands), _ptr(alns), _ptr(edits),
                                        C.byref(nc), C.byref(ne)))
        if out is None:
            if cands is not None:
                cands = cands[: nc.value]
            edits = edits[: ne.value]
        return results, cands, alns, edits, nc.value, ne.value

    def dp_phase_ms(self):
        ms = (C.c_float * 3)()
        _check(self.lib.bt2gpu_dp_phase_ms(self._h, ms))
        return [float(x) for x in ms]

    def dp_cells(self):
        return int(self.lib.bt2gpu_dp_staged_cells(self._h))

    def dp_align_batch(self, seq, qual, win, jobs, max_aln=8, cand_cap=None, edit_cap=None):
        """initRef + align + nextAlignment* for a whole batch (host buffers in, host arrays out)."""
        self.dp_stage(seq, qual, win, jobs, max_aln, cand_cap, edit_cap)
        self.dp_run(7)
        r = self.dp_fetch()
        return r[0], r[1], r[2], r[3]

    def dp_align_batch_abi(self, seq, qual, win, jobs, max_aln=8, cand_cap=None, edit_cap=None):
        """The same through the single entry point bt2gpu_dp_align_batch (what the server shim calls); pools are
        returned whole: results[i].cand_off / aln_off and alns[k].edit_off index them."""
        n = len(jobs)
        cand_cap = cand_cap if cand_cap is not None else 512 * max(n, 1)
        edit_cap = edit_cap if edit_cap is not None else 256 * max(n, 1)
        results = np.zeros(n, dtype=structs.dp_result)
        cands = np.zeros(cand_cap, dtype=structs.dp_cand)
        alns = np.zeros(n * max_aln, dtype=structs.dp_aln)
        edits = np.zeros(edit_cap, dtype=structs.edit)
        _check(self.lib.bt2gpu_dp_align_batch(
            self._h, _ptr(seq), _ptr(qual), C.c_size_t(len(seq)), _ptr(win), C.c_size_t(len(win)), _ptr(jobs), C.c_size_t(n),
            C.c_uint32(max_aln), _ptr(results), _ptr(cands), C.c_size_t(cand_cap), _ptr(alns), _ptr(edits), C.c_size_t(edit_cap)))
        return results, cands, alns, edits

    def ungapped_batch(self, seq, qual, win, jobs, edit_cap=None):
        """bt2gpu_ungapped_batch: SwAligner::ungappedAlign for a batch -> (results, edits)."""
        n = len(jobs)
        edit_cap = edit_cap if edit_cap is not None else int(jobs["rd_len"].sum()) + 16
        res = np.zeros(n, dtype=structs.ungap_result)
        edits = np.zeros(edit_cap, dtype=structs.edit)
        ne = C.c_size_t(0)
        _check(self.lib.bt2gpu_ungapped_batch(self._h, _ptr(seq), _ptr(qual), C.c_size_t(len(seq)), _ptr(win),
                                              C.c_size_t(len(win)), _ptr(jobs), C.c_size_t(n), _ptr(res), _ptr(edits),
                                              C.c_size_t(edit_cap), C.byref(ne)))
        return res, edits[: ne.value]

    # ---- FM index (seed side) ---------------------------------------------------------------
    def ungapped_batch_noalloc(self, seq, qual, win, jobs, edit_cap=None):
        """bt2gpu_ungapped_batch_noalloc: the same call with its device scratch in the library's static array (no
        cudaMalloc); fails for batches beyond BT2GPU_NOALLOC_SCRATCH_BYTES."""
        n = len(jobs)
        edit_cap = edit_cap if edit_cap is not None else int(jobs["rd_len"].sum()) + 16
        res = np.zeros(n, dtype=structs.ungap_result)
        edits = np.zeros(edit_cap, dtype=structs.edit)
        ne = C.c_size_t(0)
        _check(self.lib.bt2gpu_ungapped_batch_noalloc(self._h, _ptr(seq), _ptr(qual), C.c_size_t(len(seq)), _ptr(win),
                                                      C.c_size_t(len(win)), _ptr(jobs), C.c_size_t(n), _ptr(res), _ptr(edits),
                                                      C.c_size_t(edit_cap), C.byref(ne)))
        return res, edits[: ne.value]

    def stage_reads(self, pool, quals=True):
        """bt2gpu_reads_stage: pool = fmio.ReadPool (replaces the per-call `const Read&`); quals=False stages no
        qualities (NULL qual_pool: enough for every call but one_mm)."""
        self._pool_n = len(pool)
        _check(self.lib.bt2gpu_reads_stage(self._h, _ptr(pool.seq), _ptr(pool.qual) if quals else None, C.c_size_t(len(pool.seq)),
                                           _ptr(pool.reads), C.c_size_t(len(pool))))

    def bind_index(self, dev_index):
        """Forward + mirror DeviceIndex pair used by the seed-side calls."""
        self._ix = dev_index

    def seed_exact(self, jobs):
        out = np.zeros(len(jobs), dtype=structs.seed_hit)
        _check(self.lib.bt2gpu_seed_exact_batch(self._h, self._ix.fw, self._ix.bw, _ptr(jobs),
                                                C.c_size_t(len(jobs)), _ptr(out)))
        return out

    def seed_mm(self, jobs, mms, hit_cap=None):
        """bt2gpu_seed_mm_batch: multi-zone seeds (-N mms) -> (results[njobs], hits in job order; results[j].hit_off /
        nhits slice them)."""
        n = len(jobs)
        hit_cap = hit_cap or 256 * n + 1024
        out = np.zeros(n, dtype=structs.mmseed_result)
        hits = np.zeros(hit_cap, dtype=structs.mmseed_hit)
        nh = C.c_size_t(0)
        _check(self.lib.bt2gpu_seed_mm_batch(self._h, self._ix.fw, self._ix.bw, _ptr(jobs), C.c_size_t(n), C.c_int(mms),
                                             _ptr(out), _ptr(hits), C.c_size_t(hit_cap), C.byref(nh)))
        return out, hits[: nh.value]

    def seed_exact_stage(self, jobs):
        self._n_seed = len(jobs)
        _check(self.lib.bt2gpu_seed_exact_stage(self._h, _ptr(jobs), C.c_size_t(len(jobs))))

    def seed_exact_run(self):
        _check(self.lib.bt2gpu_seed_exact_run(self._h, self._ix.fw, self._ix.bw))

    def seed_exact_fetch(self, out=None):
        out = np.zeros(self._n_seed, dtype=structs.seed_hit) if out is None else out
        _check(self.lib.bt2gpu_seed_exact_fetch(self._h, _ptr(out)))
        return out

    def exact_sweep(self, nofw=False, norc=False, mine_max=2):
        out = np.zeros(self._pool_n, dtype=structs.sweep_result)
        _check(self.lib.bt2gpu_exact_sweep_batch(self._h, self._ix.fw, C.c_int(nofw), C.c_int(norc),
                                                 C.c_uint32(mine_max), _ptr(out)))
        return out

    def one_mm(self, jobs, local=False, repex=True, rep1mm=True, hit_cap=None):
        n = self._pool_n
        hit_cap = hit_cap or 64 * n + 1024
        out = np.zeros(n, dtype=structs.onemm_result)
        hits = np.zeros(hit_cap, dtype=structs.ee_hit)
        nh = C.c_size_t(0)
        _check(self.lib.bt2gpu_onemm_batch(self._h, self._ix.fw, self._ix.bw, _ptr(jobs), C.c_int(local),
                                           C.c_int(repex), C.c_int(rep1mm), _ptr(out), _ptr(hits),
                                           C.c_size_t(hit_cap), C.byref(nh)))
        return out, hits[: nh.value]

    def extend(self, jobs):
        out = np.zeros(len(jobs), dtype=structs.extend_result)
        _check(self.lib.bt2gpu_extend_batch(self._h, self._ix.fw, self._ix.bw, _ptr(jobs),
                                            C.c_size_t(len(jobs)), _ptr(out)))
        return out

    def locate(self, rows):
        rows = np.ascontiguousarray(rows, dtype=np.uint32)
        offs = np.zeros(len(rows), dtype=np.uint32)
        steps = np.zeros(len(rows), dtype=np.uint32)
        _check(self.lib.bt2gpu_locate_batch(self._h, self._ix.fw, _ptr(rows), C.c_size_t(len(rows)),
                                            _ptr(offs), _ptr(steps)))
        return offs, steps

    def gather_probe(self, which="fw", nloads=1 << 28):
        """Random 64-byte gather bandwidth (GB/s) over the bound index (roofline denominator)."""
        g = C.c_double(0)
        h = self._ix.fw if which == "fw" else self._ix.bw
        _check(self.lib.bt2gpu_index_gather_probe(self._h, h, C.c_uint64(nloads), C.byref(g)))
        return float(g.value)

    def probe_random_gather(self, table_bytes=8 << 30, nloads=1 << 27):
        """GB/s of independent random 64-byte gathers over a scratch table far larger than L2."""
        g = C.c_double(0)
        _check(self.lib.bt2gpu_probe_random_gather(self._h, C.c_uint64(table_bytes), C.c_uint64(nloads), C.byref(g)))
        return g.value

    def probe_dpx(self):
        """G thread-instructions/s of packed s16x2 DPX (VIADDMNMX / VIMNMX3)."""
        g = C.c_double(0)
        _check(self.lib.bt2gpu_probe_dpx(self._h, C.byref(g)))
        return g.value

    def hotpath_round0(self, ref, seq, qual, reads, minsc, nceil, seed_len=22, ival=(1.0, 1.15), maxgap=15,
                       max_aln=2, mine_max=2, nthreads=0, out=None, cand_per_job=0):
        """bt2gpu_hotpath_round0_batch: exactSweep -> seeds -> extend -> pick -> locate -> frame -> DP
        for a batch of reads, every intermediate resident in HBM (ref = DeviceRef).  Returns
        (out dict, stats dict)."""
        n = len(reads)
        if out is None:
            out = dict(sweep=np.zeros(n, dtype=structs.sweep_result), job_read=np.zeros(n, dtype=np.uint32),
                       job_start=np.zeros(n, dtype=np.int64), job_fw=np.zeros(n, dtype=np.uint8),
                       results=np.zeros(n, dtype=structs.dp_result), alns=np.zeros(n * max_aln, dtype=structs.dp_aln),
                       edits=np.zeros(64 * n + 1024, dtype=structs.edit))
        P = structs.HotpathParams(seed_len, ival[0], ival[1], maxgap, max_aln, mine_max, 0, cand_per_job)
        S = structs.HotpathStats()
        nj, ne = C.c_size_t(0), C.c_size_t(0)
        _check(self.lib.bt2gpu_hotpath_round0_batch(
            self._h, self._ix.fw, self._ix.bw, ref.handle, _ptr(seq), _ptr(qual),
            C.c_size_t(len(seq)), _ptr(reads), C.c_size_t(n), _ptr(minsc), _ptr(nceil), C.byref(P),
            _ptr(out["sweep"]), _ptr(out["job_read"]), _ptr(out["job_start"]), _ptr(out["job_fw"]), C.byref(nj),
            _ptr(out["results"]), _ptr(out["alns"]), _ptr(out["edits"]), C.c_size_t(len(out["edits"])), C.byref(ne),
            C.byref(S)))
        out["njobs"], out["nedits"] = nj.value, ne.value
        # the batch stays staged on the device: dp_fetch() can read it again (with the candidate pool)
        self._dp_n, self._dp_k = nj.value, max_aln
        self._cand_cap, self._edit_cap = 4096 * max(nj.value, 1), len(out["edits"])
        return out, S.as_dict()

    def hotpath_job_tidx(self, njobs):
        """Reference sequence of every DP job of the last hotpath_* call (job_start / refoff are offsets inside it)."""
        t = np.zeros(max(njobs, 1), dtype=np.uint32)
        _check(self.lib.bt2gpu_hotpath_job_tidx(self._h, _ptr(t)))
        return t[:njobs]

    def hotpath_pairs(self, ref, seq, qual, reads, minsc, nceil, pol, maxhalf=15, seed_len=22, ival=(1.0, 1.15), maxgap=15,
                      max_aln=2, mine_max=2, out=None, cand_per_job=0):
        """bt2gpu_hotpath_pairs_batch: the anchor pass of hotpath_round0 over reads[2i] = mate 1, reads[2i+1] = mate 2,
        then otherMate -> frameFindMateRect -> DP of the opposite mate -> peClassifyPair for every anchor alignment.
        Returns (out dict, anchor-pass stats, mate-pass stats)."""
        n = len(reads)
        if out is None:
            out = dict(sweep=np.zeros(n, dtype=structs.sweep_result), job_read=np.zeros(n, dtype=np.uint32),
                       job_start=np.zeros(n, dtype=np.int64), job_fw=np.zeros(n, dtype=np.uint8),
                       results=np.zeros(n, dtype=structs.dp_result), alns=np.zeros(n * max_aln, dtype=structs.dp_aln),
                       edits=np.zeros(64 * n + 1024, dtype=structs.edit),
                       mates=np.zeros(n, dtype=structs.mate_info), mresults=np.zeros(n, dtype=structs.dp_result),
                       malns=np.zeros(n * max_aln, dtype=structs.dp_aln), medits=np.zeros(64 * n + 1024, dtype=structs.edit))
        # both mates are searched: the seed interval gets the reference's 20 % paired boost (bt2_search.cpp:3392-3395)
        P = structs.HotpathParams(seed_len, ival[0], ival[1], maxgap, max_aln, mine_max, 1, cand_per_job)
        S, PS = structs.HotpathStats(), structs.PairsStats()
        nj, ne, nm, nme = C.c_size_t(0), C.c_size_t(0), C.c_size_t(0), C.c_size_t(0)
        _check(self.lib.bt2gpu_hotpath_pairs_batch(
            self._h, self._ix.fw, self._ix.bw, ref.handle, _ptr(seq), _ptr(qual),
            C.c_size_t(len(seq)), _ptr(reads), C.c_size_t(n), _ptr(minsc), _ptr(nceil), C.byref(P), _ptr(pol), C.c_uint32(maxhalf),
            _ptr(out["sweep"]), _ptr(out["job_read"]), _ptr(out["job_start"]), _ptr(out["job_fw"]), C.byref(nj),
            _ptr(out["results"]), _ptr(out["alns"]), _ptr(out["edits"]), C.c_size_t(len(out["edits"])), C.byref(ne),
            C.byref(S),
            _ptr(out["mates"]), C.byref(nm), _ptr(out["mresults"]), _ptr(out["malns"]), _ptr(out["medits"]),
            C.c_size_t(len(out["medits"])), C.byref(nme), C.byref(PS)))
        out["njobs"], out["nedits"], out["nmates"], out["nmedits"] = nj.value, ne.value, nm.value, nme.value
        self._dp_n, self._dp_k = nm.value, max_aln          # the mate pass is what stays staged
        self._cand_cap, self._edit_cap = 4096 * max(nm.value, 1), len(out["medits"])
        return out, S.as_dict(), PS.as_dict()

    # ---- framing / paired-end geometry (frame.cu) ----
    def frame_seed_ext(self, jobs, trim_to_ref=True):
        """DynProgFramer::frameSeedExtensionRect per job -> (rects, ok)."""
        jobs = np.ascontiguousarray(jobs, dtype=structs.frame_job)
        rects = np.zeros(len(jobs), dtype=structs.rect)
        ok = np.zeros(len(jobs), dtype=np.uint8)
        _check(self.lib.bt2gpu_frame_seed_ext_batch(self._h, C.c_int(int(trim_to_ref)), _ptr(jobs), C.c_size_t(len(jobs)),
                                                    _ptr(rects), _ptr(ok)))
        return rects, ok

    def frame_mate(self, pol, jobs):
        """PairedEndPolicy::otherMate + DynProgFramer::frameFindMateRect per anchor -> mate_frame records."""
        jobs = np.ascontiguousarray(jobs, dtype=structs.mate_job)
        out = np.zeros(len(jobs), dtype=structs.mate_frame)
        _check(self.lib.bt2gpu_frame_mate_batch(self._h, _ptr(pol), _ptr(jobs), C.c_size_t(len(jobs)), _ptr(out)))
        return out

    def frame_mate_rect(self, trim_to_ref, jobs):
        """DynProgFramer::frameFindMateRect for given otherMate windows -> (rects, ok)."""
        jobs = np.ascontiguousarray(jobs, dtype=structs.mate_rect_job)
        rects = np.zeros(len(jobs), dtype=structs.rect)
        ok = np.zeros(len(jobs), dtype=np.uint8)
        _check(self.lib.bt2gpu_frame_mate_rect_batch(self._h, C.c_int(1 if trim_to_ref else 0), _ptr(jobs),
                                                     C.c_size_t(len(jobs)), _ptr(rects), _ptr(ok)))
        return rects, ok

    def pe_classify(self, pol, jobs):
        jobs = np.ascontiguousarray(jobs, dtype=structs.pair_job)
        cls = np.zeros(len(jobs), dtype=np.int32)
        _check(self.lib.bt2gpu_pe_classify_batch(self._h, _ptr(pol), _ptr(jobs), C.c_size_t(len(jobs)), _ptr(cls)))
        return cls

    def max_gaps(self, minsc, rdlen):
        minsc = np.ascontiguousarray(minsc, dtype=np.int64)
        rdlen = np.ascontiguousarray(rdlen, dtype=np.uint32)
        a = np.zeros(len(minsc), dtype=np.int32)
        b = np.zeros(len(minsc), dtype=np.int32)
        _check(self.lib.bt2gpu_max_gaps_batch(self._h, _ptr(minsc), _ptr(rdlen), C.c_size_t(len(minsc)), _ptr(a), _ptr(b)))
        return a, b

    def joined_to_text_off(self, e, qlen, off, reject_straddle=True):
        """Ebwt::joinedToTextOff over EbwtData `e` -> (tidx, textoff, tlen, straddled)."""
        rs = np.ascontiguousarray(e.rstarts, dtype=np.uint32)
        pl = np.ascontiguousarray(e.plen, dtype=np.uint32)
        qlen = np.ascontiguousarray(qlen, dtype=np.uint32)
        off = np.ascontiguousarray(off, dtype=np.uint32)
        n = len(off)
        o = [np.zeros(n, dtype=np.uint32) for _ in range(3)]
        st = np.zeros(n, dtype=np.uint8)
        _check(self.lib.bt2gpu_joined_to_text_off_batch(self._h, _ptr(rs), C.c_uint32(len(rs)), _ptr(pl), C.c_uint32(len(pl)),
                                                        C.c_uint32(e.len), C.c_int(1 if e.fw else 0), _ptr(qlen), _ptr(off),
                                                        C.c_size_t(n), C.c_int(int(reject_straddle)),
                                                        _ptr(o[0]), _ptr(o[1]), _ptr(o[2]), _ptr(st)))
        return o[0], o[1], o[2], st

    def locate_stage(self, rows):
        rows = np.ascontiguousarray(rows, dtype=np.uint32)
        self._n_rows = len(rows)
        _check(self.lib.bt2gpu_locate_stage(self._h, _ptr(rows), C.c_size_t(len(rows))))

    def locate_run(self):
        _check(self.lib.bt2gpu_locate_run(self._h, self._ix.fw))

    def locate_fetch(self):
        offs = np.zeros(self._n_rows, dtype=np.uint32)
        steps = np.zeros(self._n_rows, dtype=np.uint32)
        _check(self.lib.bt2gpu_locate_fetch(self._h, _ptr(offs), _ptr(steps)))
        return offs, steps


class DeviceIndex:
    """Forward + mirror Ebwt resident in HBM (bt2gpu_index_create x2); immutable, shareable by
    every Context of the same device (mirrors the shared read-only Ebwt objects of
    bt2_search.cpp:4870-4892)."""

    def __init__(self, ix, device=0):
        self.lib = load()
        self.fw, self.bw = C.c_void_p(0), C.c_void_p(0)
        for e, h in ((ix.fw, self.fw), (ix.bw, self.bw)):
            d = structs.index_desc(e)
            _check(self.lib.bt2gpu_index_create(C.c_int(device), C.byref(d), C.byref(h)))
        self.device = device

    @property
    def nbytes(self):
        return int(self.lib.bt2gpu_index_bytes(self.fw)) + int(self.lib.bt2gpu_index_bytes(self.bw))

    def close(self):
        for h in (self.fw, self.bw):
            if h:
                self.lib.bt2gpu_index_destroy(h)
        self.fw, self.bw = C.c_void_p(0), C.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DeviceRef:
    """The joined reference text, 2 bits per base, resident in HBM (bt2gpu_ref_create): the device
    counterpart of BitPairReference (reference.h:180) that SwAligner::initRef reads windows from."""

    def __init__(self, text, device=0, recs=None):
        """text: the joined unambiguous text (codes 0..3).  recs: the `.3.bt2` RefRecords [(off, len, first)]
        (build.build_index's extras["recs"]); None = one gap-free sequence."""
        self.lib = load()
        self.handle = C.c_void_p(0)
        text = np.ascontiguousarray(text, dtype=np.uint8)
        if recs is None:
            _check(self.lib.bt2gpu_ref_create(C.c_int(device), _ptr(text), C.c_uint64(len(text)), C.byref(self.handle)))
        else:
            r = np.zeros(len(recs), dtype=structs.ref_record)
            for i, (off, ln, first) in enumerate(recs):
                r[i]["off"], r[i]["len"], r[i]["first"] = off, ln, first
            _check(self.lib.bt2gpu_ref_create_records(C.c_int(device), _ptr(text), C.c_uint64(len(text)), _ptr(r),
                                                      C.c_uint64(len(r)), C.byref(self.handle)))
        self.length = len(text)

    @property
    def nbytes(self):
        return int(self.lib.bt2gpu_ref_bytes(self.handle))

    def close(self):
        if self.handle:
            self.lib.bt2gpu_ref_destroy(self.handle)
        self.handle = C.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- what the current kernels do per DP cell (bench.py's roofline inputs) ----
def dpx_per_pair():
    """DPX instructions k_dp_fill issues per packed pair of cells."""
    return 5.0


def fill_bytes_per_cell(local):
    """bytes k_dp_fill stores per cell (E, F, H): byte records on the u8 end-to-end path, s16 records otherwise."""
    return 6.0 if local else 3.0


def backtrace_bytes_per_cell(local):
    return 6.0 if local else 3.0
