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
    "bt2gpu_ctx_destroy", "bt2gpu_ctx_sync", "bt2gpu_ctx_stream", "bt2gpu_ctx_launches",
    "bt2gpu_scoring_set", "bt2gpu_dp_align_batch", "bt2gpu_dp_stage", "bt2gpu_dp_run",
    "bt2gpu_dp_fetch", "bt2gpu_dp_staged_cells", "bt2gpu_dp_phase_ms",
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
