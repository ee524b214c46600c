"""numpy mirrors of the POD structs in include/bt2gpu.h (same field order, C alignment)."""
import numpy as np

dp_job = np.dtype([
    ("rd_off", "<u4"), ("rd_len", "<u4"), ("rf_off", "<u4"), ("ncol", "<u4"),
    ("minsc", "<i4"), ("nceil", "<i4"), ("refl", "<i8"),
    ("triml", "<u4"), ("corel", "<u4"), ("corer", "<u4"),
    ("fw", "u1"), ("enable8", "u1"), ("reserved", "u1", (2,)),
], align=True)
dp_result = np.dtype([
    ("best", "<i8"), ("aligned", "<i4"), ("width", "<i4"),
    ("ncand", "<u4"), ("cand_off", "<u4"), ("naln", "<u4"), ("aln_off", "<u4"),
    ("tail_tries", "<u4"), ("cural", "<u4"),
], align=True)
dp_cand = np.dtype([("row", "<u4"), ("col", "<u4"), ("score", "<i4")], align=True)
dp_aln = np.dtype([
    ("score", "<i4"), ("ns", "<i4"), ("gaps", "<i4"), ("refns", "<i4"), ("refoff", "<i8"),
    ("trim5", "<u4"), ("trim3", "<u4"), ("nedits", "<u4"), ("edit_off", "<u4"),
    ("cural", "<u4"), ("tries", "<u4"),
], align=True)
edit = np.dtype([("pos", "<u4"), ("chr", "u1"), ("qchr", "u1"), ("type", "u1"), ("pad", "u1")], align=True)
scoring = np.dtype([
    ("mmpens", "<i4", (256,)), ("npens", "<i4", (256,)),
    ("match_bonus", "<i4"), ("rd_gap_open", "<i4"), ("rd_gap_ext", "<i4"),
    ("rf_gap_open", "<i4"), ("rf_gap_ext", "<i4"), ("gapbar", "<i4"), ("monotone", "<i4"),
], align=True)

assert dp_job.itemsize == 48 and dp_result.itemsize == 40 and dp_cand.itemsize == 12
assert dp_aln.itemsize == 48 and edit.itemsize == 8 and scoring.itemsize == 2048 + 28

EDIT_READ_GAP, EDIT_REF_GAP, EDIT_MM = 1, 2, 3
INT64_MIN = -(1 << 63)

# ---- FM-index side (include/bt2gpu.h) ----
read = np.dtype([("rd_off", "<u4"), ("rd_len", "<u4")], align=True)
seed_job = np.dtype([("read", "<u4"), ("off", "<u2"), ("len", "u1"), ("fw", "u1")], align=True)
seed_hit = np.dtype([("topf", "<u4"), ("botf", "<u4"), ("topb", "<u4"), ("botb", "<u4"), ("bwops", "<u4"), ("sides", "<u4")], align=True)
mmseed_hit = np.dtype([("topf", "<u4"), ("botf", "<u4"), ("topb", "<u4"), ("botb", "<u4"), ("key", "<u8")], align=True)
mmseed_result = np.dtype([("hit_off", "<u4"), ("nhits", "<u4"), ("bwops", "<u4"), ("bwedits", "<u4"), ("ntypes", "<u4"), ("pad", "<u4")], align=True)
sweep_result = np.dtype([("top", "<u4", (2,)), ("bot", "<u4", (2,)), ("nedit", "<u2", (2,)),
                         ("done", "u1", (2,)), ("bwops", "<u4")], align=True)
onemm_job = np.dtype([("minsc", "<i4"), ("nceil", "<i4"), ("nofw", "u1"), ("norc", "u1"), ("pad", "u1", (2,))], align=True)
ee_hit = np.dtype([("top", "<u4"), ("bot", "<u4"), ("score", "<i4"), ("pos", "<u2"), ("chr", "u1"),
                   ("qchr", "u1"), ("fw", "u1"), ("pad", "u1", (3,))], align=True)
onemm_result = np.dtype([("hit_off", "<u4"), ("nhits", "<u4"), ("bwops", "<u4")], align=True)
extend_job = np.dtype([("read", "<u4"), ("topf", "<u4"), ("botf", "<u4"), ("topb", "<u4"), ("botb", "<u4"),
                       ("off", "<u2"), ("len", "u1"), ("fw", "u1")], align=True)
extend_result = np.dtype([("nlex", "<u2"), ("nrex", "<u2"), ("fmops", "<u4")], align=True)

assert read.itemsize == 8 and seed_job.itemsize == 8 and seed_hit.itemsize == 24
assert sweep_result.itemsize == 28 and onemm_job.itemsize == 12 and ee_hit.itemsize == 20
assert onemm_result.itemsize == 12 and extend_job.itemsize == 24 and extend_result.itemsize == 8


# framing / paired-end geometry (include/bt2gpu.h "framing" section; DPRect dp_framer.h:34-90)
rect = np.dtype([("refl", "<i8"), ("refr", "<i8"), ("refl_pretrim", "<i8"), ("refr_pretrim", "<i8"),
                 ("triml", "<u8"), ("trimr", "<u8"), ("corel", "<u8"), ("corer", "<u8"), ("maxgap", "<u8")], align=True)
pe_policy = np.dtype([("pol", "<i4"), ("trim_to_ref", "<i4"), ("maxfrag", "<u8"), ("minfrag", "<u8"),
                      ("local", "u1"), ("flip_ok", "u1"), ("dovetail_ok", "u1"), ("contain_ok", "u1"),
                      ("olap_ok", "u1"), ("expand_to_fit", "u1"), ("pad", "u1", (2,))], align=True)
o_pe_policy = pe_policy       # oracle/frame_oracle.c reads the same layout (trim_to_ref sits in its padding)
frame_job = np.dtype([("off", "<i8"), ("reflen", "<i8"), ("maxns", "<i8"), ("rdlen", "<u4"), ("maxrdgap", "<u4"),
                      ("maxrfgap", "<u4"), ("maxhalf", "<u4")], align=True)
mate_job = np.dtype([("off", "<i8"), ("maxalcols", "<i8"), ("reflen", "<i8"), ("maxns", "<i8"),
                     ("len1", "<u4"), ("len2", "<u4"), ("maxrdgap", "<u4"), ("maxrfgap", "<u4"), ("maxhalf", "<u4"),
                     ("is1", "u1"), ("fw", "u1"), ("pad", "u1", (2,))], align=True)
mate_frame = np.dtype([("rect", rect), ("oll", "<i8"), ("olr", "<i8"), ("orl", "<i8"), ("orr", "<i8"),
                       ("found", "u1"), ("ok", "u1"), ("oleft", "u1"), ("ofw", "u1"), ("pad", "u1", (4,))], align=True)
mate_info = np.dtype([("anchor_job", "<u4"), ("read", "<u4"), ("cls", "<i4"), ("ofw", "u1"), ("oleft", "u1"), ("pad", "u1", (2,)),
                      ("anchor_refoff", "<i8"), ("anchor_extent", "<u4"), ("mate_extent", "<u4"), ("rect", rect)], align=True)
ref_record = np.dtype([("off", "<u4"), ("len", "<u4"), ("first", "u1"), ("pad", "u1", (3,))], align=True)
pair_job = np.dtype([("off1", "<i8"), ("off2", "<i8"), ("len1", "<u4"), ("len2", "<u4"),
                     ("fw1", "u1"), ("fw2", "u1"), ("pad", "u1", (6,))], align=True)


class IndexDesc(__import__("ctypes").Structure):
    """bt2gpu_index_desc"""
    import ctypes as _C
    _fields_ = [
        ("ebwt", _C.c_void_p), ("ebwt_bytes", _C.c_uint64), ("offs", _C.c_void_p), ("n_offs", _C.c_uint64),
        ("ftab", _C.c_void_p), ("eftab", _C.c_void_p), ("fchr", _C.c_uint32 * 5), ("zoff", _C.c_uint32),
        ("len", _C.c_uint32), ("off_rate", _C.c_int32), ("ftab_chars", _C.c_int32), ("is_fw", _C.c_int32),
    ]


def index_desc(e):
    """bt2gpu_index_desc over the numpy arrays of an index.EbwtData (keeps them alive via ._keep)."""
    d = IndexDesc()
    d.ebwt = e.ebwt.ctypes.data
    d.ebwt_bytes = e.ebwt.nbytes
    if e.offs is not None:
        d.offs = e.offs.ctypes.data
        d.n_offs = len(e.offs)
    else:
        d.offs = None
        d.n_offs = 0
    d.ftab = e.ftab.ctypes.data
    d.eftab = e.eftab.ctypes.data
    for i in range(5):
        d.fchr[i] = int(e.fchr[i])
    d.zoff, d.len, d.off_rate, d.ftab_chars, d.is_fw = e.zoff, e.len, e.off_rate, e.ftab_chars, int(e.fw)
    d._keep = e
    return d


import ctypes as _ct

_HP_U64 = ("seeds", "seed_bwops", "seed_sides", "seed_hits", "sweep_bwops", "extend_fmops", "locate_rows",
           "locate_steps", "dp_jobs", "dp_cells", "aligned", "h2d_bytes", "d2h_bytes", "bt_cells", "dp_cands")
_HP_F32 = ("ms_sweep", "ms_seed", "ms_extend", "ms_locate", "ms_dp", "ms_dp_fill", "ms_dp_gather", "ms_dp_backtrace",
           "wall_sweep", "wall_inst", "wall_seed", "wall_extend", "wall_locate", "wall_frame", "wall_dp_stage", "wall_dp", "ms_total", "ms_glue")


class HotpathParams(_ct.Structure):
    """bt2gpu_hotpath_params"""
    _fields_ = [("seed_len", _ct.c_int32), ("ival_const", _ct.c_double), ("ival_coeff", _ct.c_double),
                ("maxgap", _ct.c_int32), ("max_aln", _ct.c_uint32), ("mine_max", _ct.c_uint32), ("pair_boost", _ct.c_int32),
                ("cand_per_job", _ct.c_uint32)]


class HotpathStats(_ct.Structure):
    """bt2gpu_hotpath_stats"""
    _fields_ = [(k, _ct.c_uint64) for k in _HP_U64] + [(k, _ct.c_float) for k in _HP_F32]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


_PS_U64 = ("mate_jobs", "mate_cells", "mate_aligned", "concordant", "bt_cells", "dp_cands", "d2h_bytes")
_PS_F32 = ("ms_frame", "ms_dp", "ms_dp_fill", "ms_dp_gather", "ms_dp_backtrace", "ms_total")


class PairsStats(_ct.Structure):
    """bt2gpu_pairs_stats"""
    _fields_ = [(k, _ct.c_uint64) for k in _PS_U64] + [(k, _ct.c_float) for k in _PS_F32]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}

# SwAligner::ungappedAlign (bt2gpu_ungapped_batch)
ungap_job = np.dtype([("rd_off", "<u4"), ("rd_len", "<u4"), ("rf_off", "<u4"), ("minsc", "<i4"), ("nceil", "<i4"),
                      ("fw", "u1"), ("ohang", "u1"), ("pad", "u1", (2,)), ("off", "<i8"), ("reflen", "<i8")], align=True)
ungap_result = np.dtype([("ret", "<i4"), ("score", "<i4"), ("ns", "<i4"), ("refns", "<i4"), ("refoff", "<i8"),
                         ("trim5", "<u4"), ("trim3", "<u4"), ("nedits", "<u4"), ("edit_off", "<u4")], align=True)
assert ungap_job.itemsize == 40 and ungap_result.itemsize == 40

mate_rect_job = np.dtype([("ll", "<i8"), ("lr", "<i8"), ("rl", "<i8"), ("rr", "<i8"), ("reflen", "<i8"), ("maxns", "<i8"),
                          ("rdlen", "<u4"), ("maxrdgap", "<u4"), ("maxrfgap", "<u4"), ("maxhalf", "<u4"),
                          ("anchor_left", "u1"), ("pad", "u1", (7,))], align=True)
assert mate_rect_job.itemsize == 72
