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
