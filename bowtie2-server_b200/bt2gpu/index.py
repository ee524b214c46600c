"""Host-side loader of the reference's small (32-bit) Bowtie 2 index files.

Mirrors Ebwt::readIntoMemory (bt2_io.cpp:131-614) and EbwtParams::init (bt2_idx.h:133-167):
  <base>.1.bt2 / <base>.rev.1.bt2 : u32 endian hint (1), u32 len, i32 lineRate, i32 linesPerSide,
      i32 offRate, i32 ftabChars, i32 flags, u32 nPat, plen[nPat], u32 nFrag, rstarts[3*nFrag],
      ebwt[numSides*64], u32 zOff, fchr[5], ftab[4^ftabChars+1], eftab[2*ftabChars], names...
  <base>.2.bt2 : u32 endian hint, offs[offsLen]            (forward index only is needed)
  <base>.3.bt2 / .4.bt2 : RefRecords + 2-bit reference     (reference.cpp:40-300)
Arrays are returned as numpy views in the file's own layout: the 64-byte side layout
(48 B of 2-bit BWT + u32 occ[A,C,G,T]) is uploaded to the GPU verbatim.
"""
import os

import numpy as np

EBWT_ENTIRE_REV = 4


class EbwtData:
    """One Ebwt (forward: fw=True, or mirror: fw=False) as plain arrays."""

    def __init__(self, fw):
        self.fw = fw

    # EbwtParams::init (bt2_idx.h:133-167), small index: OFF_SIZE = 4
    def derive(self):
        self.bwt_len = self.len + 1
        self.side_sz = 1 << self.line_rate
        self.side_bwt_sz = self.side_sz - 16
        self.side_bwt_len = self.side_bwt_sz * 4
        bwt_sz = self.len // 4 + 1
        self.num_sides = (bwt_sz + self.side_bwt_sz - 1) // self.side_bwt_sz
        self.ebwt_tot_len = self.num_sides * self.side_sz
        self.ftab_len = (1 << (2 * self.ftab_chars)) + 1
        self.eftab_len = 2 * self.ftab_chars
        self.offs_len = (self.bwt_len + (1 << self.off_rate) - 1) >> self.off_rate


def _read_primary(path, fw):
    raw = np.fromfile(path, dtype=np.uint8)
    u32 = lambda off, n=1: raw[off:off + 4 * n].view("<u4")  # noqa: E731
    i32 = lambda off: int(raw[off:off + 4].view("<i4")[0])   # noqa: E731
    if int(u32(0)[0]) != 1:
        raise ValueError("%s: opposite-endian or not a .bt2 file" % path)
    e = EbwtData(fw)
    e.len = int(u32(4)[0])
    e.line_rate = i32(8)
    e.off_rate = i32(16)
    e.ftab_chars = i32(20)
    flags = i32(24)
    e.entire_rev = not (flags < 0 and ((-flags) & EBWT_ENTIRE_REV) == 0)
    if e.line_rate != 6:
        raise ValueError("only 64-byte sides (lineRate 6, small index) are supported")
    e.derive()
    p = 28
    e.npat = int(u32(p)[0]); p += 4
    e.plen = u32(p, e.npat).copy(); p += 4 * e.npat
    e.nfrag = int(u32(p)[0]); p += 4
    e.rstarts = u32(p, 3 * e.nfrag).copy().reshape(-1, 3); p += 12 * e.nfrag
    e.ebwt = raw[p:p + e.ebwt_tot_len].copy(); p += e.ebwt_tot_len
    e.zoff = int(u32(p)[0]); p += 4
    e.fchr = u32(p, 5).copy(); p += 20
    e.ftab = u32(p, e.ftab_len).copy(); p += 4 * e.ftab_len
    e.eftab = u32(p, e.eftab_len).copy(); p += 4 * e.eftab_len
    names = bytes(raw[p:]).split(b"\0")[0].decode("ascii", "replace")
    e.names = [n for n in names.split("\n") if n]
    e.offs = None
    return e


def load_ebwt(base, fw=True, load_offs=None):
    """Load <base>.1.bt2 (+ .2.bt2); for the mirror index pass base + '.rev' and fw=False."""
    e = _read_primary(base + ".1.bt2", fw)
    if load_offs is None:
        load_offs = fw
    if load_offs:
        raw = np.fromfile(base + ".2.bt2", dtype="<u4")
        if int(raw[0]) != 1:
            raise ValueError("bad .2.bt2")
        e.offs = raw[1:1 + e.offs_len].copy()
        if len(e.offs) != e.offs_len:
            raise ValueError("truncated .2.bt2")
    return e


def load_reference(base):
    """BitPairReference (.3.bt2 records, .4.bt2 2-bit bases; reference.cpp:40-300).
    Returns (records[n,3] = off,len,first ; packed bytes ; unpacked length)."""
    raw = np.fromfile(base + ".3.bt2", dtype=np.uint8)
    one = int(raw[0:4].view("<u4")[0])
    if one != 1:
        raise ValueError("bad .3.bt2")
    nrec = int(raw[4:8].view("<u4")[0])
    recs = np.zeros((nrec, 3), dtype=np.uint32)
    p = 8
    for i in range(nrec):
        recs[i, 0] = raw[p:p + 4].view("<u4")[0]
        recs[i, 1] = raw[p + 4:p + 8].view("<u4")[0]
        recs[i, 2] = raw[p + 8]
        p += 9
    packed = np.fromfile(base + ".4.bt2", dtype=np.uint8)
    return recs, packed


class Index:
    """Forward + mirror Ebwt pair as used by SeedAligner (bt2_search.cpp:4870-4892)."""

    def __init__(self, fw, bw):
        self.fw, self.bw = fw, bw

    @staticmethod
    def load(base):
        return Index(load_ebwt(base, True), load_ebwt(base + ".rev", False))

    def save_npz(self, path):
        d = {}
        for tag, e in (("fw", self.fw), ("bw", self.bw)):
            d[tag + "_hdr"] = np.array([e.len, e.line_rate, e.off_rate, e.ftab_chars, e.zoff,
                                        int(e.entire_rev)], dtype=np.int64)
            for k in ("plen", "rstarts", "ebwt", "fchr", "ftab", "eftab"):
                d[tag + "_" + k] = getattr(e, k)
            if e.offs is not None:
                d[tag + "_offs"] = e.offs
        np.savez_compressed(path, **d)

    @staticmethod
    def load_npz(path):
        z = np.load(path)
        out = []
        for tag, fw in (("fw", True), ("bw", False)):
            e = EbwtData(fw)
            h = z[tag + "_hdr"]
            e.len, e.line_rate, e.off_rate, e.ftab_chars, e.zoff = (int(x) for x in h[:5])
            e.entire_rev = bool(h[5])
            e.derive()
            for k in ("plen", "rstarts", "ebwt", "fchr", "ftab", "eftab"):
                setattr(e, k, np.ascontiguousarray(z[tag + "_" + k]))
            e.npat, e.nfrag = len(e.plen), len(e.rstarts)
            e.offs = np.ascontiguousarray(z[tag + "_offs"]) if (tag + "_offs") in z.files else None
            e.names = []
            out.append(e)
        return Index(out[0], out[1])


def exists(base):
    return all(os.path.exists(base + s) for s in (".1.bt2", ".2.bt2", ".rev.1.bt2"))


def read_ref_records(base):
    """RefRecords of <base>.3.bt2 (reference.cpp:100-167): [(off, len, first)]."""
    raw = np.fromfile(base + ".3.bt2", dtype=np.uint8)
    if int(raw[0:4].view("<u4")[0]) != 1:
        raise ValueError("bad .3.bt2")
    n = int(raw[4:8].view("<u4")[0])
    recs = []
    p = 8
    for _ in range(n):
        off, ln = (int(x) for x in raw[p:p + 8].view("<u4"))
        recs.append((off, ln, int(raw[p + 8])))
        p += 9
    return recs
