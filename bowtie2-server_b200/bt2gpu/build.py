"""Index construction in the reference's on-disk layout (what bowtie2-build-s writes for the
small index: Ebwt::buildToDisk bt2_idx.h:2829-3173, Ebwt::initFromVector :940-1170,
EbwtParams::init :133-167, RefRecord/.4.bt2 reference.cpp), done with torch sorts so that a
human-scale synthetic genome can be indexed in seconds on the GPU box instead of hours on its
host cores.  Plumbing for tests and bench.py: the product consumes any .bt2 index; parity with
bowtie2-build-s output is checked byte for byte in tests/test_index_build.py.

Conventions restated from the reference:
  * joined text = concatenation of the unambiguous stretches ("fragments") of all sequences;
    the mirror index (.rev) is built over the ENTIRE joined text reversed (EBWT_ENTIRE_REV);
  * suffix order treats the end of text as GREATER than every base ('$' sorts last; row
    len is the empty suffix), rows = len + 1;
  * BWT char of the row whose suffix starts at 0 is '$': stored as A, its row is zOff, and it
    is not counted in occ[] / fchr[];
  * side = 48 bytes of BWT (base k of a byte at bits 2k..2k+1) + u32 occ[A,C,G,T] counted
    before the side;
  * ftab[k] = first row of 10-mer k; suffixes shorter than ftabChars are "absorbed" in front
    of the next 10-mer and described by eftab (lo, hi) pairs;
  * offs[i] = SA[i << offRate].
"""
import numpy as np
import torch

from .index import EbwtData, Index

OFF_MASK = 0xFFFFFFFF
_KCH = 21          # bases per 63-bit sort key (3 bits each; 7 = beyond the end)


def _key21(codes3, pos):
    """63-bit keys of the 21 bases starting at each position in `pos` (int64 tensor)."""
    k = torch.zeros_like(pos)
    for j in range(_KCH):
        k = (k << 3) | codes3[pos + j].to(torch.int64)
    return k


def _keys21_range(codes3, lo, hi):
    """_key21 for the contiguous positions lo..hi-1 by doubling (codes3 is padded with >= 42 sevens)."""
    m = hi - lo
    k1 = codes3[lo:hi + 2 * _KCH].to(torch.int64)
    k2 = (k1[:-1] << 3) | k1[1:]
    k4 = (k2[:-2] << 6) | k2[2:]
    del k2
    k8 = (k4[:-4] << 12) | k4[4:]
    k16 = (k8[:-8] << 24) | k8[8:]
    del k8
    k20 = (k16[:m + 1] << 12) | k4[16:16 + m + 1]
    del k16, k4
    return (k20[:m] << 3) | k1[20:20 + m]


_CHUNK = 1 << 27        # rows per pass of the chunked whole-array operations
_PFX_BITS = 12          # bucket histogram over the first 4 bases of a key (3 bits each)


def _sort_bucket(codes3, pos, keys, n):
    """Suffix order of the positions `pos` (int64) whose 21-base keys are `keys`: sort by key, then
    break ties 21 more bases at a time (suffixes of one bucket never tie with another bucket's)."""
    device = pos.device
    skey, o = torch.sort(keys)
    sa = pos[o]
    del o
    m = int(sa.numel())
    new = torch.ones(m, dtype=torch.bool, device=device)
    new[1:] = skey[1:] != skey[:-1]
    del skey
    if bool(new.all()):
        return sa
    starts = torch.nonzero(new).squeeze(1)
    grp = starts[torch.cumsum(new, 0) - 1]            # first local position of each run of equal keys
    del starts, new
    depth = _KCH
    while True:
        tied = torch.zeros(m, dtype=torch.bool, device=device)
        same_next = grp[1:] == grp[:-1]
        tied[1:] |= same_next
        tied[:-1] |= same_next
        tpos = torch.nonzero(tied).squeeze(1)
        if tpos.numel() == 0:
            break
        sa_t, g = sa[tpos], grp[tpos]
        p = torch.clamp(sa_t + depth, max=n + 1)             # beyond the end: all sevens
        k2 = _key21(codes3, p)
        o = torch.sort(k2, stable=True).indices
        o = o[torch.sort(g[o], stable=True).indices]
        sa[tpos] = sa_t[o]
        g2, k2 = g[o], k2[o]
        nb = torch.ones(tpos.numel(), dtype=torch.bool, device=device)
        nb[1:] = (g2[1:] != g2[:-1]) | (k2[1:] != k2[:-1])
        grp[tpos] = tpos[torch.nonzero(nb).squeeze(1)][torch.cumsum(nb, 0) - 1]
        depth += _KCH
        if depth > n + _KCH:
            break
    return sa


def suffix_array(s, device, bucket_rows=1 << 27):
    """SA of s (uint8 codes 0..3) with the end of text greater than every base; n+1 rows (int64).

    Memory-lean for human-scale texts (3.1 G rows on one 180 GB device): the 21-base keys of all
    positions are computed chunk by chunk into one int64 array, the rows are cut into buckets of at
    most ~bucket_rows suffixes by the first four bases of their key (consecutive key ranges, so the
    buckets are consecutive SA ranges), and each bucket is sorted and tie-broken on its own."""
    n = int(s.numel())
    n1 = n + 1
    codes3 = torch.cat([s.to(device), torch.full((2 * _KCH + 8,), 7, dtype=torch.uint8, device=device)])
    keys = torch.empty(n1, dtype=torch.int64, device=device)
    for lo in range(0, n1, _CHUNK):
        hi = min(n1, lo + _CHUNK)
        keys[lo:hi] = _keys21_range(codes3, lo, hi)
    if n1 <= bucket_rows:
        return _sort_bucket(codes3, torch.arange(n1, dtype=torch.int64, device=device), keys, n)
    shift = 3 * _KCH - _PFX_BITS
    hist = torch.zeros(1 << _PFX_BITS, dtype=torch.int64, device=device)
    for lo in range(0, n1, _CHUNK):
        hist += torch.bincount(keys[lo:lo + _CHUNK] >> shift, minlength=1 << _PFX_BITS)
    hist = hist.cpu().tolist()
    sa = torch.empty(n1, dtype=torch.int64, device=device)
    out, b0 = 0, 0
    while b0 < len(hist):
        b1, cnt = b0 + 1, hist[b0]
        while b1 < len(hist) and cnt + hist[b1] <= bucket_rows:
            cnt += hist[b1]
            b1 += 1
        if cnt:
            klo, khi = b0 << shift, b1 << shift          # khi = 2^63 for the last bucket: no upper bound then
            parts = []
            for lo in range(0, n1, _CHUNK):
                kc = keys[lo:lo + _CHUNK]
                m = (kc >= klo) & (kc < khi) if b1 < len(hist) else (kc >= klo)
                parts.append(torch.nonzero(m).squeeze(1) + lo)
            pos = torch.cat(parts)
            del parts
            sa[out:out + cnt] = _sort_bucket(codes3, pos, keys[pos], n)
            out += cnt
        b0 = b1
    assert out == n1
    return sa


def _kmers(s, k, device):
    """Integer of the k bases starting at every i <= n-k (first base most significant)."""
    n = int(s.numel())
    v = torch.zeros(max(n - k + 1, 0), dtype=torch.int32, device=device)
    s32 = s.to(device=device, dtype=torch.int32)
    for j in range(k):
        v = (v << 2) | s32[j:n - k + 1 + j]
    return v


def build_ebwt(s, fw, off_rate=4, ftab_chars=10, line_rate=6, device=None, want_offs=True, bucket_rows=1 << 27):
    """One Ebwt over the joined text s (torch uint8 codes 0..3), as Ebwt::buildToDisk writes it."""
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    e = EbwtData(fw)
    e.len, e.line_rate, e.off_rate, e.ftab_chars = int(s.numel()), line_rate, off_rate, ftab_chars
    e.entire_rev = not fw
    e.derive()
    n = e.len
    s = s.to(device)
    sa = suffix_array(s, device, bucket_rows=bucket_rows)
    # ---- BWT, zOff, sides ----
    rows_pad = e.num_sides * e.side_bwt_len
    bp = torch.zeros(rows_pad, dtype=torch.uint8, device=device)
    e.zoff = -1
    for lo in range(0, n + 1, _CHUNK):
        c = sa[lo:lo + _CHUNK]
        z = torch.nonzero(c == 0).squeeze(1)
        if z.numel():
            e.zoff = lo + int(z[0])
        bp[lo:lo + c.numel()] = s[torch.clamp(c - 1, min=0)]
    bp[e.zoff] = 0
    q = bp.view(-1, 4)
    packed = (q[:, 0] | (q[:, 1] << 2) | (q[:, 2] << 4) | (q[:, 3] << 6)).view(e.num_sides, e.side_bwt_sz)
    cnt = torch.stack([(bp.view(e.num_sides, e.side_bwt_len) == c).sum(1) for c in range(4)], 1)   # [sides,4]
    # the zero padding after row n lies in the last side only and occ[] counts what precedes a side: no correction
    zside = e.zoff // e.side_bwt_len
    cnt[zside, 0] -= 1                                       # the '$' is not an A
    occ = torch.cumsum(cnt, 0) - cnt                         # counted before each side
    del bp, q
    sides = torch.zeros((e.num_sides, e.side_sz), dtype=torch.uint8, device=device)
    sides[:, :e.side_bwt_sz] = packed
    del packed
    occ32 = occ.to(torch.int64)
    for c in range(4):
        for b in range(4):
            sides[:, e.side_bwt_sz + 4 * c + b] = ((occ32[:, c] >> (8 * b)) & 0xFF).to(torch.uint8)
    e.ebwt = sides.reshape(-1).cpu().numpy()
    del sides, occ, occ32, cnt
    # ---- fchr ----
    tot = np.array([int((s == c).sum()) for c in range(4)], dtype=np.int64)
    e.fchr = np.concatenate([[0], np.cumsum(tot)]).astype(np.uint32)
    # ---- ftab / eftab ----
    # count[k] = suffixes of at least ftab_chars bases that start with 10-mer k; the (at most ftab_chars + 1)
    # shorter suffixes are "absorbed" in front of the next long row's 10-mer
    nk = 1 << (2 * ftab_chars)
    count_t = torch.zeros(nk, dtype=torch.int64, device=device)
    nlong = max(n - ftab_chars + 1, 0)
    for lo in range(0, nlong, _CHUNK):
        hi = min(nlong, lo + _CHUNK)
        count_t += torch.bincount(_kmers(s[lo:hi + ftab_chars - 1], ftab_chars, device).to(torch.int64), minlength=nk)
    count = count_t.cpu().numpy().astype(np.int64)
    del count_t

    def kmer_at(p):
        v = 0
        for c in s[p:p + ftab_chars].cpu().tolist():
            v = (v << 2) | int(c)
        return v
    short_rows = []
    for lo in range(0, n + 1, _CHUNK):
        z = torch.nonzero(sa[lo:lo + _CHUNK] > n - ftab_chars).squeeze(1)
        short_rows += [lo + int(x) for x in z.cpu().tolist()]
    short_set = set(short_rows)
    absorb = np.zeros(nk + 1, dtype=np.int64)
    for r in short_rows:                                     # <= ftab_chars + 1 rows
        r2 = r + 1
        while r2 <= n and r2 in short_set:
            r2 += 1
        if r2 <= n:
            absorb[kmer_at(int(sa[r2]))] += 1
        else:
            absorb[nk] += 1
    before = np.concatenate([[0], np.cumsum(count)])         # long rows before 10-mer i
    hi = before + np.cumsum(absorb)
    lo = hi - absorb
    ftab = lo.astype(np.uint32)
    eftab = np.zeros(2 * ftab_chars, dtype=np.uint32)
    cur = 0
    for i in np.nonzero(absorb[1:])[0] + 1:
        eftab[2 * cur], eftab[2 * cur + 1] = lo[i], hi[i]
        ftab[i] = cur ^ OFF_MASK
        cur += 1
    assert absorb[0] == 0 and hi[nk] == n + 1
    e.ftab, e.eftab = ftab, eftab
    # ---- SA sample ----
    e.offs = sa[:: (1 << off_rate)].to(torch.int64).cpu().numpy().astype(np.uint32) if want_offs else None
    del sa
    return e


def fragments(seqs):
    """Unambiguous stretches of each sequence (codes 0..3; 4 = N): [(seq idx, off in seq, len)]
    plus the RefRecords of the .3.bt2 file [(off = Ns skipped before the stretch, len, first)]."""
    frags, recs = [], []
    for si, sq in enumerate(seqs):
        isn = np.concatenate([[True], sq > 3, [True]])
        starts = np.nonzero(isn[:-1] & ~isn[1:])[0]
        ends = np.nonzero(~isn[:-1] & isn[1:])[0]
        prev_end, first = 0, 1
        if len(starts) == 0:
            recs.append((len(sq), 0, 1))
        for a, b in zip(starts, ends):
            frags.append((si, int(a), int(b - a)))
            recs.append((int(a - prev_end), int(b - a), first))
            prev_end, first = int(b), 0
        if len(starts) and prev_end < len(sq):
            recs.append((len(sq) - prev_end, 0, 0))
    return frags, recs


def build_index(seqs, names=None, off_rate=4, ftab_chars=10, device=None, bucket_rows=1 << 27):
    """seqs: list of numpy uint8 code arrays (0..3, 4 = N).  Returns (Index, extras) where extras
    holds what the .3/.4.bt2 files need."""
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    names = names or ["seq%d" % i for i in range(len(seqs))]
    frags, recs = fragments(seqs)
    joined = np.concatenate([seqs[si][a:a + ln] for si, a, ln in frags]) if frags else np.zeros(0, np.uint8)
    s = torch.from_numpy(np.ascontiguousarray(joined))
    fw = build_ebwt(s, True, off_rate, ftab_chars, device=device, bucket_rows=bucket_rows)
    bw = build_ebwt(torch.flip(s, [0]), False, off_rate, ftab_chars, device=device, bucket_rows=bucket_rows)
    plen = np.array([len(x) for x in seqs], dtype=np.uint32)
    rs, pos = [], 0
    for si, a, ln in frags:
        rs.append((pos, si, a))
        pos += ln
    rs_rev, pos = [], 0
    for si, a, ln in reversed(frags):
        rs_rev.append((pos, si, a))
        pos += ln
    for e, r in ((fw, rs), (bw, rs_rev)):
        e.plen, e.npat = plen, len(seqs)
        e.rstarts, e.nfrag = np.array(r, dtype=np.uint32).reshape(-1, 3), len(r)
        e.names = list(names)
    return Index(fw, bw), dict(recs=recs, joined=joined)


def write_index(ix, extras, base):
    """Write <base>.{1,2,3,4}.bt2 and <base>.rev.{1,2}.bt2 (bt2_io.cpp:131-614 reads them back)."""
    for e, b in ((ix.fw, base), (ix.bw, base + ".rev")):
        flags = -(1 | (4 if not e.fw else 0))
        with open(b + ".1.bt2", "wb") as fh:
            fh.write(np.array([1, e.len], dtype="<u4").tobytes())
            fh.write(np.array([e.line_rate, 2, e.off_rate, e.ftab_chars, flags], dtype="<i4").tobytes())
            fh.write(np.array([e.npat], dtype="<u4").tobytes())
            fh.write(e.plen.astype("<u4").tobytes())
            fh.write(np.array([e.nfrag], dtype="<u4").tobytes())
            fh.write(e.rstarts.astype("<u4").tobytes())
            fh.write(e.ebwt.tobytes())
            fh.write(np.array([e.zoff], dtype="<u4").tobytes())
            fh.write(e.fchr.astype("<u4").tobytes())
            fh.write(e.ftab.astype("<u4").tobytes())
            fh.write(e.eftab.astype("<u4").tobytes())
            fh.write(("".join(n + "\n" for n in e.names)).encode() + b"\0")
        with open(b + ".2.bt2", "wb") as fh:
            fh.write(np.array([1], dtype="<u4").tobytes())
            fh.write(e.offs.astype("<u4").tobytes())
    with open(base + ".3.bt2", "wb") as fh:
        recs = extras["recs"]
        fh.write(np.array([1, len(recs)], dtype="<u4").tobytes())
        for off, ln, first in recs:
            fh.write(np.array([off, ln], dtype="<u4").tobytes() + bytes([first]))
    j = extras["joined"]
    pad = (-len(j)) % 4
    q = np.concatenate([j, np.zeros(pad, np.uint8)]).reshape(-1, 4)
    (q[:, 0] | (q[:, 1] << 2) | (q[:, 2] << 4) | (q[:, 3] << 6)).astype(np.uint8).tofile(base + ".4.bt2")
