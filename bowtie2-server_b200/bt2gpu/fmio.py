"""Host-side helpers of the seed path: read pools, the host half of
SeedAligner::instantiateSeeds (aligner_seed.cpp:498-587: which seed windows exist), and the
parser of the dumps written by oracle/ref_seed_dump.cpp (the reference's own results)."""
import math

import numpy as np

from . import structs
from .dpio import encode_seq


class ReadPool:
    """Reads in the pooled layout of bt2gpu_reads_stage (bytes 0..4 + Phred+33 qualities)."""

    def __init__(self, seqs, quals):
        n = len(seqs)
        self.reads = np.zeros(n, dtype=structs.read)
        lens = np.array([len(s) for s in seqs], dtype=np.uint32)
        offs = np.zeros(n, dtype=np.uint32)
        if n:
            offs[1:] = np.cumsum(lens[:-1])
        self.reads["rd_off"], self.reads["rd_len"] = offs, lens
        if n and isinstance(seqs[0], str):
            self.seq = encode_seq("".join(seqs))
            self.qual = np.frombuffer("".join(quals).encode(), dtype=np.uint8).copy()
        else:
            self.seq = np.concatenate(seqs).astype(np.uint8) if n else np.zeros(0, np.uint8)
            self.qual = np.concatenate(quals).astype(np.uint8) if n else np.zeros(0, np.uint8)
        assert len(self.seq) == len(self.qual) == int(lens.sum())

    def __len__(self):
        return len(self.reads)

    def read_codes(self, i):
        o, l = int(self.reads["rd_off"][i]), int(self.reads["rd_len"][i])
        return self.seq[o:o + l]


def seed_interval(rdlen, const=1.0, coeff=1.15):
    """msIval.f<int>(len) for the default S,1,1.15 (SimpleFunc::f, simple_func.h; bt2_search.cpp:3336)."""
    v = int(const + coeff * math.sqrt(float(rdlen)))
    return max(v, 1)


def instantiate_exact_seeds(pool, seed_len, interval_fn=seed_interval, off=0, nofw=False, norc=False):
    """Seed windows SeedAligner::instantiateSeeds creates for -N 0 (one SEED_TYPE_EXACT seed per
    strand x offset; windows containing an N are dropped, Seed::instantiate :338-345).
    Returns (jobs, per_read) with per_read[i] = (interval, nseeds)."""
    jobs, per_read = [], []
    for i in range(len(pool)):
        rd = pool.read_codes(i)
        rdlen = len(rd)
        per = interval_fn(rdlen)
        sl = min(seed_len, rdlen)
        nseeds = 1
        if rdlen - off > seed_len:
            nseeds += (rdlen - off - seed_len) // per
        per_read.append((per, nseeds))
        for fw in (1, 0):
            if (fw and nofw) or (not fw and norc):
                continue
            for k in range(nseeds):
                depth = k * per + off
                if (rd[depth:depth + sl] > 3).any():
                    continue
                jobs.append((i, depth, sl, fw))
    arr = np.zeros(len(jobs), dtype=structs.seed_job)
    if jobs:
        a = np.array(jobs, dtype=np.int64)
        arr["read"], arr["off"], arr["len"], arr["fw"] = a[:, 0], a[:, 1], a[:, 2], a[:, 3]
    return arr, per_read


def parse_seed_dump(path, limit=None):
    """-> list of dicts: seq, qual, SW, M1, H[], SD, S[], X[], L[] (see oracle/ref_seed_dump.cpp)."""
    out, cur = [], None
    with open(path) as fh:
        for line in fh:
            t = line.split()
            if not t:
                continue
            k = t[0]
            if k == "R":
                if limit is not None and len(out) >= limit:
                    break
                cur = dict(seq=t[3], qual=t[4] if len(t) > 4 else "", H=[], S=[], X=[], L=[], K=[], SD=None)
                # a quality string may contain no spaces in FASTQ; keep exact slice of the line
                cur["qual"] = line.rstrip("\n").split(" ", 4)[4]
                out.append(cur)
            elif k == "SW":
                cur["SW"] = tuple(int(x) for x in t[1:])
            elif k == "M1":
                cur["M1"] = tuple(int(x) for x in t[1:])
            elif k == "H":
                cur["H"].append((int(t[1]), int(t[2]), int(t[3]), int(t[4]), int(t[5]), t[6], t[7]))
            elif k == "SD":
                cur["SD"] = tuple(int(x) for x in t[1:])
            elif k == "S":
                v = [int(x) for x in t[1:]]
                cur["S"].append((v[0], v[1], v[2], [tuple(v[4 + 4 * j:8 + 4 * j]) for j in range(v[3])]))
            elif k == "K":     # --seedmms: the reference strings of the ranges of the S line before it (key, length) ...
                v = [int(x) for x in t[1:]]
                cur["K"].append((v[0], v[1], [(v[3 + 2 * j], v[4 + 2 * j]) for j in range(v[2])]))
            elif k == "X":
                cur["X"].append(tuple(int(x) for x in t[1:]))
            elif k == "L":
                cur["L"].append((int(t[1]), int(t[2])))
    return out
