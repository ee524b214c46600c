"""Build DP job batches (include/bt2gpu.h: bt2gpu_dp_job + pools) and parse the
P/R/A/E problem dumps written by oracle/ref_dp_dump.cpp (the reference's own results)."""
import numpy as np
from . import structs

_ASC2DNA = np.full(256, 4, dtype=np.uint8)
for _i, _c in enumerate("ACGT"):
    _ASC2DNA[ord(_c)] = _i
    _ASC2DNA[ord(_c.lower())] = _i
# asc2dnamask (alphabet.cpp:94) restricted to the IUPAC codes; N (15) -> 16 as bt2_dp.cpp:617-621
_ASC2MASK = np.zeros(256, dtype=np.uint8)
for _c, _m in dict(A=1, C=2, M=3, G=4, R=5, S=6, V=7, T=8, W=9, Y=10, H=11, K=12, D=13, B=14, N=16).items():
    _ASC2MASK[ord(_c)] = _m
_COMP = np.array([3, 2, 1, 0, 4], dtype=np.uint8)


def encode_seq(s):
    return _ASC2DNA[np.frombuffer(s.encode(), dtype=np.uint8)]


def encode_mask(s):
    return _ASC2MASK[np.frombuffer(s.encode(), dtype=np.uint8)]


def revcomp(codes):
    return _COMP[codes[::-1]]


class DpBatch:
    """Accumulates problems into the pooled layout of bt2gpu_dp_align_batch."""

    def __init__(self):
        self.seq, self.qual, self.win, self.jobs = [], [], [], []
        self.seq_len = 0
        self.win_len = 0

    def add(self, seq_codes, qual_bytes, fw, window_masks, minsc, nceil, refl, triml, corel, corer,
            enable8=True):
        """seq_codes/qual_bytes: the read as given to initRead (forward); the batch stores the
        strand initRef's `fw` selects (rdfw_/qufw_ or rdrc_/qurc_, aligner_sw.cpp:98-99)."""
        seq_codes = np.asarray(seq_codes, dtype=np.uint8)
        qual_bytes = np.asarray(qual_bytes, dtype=np.uint8)
        if not fw:
            seq_codes = revcomp(seq_codes)
            qual_bytes = qual_bytes[::-1]
        window_masks = np.asarray(window_masks, dtype=np.uint8)
        ncol = len(window_masks) - 1
        self.jobs.append((self.seq_len, len(seq_codes), self.win_len, ncol, minsc, nceil, refl,
                          triml, corel, corer, 1 if fw else 0, 1 if enable8 else 0))
        self.seq.append(seq_codes)
        self.qual.append(qual_bytes)
        self.win.append(window_masks)
        self.seq_len += len(seq_codes)
        self.win_len += len(window_masks)

    def finalize(self):
        n = len(self.jobs)
        jobs = np.zeros(n, dtype=structs.dp_job)
        for i, j in enumerate(self.jobs):
            (jobs["rd_off"][i], jobs["rd_len"][i], jobs["rf_off"][i], jobs["ncol"][i], jobs["minsc"][i],
             jobs["nceil"][i], jobs["refl"][i], jobs["triml"][i], jobs["corel"][i], jobs["corer"][i],
             jobs["fw"][i], jobs["enable8"][i]) = j
        seq = np.concatenate(self.seq) if n else np.zeros(0, np.uint8)
        qual = np.concatenate(self.qual) if n else np.zeros(0, np.uint8)
        win = np.concatenate(self.win) if n else np.zeros(0, np.uint8)
        return np.ascontiguousarray(seq), np.ascontiguousarray(qual), np.ascontiguousarray(win), jobs


def parse_dump(path, limit=None):
    """Yield dicts {seq, qual, fw, refidx, reflen, minsc, rect, window, R, A[], E} per problem."""
    cur = None
    n = 0
    pid = 0
    with open(path) as fh:
        for ln in fh:
            t = ln.rstrip("\n").split("\t")
            if t[0] == "P":
                if cur is not None:
                    yield cur
                    n += 1
                    if limit is not None and n >= limit:
                        return
                rect = [int(x) for x in t[7].split(",")]
                cur = dict(seq=t[1], qual=t[2], fw=(t[3] == "+"), refidx=int(t[4]), reflen=int(t[5]),
                           minsc=int(t[6]),
                           rect=dict(zip(("refl", "refr", "refl_pretrim", "refr_pretrim", "triml", "trimr",
                                          "corel", "corer", "maxgap"), rect)),
                           window=t[8], R=None, A=[], E=None, pid=pid)
                pid += 1
            elif t[0] == "R":
                cands = []
                if len(t) > 6 and t[6]:
                    cands = [tuple(int(x) for x in c.split(",")) for c in t[6].split(";")]
                cur["R"] = dict(aligned=int(t[1]), best=int(t[2]), sse8=int(t[3]), sse16=int(t[4]),
                                ncand=int(t[5]), cands=cands)
            elif t[0] == "A":
                edits = []
                if t[9]:
                    for e in t[9].split(";"):
                        pos, ch, qc, ty = e.split(":")
                        edits.append((int(pos), ch, qc, int(ty)))
                cur["A"].append(dict(score=int(t[1]), refoff=int(t[2]), trim5=int(t[3]), trim3=int(t[4]),
                                     ns=int(t[5]), gaps=int(t[6]), refns=int(t[7]), nedits=int(t[8]),
                                     edits=edits, cural=int(t[10]), rnd=int(t[11])))
            elif t[0] == "E":
                cur["E"] = dict(naln=int(t[1]), cural=int(t[2]), rnd=int(t[3]))
    if cur is not None:
        yield cur


def batch_from_problems(problems, scoring, enable8=True):
    b = DpBatch()
    for p in problems:
        seq = encode_seq(p["seq"])
        qual = np.frombuffer(p["qual"].encode(), dtype=np.uint8)
        r = p["rect"]
        b.add(seq, qual, p["fw"], encode_mask(p["window"]), p["minsc"], scoring.nceil(len(seq)),
              r["refl"], r["triml"], r["corel"], r["corer"], enable8)
    return b.finalize()


def results_to_py(results, cands, alns, edits, i):
    """Decode job i of a result set into the same shape parse_dump() yields (R, A[], E)."""
    r = results[i]
    cs = cands[r["cand_off"]: r["cand_off"] + r["ncand"]]
    R = dict(aligned=int(r["aligned"]), best=int(r["best"]), sse8=int(r["width"] == 8),
             sse16=int(r["width"] == 16), ncand=int(r["ncand"]),
             cands=[(int(c["row"]), int(c["col"]), int(c["score"])) for c in cs])
    A = []
    for a in alns[r["aln_off"]: r["aln_off"] + r["naln"]]:
        es = edits[a["edit_off"]: a["edit_off"] + a["nedits"]]
        A.append(dict(score=int(a["score"]), refoff=int(a["refoff"]), trim5=int(a["trim5"]),
                      trim3=int(a["trim3"]), ns=int(a["ns"]), gaps=int(a["gaps"]), refns=int(a["refns"]),
                      nedits=int(a["nedits"]),
                      edits=[(int(e["pos"]), chr(e["chr"]), chr(e["qchr"]), int(e["type"])) for e in es],
                      cural=int(a["cural"]), tries=int(a["tries"])))
    E = dict(naln=int(r["naln"]), cural=int(r["cural"]), tail_tries=int(r["tail_tries"]))
    return R, A, E
