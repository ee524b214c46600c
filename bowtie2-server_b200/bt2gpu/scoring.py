"""Host-side mirror of the reference's Scoring / SimpleFunc objects (scoring.h:96-486,
simple_func.h) -- parameter derivation only; the arithmetic that uses them runs on the GPU."""
import math
import numpy as np
from . import structs

COST_MODEL_ROUNDED_QUAL, COST_MODEL_QUAL, COST_MODEL_CONSTANT = 1, 2, 3


class SimpleFunc:
    """simple_func.h: f(x) = max(mn, min(mx, I + S * g(x))), g in {const, linear, sqrt, log}."""

    def __init__(self, kind="L", const=0.0, coeff=0.0, mn=-float("inf"), mx=float("inf")):
        self.kind, self.const, self.coeff, self.mn, self.mx = kind, const, coeff, mn, mx

    @staticmethod
    def parse(s, mn=-float("inf"), mx=float("inf")):
        t = s.split(",")
        kind = t[0].upper()
        const = float(t[1]) if len(t) > 1 else 0.0
        coeff = float(t[2]) if len(t) > 2 else 0.0
        return SimpleFunc(kind, const, coeff, mn, mx)

    def f(self, x):
        if self.kind == "C":
            x = 0.0
        elif self.kind == "L":
            x = float(x)
        elif self.kind == "S":
            x = math.sqrt(x)
        elif self.kind == "G":
            x = math.log(x)
        else:
            raise ValueError(self.kind)
        v = self.const + self.coeff * x
        return max(self.mn, min(self.mx, v))

    def f_int(self, x):
        """SimpleFunc::f<int>: truncation toward zero of the double (C++ static_cast)."""
        return int(self.f(x))


def _init_pens(kind, cmin, cmax):
    """Scoring::initPens (scoring.h:103-133)."""
    pens = np.zeros(256, dtype=np.int32)
    if kind == COST_MODEL_QUAL:
        for i in range(256):
            ii = min(i, 40)
            frac = np.float32(ii) / np.float32(40.0)
            pens[i] = cmin + int(np.float32(frac * np.float32(cmax - cmin)))
    elif kind == COST_MODEL_CONSTANT:
        pens[:] = cmax
    else:
        raise NotImplementedError("COST_MODEL_ROUNDED_QUAL")
    return pens


class Scoring:
    """Defaults are Bowtie 2's end-to-end scheme (scoring.h:28-83); local=True gives --local."""

    def __init__(self, local=False, ma=None, mmp_max=6, mmp_min=2, np_=1, rdg=(5, 3), rfg=(5, 3),
                 gapbar=4, score_min=None, n_ceil=None, mm_type=COST_MODEL_QUAL):
        self.local = local
        self.match_bonus = (2 if local else 0) if ma is None else ma
        self.mmpens = _init_pens(mm_type, mmp_min, mmp_max)
        self.npens = _init_pens(COST_MODEL_CONSTANT, np_, np_)
        self.rd_gap_const, self.rd_gap_linear = rdg
        self.rf_gap_const, self.rf_gap_linear = rfg
        self.gapbar = gapbar
        self.monotone = self.match_bonus == 0
        self.score_min = score_min or (SimpleFunc("G", 20.0, 8.0) if local else SimpleFunc("L", -0.6, -0.6))
        self.n_ceil = n_ceil or SimpleFunc("L", 0.0, 0.15)

    def read_gap_open(self):
        return self.rd_gap_const + self.rd_gap_linear

    def read_gap_extend(self):
        return self.rd_gap_linear

    def ref_gap_open(self):
        return self.rf_gap_const + self.rf_gap_linear

    def ref_gap_extend(self):
        return self.rf_gap_linear

    def nceil(self, rdlen):
        return self.n_ceil.f_int(float(rdlen))

    def to_struct(self):
        s = np.zeros(1, dtype=structs.scoring)
        s["mmpens"][0] = self.mmpens
        s["npens"][0] = self.npens
        s["match_bonus"] = self.match_bonus
        s["rd_gap_open"] = self.read_gap_open()
        s["rd_gap_ext"] = self.read_gap_extend()
        s["rf_gap_open"] = self.ref_gap_open()
        s["rf_gap_ext"] = self.ref_gap_extend()
        s["gapbar"] = self.gapbar
        s["monotone"] = 1 if self.monotone else 0
        return s
