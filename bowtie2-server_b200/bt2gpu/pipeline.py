"""Batch-granular host glue that chains the hot-path entry points the way the reference's
per-read driver does (multiseedSearchWorker bt2_search.cpp:3440-4060 -> SwDriver::extendSeeds
aligner_sw_driver.cpp:756-1297): exactSweep -> exact seeds (round 0) -> extend -> locate ->
frame a seed-extension window (DynProgFramer::frameSeedExtensionRect dp_framer.cpp:81-129) ->
fill / gather / backtrace.

This is measurement scaffolding for bench.py, not the reference's driver: hit selection is a
deterministic stand-in (per read: narrowest BW range, then longest exact extension) for
SeedResults::rankSeedHits + SwDriver::prioritizeSATups (which draw from the per-read RNG), and
one window is extended per read.  Every GPU call goes through the C ABI with host buffers.
"""
import numpy as np

from . import structs, fmio

_COMP = np.array([3, 2, 1, 0, 4], dtype=np.uint8)


class HotPath:
    def __init__(self, ctx, dev_index, genome, sc, seed_len=22, maxgap=15, max_aln=2):
        self.ctx, self.dix, self.genome, self.sc = ctx, dev_index, genome, sc
        self.seed_len, self.maxgap, self.max_aln = seed_len, maxgap, max_aln
        ctx.set_scoring(sc)
        ctx.bind_index(dev_index)
        self.kernel_ms = {}
        self.counters = {}

    def _t(self, name):
        self.kernel_ms[name] = self.kernel_ms.get(name, 0.0) + self.ctx.kernel_ms()

    def make_seed_jobs(self, reads):
        """instantiateSeeds for equal-length reads (aligner_seed.cpp:498-587), vectorised."""
        n, rdlen = reads.shape
        per = fmio.seed_interval(rdlen)
        sl = min(self.seed_len, rdlen)
        nseeds = 1 + ((rdlen - self.seed_len) // per if rdlen > self.seed_len else 0)
        offs = np.arange(nseeds) * per
        isn = np.concatenate([np.zeros((n, 1), np.int32), np.cumsum(reads > 3, axis=1, dtype=np.int32)], axis=1)
        clean = (isn[:, offs + sl] - isn[:, offs]) == 0                      # [n, nseeds]
        rd, k = np.nonzero(clean)
        m = len(rd)
        jobs = np.zeros(2 * m, dtype=structs.seed_job)
        for half, fw in ((slice(0, m), 1), (slice(m, 2 * m), 0)):
            jobs["read"][half], jobs["off"][half], jobs["len"][half], jobs["fw"][half] = rd, offs[k], sl, fw
        # reference order: all fw seeds of a read, then its rc seeds; group by read
        order = np.lexsort((jobs["off"], 1 - jobs["fw"].astype(np.int8), jobs["read"]))
        return jobs[order]

    def step(self, reads, quals):
        """reads, quals: [n, rdlen] uint8 (codes 0..4, Phred+33).  Returns the DP results."""
        ctx, G = self.ctx, len(self.genome)
        n, rdlen = reads.shape
        self.kernel_ms, c = {}, {}
        pool = fmio.ReadPool.__new__(fmio.ReadPool)
        pool.seq, pool.qual = reads.reshape(-1), quals.reshape(-1)
        pool.reads = np.zeros(n, dtype=structs.read)
        pool.reads["rd_off"] = np.arange(n, dtype=np.int64) * rdlen
        pool.reads["rd_len"] = rdlen
        ctx.stage_reads(pool)
        # 1. exactSweep
        sw = ctx.exact_sweep(False, False, 2)
        self._t("sweep")
        c["sweep_bwops"] = int(sw["bwops"].sum(dtype=np.int64))
        # 2. exact seeds, round 0
        sjobs = self.make_seed_jobs(reads)
        sh = ctx.seed_exact(sjobs)
        self._t("seed")
        c["seeds"] = len(sjobs)
        c["seed_bwops"] = int(sh["bwops"].sum(dtype=np.int64))
        c["seed_sides"] = int(sh["sides"].sum(dtype=np.int64))
        hit = np.nonzero(sh["botf"] > sh["topf"])[0]
        c["seed_hits"] = len(hit)
        # 3. extend every hit
        xj = np.zeros(len(hit), dtype=structs.extend_job)
        for k in ("topf", "botf", "topb", "botb"):
            xj[k] = sh[k][hit]
        for k in ("read", "off", "len", "fw"):
            xj[k] = sjobs[k][hit]
        xr = ctx.extend(xj)
        self._t("extend")
        c["extend_fmops"] = int(xr["fmops"].sum(dtype=np.int64))
        # 4. pick one hit per read: narrowest range, then longest exact extension, then first
        size = (xj["botf"] - xj["topf"]).astype(np.int64)
        ext = xr["nlex"].astype(np.int64) + xr["nrex"].astype(np.int64)
        order = np.lexsort((np.arange(len(hit)), -ext, size, xj["read"]))
        first = np.ones(len(order), dtype=bool)
        first[1:] = xj["read"][order][1:] != xj["read"][order][:-1]
        pick = order[first]
        # 5. locate the first row of each picked range
        offs, steps = ctx.locate(xj["topf"][pick])
        self._t("locate")
        c["locate_rows"] = len(pick)
        c["locate_steps"] = int(steps.sum(dtype=np.int64))
        # 6. frame the seed-extension window around the located diagonal
        rid = xj["read"][pick].astype(np.int64)
        fw = xj["fw"][pick].astype(bool)
        soff = xj["off"][pick].astype(np.int64)
        slen = xj["len"][pick].astype(np.int64)
        rdoff = np.where(fw, soff, rdlen - soff - slen)          # seed offset in the oriented read
        start = offs.astype(np.int64) - rdoff
        mg = self.maxgap
        ncol = rdlen + 4 * mg
        refl = start - 2 * mg
        cols = refl[:, None] + np.arange(ncol + 1)[None, :]
        inside = (cols >= 0) & (cols < G)
        codes = np.where(inside, self.genome[np.clip(cols, 0, G - 1)], 4).astype(np.uint8)
        win = (1 << codes).astype(np.uint8)
        m = len(pick)
        oriented = np.where(fw[:, None], reads[rid], _COMP[reads[rid][:, ::-1]])
        oq = np.where(fw[:, None], quals[rid], quals[rid][:, ::-1])
        jobs = np.zeros(m, dtype=structs.dp_job)
        jobs["rd_off"] = np.arange(m, dtype=np.int64) * rdlen
        jobs["rd_len"] = rdlen
        jobs["rf_off"] = np.arange(m, dtype=np.int64) * (ncol + 1)
        jobs["ncol"] = ncol
        jobs["minsc"] = self.sc.score_min.f_int(rdlen) if self.sc.monotone else max(1, self.sc.score_min.f_int(rdlen))
        jobs["nceil"] = self.sc.nceil(rdlen)
        jobs["refl"] = refl
        jobs["corel"], jobs["corer"] = mg, 3 * mg
        jobs["fw"] = fw
        jobs["enable8"] = 1
        # 7. fill + gather + nextAlignment loop
        dseq, dqual, dwin = (np.ascontiguousarray(oriented.reshape(-1)), np.ascontiguousarray(oq.reshape(-1)),
                             np.ascontiguousarray(win.reshape(-1)))
        ctx.dp_stage(dseq, dqual, dwin, jobs, max_aln=self.max_aln, cand_cap=256 * max(m, 1), edit_cap=64 * max(m, 1))
        ctx.dp_run(7)
        r = ctx.dp_fetch(want_cands=False)
        res = (r[0], None, r[2], r[3])
        self._t("dp")
        ph = ctx.dp_phase_ms() if m else [0.0, 0.0, 0.0]
        self.kernel_ms["dp_fill"], self.kernel_ms["dp_gather"], self.kernel_ms["dp_backtrace"] = ph
        c["dp_jobs"] = m
        c["dp_cells"] = ctx.dp_cells()
        c["aligned"] = int((res[0]["naln"] > 0).sum())
        c["h2d_bytes"] = int(pool.seq.nbytes + pool.qual.nbytes + pool.reads.nbytes + sjobs.nbytes + xj.nbytes
                             + 4 * len(pick) + dseq.nbytes + dqual.nbytes + dwin.nbytes + jobs.nbytes)
        c["d2h_bytes"] = int(sw.nbytes + sh.nbytes + xr.nbytes + 8 * len(pick) + res[0].nbytes + res[2].nbytes
                             + res[3].nbytes)
        self.counters = c
        self.last = dict(rid=rid, fw=fw, start=start, results=res, jobs=jobs, pick_offs=offs)
        return res
