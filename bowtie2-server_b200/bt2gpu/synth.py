"""Synthetic workloads restating the read model of the reference's example/reads/simulate.pl
(SNP rate 0.0012 :124-131, indels 0.0005/base each with length 1+floor(Exp(3)) :140-168,
per-read quality ceiling 40/11 with 2% overridden to 4 :209-217, per-base error 10^(-q/10)
drawn from "ACGTNNNNNN" :219-236), vectorised with a fixed numpy PRNG seed."""
import numpy as np

from . import structs


def random_genome(rng, n):
    return rng.integers(0, 4, size=n, dtype=np.uint8)


def simulate_reads(rng, genome, nreads, rdlen, pos=None):
    """Return (reads[n, rdlen] codes, quals[n, rdlen] phred+33, pos[n], fw[n]).  `pos`: given origins."""
    G = len(genome)
    if pos is None:
        pos = rng.integers(0, G - rdlen - 64, size=nreads)
    idx = pos[:, None] + np.arange(rdlen + 48)[None, :]
    frag = genome[idx]                                   # true fragment with slack for deletions
    reads = frag[:, :rdlen].copy()
    # SNPs
    snp = rng.random((nreads, rdlen)) < 0.0012
    reads[snp] = (reads[snp] + rng.integers(1, 4, size=int(snp.sum()))) % 4
    # indels (one event per affected read is enough at these rates)
    p_indel = 0.0005 * rdlen
    has_del = rng.random(nreads) < p_indel
    has_ins = rng.random(nreads) < p_indel
    for i in np.nonzero(has_del)[0]:
        p = int(rng.integers(8, rdlen - 8)); g = 1 + int(rng.exponential(3.0))
        g = min(g, 40)
        reads[i] = np.concatenate([frag[i, :p], frag[i, p + g:]])[:rdlen]
    for i in np.nonzero(has_ins)[0]:
        p = int(rng.integers(8, rdlen - 8)); g = 1 + int(rng.exponential(3.0))
        g = min(g, 20)
        reads[i] = np.concatenate([reads[i, :p], rng.integers(0, 4, size=g).astype(np.uint8), reads[i, p:]])[:rdlen]
    # qualities
    ceil = np.where(rng.random(nreads) < 0.8, 40, 11)
    q = (rng.random((nreads, rdlen)) * ceil[:, None]).astype(np.int64)
    q[rng.random((nreads, rdlen)) < 0.02] = 4
    # sequencing errors
    err = rng.random((nreads, rdlen)) < np.power(10.0, -q / 10.0)
    alphabet = np.array([0, 1, 2, 3, 4, 4, 4, 4, 4, 4], dtype=np.uint8)
    reads[err] = alphabet[rng.integers(0, 10, size=int(err.sum()))]
    fw = rng.random(nreads) < 0.5
    return reads, (q + 33).astype(np.uint8), pos, fw


def simulate_pairs(rng, genome, npairs, rdlen, frag_av=250.0, frag_sd=45.0):
    """simulate.pl --fr pairs (:191-207): fragment length N(frag_av, frag_sd) raised to at least 2 * rdlen, fragment
    strand by coin flip, mate 1 = the fragment's first rdlen bases, mate 2 = the reverse complement of its last rdlen.
    Returns (reads[2n, rdlen] with mate 1 / mate 2 interleaved, quals, pos[2n] = leftmost genome offset of each mate,
    fw[2n] = mate given in genome orientation)."""
    G = len(genome)
    flen = np.maximum(rng.normal(frag_av, frag_sd, size=npairs).astype(np.int64), 2 * rdlen)
    fs = rng.integers(0, G - flen - rdlen - 64, size=npairs)
    posl, posr = fs, fs + flen - rdlen
    rl, ql, _, _ = simulate_reads(rng, genome, npairs, rdlen, pos=posl)
    rr, qr, _, _ = simulate_reads(rng, genome, npairs, rdlen, pos=posr)
    comp = np.array([3, 2, 1, 0, 4], dtype=np.uint8)
    frag_fw = rng.random(npairs) < 0.5
    reads = np.empty((2 * npairs, rdlen), dtype=np.uint8)
    quals = np.empty((2 * npairs, rdlen), dtype=np.uint8)
    pos = np.empty(2 * npairs, dtype=np.int64)
    fw = np.empty(2 * npairs, dtype=bool)
    rc_r, rcq_r = comp[rr[:, ::-1]], qr[:, ::-1]
    # Watson fragment: mate 1 = left piece as is, mate 2 = rc(right piece); Crick fragment: mate 1 = rc(right piece), mate 2 = left piece
    reads[0::2] = np.where(frag_fw[:, None], rl, rc_r); quals[0::2] = np.where(frag_fw[:, None], ql, rcq_r)
    reads[1::2] = np.where(frag_fw[:, None], rc_r, rl); quals[1::2] = np.where(frag_fw[:, None], rcq_r, ql)
    pos[0::2] = np.where(frag_fw, posl, posr); pos[1::2] = np.where(frag_fw, posr, posl)
    fw[0::2] = frag_fw; fw[1::2] = ~frag_fw
    return reads, quals, pos, fw


def dp_extend_batch(rng, genome, nreads, rdlen, sc, maxgap=15):
    """One seed-extension DP problem per read: the window DynProgFramer::frameSeedExtensionRect
    (dp_framer.cpp:81-129) frames around the read's true diagonal:
    refl = off - 2*maxgap, refr = off + rdlen - 1 + 2*maxgap, core diagonals [maxgap, 3*maxgap].
    Returns the pooled arrays of bt2gpu_dp_align_batch; all reads are aligned as given ('+')."""
    reads, quals, pos, _ = simulate_reads(rng, genome, nreads, rdlen)
    G = len(genome)
    ncol = rdlen + 4 * maxgap
    refl = pos - 2 * maxgap
    cols = refl[:, None] + np.arange(ncol + 1)[None, :]
    inside = (cols >= 0) & (cols < G)
    codes = np.where(inside, genome[np.clip(cols, 0, G - 1)], 4).astype(np.uint8)
    win = (1 << codes).astype(np.uint8)
    jobs = np.zeros(nreads, dtype=structs.dp_job)
    jobs["rd_off"] = np.arange(nreads, dtype=np.int64) * rdlen
    jobs["rd_len"] = rdlen
    jobs["rf_off"] = np.arange(nreads, dtype=np.int64) * (ncol + 1)
    jobs["ncol"] = ncol
    jobs["minsc"] = sc.score_min.f_int(rdlen) if sc.monotone else max(1, sc.score_min.f_int(rdlen))
    jobs["nceil"] = sc.nceil(rdlen)
    jobs["refl"] = refl
    jobs["triml"] = 0
    jobs["corel"] = maxgap
    jobs["corer"] = 3 * maxgap
    jobs["fw"] = 1
    jobs["enable8"] = 1
    return (np.ascontiguousarray(reads.reshape(-1)), np.ascontiguousarray(quals.reshape(-1)),
            np.ascontiguousarray(win.reshape(-1)), jobs)


def problems_to_plines(seq, qual, win, jobs, path, reflen=1 << 30):
    """Write a batch as the self-contained 'P' lines oracle/ref_dp_dump.cpp reads."""
    m2c = np.frombuffer(b"?ACMGRSVTWYHKDBNN", dtype=np.uint8)
    c2a = np.frombuffer(b"ACGTN", dtype=np.uint8)
    with open(path, "w") as fh:
        for j in jobs:
            s = c2a[seq[j["rd_off"]: j["rd_off"] + j["rd_len"]]].tobytes().decode()
            q = qual[j["rd_off"]: j["rd_off"] + j["rd_len"]].tobytes().decode()
            w = m2c[win[j["rf_off"]: j["rf_off"] + j["ncol"] + 1]].tobytes().decode()
            refl = int(j["refl"]); refr = refl + int(j["ncol"]) - 1
            rect = "%d,%d,%d,%d,%d,0,%d,%d,%d" % (refl, refr, refl, refr, int(j["triml"]), int(j["corel"]),
                                                int(j["corer"]), int(j["corel"]))
            fh.write("P\t%s\t%s\t%s\t0\t%d\t%d\t%s\t%s\n" % (s, q, "+" if j["fw"] else "-", reflen,
                                                        int(j["minsc"]), rect, w))
