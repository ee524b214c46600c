/*
 * fm_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * Plain-C restatement of the reference's FM-index hot path (Bowtie 2 v2.5.4 as in
 * sfiligoi/bowtie2-server), used only as the checker in tests/, smoke() and the
 * cpu_baseline leg of bench.py.  Parity is PINNED: tests/test_oracle_fm.py checks every
 * function below against dumps of the unmodified reference (oracle/ref_seed_dump.cpp,
 * tests/golden/fm_*.txt).  Citations are file:line in the reference tree.
 *
 * Data model = the C ABI's (include/bt2gpu.h): bt2gpu_index_desc with HOST pointers, read
 * pools of bytes 0..4 + Phred+33 qualities.
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <stdlib.h>
#include "../include/bt2gpu.h"

#define OFF_MASK 0xffffffffu
#define SIDE_SZ 64
#define SIDE_BWT_LEN 192

typedef bt2gpu_index_desc Idx;

/* SideLocus (bt2_idx.h:306-484) */
typedef struct { uint32_t side_off; int by, bp; int valid; } Locus;

static void loc_from_row(Locus *l, uint32_t row) {          /* initFromRow, bt2_idx.h:369-396 */
	uint32_t side = row / SIDE_BWT_LEN;
	int char_off = (int)(row % SIDE_BWT_LEN);
	l->side_off = side * SIDE_SZ;
	l->by = char_off >> 2;
	l->bp = char_off & 3;
	l->valid = 1;
}

/* INIT_LOCS (aligner_seed.h:1847-1855) + initFromTopBot (bt2_idx.h:327-352): whether bot
 * shares top's side changes nothing numerically, so both loci come from their rows. */
static void init_locs(uint32_t top, uint32_t bot, Locus *tl, Locus *bl) {
	loc_from_row(tl, top);
	if(bot - top == 1) bl->valid = 0;
	else loc_from_row(bl, bot);
}

static int row_l(const Idx *ix, const Locus *l) {           /* rowL, bt2_idx.h:2247 */
	return (ix->ebwt[l->side_off + l->by] >> (2 * l->bp)) & 3;
}

/* countUpToEx (bt2_idx.h:2038-2080): occurrences of A,C,G,T among the first 4*by+bp
 * bases of the side (the reference's word/LUT split is an implementation detail). */
static void count_upto_ex(const Idx *ix, const Locus *l, uint32_t *arrs) {
	const uint8_t *side = ix->ebwt + l->side_off;
	int n = 4 * l->by + l->bp;
	for(int i = 0; i < n; i++) arrs[(side[i >> 2] >> (2 * (i & 3))) & 3]++;
}

/* '$' is stored as an A and must not be counted (bt2_idx.h:1766-1774, 1891-1899) */
static int z_fix(const Idx *ix, const Locus *l) {
	uint32_t zside = ix->zoff / SIDE_BWT_LEN, zchar = ix->zoff % SIDE_BWT_LEN;
	uint32_t zbyte = zside * SIDE_SZ + (zchar >> 2);
	int zbp = (int)(zchar & 3);
	if(l->side_off <= zbyte && l->side_off + (uint32_t)l->by >= zbyte) {
		if(l->side_off + (uint32_t)l->by > zbyte || (l->side_off + (uint32_t)l->by == zbyte && l->bp > zbp)) return 1;
	}
	return 0;
}

/* countBt2SideEx (bt2_idx.h:1887-1919) */
static void count_side_ex(const Idx *ix, const Locus *l, uint32_t *arrs) {
	arrs[0] = arrs[1] = arrs[2] = arrs[3] = 0;
	count_upto_ex(ix, l, arrs);
	if(z_fix(ix, l)) arrs[0]--;
	const uint32_t *acgt = (const uint32_t *)(ix->ebwt + l->side_off + SIDE_SZ - 16);
	for(int c = 0; c < 4; c++) arrs[c] += acgt[c] + ix->fchr[c];
}

/* countBt2Side (bt2_idx.h:1758-1793) = mapLF(l, c) (:2344) */
static uint32_t map_lf_c(const Idx *ix, const Locus *l, int c) {
	uint32_t a[4];
	count_side_ex(ix, l, a);
	return a[c];
}

/* mapLF1(row, l, c) (bt2_idx.h:2420-2443) */
static uint32_t map_lf1_c(const Idx *ix, uint32_t row, const Locus *l, int c) {
	if(row_l(ix, l) != c || row == ix->zoff) return OFF_MASK;
	return map_lf_c(ix, l, c);
}

/* mapLF1(row&, l) (bt2_idx.h:2451-2472) */
static int map_lf1(const Idx *ix, uint32_t *row, const Locus *l) {
	if(*row == ix->zoff) return -1;
	int c = row_l(ix, l);
	*row = map_lf_c(ix, l, c);
	return c;
}

/* mapBiLFEx (bt2_idx.h:2372-2413) */
static void map_bi_lf_ex(const Idx *ix, const Locus *tl, const Locus *bl,
                         uint32_t *t, uint32_t *b, uint32_t *tp, uint32_t *bp) {
	count_side_ex(ix, tl, t);
	count_side_ex(ix, bl, b);
	bp[0] = tp[0] + (b[0] - t[0]);
	tp[1] = bp[0];
	bp[1] = tp[1] + (b[1] - t[1]);
	tp[2] = bp[1];
	bp[2] = tp[2] + (b[2] - t[2]);
	tp[3] = bp[2];
	bp[3] = tp[3] + (b[3] - t[3]);
}

/* ftabHi / ftabLo / ftabLoHi (bt2_idx.h:1425-1554) */
static uint32_t ftab_hi(const Idx *ix, uint32_t i) {
	if(ix->ftab[i] <= ix->len) return ix->ftab[i];
	return ix->eftab[(ix->ftab[i] ^ OFF_MASK) * 2 + 1];
}
static uint32_t ftab_lo(const Idx *ix, uint32_t i) {
	if(ix->ftab[i] <= ix->len) return ix->ftab[i];
	return ix->eftab[(ix->ftab[i] ^ OFF_MASK) * 2];
}
/* ftabSeqToInt (bt2_idx.h:1374-1398) */
static uint32_t ftab_seq_to_int(const Idx *ix, const uint8_t *seq, size_t off, int rev) {
	int fc = ix->ftab_chars;
	size_t lo = off, hi = lo + fc;
	uint32_t v = 0;
	for(int i = 0; i < fc; i++) {
		int fwex = ix->is_fw;
		if(rev) fwex = !fwex;
		int c = fwex ? seq[lo + i] : seq[hi - i - 1];
		if(c > 3) return OFF_MASK;
		v = (v << 2) | (uint32_t)c;
	}
	return v;
}
static int ftab_lo_hi_seq(const Idx *ix, const uint8_t *seq, size_t off, int rev, uint32_t *top, uint32_t *bot) {
	uint32_t fi = ftab_seq_to_int(ix, seq, off, rev);
	if(fi == OFF_MASK) return 0;
	*top = ftab_hi(ix, fi);
	*bot = ftab_lo(ix, fi + 1);
	return 1;
}

static uint8_t comp_dna(uint8_t c) { return c > 3 ? 4 : (uint8_t)(3 - c); }

/* the four orientations of a read (read.h:125-131): patFw, patRc, patFwRev, patRcRev */
static void read_views(const uint8_t *fwd, size_t len, uint8_t *pat_rc, uint8_t *pat_fw_rev, uint8_t *pat_rc_rev) {
	for(size_t i = 0; i < len; i++) {
		pat_rc[i] = comp_dna(fwd[len - 1 - i]);
		pat_fw_rev[i] = fwd[len - 1 - i];
		pat_rc_rev[i] = comp_dna(fwd[i]);
	}
}

/* ----------------------------------------------------------------------------------------
 * Exact seeds: SeedAligner::searchAllSeeds -> startSearchSeedBi -> searchSeedBi for
 * SEED_TYPE_EXACT (aligner_seed.cpp:597-718, 1633-1714, 1854-2033).  steps[k] = -(len-k):
 * right-to-left over the forward index; the mirror range rides along (mapBiLFEx).
 * -------------------------------------------------------------------------------------- */
static void seed_exact_one(const Idx *fw, const Idx *bw, const uint8_t *seq, int len, bt2gpu_seed_hit *out) {
	uint32_t topf = 0, botf = 0, topb = 0, botb = 0, bwops = 0;
	int step = 0;
	memset(out, 0, sizeof(*out));
	int ftab_len = fw->ftab_chars;
	/* CacheAndSeed (aligner_seed.cpp:84-112): maxjump == len for an N-free exact seed */
	if(ftab_len > 1 && ftab_len <= len) {
		uint32_t fwi0 = ftab_seq_to_int(fw, seq, (size_t)(len - ftab_len), 0);
		uint32_t bwi0 = ftab_seq_to_int(bw, seq, (size_t)(len - ftab_len), 0);
		topf = ftab_hi(fw, fwi0); botf = ftab_lo(fw, fwi0 + 1);
		if(botf - topf == 0) return;
		topb = ftab_hi(bw, bwi0); botb = topb + (botf - topf);
		step += ftab_len;
	} else {
		int c = seq[len - 1];
		topf = topb = fw->fchr[c];
		botf = botb = fw->fchr[c + 1];
		if(botf - topf == 0) return;
		step++;
	}
	for(; step < len; step++) {
		int off = len - 1 - step;                        /* abs(steps[step]) - 1 */
		int c = seq[off];
		Locus tl, bl;
		if(botf - topf == 1) { loc_from_row(&tl, topf); bl.valid = 0; }   /* nextLocsBi :1364 */
		else init_locs(topf, botf, &tl, &bl);
		uint32_t t[4] = {0, 0, 0, 0}, b[4] = {0, 0, 0, 0}, tp[4], bp[4];
		tp[0] = tp[1] = tp[2] = tp[3] = topb;
		bp[0] = bp[1] = bp[2] = bp[3] = botb;
		if(bl.valid) {
			bwops++;
			out->sides += (tl.side_off == bl.side_off) ? 1 : 2;
			map_bi_lf_ex(fw, &tl, &bl, t, b, tp, bp);
		} else {
			bwops++;
			out->sides++;
			t[c] = map_lf1_c(fw, topf, &tl, c);
			if(t[c] == OFF_MASK) { out->bwops = bwops; return; }
			b[c] = t[c] + 1;
		}
		if(b[c] == t[c]) { out->bwops = bwops; return; }
		topf = t[c]; botf = b[c]; topb = tp[c]; botb = bp[c];
	}
	out->topf = topf; out->botf = botf; out->topb = topb; out->botb = botb; out->bwops = bwops;
}

/* instantiateSeq (aligner_seed.cpp:471-491): window [off, off+len) of the read as it aligns
 * to the Watson strand. */
static void seed_window(const uint8_t *rd, int off, int len, int fw, uint8_t *seq) {
	for(int i = 0; i < len; i++) seq[i] = fw ? rd[off + i] : comp_dna(rd[off + len - i - 1]);
}

int bt2oracle_seed_exact_batch(const Idx *fw, const Idx *bw, const uint8_t *seq_pool, const bt2gpu_read *reads,
                               const bt2gpu_seed_job *jobs, size_t njobs, bt2gpu_seed_hit *out) {
	uint8_t seq[256];
	for(size_t i = 0; i < njobs; i++) {
		const bt2gpu_read *r = &reads[jobs[i].read];
		seed_window(seq_pool + r->rd_off, jobs[i].off, jobs[i].len, jobs[i].fw, seq);
		seed_exact_one(fw, bw, seq, jobs[i].len, &out[i]);
	}
	return 0;
}

/* ----------------------------------------------------------------------------------------
 * SeedAligner::exactSweep (aligner_seed.cpp:750-968)
 * -------------------------------------------------------------------------------------- */
static void sweep_one(const Idx *ix, const uint8_t *rdfw, size_t len, int nofw, int norc, size_t mine_max,
                      bt2gpu_sweep_result *out) {
	uint8_t *rc = (uint8_t *)malloc(len + 1);
	for(size_t i = 0; i < len; i++) rc[i] = comp_dna(rdfw[len - 1 - i]);
	const int ftab_len = ix->ftab_chars;
	Locus tloc[2], bloc[2];
	uint32_t top[2] = {0, 0}, bot[2] = {0, 0};
	size_t dep[2] = {0, 0}, nedit[2] = {0, 0};
	int do_init[2] = {1, 1}, done[2] = {nofw, norc};
	uint64_t bwops = 0;
	while((dep[0] < len && !done[0]) || (dep[1] < len && !done[1])) {
		for(int fwi = 0; fwi < 2; fwi++) {
			if(!(dep[fwi] < len && !done[fwi])) continue;
			const uint8_t *seq = fwi == 0 ? rdfw : rc;
			if(do_init[fwi]) {
				/* exactSweepInit :750-790 */
				top[fwi] = bot[fwi] = 0;
				size_t left = len - dep[fwi];
				int do_ftab = ftab_len > 1 && left >= (size_t)ftab_len;
				if(do_ftab) {
					size_t endi = len - dep[fwi] - 1;
					for(int i = 0; i < ftab_len; i++) if(seq[endi - i] > 3) { do_ftab = 0; break; }
				}
				if(do_ftab) {
					ftab_lo_hi_seq(ix, seq, left - ftab_len, 0, &top[fwi], &bot[fwi]);
					dep[fwi] += (size_t)ftab_len;
				} else {
					int c = seq[len - dep[fwi] - 1];
					if(c < 4) { top[fwi] = ix->fchr[c]; bot[fwi] = ix->fchr[c + 1]; }
					dep[fwi]++;
				}
				/* exactSweepStep :827-846 */
				if(bot[fwi] <= top[fwi]) {
					nedit[fwi]++;
					if(nedit[fwi] >= mine_max) done[fwi] = 1;
					continue;
				}
				init_locs(top[fwi], bot[fwi], &tloc[fwi], &bloc[fwi]);
				do_init[fwi] = 0;
			}
			if(dep[fwi] < len) {
				/* exactSweepMapLF :792-824 */
				int c = seq[len - dep[fwi] - 1];
				if(c > 3) top[fwi] = bot[fwi] = 0;
				else if(bloc[fwi].valid) {
					bwops += 2;
					top[fwi] = map_lf_c(ix, &tloc[fwi], c);
					bot[fwi] = map_lf_c(ix, &bloc[fwi], c);
				} else {
					bwops++;
					top[fwi] = map_lf1_c(ix, top[fwi], &tloc[fwi], c);
					if(top[fwi] == OFF_MASK) top[fwi] = bot[fwi] = 0;
					else bot[fwi] = top[fwi] + 1;
				}
				if(bot[fwi] <= top[fwi]) {
					nedit[fwi]++;
					if(nedit[fwi] >= mine_max) done[fwi] = 1;
					do_init[fwi] = 1;
				} else init_locs(top[fwi], bot[fwi], &tloc[fwi], &bloc[fwi]);
				dep[fwi]++;
			}
		}
	}
	for(int fwi = 0; fwi < 2; fwi++) {
		out->top[fwi] = top[fwi]; out->bot[fwi] = bot[fwi];
		out->nedit[fwi] = (uint16_t)nedit[fwi];
		out->done[fwi] = (uint8_t)done[fwi];
	}
	out->bwops = (uint32_t)bwops;
	free(rc);
}

int bt2oracle_exact_sweep_batch(const Idx *fw, const uint8_t *seq_pool, const bt2gpu_read *reads, size_t nreads,
                                int nofw, int norc, uint32_t mine_max, bt2gpu_sweep_result *out) {
	for(size_t i = 0; i < nreads; i++) {
		memset(&out[i], 0, sizeof(out[i]));
		sweep_one(fw, seq_pool + reads[i].rd_off, reads[i].rd_len, nofw, norc, mine_max, &out[i]);
	}
	return 0;
}

/* ----------------------------------------------------------------------------------------
 * SeedAligner::oneMmSearch (aligner_seed.cpp:973-1323)
 * -------------------------------------------------------------------------------------- */
typedef struct {
	const int32_t *mmpens, *npens; int match;      /* Scoring (scoring.h:241-300) */
} Sc;

static int sc_score(const Sc *sc, int rdc, int refm, int q) {   /* Scoring::score scoring.h:241 */
	if(q < 0) q = 0; if(q > 255) q = 255;
	if(rdc > 3 || refm > 15) return -sc->npens[q];
	if((refm >> rdc) & 1) return sc->match;
	return -sc->mmpens[q];
}

typedef struct { bt2gpu_ee_hit *hits; size_t cap, n; int overflow; } HitSink;
static void sink_add(HitSink *s, uint32_t top, uint32_t bot, int64_t score, int pos, int chr, int qchr, int fw) {
	if(s->n >= s->cap) { s->overflow = 1; return; }
	bt2gpu_ee_hit *h = &s->hits[s->n++];
	memset(h, 0, sizeof(*h));
	h->top = top; h->bot = bot; h->score = (int32_t)score;
	h->pos = (uint16_t)pos; h->chr = (uint8_t)chr; h->qchr = (uint8_t)qchr; h->fw = (uint8_t)fw;
}

static uint32_t onemm_one(const Idx *ebwt_fw, const Idx *ebwt_bw, const Sc *sc, const uint8_t *rdfw,
                          const uint8_t *qfw, size_t len, int64_t minsc, int nceil, int nofw, int norc,
                          int local, int repex, int rep1mm, HitSink *sink) {
	uint64_t bwops = 0;
	size_t ns = 0;
	for(size_t i = 0; i < len; i++) if(rdfw[i] > 3) ns++;
	if(ns > 1) return 0;
	if(ns == 1 && !rep1mm) return 0;
	uint8_t *buf = (uint8_t *)malloc(4 * len + 4);
	uint8_t *pat_rc = buf, *pat_fw_rev = buf + len, *pat_rc_rev = buf + 2 * len, *qual_rev = buf + 3 * len;
	read_views(rdfw, len, pat_rc, pat_fw_rev, pat_rc_rev);
	for(size_t i = 0; i < len; i++) qual_rev[i] = qfw[len - 1 - i];
	size_t half_fw = len >> 1, half_bw = len >> 1;
	if(len & 1) half_bw++;
	Locus tloc, bloc;
	uint32_t t[4] = {0, 0, 0, 0}, b[4] = {0, 0, 0, 0}, tp[4] = {0, 0, 0, 0}, bp[4] = {0, 0, 0, 0};
	uint32_t top = 0, bot = 0, topp = 0, botp = 0;
	for(int fwi = 0; fwi < 2; fwi++) {
		int fw = (fwi == 0);
		if(fw && nofw) continue;
		if(!fw && norc) continue;
		int lim = rep1mm ? 2 : 1;
		for(int ebwtfwi = 0; ebwtfwi < lim; ebwtfwi++) {
			int ebwtfw = (ebwtfwi == 0);
			const Idx *ebwt = ebwtfw ? ebwt_fw : ebwt_bw;
			const Idx *ebwtp = ebwtfw ? ebwt_bw : ebwt_fw;
			const uint8_t *seq = fw ? (ebwtfw ? rdfw : pat_fw_rev) : (ebwtfw ? pat_rc : pat_rc_rev);
			const uint8_t *qual = fw ? (ebwtfw ? qfw : qual_rev) : (ebwtfw ? qual_rev : qfw);
			int ftab_len = ebwt->ftab_chars;
			size_t nea = ebwtfw ? half_fw : half_bw;
			int skip = 0;
			for(size_t dep = 0; dep < nea; dep++) if(seq[len - dep - 1] > 3) { skip = 1; break; }
			if(skip) continue;
			size_t dep = 0;
			if(ftab_len > 1 && (size_t)ftab_len <= nea) {
				int rev = !ebwtfw;
				ftab_lo_hi_seq(ebwt, seq, len - ftab_len, rev, &top, &bot);
				if(rep1mm) ftab_lo_hi_seq(ebwtp, seq, len - ftab_len, rev, &topp, &botp);
				if(bot - top == 0) continue;
				int c = seq[len - ftab_len];
				t[c] = top; b[c] = bot; tp[c] = topp; bp[c] = botp;
				dep = (size_t)ftab_len;
			} else {
				int c = seq[len - 1];
				top = topp = tp[c] = ebwt->fchr[c];
				bot = botp = bp[c] = ebwt->fchr[c + 1];
				if(bot - top == 0) continue;
				dep = 1;
			}
			init_locs(top, bot, &tloc, &bloc);
			int do_continue = 0;
			for(; dep < nea; dep++) {                                   /* near half :1078-1116 */
				int rdc = seq[len - dep - 1];
				tp[0] = tp[1] = tp[2] = tp[3] = topp;
				bp[0] = bp[1] = bp[2] = bp[3] = botp;
				if(bloc.valid) {
					bwops++;
					map_bi_lf_ex(ebwt, &tloc, &bloc, t, b, tp, bp);
					top = t[rdc]; bot = b[rdc];
					if(bot <= top) { do_continue = 1; break; }
					topp = tp[rdc]; botp = bp[rdc];
				} else {
					bwops++;
					top = map_lf1_c(ebwt, top, &tloc, rdc);
					if(top == OFF_MASK) { do_continue = 1; break; }
					bot = top + 1;
					t[rdc] = top; b[rdc] = bot; tp[rdc] = topp; bp[rdc] = botp;
				}
				init_locs(top, bot, &tloc, &bloc);
			}
			if(do_continue) continue;
			for(; dep < len; dep++) {                                   /* far half :1121-1318 */
				int rdc = seq[len - dep - 1];
				int quc = qual[len - dep - 1];
				if(rdc > 3 && nceil == 0) break;
				tp[0] = tp[1] = tp[2] = tp[3] = topp;
				bp[0] = bp[1] = bp[2] = bp[3] = botp;
				int clo = 0, chi = 3, match = 1;
				if(bloc.valid) {
					bwops++;
					map_bi_lf_ex(ebwt, &tloc, &bloc, t, b, tp, bp);
					match = rdc < 4;
					if(rdc < 4) { top = t[rdc]; bot = b[rdc]; topp = tp[rdc]; botp = bp[rdc]; }
					/* rdc == 4: the reference indexes t[4]; the values are never used because
					   match == false ends the far loop below (:1298-1316) */
				} else {
					bwops++;
					clo = map_lf1(ebwt, &top, &tloc);
					match = (clo == rdc);
					if(clo < 0) break;
					t[clo] = top; b[clo] = bot = top + 1;
					bp[clo] = botp; tp[clo] = topp;
					chi = clo;
				}
				if(rep1mm && (ns == 0 || rdc > 3)) {
					for(int j = clo; j <= chi; j++) {
						if(j == rdc || b[j] == t[j]) continue;
						size_t depm = dep + 1;
						uint32_t topm = t[j], botm = b[j], topmp = tp[j], botmp = bp[j];
						uint32_t tm[4], bm[4], tmp[4], bmp[4];
						Locus tlocm, blocm;
						init_locs(topm, botm, &tlocm, &blocm);
						for(; depm < len; depm++) {
							int rdcm = seq[len - depm - 1];
							tmp[0] = tmp[1] = tmp[2] = tmp[3] = topmp;
							bmp[0] = bmp[1] = bmp[2] = bmp[3] = botmp;
							if(blocm.valid) {
								bwops++;
								map_bi_lf_ex(ebwt, &tlocm, &blocm, tm, bm, tmp, bmp);
								/* rdcm < 4 here: the only N of the read (if any) is at `dep` (:1164) */
								topm = tm[rdcm]; botm = bm[rdcm];
								topmp = tmp[rdcm]; botmp = bmp[rdcm];
								if(botm <= topm) break;
							} else {
								bwops++;
								topm = map_lf1_c(ebwt, topm, &tlocm, rdcm);
								if(topm == OFF_MASK) break;
								botm = topm + 1;
							}
							init_locs(topm, botm, &tlocm, &blocm);
						}
						if(depm == len) {
							size_t off5p = dep;
							if(fw == ebwtfw) off5p = len - off5p - 1;
							int64_t score = (int64_t)(len - 1) * sc->match;
							int pen = sc_score(sc, rdc, 1 << j, quc - 33);
							score += pen;
							int valid = 1;
							if(local) {
								int64_t lf = 0, lb = 0;
								for(size_t i = 0; i < len; i++) {
									if(i == dep) { if(lf + pen <= 0) { valid = 0; break; } lf += pen; }
									else lf += sc->match;
									if(len - i - 1 == dep) { if(lb + pen <= 0) { valid = 0; break; } lb += pen; }
									else lb += sc->match;
								}
							}
							if(valid) valid = score >= minsc;
							if(valid) {
								uint32_t toprep = ebwtfw ? topm : topmp, botrep = ebwtfw ? botm : botmp;
								sink_add(sink, toprep, botrep, score, (int)off5p, "ACGT"[j], "ACGTN"[rdc], fw);
							}
						}
					}
				}
				if(bot > top && match) {
					if(dep == len - 1) {
						if(ebwtfw && repex) {
							int64_t score = (int64_t)len * sc->match;
							sink_add(sink, top, bot, score, 0xffff, 0, 0, fw);
						}
						break;
					} else init_locs(top, bot, &tloc, &bloc);
				} else break;
			}
		}
	}
	free(buf);
	return (uint32_t)bwops;
}

int bt2oracle_onemm_batch(const Idx *fw, const Idx *bw, const bt2gpu_scoring *scs, const uint8_t *seq_pool,
                          const uint8_t *qual_pool, const bt2gpu_read *reads, size_t nreads,
                          const bt2gpu_onemm_job *jobs, int local, int repex, int rep1mm,
                          bt2gpu_onemm_result *out, bt2gpu_ee_hit *hits, size_t hit_cap, size_t *nhits) {
	Sc sc = { scs->mmpens, scs->npens, scs->match_bonus };
	HitSink sink = { hits, hit_cap, 0, 0 };
	for(size_t i = 0; i < nreads; i++) {
		out[i].hit_off = (uint32_t)sink.n;
		out[i].bwops = onemm_one(fw, bw, &sc, seq_pool + reads[i].rd_off, qual_pool + reads[i].rd_off,
		                         reads[i].rd_len, jobs[i].minsc, jobs[i].nceil, jobs[i].nofw, jobs[i].norc, local, repex, rep1mm, &sink);
		out[i].nhits = (uint32_t)sink.n - out[i].hit_off;
	}
	*nhits = sink.n;
	return sink.overflow;
}

/* ----------------------------------------------------------------------------------------
 * SwDriver::extend (aligner_sw_driver.cpp:299-484)
 * -------------------------------------------------------------------------------------- */
static void extend_dir(const Idx *ebwt, const uint8_t *seq, size_t rdlen, uint32_t top, uint32_t bot,
                       uint32_t topp, uint32_t botp, int fw, size_t off, size_t len, size_t lim, int left,
                       uint32_t *fmops, size_t *nex) {
	uint32_t t[4], b[4], tp[4], bp[4];
	Locus tloc, bloc;
	tp[0] = tp[1] = tp[2] = tp[3] = topp;
	bp[0] = bp[1] = bp[2] = bp[3] = botp;
	init_locs(top, bot, &tloc, &bloc);
	for(size_t ii = 0; ii < lim; ii++) {
		size_t i;
		if(left) i = fw ? off - ii - 1 : rdlen - off - len - 1 - ii;
		else     i = fw ? ii + len + off : rdlen - off + ii;
		int rdc = seq[i];
		if(bloc.valid) {
			(*fmops)++;
			map_bi_lf_ex(ebwt, &tloc, &bloc, t, b, tp, bp);
			int nonz = -1, abort = 0;
			size_t orig_sz = bot - top;
			for(int j = 0; j < 4; j++) {
				if(b[j] > t[j]) {
					if(nonz >= 0) { abort = 1; break; }
					nonz = j; top = t[j]; bot = b[j];
				}
			}
			if(abort || (nonz != rdc && rdc <= 3) || bot - top < orig_sz) break;
		} else {
			(*fmops)++;
			int c = map_lf1(ebwt, &top, &tloc);
			if(c != rdc && rdc <= 3) break;
			bot = top + 1;
		}
		if(++(*nex) == 255) break;
		init_locs(top, bot, &tloc, &bloc);
	}
}

int bt2oracle_extend_batch(const Idx *fw, const Idx *bw, const uint8_t *seq_pool, const bt2gpu_read *reads,
                           const bt2gpu_extend_job *jobs, size_t njobs, bt2gpu_extend_result *out) {
	for(size_t k = 0; k < njobs; k++) {
		const bt2gpu_extend_job *j = &jobs[k];
		const uint8_t *rd = seq_pool + reads[j->read].rd_off;
		size_t rdlen = reads[j->read].rd_len;
		uint8_t *rc = (uint8_t *)malloc(rdlen + 1);
		for(size_t i = 0; i < rdlen; i++) rc[i] = comp_dna(rd[rdlen - 1 - i]);
		const uint8_t *seq = j->fw ? rd : rc;
		size_t nlex = 0, nrex = 0; uint32_t fmops = 0;
		size_t lim = j->fw ? j->off : rdlen - j->len - j->off;
		if(lim > 0) extend_dir(fw, seq, rdlen, j->topf, j->botf, j->topb, j->botb, j->fw, j->off, j->len, lim, 1, &fmops, &nlex);
		lim = j->fw ? rdlen - j->len - j->off : j->off;
		if(lim > 0 && bw) extend_dir(bw, seq, rdlen, j->topb, j->botb, j->topf, j->botf, j->fw, j->off, j->len, lim, 0, &fmops, &nrex);
		out[k].nlex = (uint16_t)nlex; out[k].nrex = (uint16_t)nrex; out[k].fmops = fmops;
		free(rc);
	}
	return 0;
}

/* ----------------------------------------------------------------------------------------
 * Ebwt::getOffset (bt2_idx.cpp:145-171)
 * -------------------------------------------------------------------------------------- */
int bt2oracle_locate_batch(const Idx *ix, const uint32_t *rows, size_t n, uint32_t *offs_out, uint32_t *steps_out) {
	const uint32_t off_mask = OFF_MASK << ix->off_rate;
	for(size_t i = 0; i < n; i++) {
		uint32_t row = rows[i], jumps = 0, res;
		if(row == ix->zoff) res = 0;
		else if((row & off_mask) == row) res = ix->offs[row >> ix->off_rate];
		else {
			Locus l;
			loc_from_row(&l, row);
			for(;;) {
				uint32_t newrow = map_lf_c(ix, &l, row_l(ix, &l));     /* mapLF(l) :2313 */
				jumps++;
				row = newrow;
				if(row == ix->zoff) { res = jumps; break; }
				else if((row & off_mask) == row) { res = jumps + ix->offs[row >> ix->off_rate]; break; }
				loc_from_row(&l, row);
			}
		}
		offs_out[i] = res;
		if(steps_out) steps_out[i] = jumps;
	}
	return 0;
}

/* ----------------------------------------------------------------------------------------
 * Multi-zone seeds, -N 1 / -N 2: Seed::mmSeeds (aligner_seed.cpp:365-448),
 * Seed::instantiate (:213-358), startSearchSeedBi (:1633-1714), nextLocsBi (:1364-1432),
 * searchSeedBi with its mismatch recursion (:1854-2033), reportHit (:1575-1630), and the order
 * in which searchAllSeeds (:597-718) runs the seed types of one offset (lock step, recursions
 * depth first).  -N 0 goes through the same code (one exact seed): it must agree with
 * seed_exact_one above.  Pinned by tests/test_oracle_fm.py against ref_seed_dump --seedmms.
 * The CUDA kernel for these seeds does not exist yet; this is the checker it will be held to.
 * -------------------------------------------------------------------------------------- */
#define MAX_I 0x7fffffff
#define MM_MAXLEN 32
enum { SEED_EXACT = 1, SEED_L2R = 2, SEED_R2L = 3, SEED_INSIDE_OUT = 4 };   /* aligner_seed.h SEED_TYPE_* */

typedef struct { int edits, mms, ins, dels, penalty, editsCeil, mmsCeil, insCeil, delsCeil, penaltyCeil; } Cons;
static Cons cons_init(void) { Cons c = { MAX_I, MAX_I, MAX_I, MAX_I, MAX_I, MAX_I, MAX_I, MAX_I, MAX_I, MAX_I }; return c; }
static Cons cons_exact(void) { Cons c = cons_init(); c.edits = c.mms = c.ins = c.dels = c.penalty = 0; return c; }
static Cons cons_mm(int n) { Cons c = cons_init(); c.mms = n; c.edits = c.dels = c.ins = 0; return c; }
static int cons_must_match(const Cons *c) {                    /* aligner_seed.h:69-74 */
	return (c->mms == 0 && c->edits == 0) || c->penalty == 0 || (c->mms == 0 && c->dels == 0 && c->ins == 0);
}
static int cons_can(const Cons *c, int pen) { return (c->mms > 0 || c->edits > 0) && c->penalty >= pen; }   /* canMismatch / canN */
static void cons_charge(Cons *c, int pen) { if(c->mms == 0) c->edits--; else c->mms--; c->penalty -= pen; }
static int cons_acceptable(const Cons *c) {
	return c->edits <= c->editsCeil && c->mms <= c->mmsCeil && c->ins <= c->insCeil && c->dels <= c->delsCeil && c->penalty <= c->penaltyCeil;
}
static int cons_can_gap(const Cons *c) { return ((c->ins > 0 || c->dels > 0) || c->edits > 0) && c->penalty > 0; }

typedef struct { int type; Cons zones[3]; } SeedPol;
static int mm_seed_pols(int mms, SeedPol *p) {                 /* Seed::mmSeeds */
	int n = 0;
	if(mms == 0) {
		p[n].type = SEED_EXACT; p[n].zones[0] = p[n].zones[1] = p[n].zones[2] = cons_exact(); n++;
	} else if(mms == 1) {
		p[n].type = SEED_L2R; p[n].zones[0] = cons_exact(); p[n].zones[1] = cons_mm(1); p[n].zones[2] = cons_exact(); n++;
		p[n].type = SEED_R2L; p[n].zones[0] = cons_exact(); p[n].zones[1] = cons_mm(1); p[n].zones[1].mmsCeil = 0; p[n].zones[2] = cons_exact(); n++;
	} else {
		p[n].type = SEED_L2R; p[n].zones[0] = cons_exact(); p[n].zones[1] = cons_mm(2); p[n].zones[2] = cons_exact(); n++;
		p[n].type = SEED_R2L; p[n].zones[0] = cons_exact(); p[n].zones[1] = cons_mm(2); p[n].zones[1].mmsCeil = 1; p[n].zones[2] = cons_exact(); n++;
		p[n].type = SEED_INSIDE_OUT; p[n].zones[0] = cons_exact(); p[n].zones[1] = cons_mm(1); p[n].zones[1].mmsCeil = 0;
		p[n].zones[2] = cons_mm(1); p[n].zones[2].mmsCeil = 0; n++;
	}
	return n;
}

typedef struct {
	int n, steps[MM_MAXLEN], z1[MM_MAXLEN], z2[MM_MAXLEN];     /* InstantiatedSeed::steps / zones (first, second) */
	Cons cons[3], overall;
	int maxjump;
} InstSeed;

/* Seed::instantiate; returns 0 when the Ns of the seed window disqualify it */
static int seed_instantiate(const SeedPol *pol, int seedlen, const uint8_t *seq, const uint8_t *qual, const Sc *sc, InstSeed *is) {
	is->n = seedlen;
	switch(pol->type) {
	case SEED_EXACT:
		for(int k = 0; k < seedlen; k++) { is->steps[k] = -(seedlen - k); is->z1[k] = is->z2[k] = 0; }
		break;
	case SEED_L2R:
		for(int k = 0; k < seedlen; k++) { is->steps[k] = k + 1; is->z1[k] = is->z2[k] = (k < (seedlen + 1) / 2) ? 0 : 1; }
		is->z1[seedlen - 1] = is->z2[seedlen - 1] = -1;
		break;
	case SEED_R2L:
		for(int k = 0; k < seedlen; k++) {
			is->steps[k] = -(seedlen - k);
			is->z1[k] = (k < seedlen / 2) ? 0 : 1;
			is->z2[k] = (k < (seedlen + 1) / 2 + 1) ? 0 : 1;
		}
		is->z1[seedlen - 1] = is->z2[seedlen - 1] = -1;
		break;
	default: {
		int step = 0;
		for(int k = (seedlen + 3) / 4; k < seedlen - (seedlen / 4); k++) { is->z1[step] = is->z2[step] = 0; is->steps[step++] = k + 1; }
		for(int k = seedlen - (seedlen / 4); k < seedlen; k++) { is->z1[step] = is->z2[step] = 1; is->steps[step++] = k + 1; }
		is->z1[step - 1] = is->z2[step - 1] = -1;
		for(int k = ((seedlen + 3) / 4) - 1; k >= 0; k--) { is->z1[step] = is->z2[step] = 2; is->steps[step++] = -(k + 1); }
		is->z1[step - 1] = is->z2[step - 1] = -2;
		break;
	}
	}
	for(int i = 0; i < 3; i++) is->cons[i] = pol->zones[i];
	is->overall = cons_init();                                 /* mmSeeds: oall.init() */
	int streak = 1, ltr = is->steps[0] > 0;
	is->maxjump = 0;
	for(int i = 0; i < is->n; i++) {
		const int off = abs(is->steps[i]) - 1;
		Cons *cons = &is->cons[abs(is->z1[i])];
		const int c = seq[off], q = qual[off];
		if(ltr != (is->steps[i] > 0) || is->z1[i] != 0 || is->z2[i] != 0) streak = 0;
		if(c == 4) {
			if(cons_can(cons, sc->npens[q])) cons_charge(cons, sc->npens[q]);
			else return 0;
		}
		if(streak) is->maxjump++;
	}
	return 1;
}

typedef struct { uint32_t topf, botf, topb, botb; uint64_t key; uint32_t job; } MmHit;
typedef struct { MmHit *hits; size_t cap, n; int overflow; uint64_t bwops, bwedits; const Idx *fw, *bw; const Sc *sc; uint32_t job; } MmCtx;

typedef struct {
	const InstSeed *s; const uint8_t *seq, *qual;
	int step, depth;
	uint32_t topf, botf, topb, botb;
	Locus tloc, bloc;
	Cons cv[3], overall;
	int nedits, epos[4], echr[4];
	int hasi0; uint32_t fwi0, bwi0;
} MmParam;

static void mm_report(MmCtx *x, const MmParam *p) {            /* reportHit: the reference string = seed with its edits applied */
	uint8_t rf[MM_MAXLEN];
	memcpy(rf, p->seq, (size_t)p->s->n);
	for(int e = 0; e < p->nedits; e++) rf[p->epos[e]] = (uint8_t)p->echr[e];
	uint64_t key = 0;
	for(int i = 0; i < p->s->n; i++) key = (key << 2) | rf[i];
	if(x->n >= x->cap) { x->overflow = 1; return; }
	MmHit *h = &x->hits[x->n++];
	h->topf = p->topf; h->botf = p->botf; h->topb = p->topb; h->botb = p->botb; h->key = key; h->job = x->job;
}

static void mm_next_locs(const InstSeed *s, Locus *tl, Locus *bl, uint32_t topf, uint32_t botf, uint32_t topb, uint32_t botb, int step) {
	if(step == s->n) return;
	if(s->steps[step] > 0) {                                   /* left to right: BWT' */
		if(botb - topb == 1) { loc_from_row(tl, topb); bl->valid = 0; }
		else init_locs(topb, botb, tl, bl);
	} else {
		if(botf - topf == 1) { loc_from_row(tl, topf); bl->valid = 0; }
		else init_locs(topf, botf, tl, bl);
	}
}

static int mm_start(MmCtx *x, MmParam *p) {                    /* startSearchSeedBi; 1 = already done */
	const InstSeed *s = p->s;
	if(p->step == s->n) { mm_report(x, p); return 1; }
	if(p->step == 0) {
		const int off = abs(s->steps[0]) - 1;
		const int ftab_len = x->fw->ftab_chars;
		if(p->hasi0) {
			p->topf = ftab_hi(x->fw, p->fwi0); p->botf = ftab_lo(x->fw, p->fwi0 + 1);
			if(p->botf - p->topf == 0) return 1;
			p->topb = ftab_hi(x->bw, p->bwi0); p->botb = p->topb + (p->botf - p->topf);
			p->step += ftab_len;
		} else if(s->maxjump > 0) {
			const int c = p->seq[off];
			p->topf = p->topb = x->fw->fchr[c];
			p->botf = p->botb = x->fw->fchr[c + 1];
			if(p->botf - p->topf == 0) return 1;
			p->step++;
		} else {
			p->topf = p->topb = 0;
			p->botf = p->botb = x->fw->fchr[4];
		}
		if(p->step == s->n) { mm_report(x, p); return 1; }
		mm_next_locs(s, &p->tloc, &p->bloc, p->topf, p->botf, p->topb, p->botb, p->step);
	}
	return 0;
}

static void mm_param_init(MmCtx *x, MmParam *p, const InstSeed *s, const uint8_t *seq, const uint8_t *qual) {
	memset(p, 0, sizeof(*p));
	p->s = s; p->seq = seq; p->qual = qual;
	for(int i = 0; i < 3; i++) p->cv[i] = s->cons[i];
	p->overall = s->overall;
	/* CacheAndSeed (aligner_seed.cpp:84-112) */
	int off = abs(s->steps[0]) - 1;
	const int ltr = s->steps[0] > 0, ftab_len = x->fw->ftab_chars;
	p->hasi0 = (ftab_len > 1 && ftab_len <= s->maxjump);
	if(p->hasi0) {
		if(!ltr) off = off - ftab_len + 1;
		p->fwi0 = ftab_seq_to_int(x->fw, seq, (size_t)off, 0);
		p->bwi0 = ftab_seq_to_int(x->bw, seq, (size_t)off, 0);
	}
}

static void mm_search(MmCtx *x, MmParam *pv, size_t nparams) { /* searchSeedBi */
	int done[8];
	size_t nleft = nparams;
	for(size_t n = 0; n < nparams; n++) { done[n] = mm_start(x, &pv[n]); if(done[n]) nleft--; }
	while(nleft > 0) {
		for(size_t n = 0; n < nparams; n++) {
			if(done[n]) continue;
			MmParam *p = &pv[n];
			const InstSeed *s = p->s;
			if(p->step >= s->n) { done[n] = 1; nleft--; continue; }
			const int i = p->step++;
			/* SeedAlignerSearchState::setOff (:1744-1770) */
			uint32_t t[4] = {0, 0, 0, 0}, b[4] = {0, 0, 0, 0}, tp[4], bp[4], *tf, *tb, *bf, *bb, ntop;
			const int ltr = s->steps[i] > 0, off = abs(s->steps[i]) - 1;
			const Idx *ebwt;
			if(ltr) { ebwt = x->bw; for(int k = 0; k < 4; k++) { tp[k] = p->topf; bp[k] = p->botf; } tf = tp; tb = t; bf = bp; bb = b; ntop = p->topb; }
			else    { ebwt = x->fw; for(int k = 0; k < 4; k++) { tp[k] = p->topb; bp[k] = p->botb; } tf = t; tb = tp; bf = b; bb = bp; ntop = p->topf; }
			if(p->bloc.valid) { x->bwops++; map_bi_lf_ex(ebwt, &p->tloc, &p->bloc, t, b, tp, bp); }
			const int c = p->seq[off];
			const int leave_zone = s->z1[i] < 0;
			Cons *cons = &p->cv[abs(s->z1[i])];
			if(!(cons_must_match(cons) && !cons_must_match(&p->overall)) || c == 4) {
				int bail = 0;
				if(!p->bloc.valid) {
					x->bwops++;
					const int cc = map_lf1(ebwt, &ntop, &p->tloc);       /* advances ntop, as the reference's call does */
					if(cc < 0) bail = 1;
					else { t[cc] = ntop; b[cc] = ntop + 1; }
				}
				if(!bail) {
					const int q = p->qual[off], pen = x->sc->mmpens[q];     /* the raw quality character indexes the table (:1934) */
					if((cons_can(cons, pen) && cons_can(&p->overall, pen)) || c == 4) {
						const Cons save_c = *cons, save_o = p->overall;
						if(c != 4) { cons_charge(cons, pen); cons_charge(&p->overall, pen); }
						if(!leave_zone || (cons_acceptable(cons) && cons_acceptable(&p->overall))) {
							for(int j = 0; j < 4; j++) {
								if(j == c || b[j] == t[j]) continue;
								MmParam p2 = *p;
								p2.topf = tf[j]; p2.botf = bf[j]; p2.topb = tb[j]; p2.botb = bb[j];
								p2.epos[p2.nedits] = off; p2.echr[p2.nedits] = j; p2.nedits++;
								mm_next_locs(s, &p2.tloc, &p2.bloc, p2.topf, p2.botf, p2.topb, p2.botb, i + 1);
								x->bwedits++;
								p2.step = i + 1; p2.depth = p->depth + 1;
								mm_search(x, &p2, 1);
							}
						}
						*cons = save_c; p->overall = save_o;
					}
					if(cons_can_gap(cons) && cons_can_gap(&p->overall)) { x->overflow = 2; return; }   /* `throw 1; // TODO` in the reference */
				}
			}
			if(c == 4) { done[n] = 1; nleft--; continue; }
			if(leave_zone && (!cons_acceptable(cons) || !cons_acceptable(&p->overall))) { done[n] = 1; nleft--; continue; }
			if(!p->bloc.valid) {
				x->bwops++;
				t[c] = map_lf1_c(ebwt, ntop, &p->tloc, c);
				if(t[c] == OFF_MASK) { done[n] = 1; nleft--; continue; }
				b[c] = t[c] + 1;
			}
			if(b[c] == t[c]) { done[n] = 1; nleft--; continue; }
			p->topf = tf[c]; p->botf = bf[c]; p->topb = tb[c]; p->botb = bb[c];
			if(i + 1 == s->n) { mm_report(x, p); done[n] = 1; nleft--; continue; }
			mm_next_locs(s, &p->tloc, &p->bloc, p->topf, p->botf, p->topb, p->botb, i + 1);
		}
	}
}

/* One job = one seed offset of one read strand (what instantiateSeeds lays out); its hits come back in the order of the
 * reportHit calls.  jobs[].len <= 32.  out_n[j] = hits of job j (they are contiguous in `hits`).  Returns 0, or 1 when
 * `cap` is too small. */
int bt2oracle_seed_mm_batch(const Idx *fw, const Idx *bw, const bt2gpu_scoring *scs, const uint8_t *seq_pool, const uint8_t *qual_pool,
                            const bt2gpu_read *reads, const bt2gpu_seed_job *jobs, size_t njobs, int mms,
                            uint32_t *hit_job, uint32_t *hit_ranges /* 4 per hit */, uint64_t *hit_keys, size_t cap, size_t *nhits,
                            uint32_t *out_n, uint64_t *bwops)
{
	Sc sc; sc.mmpens = scs->mmpens; sc.npens = scs->npens; sc.match = scs->match_bonus;
	MmCtx x; memset(&x, 0, sizeof(x));
	x.hits = (MmHit*)malloc((cap + 1) * sizeof(MmHit)); x.cap = cap; x.fw = fw; x.bw = bw; x.sc = &sc;
	SeedPol pols[3];
	const int npol = mm_seed_pols(mms, pols);
	for(size_t j = 0; j < njobs; j++) {
		const bt2gpu_seed_job *jb = &jobs[j];
		const bt2gpu_read *rd = &reads[jb->read];
		uint8_t seq[MM_MAXLEN], qual[MM_MAXLEN];
		const int len = jb->len;
		/* instantiateSeq (aligner_seed.cpp:471-491): the window as it aligns to the Watson strand, qualities alongside */
		seed_window(seq_pool + rd->rd_off, jb->off, len, jb->fw, seq);
		for(int i = 0; i < len; i++) qual[i] = jb->fw ? qual_pool[rd->rd_off + jb->off + i] : qual_pool[rd->rd_off + jb->off + len - 1 - i];
		InstSeed is[3]; MmParam pv[3]; size_t np = 0;
		for(int k = 0; k < npol; k++)
			if(seed_instantiate(&pols[k], len, seq, qual, &sc, &is[np])) { mm_param_init(&x, &pv[np], &is[np], seq, qual); np++; }
		x.job = (uint32_t)j;
		const size_t n0 = x.n;
		if(np > 0) mm_search(&x, pv, np);
		out_n[j] = (uint32_t)(x.n - n0);
		if(x.overflow) break;
	}
	for(size_t i = 0; i < x.n; i++) {
		hit_job[i] = x.hits[i].job; hit_keys[i] = x.hits[i].key;
		hit_ranges[4 * i] = x.hits[i].topf; hit_ranges[4 * i + 1] = x.hits[i].botf; hit_ranges[4 * i + 2] = x.hits[i].topb; hit_ranges[4 * i + 3] = x.hits[i].botb;
	}
	*nhits = x.n; *bwops = x.bwops;
	const int rc = x.overflow;
	free(x.hits);
	return rc;
}
