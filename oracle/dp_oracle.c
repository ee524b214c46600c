/*
 * oracle/dp_oracle.c — TEST INFRASTRUCTURE ONLY (CPU restatement, plain C).
 *
 * Restates, as straight scalar recurrences, what the reference's striped SSE
 * Smith-Waterman path computes (sfiligoi/bowtie2-server, Bowtie 2 v2.5.4):
 *   SwAligner::align            aligner_sw.cpp:500-729
 *   alignNucleotides*Sse{U8,I16} aligner_swsse_ee_u8.cpp:775-1146, _ee_i16.cpp:780-1171,
 *                               _loc_u8.cpp:927-1336, _loc_i16.cpp:938-1386
 *   gatherCells*                ee_u8.cpp:1176-1208, loc_u8.cpp:1389-1499 (+ i16 twins)
 *   backtraceNucleotides*       ee_u8.cpp:1283-1877 (+ 3 twins that differ only in
 *                               floorsc/offsetsc, see diff notes in DESIGN.md)
 *   SwAligner::nextAlignment    aligner_sw.cpp:737-1146
 * Parity is PINNED: tests/test_oracle_dp.py checks this file against dumps of
 * the compiled reference itself (oracle/ref_dp_dump.cpp -> tests/golden/).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
 * load this library; the product (libbt2gpu.so) never does.
 *
 * Why a scalar recurrence equals the striped + lazy-F code: every SSE op used
 * is a saturating add/sub or a max, all monotone, so the lazy-F fix-up loop
 * (ee_u8.cpp:1025-1076) converges to the unique fixed point of
 *     F(i,j) = max(F(i-1,j) (-) rfe, H(i-1,j) (-) rfo) (-) veto(i)
 *     H(i,j) = max(H(i-1,j-1) (+) s(i,j), E(i,j), F(i,j))
 *     E(i,j+1) = max(E(i,j) (-) rde, H(i,j) (-) rdo (-) veto(i))
 * (re-opening from an H raised by F is dominated by extending F because
 * rfe <= rfo, asserted at ee_u8.cpp:836).  Values are kept here in the SCORE
 * domain (raw - 0xff, raw - 0x7fff, raw, raw + 0x8000) with the variant's
 * floor; the stripe padding rows (NWORDS_PER_REG*seglen - len pseudo-rows with
 * score 0 and no gap barrier, ee_u8.cpp:108-139) are materialised because in
 * local mode they leak into the per-column maximum (loc_u8.cpp:1144,1268-1295).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "../include/bt2gpu.h"

#define MIN_I64 INT64_MIN

static const int firsts5[32] = {
	-1, 0, 1, 0, 2, 0, 1, 0, 3, 0, 1, 0, 2, 0, 1, 0,
	 4, 0, 1, 0, 2, 0, 1, 0, 3, 0, 1, 0, 2, 0, 1, 0 };  /* mask.cpp:31 */
static const int alts5[32] = {
	0, 1, 1, 2, 1, 2, 2, 3, 1, 2, 2, 3, 2, 3, 3, 4,
	1, 2, 2, 3, 2, 3, 3, 4, 2, 3, 3, 4, 3, 4, 4, 5 };      /* mask.cpp:23 */
static const char mask2dna[] = "?ACMGRSVTWYHKDBNN";          /* alphabet.cpp:71 */

/* Scoring::score(rdc, refm, q) scoring.h:241 */
static int sc_score(const bt2gpu_scoring *sc, int rdc, int refm, int q) {
	if(q < 0) q = 0; if(q > 255) q = 255;
	if(rdc > 3 || refm > 15) return -sc->npens[q];
	if((refm & (1 << rdc)) != 0) return sc->match_bonus;
	return -sc->mmpens[q];
}
/* Scoring::mm(rdc, refm, q) scoring.h:232 */
static int sc_mm(const bt2gpu_scoring *sc, int rdc, int refm, int q) {
	if(q < 0) q = 0; if(q > 255) q = 255;
	return (rdc > 3 || refm > 15) ? sc->npens[q] : sc->mmpens[q];
}
/* matchesEx mask.h:37 */
static int matchesEx(int i, int j) {
	if(j >= 16 || i > 3) return -1;
	return (((1 << i) & j) != 0) ? 1 : 0;
}

typedef struct {
	const bt2gpu_scoring *sc;
	const uint8_t *rd, *qu, *rf;   /* rf has ncol+1 masks */
	int nrow, ncol, nrowp;         /* nrowp = rows incl. stripe padding */
	int64_t minsc;
	int local, w16;                /* w16: 16-bit variant */
	int64_t fl;                    /* floor in score domain */
	int32_t *E, *F, *H;            /* [col*nrowp + row], score domain */
	int32_t *colmax;               /* local: per-column max incl. padding rows */
	int colstop, lastsolcol;
	uint8_t *rep;                  /* reportedThrough bit (masks_ bit 0), [row*ncol+col] */
	uint16_t *masks;               /* remaining mask bits, kept for faithfulness */
} dpmat;

static inline int64_t subf(int64_t a, int64_t p, int64_t fl) { a -= p; return a < fl ? fl : a; }
static inline int64_t max2(int64_t a, int64_t b) { return a > b ? a : b; }

/*
 * Fill. Returns `best` and sets *flag like alignNucleotides*Sse*().
 * nwords = NWORDS_PER_REG of the emulated build (16 for u8, 8 for i16; SSE2).
 */
static int64_t dp_fill(dpmat *m, int *flag) {
	const bt2gpu_scoring *sc = m->sc;
	const int nrow = m->nrow, ncol = m->ncol;
	const int nwords = m->w16 ? 8 : 16;
	const int seglen = (nrow + nwords - 1) / nwords;
	const int nrowp = m->local ? seglen * nwords : nrow;
	m->nrowp = nrowp;
	const int64_t fl = m->fl;
	const int64_t rdo = sc->rd_gap_open, rde = sc->rd_gap_ext;
	const int64_t rfo = sc->rf_gap_open, rfe = sc->rf_gap_ext;
	size_t ncell = (size_t)nrowp * (size_t)ncol;
	m->E = (int32_t*)malloc(sizeof(int32_t) * ncell);
	m->F = (int32_t*)malloc(sizeof(int32_t) * ncell);
	m->H = (int32_t*)malloc(sizeof(int32_t) * ncell);
	m->colmax = (int32_t*)calloc((size_t)ncol + 1, sizeof(int32_t));
	/* local u8: bias (loc_u8.cpp:98-110) and saturating add ceiling */
	int bias = 0;
	if(m->local && !m->w16) {
		for(int refc = 0; refc < 5; refc++) for(int i = 0; i < nrow; i++) {
			int s = sc_score(sc, m->rd[i], 1 << refc, (int)m->qu[i] - 33);
			if(s < 0 && s < bias) bias = s;
		}
		bias = -bias;
	}
	const int64_t ceil_u8 = 255, ceil_i16 = 65535;
	int64_t vmax = m->local ? 0 : MIN_I64; /* local: max of column maxima; e2e: last-row max */
	int64_t lrmax = fl; int found = 0;
	m->colstop = m->local ? ncol : ncol - 1; /* loc_u8.cpp:1050 / ee_u8.cpp:906 */
	m->lastsolcol = 0;
	*flag = 0;
	int64_t *Ecur = (int64_t*)malloc(sizeof(int64_t) * (size_t)nrowp);
	int64_t *Hprev = (int64_t*)malloc(sizeof(int64_t) * (size_t)nrowp);
	for(int i = 0; i < nrowp; i++) { Ecur[i] = fl; Hprev[i] = fl; } /* ee_u8.cpp:877-882 */
	int sat = 0;
	for(int j = 0; j < ncol; j++) {
		const int refm = m->rf[j];
		const int refc = firsts5[refm & 31]; /* ee_u8.cpp:916 profile row */
		int64_t Fprev = fl, Hup = fl, cmax = m->local ? 0 : fl;
		int64_t hdiag = m->local ? 0 : 0; /* H(-1,j-1): 0xff / 0x7fff -> score 0; local 0 */
		for(int i = 0; i < nrowp; i++) {
			int gapbar_row = 0; int64_t s = 0;
			if(i < nrow) {
				s = sc_score(sc, m->rd[i], 1 << refc, (int)m->qu[i] - 33);
				if(i < sc->gapbar || (nrow - i - 1) < sc->gapbar) gapbar_row = 1;
			}
			/* F(i,j) */
			int64_t f;
			if(i == 0) f = fl;
			else {
				f = max2(subf(Fprev, rfe, fl), subf(Hup, rfo, fl));
				if(gapbar_row) f = fl;
			}
			/* H(i,j) */
			int64_t d = hdiag;
			if(!m->local) d = subf(d, -s, fl);          /* s <= 0 */
			else if(!m->w16) { d = d + s + bias; if(d > ceil_u8) d = ceil_u8; d -= bias; if(d < 0) d = 0; }
			else { d = d + s; if(d > ceil_i16) d = ceil_i16; if(d < 0) d = 0; }
			int64_t e = Ecur[i];
			int64_t h = max2(d, max2(e, f));
			size_t idx = (size_t)j * nrowp + i;
			m->E[idx] = (int32_t)e; m->F[idx] = (int32_t)f; m->H[idx] = (int32_t)h;
			if(h > cmax) cmax = h;
			/* E(i,j+1) */
			int64_t eo = subf(h, rdo, fl);
			if(gapbar_row) eo = fl;
			Ecur[i] = max2(subf(e, rde, fl), eo);
			/* advance */
			hdiag = Hprev[i]; Hprev[i] = h; Fprev = f; Hup = h;
		}
		if(!m->local) {
			int64_t lr = m->H[(size_t)j * nrowp + (nrow - 1)];
			found = 1;
			if(lr > lrmax) lrmax = lr;
		} else {
			m->colmax[j] = (int32_t)cmax;
			if(cmax > vmax) vmax = cmax;
			if(!m->w16 && cmax + bias >= 255) { sat = 1; break; }   /* loc_u8.cpp:1278 */
			if(cmax < m->minsc) {
				int64_t ncolleft = ncol - j - 1;
				if(cmax + ncolleft * (int64_t)sc->match_bonus < m->minsc) { /* loc_u8.cpp:1284-1291 */
					m->colstop = j + 1;
					break;
				}
			} else m->lastsolcol = j;
		}
	}
	free(Ecur); free(Hprev);
	if(!m->local) {
		if(!found) { *flag = -1; return MIN_I64; }
		int64_t score = lrmax;
		if(score < m->minsc) { *flag = -1; return score; }
		if(lrmax == fl) { *flag = -2; return MIN_I64; }
		return score;
	}
	if(!m->w16) {
		if(sat || vmax + bias >= 255) { *flag = -2; return MIN_I64; }
		if(vmax == 0 || vmax < m->minsc) { *flag = -1; return vmax; }
		return vmax;
	}
	if(vmax == 0) { *flag = -1; return MIN_I64; }           /* ret == MIN_I16 */
	if(vmax < m->minsc) { *flag = -1; return vmax; }
	if(vmax == ceil_i16) { *flag = -2; return MIN_I64; }
	return vmax;
}

static void dp_free(dpmat *m) {
	free(m->E); free(m->F); free(m->H); free(m->colmax); free(m->rep); free(m->masks);
	m->E = m->F = m->H = m->colmax = NULL; m->rep = NULL; m->masks = NULL;
}

static int cand_cmp(const void *a, const void *b) {
	const bt2gpu_dp_cand *x = (const bt2gpu_dp_cand*)a, *y = (const bt2gpu_dp_cand*)b;
	if(x->score != y->score) return x->score > y->score ? -1 : 1; /* aligner_sw_nuc.h:134-142 */
	if(x->row != y->row) return x->row > y->row ? -1 : 1;
	if(x->col != y->col) return x->col > y->col ? -1 : 1;
	return 0;
}

/* gatherCells*: returns number of candidates appended (sorted), or -1 on overflow */
static long dp_gather(dpmat *m, bt2gpu_dp_cand *out, size_t cap) {
	size_t n = 0;
	const int nrow = m->nrow;
	if(!m->local) {
		for(int j = 0; j < m->ncol; j++) {                  /* ee_u8.cpp:1191-1202 */
			int64_t s = m->H[(size_t)j * m->nrowp + (nrow - 1)];
			if(s >= m->minsc) {
				if(n >= cap) return -1;
				out[n].row = (uint32_t)(nrow - 1); out[n].col = (uint32_t)j; out[n].score = (int32_t)s; n++;
			}
		}
	} else {
		uint64_t bonus = (uint64_t)m->sc->match_bonus;
		uint64_t minrow = (((uint64_t)m->minsc + bonus - 1) / bonus) - 1; /* loc_u8.cpp:1408 */
		int ncol = m->lastsolcol + 1;
		for(int j = 0; j < ncol; j++) {
			if((int64_t)m->colmax[j] < m->minsc) continue;   /* loc_u8.cpp:1436 */
			for(int i = 0; i < nrow; i++) {
				if((uint64_t)i < minrow) continue;
				int64_t s = m->H[(size_t)j * m->nrowp + i];
				if(s < m->minsc) continue;
				int readc = m->rd[i], refc = m->rf[j];
				int match = ((refc & (1 << readc)) != 0), matchSucc = 0;
				if(i < nrow - 1) {
					int readcSucc = m->rd[i + 1], refcSucc = m->rf[j + 1];
					matchSucc = ((refcSucc & (1 << readcSucc)) != 0);
				}
				if(match && !matchSucc) {
					if(n >= cap) return -1;
					out[n].row = (uint32_t)i; out[n].col = (uint32_t)j; out[n].score = (int32_t)s; n++;
				}
			}
		}
	}
	qsort(out, n, sizeof(bt2gpu_dp_cand), cand_cmp);
	return (long)n;
}

/* ---- backtrace (ee_u8.cpp:1283-1877; loc/i16 twins differ in floorsc only) -- */

typedef struct { int64_t score; int ns, gaps; } alnscore;
typedef struct {
	size_t nedsz, celsz; int row, col, gaps, readGaps, refGaps; alnscore score; int ct;
} btframe;

enum { CT_E = 0, CT_F = 1, CT_H = 2 };
enum { BT_DIAG = 1, BT_REF_OPEN, BT_READ_OPEN, BT_RDGAP_EXT, BT_RFGAP_EXT };

#define MSK(r, c) m->masks[(size_t)(r) * m->ncol + (c)]
#define REP(r, c) m->rep[(size_t)(r) * m->ncol + (c)]
#define HH(r, c) ((int64_t)m->H[(size_t)(c) * m->nrowp + (r)])
#define EE(r, c) ((int64_t)m->E[(size_t)(c) * m->nrowp + (r)])
#define FF(r, c) ((int64_t)m->F[(size_t)(c) * m->nrowp + (r)])

/*
 * Returns 1 on success.  Edits are appended to ed[*ned...] in the order the
 * reference leaves them after res.reverse() (ascending DP row).
 */
static int dp_backtrace(
	dpmat *m, const bt2gpu_dp_job *job, int64_t escore, int row, int col,
	bt2gpu_edit *ed, size_t edcap, size_t *nedp, bt2gpu_dp_aln *out, int *overflow)
{
	const bt2gpu_scoring *sc = m->sc;
	/* floorsc: MIN_I64 in both end-to-end variants, 0 in both local variants */
	const int64_t floorsc = m->local ? 0 : MIN_I64;
	alnscore score; score.score = 0; score.ns = 0; score.gaps = 0;
	const int origCol = col;
	int gaps = 0, readGaps = 0, refGaps = 0;
	size_t ned = 0;
	const int nrow = m->nrow;
	size_t trimEnd = (size_t)(nrow - row - 1), trimBeg = 0;
	int ct = CT_H;
	size_t stackcap = 64, nstack = 0;
	btframe *stack = (btframe*)malloc(sizeof(btframe) * stackcap);
	size_t cellcap = (size_t)nrow + m->ncol + 8, ncells = 0;
	int *cells = (int*)malloc(sizeof(int) * 2 * cellcap);
	int ok = 0;
	while(row >= 0) {
		int readc = m->rd[row];
		int refm = m->rf[col];
		int readq = m->qu[row];
		int empty = 0, reportedThru, canMoveThru = 1, branch = 0;
		int cur = 0;
		reportedThru = REP(row, col);
		if(reportedThru) {
			canMoveThru = 0;
		} else {
			empty = 0;
			if(row > 0) {
				int rowFromEnd = nrow - row - 1;
				int gapsAllowed = !(row < sc->gapbar || rowFromEnd < sc->gapbar);
				if(ct == CT_E) {
					int64_t sc_cur = EE(row, col);
					int origMask = 0, mask = 0;
					int64_t sc_h_left = HH(row, col - 1);
					if(sc_h_left > floorsc && sc_h_left - sc->rd_gap_open == sc_cur) mask |= 1;
					int64_t sc_e_left = EE(row, col - 1);
					if(sc_e_left > floorsc && sc_e_left - sc->rd_gap_ext == sc_cur) mask |= 2;
					origMask = mask;
					if(MSK(row, col) & (1 << 7)) mask = (MSK(row, col) >> 8) & 3;
					if(mask == 3) { cur = BT_READ_OPEN; MSK(row, col) = (MSK(row, col) & ~(7 << 7)) | (1 << 7 | 2 << 8); branch = 1; }
					else if(mask == 2) { cur = BT_RDGAP_EXT; MSK(row, col) = (MSK(row, col) & ~(7 << 7)) | (1 << 7); }
					else if(mask == 1) { cur = BT_READ_OPEN; MSK(row, col) = (MSK(row, col) & ~(7 << 7)) | (1 << 7); }
					else { empty = 1; canMoveThru = (origMask == 0); }
				} else if(ct == CT_F) {
					int64_t sc_h_up = HH(row - 1, col), sc_f_up = FF(row - 1, col), sc_cur = FF(row, col);
					int origMask = 0, mask = 0;
					if(sc_h_up > floorsc && sc_h_up - sc->rf_gap_open == sc_cur) mask |= 1;
					if(sc_f_up > floorsc && sc_f_up - sc->rf_gap_ext == sc_cur) mask |= 2;
					origMask = mask;
					if(MSK(row, col) & (1 << 10)) mask = (MSK(row, col) >> 11) & 3;
					if(mask == 3) { cur = BT_REF_OPEN; MSK(row, col) = (MSK(row, col) & ~(7 << 10)) | (1 << 10 | 2 << 11); branch = 1; }
					else if(mask == 2) { cur = BT_RFGAP_EXT; MSK(row, col) = (MSK(row, col) & ~(7 << 10)) | (1 << 10); }
					else if(mask == 1) { cur = BT_REF_OPEN; MSK(row, col) = (MSK(row, col) & ~(7 << 10)) | (1 << 10); }
					else { empty = 1; canMoveThru = (origMask == 0); }
				} else {
					int64_t sc_cur = HH(row, col);
					int64_t sc_f_up = FF(row - 1, col), sc_h_up = HH(row - 1, col);
					int64_t sc_h_left = col > 0 ? HH(row, col - 1) : floorsc;
					int64_t sc_e_left = col > 0 ? EE(row, col - 1) : floorsc;
					int64_t sc_h_upleft = col > 0 ? HH(row - 1, col - 1) : floorsc;
					int64_t sc_diag = sc_score(sc, readc, refm, readq - 33);
					int origMask = 0, mask = 0;
					if(gapsAllowed) {
						if(sc_h_up   > floorsc && sc_cur == sc_h_up   - sc->rf_gap_open) mask |= 1;
						if(sc_h_left > floorsc && sc_cur == sc_h_left - sc->rd_gap_open) mask |= 2;
						if(sc_f_up   > floorsc && sc_cur == sc_f_up   - sc->rf_gap_ext)  mask |= 4;
						if(sc_e_left > floorsc && sc_cur == sc_e_left - sc->rd_gap_ext)  mask |= 8;
					}
					if(sc_h_upleft > floorsc && sc_cur == sc_h_upleft + sc_diag) mask |= 16;
					origMask = mask;
					if(MSK(row, col) & (1 << 1)) mask = (MSK(row, col) >> 2) & 31;
					int opts = alts5[mask], select = -1;
					if(opts == 1) {
						select = firsts5[mask];
						MSK(row, col) = (MSK(row, col) & ~(31 << 1)) | (1 << 1);
					} else if(opts > 1) {
						if(mask & 16) select = 4;          /* ee_u8.cpp:1510-1520 */
						else if(mask & 1) select = 0;
						else if(mask & 4) select = 2;
						else if(mask & 2) select = 1;
						else if(mask & 8) select = 3;
						mask &= ~(1 << select);
						MSK(row, col) = (MSK(row, col) & ~(31 << 1)) | (1 << 1 | mask << 2);
						branch = 1;
					}
					if(select != -1) {
						if(select == 4) cur = BT_DIAG;
						else if(select == 0) cur = BT_REF_OPEN;
						else if(select == 1) cur = BT_READ_OPEN;
						else if(select == 2) cur = BT_RFGAP_EXT;
						else cur = BT_RDGAP_EXT;
					} else { empty = 1; canMoveThru = (origMask == 0); }
				}
			}
		}
		REP(row, col) = 1;
		if(!canMoveThru) {
			if(nstack > 0) {
				btframe *f = &stack[nstack - 1];
				ncells = f->celsz; ned = f->nedsz; row = f->row; col = f->col;
				gaps = f->gaps; readGaps = f->readGaps; refGaps = f->refGaps;
				score = f->score; ct = f->ct; nstack--;
				continue;
			}
			goto done;
		}
		if(empty || row == 0) {
			cells[2 * ncells] = row; cells[2 * ncells + 1] = col; ncells++;
			trimBeg = (size_t)row;
			break;
		}
		if(branch) {
			if(nstack == stackcap) { stackcap *= 2; stack = (btframe*)realloc(stack, sizeof(btframe) * stackcap); }
			btframe *f = &stack[nstack++];
			f->nedsz = ned; f->celsz = ncells; f->row = row; f->col = col; f->gaps = gaps;
			f->readGaps = readGaps; f->refGaps = refGaps; f->score = score; f->ct = ct;
		}
		cells[2 * ncells] = row; cells[2 * ncells + 1] = col; ncells++;
		if(ned + 2 >= edcap) { *overflow = 1; goto done; }
		switch(cur) {
		case BT_DIAG: {
			int mm = matchesEx(m->rd[row], m->rf[col]);
			ct = CT_H;
			if(mm != 1) {
				ed[ned].pos = (uint32_t)row; ed[ned].chr = (uint8_t)mask2dna[m->rf[col]];
				ed[ned].qchr = (uint8_t)"ACGTN"[m->rd[row]]; ed[ned].type = BT2GPU_EDIT_MM; ed[ned].pad = 0; ned++;
				score.score -= sc_mm(sc, m->rd[row], m->rf[col], (int)m->qu[row] - 33);
			} else score.score += sc->match_bonus;
			if(mm == -1) score.ns++;
			row--; col--;
			break; }
		case BT_REF_OPEN: case BT_RFGAP_EXT:
			ed[ned].pos = (uint32_t)row; ed[ned].chr = '-'; ed[ned].qchr = (uint8_t)"ACGTN"[m->rd[row]];
			ed[ned].type = BT2GPU_EDIT_REF_GAP; ed[ned].pad = 0; ned++;
			row--;
			if(cur == BT_REF_OPEN) { ct = CT_H; score.score -= sc->rf_gap_open; }
			else { ct = CT_F; score.score -= sc->rf_gap_ext; }
			gaps++; refGaps++;
			break;
		case BT_READ_OPEN: case BT_RDGAP_EXT:
			ed[ned].pos = (uint32_t)row + 1; ed[ned].chr = (uint8_t)mask2dna[m->rf[col]]; ed[ned].qchr = '-';
			ed[ned].type = BT2GPU_EDIT_READ_GAP; ed[ned].pad = 0; ned++;
			col--;
			if(cur == BT_READ_OPEN) { ct = CT_H; score.score -= sc->rd_gap_open; }
			else { ct = CT_E; score.score -= sc->rd_gap_ext; }
			gaps++; readGaps++;
			break;
		default: goto done;
		}
	}
	{
		/* core-diagonal test ee_u8.cpp:1764-1795 */
		int overlapped = 0;
		for(size_t i = 0; i < ncells; i++) {
			int64_t diagi = (int64_t)cells[2 * i + 1] - (int64_t)cells[2 * i] + (int64_t)job->triml;
			if(diagi >= 0 && (uint64_t)diagi >= job->corel && (uint64_t)diagi <= job->corer) { overlapped = 1; break; }
		}
		if(!overlapped) goto done;
		int readC = m->rd[row], refNmask = m->rf[col];
		int mm = matchesEx(readC, refNmask);
		if(mm != 1) {
			ed[ned].pos = (uint32_t)row; ed[ned].chr = (uint8_t)mask2dna[refNmask];
			ed[ned].qchr = (uint8_t)"ACGTN"[readC]; ed[ned].type = BT2GPU_EDIT_MM; ed[ned].pad = 0; ned++;
			score.score -= sc_mm(sc, readC, refNmask, (int)m->qu[row] - 33);
		} else score.score += sc->match_bonus;
		if(mm == -1) score.ns++;
		if(score.ns > job->nceil) goto done;
		/* res.reverse() */
		for(size_t i = 0; i < ned / 2; i++) { bt2gpu_edit t = ed[i]; ed[i] = ed[ned - 1 - i]; ed[ned - 1 - i] = t; }
		(void)escore;
		out->score = (int32_t)score.score; out->ns = score.ns; out->gaps = gaps;
		out->refoff = (int64_t)col + job->refl;            /* off + rfi_(0) + rect.refl */
		out->trim5 = (uint32_t)(job->fw ? trimBeg : trimEnd);
		out->trim3 = (uint32_t)(job->fw ? trimEnd : trimBeg);
		int refns = 0;
		for(int i = col; i <= origCol; i++) if(m->rf[i] > 15) refns++;
		out->refns = refns;
		/* AlnRes::setShape: shift edit positions by trimBeg (aligner_result.cpp:101-108) */
		if(trimBeg > 0) for(size_t i = 0; i < ned; i++) ed[i].pos -= (uint32_t)trimBeg;
		/* nextAlignment: invertEdits if !fw (aligner_sw.cpp:1137-1142, edit.cpp:50-78) */
		if(!job->fw) {
			size_t rdexrows = (size_t)nrow - trimBeg - trimEnd;
			for(size_t i = 0; i < ned / 2; i++) { bt2gpu_edit t = ed[i]; ed[i] = ed[ned - 1 - i]; ed[ned - 1 - i] = t; }
			for(size_t i = 0; i < ned; i++)
				ed[i].pos = (uint32_t)(rdexrows - ed[i].pos - (ed[i].type == BT2GPU_EDIT_READ_GAP ? 0 : 1));
		}
		out->nedits = (uint32_t)ned;
		*nedp = ned;
		ok = 1;
	}
done:
	free(stack); free(cells);
	return ok;
}

/* ---- SwAligner::align + nextAlignment loop -------------------------------- */

static int dp_one(
	const bt2gpu_scoring *sc, const uint8_t *seq_pool, const uint8_t *qual_pool,
	const uint8_t *win_pool, const bt2gpu_dp_job *job, uint32_t max_aln,
	bt2gpu_dp_result *res, bt2gpu_dp_cand *cands, size_t cand_cap, size_t *ncand_tot,
	bt2gpu_dp_aln *alns, bt2gpu_edit *edits, size_t edit_cap, size_t *nedit_tot)
{
	dpmat m; memset(&m, 0, sizeof(m));
	m.sc = sc; m.rd = seq_pool + job->rd_off; m.qu = qual_pool + job->rd_off;
	m.rf = win_pool + job->rf_off; m.nrow = (int)job->rd_len; m.ncol = (int)job->ncol;
	m.minsc = job->minsc; m.local = !sc->monotone;
	int flag = 0, sse8succ = 0, sse16succ = 0;
	int64_t best = MIN_I64;
	res->cand_off = (uint32_t)*ncand_tot; res->ncand = 0; res->naln = 0;
	res->tail_tries = 0; res->cural = 0; res->width = 0; res->aligned = 0;
	if(sc->monotone) {
		if(job->enable8 && job->minsc >= -254) {            /* aligner_sw.cpp:518 */
			m.w16 = 0; m.fl = -255;
			best = dp_fill(&m, &flag); sse8succ = (flag == 0);
		} else {
			m.w16 = 1; m.fl = -65535;
			best = dp_fill(&m, &flag); sse16succ = (flag == 0);
		}
	} else {
		flag = -2; m.fl = 0;
		if(job->enable8) { m.w16 = 0; best = dp_fill(&m, &flag); }   /* :570-586 */
		if(flag == -2) {
			dp_free(&m);
			flag = 0; m.w16 = 1;
			best = dp_fill(&m, &flag); sse16succ = (flag == 0);  /* :587-605 */
		} else sse8succ = (flag == 0);
	}
	res->best = best;
	res->width = sse8succ ? 8 : (sse16succ ? 16 : 0);
	if(best == MIN_I64 || best < job->minsc) { dp_free(&m); return 0; } /* :681 */
	long nc = dp_gather(&m, cands + *ncand_tot, cand_cap - *ncand_tot);
	if(nc < 0) { dp_free(&m); return 1; }
	res->ncand = (uint32_t)nc; *ncand_tot += (size_t)nc;
	res->aligned = nc > 0;
	if(nc == 0) { dp_free(&m); return 0; }
	const bt2gpu_dp_cand *cd = cands + res->cand_off;
	/* nextAlignment loop (aligner_sw.cpp:737-1146) */
	m.rep = (uint8_t*)calloc((size_t)m.nrow * m.ncol, 1);
	m.masks = (uint16_t*)calloc((size_t)m.nrow * m.ncol, sizeof(uint16_t));
	size_t cural = 0, ncandsz = (size_t)nc;
	size_t SQ = (size_t)m.nrow >> 4; if(SQ == 0) SQ = 1;
	size_t ndone = 0; uint32_t *done_rc = (uint32_t*)malloc(sizeof(uint32_t) * 2 * (ncandsz + 1));
	uint32_t tries = 0; int rc = 0;
	while(res->naln < max_aln && cural < ncandsz) {
		int got = 0;
		bt2gpu_dp_aln *out = alns + res->naln;
		while(cural < ncandsz) {
			if(cd[cural].score < job->minsc) { cural++; continue; }       /* :762 */
			int row = (int)cd[cural].row, col = (int)cd[cural].col;
			if(m.rep[(size_t)row * m.ncol + col]) { cural++; continue; } /* :774-789 */
			if(m.local) {                                                 /* :946-971 */
				int dom = 0;
				for(size_t i = 0; i < ndone; i++) {
					size_t rowlo = done_rc[2 * i], collo = done_rc[2 * i + 1];
					size_t rowhi = (size_t)row, colhi = (size_t)col, t;
					if(colhi < collo) { t = colhi; colhi = collo; collo = t; }
					if(rowhi < rowlo) { t = rowhi; rowhi = rowlo; rowlo = t; }
					if(colhi - collo <= SQ && rowhi - rowlo <= SQ) { dom = 1; break; }
				}
				if(dom) { cural++; continue; }
			}
			tries++;   /* one rnd.nextU32() re-seed per attempted backtrace */
			size_t ned = 0; int overflow = 0;
			out->edit_off = (uint32_t)*nedit_tot;
			int ok = dp_backtrace(&m, job, cd[cural].score, row, col,
			                      edits + *nedit_tot, edit_cap - *nedit_tot, &ned, out, &overflow);
			if(overflow) { rc = 2; goto out; }   /* 1 = candidate pool, 2 = edit pool */
			if(m.local) { done_rc[2 * ndone] = (uint32_t)row; done_rc[2 * ndone + 1] = (uint32_t)col; ndone++; }
			if(ok) { *nedit_tot += ned; got = 1; break; }
			cural++;
		}
		if(!got) break;                      /* cural_ == btncand_.size() -> false */
		cural++;                             /* :1143 */
		out->cural = (uint32_t)cural; out->tries = tries; tries = 0;
		res->naln++;
	}
out:
	res->tail_tries = tries; res->cural = (uint32_t)cural;
	free(done_rc); dp_free(&m);
	return rc;
}

/* Same contract as bt2gpu_dp_align_batch (include/bt2gpu.h), computed on the CPU. */
int bt2oracle_dp_align_batch(
	const bt2gpu_scoring *sc,
	const uint8_t *seq_pool, const uint8_t *qual_pool, size_t pool_len,
	const uint8_t *win_pool, size_t win_len,
	const bt2gpu_dp_job *jobs, size_t njobs, uint32_t max_aln,
	bt2gpu_dp_result *results, bt2gpu_dp_cand *cands, size_t cand_cap,
	bt2gpu_dp_aln *alns, bt2gpu_edit *edits, size_t edit_cap,
	size_t *ncands, size_t *nedits)
{
	(void)pool_len; (void)win_len;
	size_t nc = 0, ne = 0;
	for(size_t i = 0; i < njobs; i++) {
		results[i].aln_off = (uint32_t)(i * max_aln);
		int rc = dp_one(sc, seq_pool, qual_pool, win_pool, &jobs[i], max_aln, &results[i],
		                cands, cand_cap, &nc, alns + i * max_aln, edits, edit_cap, &ne);
		if(rc) return rc;
	}
	if(ncands) *ncands = nc;
	if(nedits) *nedits = ne;
	return 0;
}

/* DP cells of a batch: rows x cols as bowtie2-dp counts them (bt2_dp.cpp:744). */
uint64_t bt2oracle_dp_cells(const bt2gpu_dp_job *jobs, size_t njobs) {
	uint64_t c = 0;
	for(size_t i = 0; i < njobs; i++) c += (uint64_t)jobs[i].rd_len * jobs[i].ncol;
	return c;
}

/* ---- SwAligner::ungappedAlign (aligner_sw.cpp:286-494) ------------------------------------
 * Pinned against tests/golden/ungapped_{ee,loc}.txt (oracle/make_ungapped_goldens.py: dumps of
 * the unmodified reference through ref_dp_dump --ungapped). */
static int ungap_one(const bt2gpu_scoring *sc, const uint8_t *rd, const uint8_t *qu, const uint8_t *rf,
                     const bt2gpu_ungap_job *j, bt2gpu_ungap_result *r, bt2gpu_edit *ed, size_t cap, size_t *ne)
{
	const int64_t len = j->rd_len;
	int ns = 0;
	memset(r, 0, sizeof *r);
	int64_t rfi = j->off, rff = rfi + len;
	int64_t leftNs = 0, rightNs = 0;
	if(rfi < 0) { if(j->ohang) leftNs = -rfi; else return 0; }                    /* :307-314 */
	if(rff > j->reflen) { if(j->ohang) rightNs = rff - j->reflen; else return 0; } /* :315-322 */
	if(leftNs + rightNs > (int64_t)j->nceil) return 0;                             /* :323-325 */
	int64_t score = 0, rowi = 0, rowf = len - 1;
	if(sc->monotone) {                                                             /* :387-398 */
		for(int64_t i = 0; i < len; i++) {
			const int rdc = rd[i], refm = 1 << rf[i], q = qu[i] - 33;
			if(rdc > 3 || refm > 15) ns++;
			score += sc_score(sc, rdc, refm, q);
			if(score < j->minsc || ns > j->nceil) return 0;
		}
	} else {                                                                       /* :399-435 */
		int64_t scoreMax = 0, lastfloor = 0, sols = 0;
		rowi = -1;
		for(int64_t i = 0; i < len; i++) {
			const int rdc = rd[i], refm = 1 << rf[i], q = qu[i] - 33;
			if(rdc > 3 || refm > 15) ns++;
			score += sc_score(sc, rdc, refm, q);
			if(score >= j->minsc && score >= scoreMax) {
				scoreMax = score; rowf = i;
				if(rowi != lastfloor) { rowi = lastfloor; sols++; }
			}
			if(score <= 0) { score = 0; lastfloor = i + 1; }
		}
		if(ns > j->nceil || scoreMax < j->minsc || rowi < 0) return 0;   /* rowi < 0: minsc <= 0 and no cell reached it */
		if(sols > 1) { r->ret = -1; return 0; }
		score = scoreMax;
	}
	const size_t e0 = *ne;
	uint32_t n = 0; int refns = 0;
	for(int64_t i = rowi; i <= rowf; i++) {                                        /* :441-455 */
		if(rf[i] > 3 || rd[i] != rf[i]) {
			if(e0 + n >= cap) return 1;
			bt2gpu_edit *e = &ed[e0 + n++];
			e->pos = (uint32_t)i; e->chr = (uint8_t)"ACGTN"[rf[i]]; e->qchr = (uint8_t)"ACGTN"[rd[i]];
			e->type = BT2GPU_EDIT_MM; e->pad = 0;
			if(rf[i] > 3) refns++;
		}
	}
	const int64_t trimEnd = (len - 1) - rowf;
	/* setShape's trim shift (aligner_result.cpp:101-108) then, for the '-' strand, invertEdits (:489) */
	for(uint32_t k = 0; k < n; k++) ed[e0 + k].pos -= (uint32_t)rowi;
	if(!j->fw) {
		const uint32_t ext = (uint32_t)(rowf - rowi + 1);
		for(uint32_t k = 0; k < n / 2; k++) { bt2gpu_edit t = ed[e0 + k]; ed[e0 + k] = ed[e0 + n - 1 - k]; ed[e0 + n - 1 - k] = t; }
		for(uint32_t k = 0; k < n; k++) ed[e0 + k].pos = ext - ed[e0 + k].pos - 1;
	}
	*ne = e0 + n;
	r->ret = 1; r->score = (int32_t)score; r->ns = ns; r->refns = refns;
	r->refoff = j->off + rowi;
	r->trim5 = (uint32_t)(j->fw ? rowi : trimEnd); r->trim3 = (uint32_t)(j->fw ? trimEnd : rowi);
	r->nedits = n; r->edit_off = (uint32_t)e0;
	return 0;
}

int bt2oracle_ungapped_batch(const bt2gpu_scoring *sc, const uint8_t *seq_pool, const uint8_t *qual_pool,
                             const uint8_t *win_pool, const bt2gpu_ungap_job *jobs, size_t njobs,
                             bt2gpu_ungap_result *results, bt2gpu_edit *edits, size_t edit_cap, size_t *nedits)
{
	size_t ne = 0;
	for(size_t i = 0; i < njobs; i++) {
		const bt2gpu_ungap_job *j = &jobs[i];
		if(ungap_one(sc, seq_pool + j->rd_off, qual_pool + j->rd_off, win_pool + j->rf_off, j, &results[i], edits, edit_cap, &ne))
			return 1;
	}
	*nedits = ne;
	return 0;
}
