/* oracle/frame_oracle.c — TEST INFRASTRUCTURE, not product.
 *
 * CPU restatement of the window framing and paired-end geometry that sit between the two
 * kernels of the hot path (SURVEY.md §8a rows K2-2, DRV-3, IDX-6, K2-1 gap budgets):
 *   bt2o_frame_seed_ext   DynProgFramer::frameSeedExtensionRect      dp_framer.cpp:81-129
 *   bt2o_frame_mate       DynProgFramer::frameFindMateRect           dp_framer.h:155-197,
 *                         ...AnchorLeftRect / AnchorRightRect        dp_framer.cpp:177-357
 *   bt2o_pe_other_mate    PairedEndPolicy::otherMate                 pe.cpp:161-353 (+ pePolicyMateDir pe.h:130-164)
 *   bt2o_pe_classify      PairedEndPolicy::peClassifyPair            pe.cpp:37-136
 *   bt2o_max_read_gaps / bt2o_max_ref_gaps   Scoring::maxReadGaps / maxRefGaps   scoring.cpp:42-105
 *   bt2o_joined_to_text_off   Ebwt::joinedToTextOff                  bt2_idx.cpp:54-124
 * Parity: pinned against the unmodified reference (oracle/ref_frame_dump.cpp links the
 * reference's pe.o / dp_framer.o / scoring.o; vectors committed as tests/golden/frame_vectors.txt,
 * checked by tests/test_oracle_frame.py).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg may call this.
 *
 * The reference mixes int64_t and size_t in these expressions; all of it is two's-complement
 * wrap-around arithmetic, restated here on uint64_t and cast back. */
#include <stdint.h>
#include <stddef.h>

typedef struct {
	int64_t  refl, refr, refl_pretrim, refr_pretrim;
	uint64_t triml, trimr, corel, corer, maxgap;
} bt2o_rect;

typedef struct {
	int32_t  pol;                /* PE_POLICY_FF=1, RR=2, FR=3, RF=4 (pe.h:38-54) */
	uint64_t maxfrag, minfrag;
	uint8_t  local, flip_ok, dovetail_ok, contain_ok, olap_ok, expand_to_fit;
} bt2o_pe_policy;

enum { PE_FF = 1, PE_RR, PE_FR, PE_RF };
enum { ALS_NORMAL = 1, ALS_OVERLAP, ALS_CONTAIN, ALS_DOVETAIL, ALS_DISCORD };

static uint64_t umax(uint64_t a, uint64_t b) { return a > b ? a : b; }
static uint64_t umin(uint64_t a, uint64_t b) { return a < b ? a : b; }
static int64_t  smax(int64_t a, int64_t b) { return a > b ? a : b; }
static int64_t  smin(int64_t a, int64_t b) { return a < b ? a : b; }

/* the common tail of the three framers: trimming against the reference ends (dp_framer.cpp:100-119) */
static int finish_rect(int64_t refl, int64_t refr, uint64_t rdlen, int64_t reflen, int64_t maxns,
                       int trim_to_ref, uint64_t maxgap, uint64_t corel, uint64_t corer, bt2o_rect *r)
{
	uint64_t triml = 0, trimr = 0;
	if(trim_to_ref) maxns = 0;
	else if(maxns == (int64_t)rdlen) maxns--;
	if(refr >= reflen + maxns) trimr = (uint64_t)(refr - (reflen + maxns - 1));
	if(refl < -maxns) triml = (uint64_t)(-refl) - (uint64_t)maxns;
	r->refl_pretrim = refl; r->refr_pretrim = refr;
	r->refl = (int64_t)((uint64_t)refl + triml);
	r->refr = (int64_t)((uint64_t)refr - trimr);
	r->triml = triml; r->trimr = trimr; r->maxgap = maxgap; r->corel = corel; r->corer = corer;
	return !(r->refr < r->refl);                 /* !rect.entirelyTrimmed() (dp_framer.h:100) */
}

int bt2o_frame_seed_ext(int64_t off, uint64_t rdlen, int64_t reflen, uint64_t maxrdgap, uint64_t maxrfgap,
                        int64_t maxns, uint64_t maxhalf, int trim_to_ref, bt2o_rect *r)
{
	const uint64_t maxgap = umin(umax(maxrdgap, maxrfgap), maxhalf);
	const int64_t refl = (int64_t)((uint64_t)off - 2 * maxgap);
	const int64_t refr = (int64_t)((uint64_t)off + (rdlen - 1) + 2 * maxgap);
	return finish_rect(refl, refr, rdlen, reflen, maxns, trim_to_ref, maxgap, maxgap, maxgap + 2 * maxgap, r);
}

int bt2o_frame_mate(int anchor_left, int64_t ll, int64_t lr, int64_t rl, int64_t rr, uint64_t rdlen, int64_t reflen,
                    uint64_t maxrdgap, uint64_t maxrfgap, int64_t maxns, uint64_t maxhalf, int trim_to_ref, bt2o_rect *r)
{
	const uint64_t maxgap = umax(umax(maxrdgap, maxrfgap), maxhalf);
	int64_t refl, refr;
	if(anchor_left) {        /* opposite mate to the right: built on where its RHS may fall (dp_framer.cpp:177-249) */
		const int64_t st_left = (int64_t)((uint64_t)rl - (rdlen - 1));
		refl = (int64_t)((uint64_t)st_left - maxgap);
		refr = (int64_t)((uint64_t)rr + maxgap);
	} else {                 /* opposite mate to the left: built on where its LHS may fall (dp_framer.cpp:291-357) */
		const int64_t en_right = (int64_t)((uint64_t)lr + (rdlen - 1));
		refl = (int64_t)((uint64_t)ll - maxgap);
		refr = (int64_t)((uint64_t)en_right + maxgap);
	}
	const uint64_t width = (uint64_t)(refr - refl + 1);
	return finish_rect(refl, refr, rdlen, reflen, maxns, trim_to_ref, maxgap, maxgap, width - maxgap - 1, r);
}

int bt2o_pe_other_mate(const bt2o_pe_policy *p, int is1, int fw, int64_t off, int64_t maxalcols, uint64_t reflen,
                       uint64_t len1, uint64_t len2, int *oleft, int64_t *oll_, int64_t *olr_, int64_t *orl_, int64_t *orr_,
                       int *ofw)
{
	(void)reflen;
	int left, mfw;
	switch(p->pol) {          /* pePolicyMateDir (pe.h:130-164) */
		case PE_FF: left = (is1 != fw); mfw = fw; break;
		case PE_RR: left = (is1 == fw); mfw = fw; break;
		case PE_FR: left = !fw; mfw = !fw; break;
		case PE_RF: left = fw; mfw = !fw; break;
		default: return -1;
	}
	*oleft = left; *ofw = mfw;
	const uint64_t alen = is1 ? len1 : len2;
	uint64_t maxfrag = p->maxfrag, minfrag = p->minfrag;
	if(minfrag < 1) minfrag = 1;
	if(len1 > maxfrag && p->expand_to_fit) maxfrag = len1;
	if(len2 > maxfrag && p->expand_to_fit) maxfrag = len2;
	if(!p->expand_to_fit && (len1 > maxfrag || len2 > maxfrag)) return 0;
	const uint64_t o = (uint64_t)off;
	int64_t oll, olr, orl, orr;
	if(left) {
		oll = (int64_t)(o + alen - maxfrag);
		olr = (int64_t)(o + alen - minfrag);
		orl = oll;
		orr = (int64_t)(o + maxfrag - 1);
		if(!p->olap_ok) {
			orr = smin(orr, off - 1);
			if(orr < olr) olr = orr;
		} else if(!p->dovetail_ok) {
			orr = smin(orr, (int64_t)(o + alen - 1));
		} else if(!p->flip_ok && maxalcols != -1) {
			orr = smin(orr, (int64_t)(o + alen - 1 + (uint64_t)(maxalcols - 1)));
		}
	} else {
		orr = (int64_t)(o + (maxfrag - 1));
		orl = (int64_t)(o + (minfrag - 1));
		oll = (int64_t)(o + alen - maxfrag);
		olr = orr;
		if(!p->olap_ok) {
			oll = smax(oll, (int64_t)(o + alen));
			if(oll > orl) orl = oll;
		} else if(!p->dovetail_ok) {
			oll = smax(oll, off);
		} else if(!p->flip_ok && maxalcols != -1) {
			oll = smax(oll, off - maxalcols + 1);
		}
	}
	*oll_ = oll; *olr_ = olr; *orl_ = orl; *orr_ = orr;
	return 1;
}

int bt2o_pe_classify(const bt2o_pe_policy *p, int64_t off1, uint64_t len1, int fw1, int64_t off2, uint64_t len2, int fw2)
{
	uint64_t maxfrag = p->maxfrag, minfrag = p->minfrag;
	if(len1 > maxfrag && p->expand_to_fit) maxfrag = len1;
	if(len2 > maxfrag && p->expand_to_fit) maxfrag = len2;
	if(minfrag < 1) minfrag = 1;
	int one_left = 0;
	if(p->pol == PE_FF)      { if(fw1 != fw2) return ALS_DISCORD; one_left = fw1; }
	else if(p->pol == PE_RR) { if(fw1 != fw2) return ALS_DISCORD; one_left = !fw1; }
	else if(p->pol == PE_FR) { if(fw1 == fw2) return ALS_DISCORD; one_left = fw1; }
	else if(p->pol == PE_RF) { if(fw1 == fw2) return ALS_DISCORD; one_left = !fw1; }
	const int64_t fraglo = smin(off1, off2);
	const int64_t fraghi = smax((int64_t)((uint64_t)off1 + len1), (int64_t)((uint64_t)off2 + len2));
	const uint64_t frag = (uint64_t)(fraghi - fraglo);
	if(frag > maxfrag || frag < minfrag) return ALS_DISCORD;
	const int64_t lo1 = off1, hi1 = (int64_t)((uint64_t)off1 + len1 - 1);
	const int64_t lo2 = off2, hi2 = (int64_t)((uint64_t)off2 + len2 - 1);
	int containment = 0;
	if((lo1 >= lo2 && hi1 <= hi2) || (lo2 >= lo1 && hi2 <= hi1)) containment = 1;
	int type = ALS_NORMAL, olap = 0;
	if((lo1 <= lo2 && hi1 >= lo2) || (lo1 <= hi2 && hi1 >= hi2) || containment) {
		olap = 1;
		if(!p->olap_ok) return ALS_DISCORD;
		type = ALS_OVERLAP;
	}
	if(!olap) {
		if((one_left && lo2 < lo1) || (!one_left && lo1 < lo2)) return ALS_DISCORD;
	}
	if(containment) {
		if(!p->contain_ok) return ALS_DISCORD;
		type = ALS_CONTAIN;
	}
	if((one_left && (hi1 > hi2 || lo2 < lo1)) || (!one_left && (hi2 > hi1 || lo1 < lo2))) {
		if(!p->dovetail_ok) return ALS_DISCORD;
		type = ALS_DOVETAIL;
	}
	return type;
}

/* Scoring::maxReadGaps / maxRefGaps (scoring.cpp:42-105): match = Scoring::match(30) */
int bt2o_max_read_gaps(int64_t minsc, uint64_t rdlen, int64_t match, int64_t gap_open, int64_t gap_ext)
{
	int64_t sc = (int64_t)(rdlen * (uint64_t)match);
	int first = 1, num = 0;
	while(sc >= minsc) {
		if(first) { first = 0; sc -= gap_open; } else sc -= gap_ext;
		num++;
	}
	return num - 1;
}

int bt2o_max_ref_gaps(int64_t minsc, uint64_t rdlen, int64_t match, int64_t gap_open, int64_t gap_ext)
{
	int64_t sc = (int64_t)(rdlen * (uint64_t)match);
	int first = 1, num = 0;
	while(sc >= minsc) {
		sc -= match;
		if(first) { first = 0; sc -= gap_open; } else sc -= gap_ext;
		num++;
	}
	return num - 1;
}

/* Ebwt::joinedToTextOff (bt2_idx.cpp:54-124).  rstarts = nfrag triples (joined offset, text id, offset in
 * the text); len = EbwtParams::_len; fw = the index's fw_ flag.  tidx = 0xffffffff when the hit straddles
 * a fragment boundary and reject_straddle is set. */
void bt2o_joined_to_text_off(const uint32_t *rstarts, uint32_t nfrag, const uint32_t *plen, uint32_t len, int fw,
                             uint32_t qlen, uint32_t off, int reject_straddle,
                             uint32_t *tidx, uint32_t *textoff, uint32_t *tlen, int *straddled)
{
	uint32_t top = 0, bot = nfrag, elt;
	for(;;) {
		elt = top + ((bot - top) >> 1);
		const uint32_t lower = rstarts[elt * 3];
		const uint32_t upper = (elt == nfrag - 1) ? len : rstarts[(elt + 1) * 3];
		const uint32_t fraglen = upper - lower;
		if(lower <= off) {
			if(upper > off) {
				if(off + qlen > upper) {
					*straddled = 1;
					if(reject_straddle) { *tidx = 0xffffffffu; return; }
				}
				*tidx = rstarts[elt * 3 + 1];
				uint32_t fragoff = off - rstarts[elt * 3];
				if(!fw) { fragoff = fraglen - fragoff - 1; fragoff -= (qlen - 1); }
				*textoff = fragoff + rstarts[elt * 3 + 2];
				break;
			} else top = elt;
		} else bot = elt;
	}
	*tlen = plen[*tidx];
}
