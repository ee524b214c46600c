// frame.cuh — window framing and paired-end geometry between the FM kernels and the DP kernels.
// Device-side (also host-callable) integer logic, one call per DP problem:
//   frame_seed_ext   DynProgFramer::frameSeedExtensionRect           (dp_framer.cpp:81-129)
//   frame_mate       DynProgFramer::frameFindMateRect                (dp_framer.h:155-197, dp_framer.cpp:177-357)
//   other_mate       PairedEndPolicy::otherMate + pePolicyMateDir    (pe.cpp:161-353, pe.h:130-164)
//   pe_classify      PairedEndPolicy::peClassifyPair                 (pe.cpp:37-136)
//   max_gaps         Scoring::maxReadGaps / maxRefGaps               (scoring.cpp:42-105)
//   joined_to_text   Ebwt::joinedToTextOff                           (bt2_idx.cpp:54-124)
// The reference evaluates these with int64_t / size_t mixed; everything below is done on 64-bit
// two's-complement values so that even degenerate inputs wrap the same way.
#pragma once
#include <stdint.h>
#include "../../include/bt2gpu.h"

namespace bt2 {

#define BT2_HD __host__ __device__ __forceinline__

typedef long long i64;
typedef unsigned long long u64_t;

// A rectangle [refl, refr] (inclusive, before trimming) cut back to what the reference sequence
// plus `maxns` overhanging positions allows.  Returns false if nothing is left.
BT2_HD bool rect_trim(i64 refl, i64 refr, u64_t rdlen, i64 reflen, i64 maxns, bool trim_to_ref,
                      u64_t maxgap, u64_t corer, bt2gpu_rect &r)
{
	if(trim_to_ref) maxns = 0;
	else if(maxns == (i64)rdlen) maxns--;
	const i64 lim_r = reflen + maxns;                  // first column not allowed on the right
	const u64_t trimr = (refr >= lim_r) ? (u64_t)(refr - (lim_r - 1)) : 0;
	const u64_t triml = (refl < -maxns) ? (u64_t)(-refl) - (u64_t)maxns : 0;
	r.refl_pretrim = refl; r.refr_pretrim = refr;
	r.refl = (i64)((u64_t)refl + triml);
	r.refr = (i64)((u64_t)refr - trimr);
	r.triml = triml; r.trimr = trimr;
	r.maxgap = maxgap; r.corel = maxgap; r.corer = corer;
	return r.refr >= r.refl;
}

// Seed extension: the read laid gap-free at `off`; 2*maxgap columns either side; the middle
// 2*maxgap+1 diagonals are the ones this window covers exhaustively.
BT2_HD bool frame_seed_ext(i64 off, u64_t rdlen, i64 reflen, u64_t maxrdgap, u64_t maxrfgap, i64 maxns, u64_t maxhalf,
                           bool trim_to_ref, bt2gpu_rect &r)
{
	u64_t g = maxrdgap > maxrfgap ? maxrdgap : maxrfgap;
	if(g > maxhalf) g = maxhalf;
	const u64_t o = (u64_t)off;
	return rect_trim((i64)(o - 2 * g), (i64)(o + (rdlen - 1) + 2 * g), rdlen, reflen, maxns, trim_to_ref, g, 3 * g, r);
}

// Mate search: the window spans every placement whose right end (anchor on the left) or left end
// (anchor on the right) falls in the range otherMate allows, padded by maxgap both sides.
BT2_HD bool frame_mate(bool anchor_left, i64 ll, i64 lr, i64 rl, i64 rr, u64_t rdlen, i64 reflen, u64_t maxrdgap,
                       u64_t maxrfgap, i64 maxns, u64_t maxhalf, bool trim_to_ref, bt2gpu_rect &r)
{
	u64_t g = maxrdgap > maxrfgap ? maxrdgap : maxrfgap;
	if(g < maxhalf) g = maxhalf;
	const u64_t lo = anchor_left ? (u64_t)rl - (rdlen - 1) : (u64_t)ll;
	const u64_t hi = anchor_left ? (u64_t)rr : (u64_t)lr + (rdlen - 1);
	const i64 refl = (i64)(lo - g), refr = (i64)(hi + g);
	const u64_t width = (u64_t)(refr - refl + 1);
	return rect_trim(refl, refr, rdlen, reflen, maxns, trim_to_ref, g, width - g - 1, r);
}

struct MateWin { i64 oll, olr, orl, orr; bool oleft, ofw; };

BT2_HD bool other_mate(const bt2gpu_pe_policy &p, bool is1, bool fw, i64 off, i64 maxalcols, u64_t len1, u64_t len2, MateWin &w)
{
	switch(p.pol) {
		case BT2GPU_PE_FF: w.oleft = (is1 != fw); w.ofw = fw; break;
		case BT2GPU_PE_RR: w.oleft = (is1 == fw); w.ofw = fw; break;
		case BT2GPU_PE_FR: w.oleft = !fw; w.ofw = !fw; break;
		default:           w.oleft = fw;  w.ofw = !fw; break;      // BT2GPU_PE_RF (policy validated by the caller)
	}
	const u64_t alen = is1 ? len1 : len2;
	u64_t maxfrag = p.maxfrag, minfrag = p.minfrag ? p.minfrag : 1;
	const u64_t longer = len1 > len2 ? len1 : len2;
	if(longer > maxfrag) {
		if(!p.expand_to_fit) return false;
		maxfrag = longer;
	}
	const u64_t o = (u64_t)off;
	const bool cap_olap = !p.olap_ok, cap_dove = !cap_olap && !p.dovetail_ok,
	           cap_flip = !cap_olap && !cap_dove && !p.flip_ok && maxalcols != -1;
	if(w.oleft) {
		w.oll = (i64)(o + alen - maxfrag);
		w.olr = (i64)(o + alen - minfrag);
		w.orl = w.oll;
		i64 orr = (i64)(o + maxfrag - 1);
		i64 cap = orr;
		if(cap_olap) cap = off - 1;
		else if(cap_dove) cap = (i64)(o + alen - 1);
		else if(cap_flip) cap = (i64)(o + alen - 1 + (u64_t)(maxalcols - 1));
		if(cap < orr) orr = cap;
		if(cap_olap && orr < w.olr) w.olr = orr;
		w.orr = orr;
	} else {
		w.orr = (i64)(o + (maxfrag - 1));
		w.orl = (i64)(o + (minfrag - 1));
		w.olr = w.orr;
		i64 oll = (i64)(o + alen - maxfrag);
		i64 flo = oll;
		if(cap_olap) flo = (i64)(o + alen);
		else if(cap_dove) flo = off;
		else if(cap_flip) flo = off - maxalcols + 1;
		if(flo > oll) oll = flo;
		if(cap_olap && oll > w.orl) w.orl = oll;
		w.oll = oll;
	}
	return true;
}

BT2_HD int pe_classify(const bt2gpu_pe_policy &p, i64 off1, u64_t len1, bool fw1, i64 off2, u64_t len2, bool fw2)
{
	u64_t maxfrag = p.maxfrag;
	const u64_t minfrag = p.minfrag ? p.minfrag : 1;
	if(p.expand_to_fit) { if(len1 > maxfrag) maxfrag = len1; if(len2 > maxfrag) maxfrag = len2; }
	const bool same = (fw1 == fw2);
	const bool want_same = (p.pol == BT2GPU_PE_FF || p.pol == BT2GPU_PE_RR);
	if(same != want_same) return BT2GPU_PE_ALS_DISCORD;
	const bool one_left = (p.pol == BT2GPU_PE_FF || p.pol == BT2GPU_PE_FR) ? fw1 : !fw1;
	const i64 end1 = (i64)((u64_t)off1 + len1), end2 = (i64)((u64_t)off2 + len2);
	const u64_t frag = (u64_t)((end1 > end2 ? end1 : end2) - (off1 < off2 ? off1 : off2));
	if(frag > maxfrag || frag < minfrag) return BT2GPU_PE_ALS_DISCORD;
	const i64 lo1 = off1, hi1 = end1 - 1, lo2 = off2, hi2 = end2 - 1;
	const bool contain = (lo1 >= lo2 && hi1 <= hi2) || (lo2 >= lo1 && hi2 <= hi1);
	const bool olap = contain || (lo1 <= lo2 && hi1 >= lo2) || (lo1 <= hi2 && hi1 >= hi2);
	int type = BT2GPU_PE_ALS_NORMAL;
	if(olap) {
		if(!p.olap_ok) return BT2GPU_PE_ALS_DISCORD;
		type = BT2GPU_PE_ALS_OVERLAP;
	} else if(one_left ? (lo2 < lo1) : (lo1 < lo2)) return BT2GPU_PE_ALS_DISCORD;
	if(contain) {
		if(!p.contain_ok) return BT2GPU_PE_ALS_DISCORD;
		type = BT2GPU_PE_ALS_CONTAIN;
	}
	const bool dovetail = one_left ? (hi1 > hi2 || lo2 < lo1) : (hi2 > hi1 || lo1 < lo2);
	if(dovetail) {
		if(!p.dovetail_ok) return BT2GPU_PE_ALS_DISCORD;
		type = BT2GPU_PE_ALS_DOVETAIL;
	}
	return type;
}

// Largest number of read gaps (ref_gaps = false) or reference gaps (true) a valid alignment can
// hold: gaps are charged against the all-match score until it drops below minsc.  Closed form of
// the reference's subtract-in-a-loop.
BT2_HD int max_gaps(i64 minsc, u64_t rdlen, i64 match, i64 open, i64 ext, bool ref_gaps)
{
	const i64 perfect = (i64)(rdlen * (u64_t)match);
	if(perfect < minsc) return -1;
	const i64 first = open + (ref_gaps ? match : 0), next = ext + (ref_gaps ? match : 0);
	const i64 room = perfect - minsc;           // >= 0
	if(room < first) return 0;
	if(next <= 0) return 0x7fffffff;
	return (int)(1 + (room - first) / next);
}

// Joined-text offset -> (text id, offset inside it, its length).  rstarts: nfrag triples (joined
// offset, text id, offset of the fragment in its text), ascending.  Returns false if the hit runs
// over the end of its fragment and such hits are to be rejected.
BT2_HD bool joined_to_text(const uint32_t *rstarts, uint32_t nfrag, const uint32_t *plen, uint32_t len, bool fw,
                           uint32_t qlen, uint32_t off, bool reject_straddle,
                           uint32_t &tidx, uint32_t &textoff, uint32_t &tlen, bool &straddled)
{
	uint32_t lo = 0, hi = nfrag;                // last fragment starting at or before off
	while(hi - lo > 1) {
		const uint32_t mid = lo + ((hi - lo) >> 1);
		if(rstarts[mid * 3] <= off) lo = mid; else hi = mid;
	}
	const uint32_t start = rstarts[lo * 3];
	const uint32_t end = (lo == nfrag - 1) ? len : rstarts[(lo + 1) * 3];
	straddled = false;
	if(off + qlen > end) {
		straddled = true;
		if(reject_straddle) { tidx = 0xffffffffu; textoff = 0; tlen = 0; return false; }
	}
	tidx = rstarts[lo * 3 + 1];
	uint32_t fragoff = off - start;
	if(!fw) fragoff = (end - start) - fragoff - 1 - (qlen - 1);
	textoff = fragoff + rstarts[lo * 3 + 2];
	tlen = plen[tidx];
	return true;
}

} // namespace bt2
