// frame_gpu.cpp — reference-side shim: the three geometry steps between the FM kernels and the DP that have a
// one-to-one entry point in the C ABI, run on the GPU (batch of one) inside the unmodified reference:
//   DynProgFramer::frameSeedExtensionRect (dp_framer.cpp:81-129)  -> bt2gpu_frame_seed_ext_batch
//   PairedEndPolicy::peClassifyPair       (pe.cpp:37-136)         -> bt2gpu_pe_classify_batch
//   Ebwt::joinedToTextOff                 (bt2_idx.cpp:54-124)    -> bt2gpu_joined_to_text_off_batch
// otherMate + frameFindMateRect stay the reference's here: the ABI fuses them into one call
// (bt2gpu_frame_mate_batch, as SwDriver::extendSeedsPaired uses them, aligner_sw_driver.cpp:1994-2023) and
// frameFindMateRect is an inline of dp_framer.h, so there is no link-level seam for it.
#include <cstdio>
#include "dp_framer.h"
#include "pe.h"
#include "bt2_idx.h"
#include "gpu_shim.h"

bool DynProgFramer::frameSeedExtensionRect(
	int64_t off, size_t rdlen, int64_t reflen, size_t maxrdgap, size_t maxrfgap, int64_t maxns, size_t maxhalf,
	DPRect& rect)
{
	bt2gpu_frame_job j;
	j.off = off; j.reflen = reflen; j.maxns = maxns;
	j.rdlen = (uint32_t)rdlen; j.maxrdgap = (uint32_t)maxrdgap; j.maxrfgap = (uint32_t)maxrfgap; j.maxhalf = (uint32_t)maxhalf;
	bt2gpu_rect r; uint8_t ok = 0;
	if(bt2gpu_frame_seed_ext_batch(bt2shim::ctx(), trimToRef_ ? 1 : 0, &j, 1, &r, &ok))
		bt2shim::fail("bt2gpu_frame_seed_ext_batch");
	bt2shim::note(bt2shim::C_FRAME, 1);
	rect.refl = r.refl; rect.refr = r.refr;
	rect.refl_pretrim = r.refl_pretrim; rect.refr_pretrim = r.refr_pretrim;
	rect.triml = (size_t)r.triml; rect.trimr = (size_t)r.trimr;
	rect.corel = (size_t)r.corel; rect.corer = (size_t)r.corer;
	rect.maxgap = (size_t)r.maxgap;
	return ok != 0;
}

int PairedEndPolicy::peClassifyPair(int64_t off1, size_t len1, bool fw1, int64_t off2, size_t len2, bool fw2) const {
	bt2gpu_pe_policy p;
	p.pol = pol_; p.trim_to_ref = 0; p.maxfrag = maxfrag_; p.minfrag = minfrag_;
	p.local = local_; p.flip_ok = flippingOk_; p.dovetail_ok = dovetailOk_; p.contain_ok = containOk_;
	p.olap_ok = olapOk_; p.expand_to_fit = expandToFit_; p.pad[0] = p.pad[1] = 0;
	bt2gpu_pair_job j;
	j.off1 = off1; j.off2 = off2; j.len1 = (uint32_t)len1; j.len2 = (uint32_t)len2;
	j.fw1 = fw1 ? 1 : 0; j.fw2 = fw2 ? 1 : 0;
	for(int i = 0; i < 6; i++) j.pad[i] = 0;
	int32_t cls = 0;
	if(bt2gpu_pe_classify_batch(bt2shim::ctx(), &p, &j, 1, &cls)) bt2shim::fail("bt2gpu_pe_classify_batch");
	bt2shim::note(bt2shim::C_PECLASS, 1);
	return (int)cls;             // BT2GPU_PE_ALS_* == PE_ALS_* (pe.h:62-95)
}

void Ebwt::joinedToTextOff(
	TIndexOffU qlen, TIndexOffU off, TIndexOffU& tidx, TIndexOffU& textoff, TIndexOffU& tlen, bool rejectStraddle,
	bool& straddled) const
{
	uint32_t ti = 0, to = 0, tl = 0; uint8_t st = 0;
	if(bt2gpu_joined_to_text_off_batch(bt2shim::ctx(), rstarts(), _nFrag, plen(), _nPat, _eh._len, fw_ ? 1 : 0,
	                                   &qlen, &off, 1, rejectStraddle ? 1 : 0, &ti, &to, &tl, &st))
		bt2shim::fail("bt2gpu_joined_to_text_off_batch");
	bt2shim::note(bt2shim::C_JOINED, 1);
	if(st) straddled = true;     // the reference only ever raises it (:84)
	tidx = ti;
	if(ti != OFF_MASK) { textoff = to; tlen = tl; }
}

// --- mate window: PairedEndPolicy::otherMate (pe.cpp:161-353) and the two non-inline halves of
// DynProgFramer::frameFindMateRect (dp_framer.cpp:177-357; the inline dispatcher dp_framer.h:155-197 stays).
static void fillPolicy(bt2gpu_pe_policy& p, int pol, bool local, bool flip, bool dove, bool cont, bool olap, bool expand,
                       size_t maxfrag, size_t minfrag)
{
	p.pol = pol; p.trim_to_ref = 0; p.maxfrag = maxfrag; p.minfrag = minfrag;
	p.local = local; p.flip_ok = flip; p.dovetail_ok = dove; p.contain_ok = cont; p.olap_ok = olap; p.expand_to_fit = expand;
	p.pad[0] = p.pad[1] = 0;
}

bool PairedEndPolicy::otherMate(
	bool is1, bool fw, int64_t off, int64_t maxalcols, size_t reflen, size_t len1, size_t len2,
	bool& oleft, int64_t& oll, int64_t& olr, int64_t& orl, int64_t& orr, bool& ofw) const
{
	bt2gpu_pe_policy p;
	fillPolicy(p, pol_, local_, flippingOk_, dovetailOk_, containOk_, olapOk_, expandToFit_, maxfrag_, minfrag_);
	bt2gpu_mate_job j;
	j.off = off; j.maxalcols = maxalcols; j.reflen = (int64_t)reflen; j.maxns = 0;
	j.len1 = (uint32_t)len1; j.len2 = (uint32_t)len2; j.maxrdgap = j.maxrfgap = 0; j.maxhalf = 0;
	j.is1 = is1 ? 1 : 0; j.fw = fw ? 1 : 0; j.pad[0] = j.pad[1] = 0;
	bt2gpu_mate_frame f;
	if(bt2gpu_frame_mate_batch(bt2shim::ctx(), &p, &j, 1, &f)) bt2shim::fail("bt2gpu_frame_mate_batch");
	bt2shim::note(bt2shim::C_OTHERMATE, 1);
	oleft = f.oleft != 0; ofw = f.ofw != 0;
	if(!f.found) return false;
	oll = f.oll; olr = f.olr; orl = f.orl; orr = f.orr;
	return true;
}

static bool mateRect(bool anchorLeft, bool trimToRef, int64_t ll, int64_t lr, int64_t rl, int64_t rr, size_t rdlen,
                     int64_t reflen, size_t maxrdgap, size_t maxrfgap, int64_t maxns, size_t maxhalf, DPRect& rect)
{
	bt2gpu_mate_rect_job j;
	j.ll = ll; j.lr = lr; j.rl = rl; j.rr = rr; j.reflen = reflen; j.maxns = maxns;
	j.rdlen = (uint32_t)rdlen; j.maxrdgap = (uint32_t)maxrdgap; j.maxrfgap = (uint32_t)maxrfgap; j.maxhalf = (uint32_t)maxhalf;
	j.anchor_left = anchorLeft ? 1 : 0;
	for(int i = 0; i < 7; i++) j.pad[i] = 0;
	bt2gpu_rect r; uint8_t ok = 0;
	if(bt2gpu_frame_mate_rect_batch(bt2shim::ctx(), trimToRef ? 1 : 0, &j, 1, &r, &ok)) bt2shim::fail("bt2gpu_frame_mate_rect_batch");
	bt2shim::note(bt2shim::C_MATERECT, 1);
	rect.refl = r.refl; rect.refr = r.refr; rect.refl_pretrim = r.refl_pretrim; rect.refr_pretrim = r.refr_pretrim;
	rect.triml = (size_t)r.triml; rect.trimr = (size_t)r.trimr; rect.corel = (size_t)r.corel; rect.corer = (size_t)r.corer;
	rect.maxgap = (size_t)r.maxgap;
	return ok != 0;
}

bool DynProgFramer::frameFindMateAnchorLeftRect(
	int64_t ll, int64_t lr, int64_t rl, int64_t rr, size_t rdlen, int64_t reflen, size_t maxrdgap, size_t maxrfgap,
	int64_t maxns, size_t maxhalf, DPRect& rect) const
{
	return mateRect(true, trimToRef_, ll, lr, rl, rr, rdlen, reflen, maxrdgap, maxrfgap, maxns, maxhalf, rect);
}

bool DynProgFramer::frameFindMateAnchorRightRect(
	int64_t ll, int64_t lr, int64_t rl, int64_t rr, size_t rdlen, int64_t reflen, size_t maxrdgap, size_t maxrfgap,
	int64_t maxns, size_t maxhalf, DPRect& rect) const
{
	return mateRect(false, trimToRef_, ll, lr, rl, rr, rdlen, reflen, maxrdgap, maxrfgap, maxns, maxhalf, rect);
}
