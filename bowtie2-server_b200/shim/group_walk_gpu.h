// group_walk_gpu.h — reference-side shim for the one hot-path step that has no link-level seam: resolving a BW row
// to its text offset.  The reference does it in GroupWalk2S::advanceElement (group_walk.h:1105-1216), a header
// template inlined into SwDriver; per element it equals Ebwt::getOffset(topf + elt) (bt2_idx.cpp:145-171): LF-walk
// to the next sampled row, then offs[] (the group walk only shares steps between the rows of a range and resolves
// some of them earlier; SwDriver reads the offsets through advanceElement alone).
//
// This header is FORCE-INCLUDED (-include) when the unmodified aligner_sw_driver.cpp and bt2_search.cpp are
// compiled for the GPU-bound binary (oracle/Makefile, target gpu): it pulls in the reference's group_walk.h, then
// defines a class with GroupWalk2S's interface over bt2gpu_locate_batch and renames every later mention of
// GroupWalk2S to it -- what a maintainer would write as `#ifdef BT2_GPU` inside group_walk.h.
#ifndef BT2GPU_GROUP_WALK_GPU_H_
#define BT2GPU_GROUP_WALK_GPU_H_

#include "group_walk.h"
#include "gpu_shim.h"

template<typename T, int S>
class GpuGroupWalk2S {
public:
	GpuGroupWalk2S() { reset(); }
	void reset() { elt_ = rep_ = 0; inited_ = false; }

	void init(const Ebwt& ebwtFw, const BitPairReference& ref, SARangeWithOffs<T>& sa, RandomSource& rnd, WalkMetrics& met) {
		reset();
		inited_ = true;
		elt_ += sa.size();
	}

	bool advanceElement(TIndexOffU elt, const Ebwt& ebwtFw, const BitPairReference& ref, SARangeWithOffs<T>& sa,
	                    GroupWalkState& gws, WalkResult& res, WalkMetrics& met, PerReadMetrics& prm)
	{
		if(sa.offs[elt] == OFF_MASK) {
			bt2shim::LocateCall lc;
			lc.fw = &ebwtFw; lc.row = (uint32_t)(sa.topf + elt); lc.off = 0; lc.steps = 0;
			bt2shim::locate(lc);
			sa.offs[elt] = (TIndexOffU)lc.off;
			met.bwops += lc.steps; prm.nExFmops += lc.steps;
			met.resolves++;
		}
		met.reports++;
		res.init(0, false, 0, elt, sa.topf + elt, (TIndexOffU)sa.len, sa.offs[elt]);
		rep_++;
		return true;
	}

	bool done() const { return rep_ == elt_; }
	bool initialized() const { return inited_; }
	bool repOk(const SARangeWithOffs<T>& sa) const { return true; }

protected:
	bool   inited_;
	size_t elt_;    // # elements in the range
	size_t rep_;    // # elements reported
};

#define GroupWalk2S GpuGroupWalk2S

#endif
