// seed_aligner_gpu.cpp — reference-side shim: SeedAligner::exactSweep / oneMmSearch / searchAllSeeds bodies that
// run the FM-index searches on the GPU through the C ABI (include/bt2gpu.h) instead of Ebwt::mapLF*/mapBiLFEx.
//
// Compiled against the reference's own headers and linked into the unmodified reference in place of the three
// member functions it replaces (aligner_seed.cpp:854-968, :973-1323, :597-718).  instantiateSeeds, the
// AlignmentCache protocol, SeedResults and every caller are the reference's.  Batch-of-one binding: each call
// stages its one read and sends that read's searches; the batched runtime is bt2gpu_hotpath_*.
// No CPU fallback: seed policies the CUDA path does not cover (-N 1/2 multi-zone seeds) throw.
#include <cstdio>
#include <vector>
#include "aligner_seed.h"
#include "gpu_shim.h"

/**
 * SeedAligner::exactSweep (aligner_seed.cpp:854-968): same contract — returns the number of BW elements of the exact
 * end-to-end hits, sets mineFw/mineRc as exactSweepStep / the final loop do, installs the exact hits in `hits`.
 */
size_t SeedAligner::exactSweep(
	const Ebwt& ebwt, const Read& read, const Scoring& sc, bool nofw, bool norc, size_t mineMax,
	size_t& mineFw, size_t& mineRc, bool repex, SeedResults& hits, SeedSearchMetrics& met)
{
	bt2shim::SweepCall sw;
	sw.fw = &ebwt; sw.sc = &sc; sw.rd = &read; sw.nofw = nofw; sw.norc = norc; sw.mineMax = (uint32_t)mineMax;
	bt2shim::sweep(sw);
	const bt2gpu_sweep_result& r = sw.res;
	bwops_ += r.bwops;
	const size_t len = read.length();
	size_t nelt = 0;
	for(int fwi = 0; fwi < 2; fwi++) {
		const bool fw = (fwi == 0);
		if(fw ? nofw : norc) continue;
		// stopped at the bound (exactSweepStep :837-841) or ran to the end of the read (:944): either way the
		// reference leaves nedit in mineFw/mineRc
		if(r.done[fwi]) { (fw ? mineFw : mineRc) = r.nedit[fwi]; continue; }
		(fw ? mineFw : mineRc) = r.nedit[fwi];
		if(r.nedit[fwi] == 0 && r.bot[fwi] > r.top[fwi]) {
			if(repex) {
				int64_t score = len * sc.match();
				if(fw) hits.addExactEeFw(r.top[fwi], r.bot[fwi], NULL, NULL, fw, score);
				else   hits.addExactEeRc(r.top[fwi], r.bot[fwi], NULL, NULL, fw, score);
			}
			nelt += (r.bot[fwi] - r.top[fwi]);
		}
	}
	return nelt;
}

/**
 * SeedAligner::oneMmSearch (aligner_seed.cpp:973-1323): end-to-end alignments with exactly one mismatch (and, with
 * repex, the exact ones), installed in `hits` in the reference's order of add1mmEe / addExactEe* calls.
 * The return value ("found something", ignored by the reference's only caller, bt2_search.cpp:3654) is true iff
 * a hit was installed.
 */
bool SeedAligner::oneMmSearch(
	const Ebwt* ebwtFw, const Ebwt* ebwtBw, const Read& read, const Scoring& sc, int64_t minsc, bool nofw, bool norc,
	bool local, bool repex, bool rep1mm, SeedResults& hits, SeedSearchMetrics& met)
{
	const size_t len = read.length();
	size_t ns = read.ns();
	if(ns > 1) return false;                 // :990-997
	else if(ns == 1 && !rep1mm) return false;
	if(ebwtBw == NULL) { fprintf(stderr, "bt2gpu shim: oneMmSearch needs the mirror index\n"); throw 1; }
	bt2shim::OneMmCall om;
	om.fw = ebwtFw; om.bw = ebwtBw; om.sc = &sc; om.rd = &read;
	om.job.minsc = (int32_t)minsc; om.job.nceil = sc.nCeil.f<int>((double)len);
	om.job.nofw = nofw ? 1 : 0; om.job.norc = norc ? 1 : 0; om.job.pad[0] = om.job.pad[1] = 0;
	om.local = local; om.repex = repex; om.rep1mm = rep1mm;
	bt2shim::oneMm(om);
	const bt2gpu_onemm_result& res = om.res;
	bwops_ += res.bwops;
	for(uint32_t k = 0; k < res.nhits; k++) {
		const bt2gpu_ee_hit& h = om.hits[res.hit_off + k];
		const bool fw = h.fw != 0;
		if(h.pos == 0xffff) {
			int64_t score = len * sc.match();
			if(fw) hits.addExactEeFw(h.top, h.bot, NULL, NULL, fw, score);
			else   hits.addExactEeRc(h.top, h.bot, NULL, NULL, fw, score);
		} else {
			Edit e((uint32_t)h.pos, (int)h.chr, (int)h.qchr, EDIT_TYPE_MM, true);   // = Edit(off5p, j, rdc, MM, false), :1226
			hits.add1mmEe(h.top, h.bot, &e, NULL, fw, h.score);
		}
	}
	return res.nhits > 0;
}

/**
 * SeedAligner::instantiateSeeds (aligner_seed.cpp:498-587): same contract -- the seed offsets, SeedResults::reset, the
 * seed sequences / qualities of both strands (instantiateSeq: the alignment cache keys on them), one entry per seed
 * that Seed::instantiate accepts (Ns), the same counts and metrics.
 * The steps / zones / constraints Seed::instantiate lays out are read by searchSeedBi only, which the GPU replaces:
 * they are computed into one scratch object per host thread instead of into the read's SeedResults (whose buffers,
 * with thousands of workers per process, are never in cache: the stores were a third of a host thread's time).  The
 * SeedResults entry keeps what searchAllSeeds below reads (offset, index, strand) and maxjump.
 */
pair<int, int> SeedAligner::instantiateSeeds(
	const EList<Seed>& seeds, size_t off, int per, const Read& read, const Scoring& pens, bool nofw, bool norc,
	AlignmentCacheIface& cache, SeedResults& sr, SeedSearchMetrics& met, pair<int, int>& instFw, pair<int, int>& instRc)
{
	offIdx2off_.clear();
	const int len = seeds[0].len;
	int nseeds = 1;
	if((int)read.length() - (int)off > len) nseeds += ((int)read.length() - (int)off - len) / per;
	for(int i = 0; i < nseeds; i++) offIdx2off_.push_back(per * i + (int)off);
	pair<int, int> ret(0, 0);
	sr.reset(read, offIdx2off_, nseeds);
	static thread_local InstantiatedSeed scratch;
	for(int fwi = 0; fwi < 2; fwi++) {
		const bool fw = (fwi == 0);
		if((fw && nofw) || (!fw && norc)) continue;
		for(int i = 0; i < nseeds; i++) {
			const int depth = i * per + (int)off;
			instantiateSeq(read, sr.seqs(fw)[i], sr.quals(fw)[i], std::min<int>(len, (int)read.length()), depth, fw);
			EList<InstantiatedSeed>& iss = sr.instantiatedSeeds(fw, i);
			for(int j = 0; j < (int)seeds.size(); j++) {
				if(seeds[j].instantiate(read, sr.seqs(fw)[i], sr.quals(fw)[i], pens, depth, i, j, fw, scratch)) {
					iss.expand();
					InstantiatedSeed& is = iss.back();
					is.seedoff = scratch.seedoff; is.seedoffidx = scratch.seedoffidx; is.fw = scratch.fw; is.maxjump = scratch.maxjump;
					ret.first++;
					if(fwi == 0) instFw.first++; else instRc.first++;
				} else met.filteredseed++;
			}
		}
	}
	return ret;
}

/**
 * SeedAligner::searchAllSeeds (aligner_seed.cpp:597-718).  Exact seeds (-N 0): every instantiated seed of the read
 * goes to the GPU in one bt2gpu_seed_exact_batch call (ftab jump + mapBiLFEx / mapLF1 steps, startSearchSeedBi :1633,
 * searchSeedBi :1854); the hits then go through the reference's own cache protocol (reportHit :1575 -> addOnTheFly,
 * beginAlign / addAllCached / finishAlign, SeedResults::add) in the reference's order: forward seeds by offset,
 * then reverse-complement seeds by offset.  The searches themselves never touch the cache, so running them all
 * before the first beginAlign leaves the cache in the same state as the reference's batches of eight.
 */
void SeedAligner::searchAllSeeds(
	const EList<Seed>& seeds, const Ebwt* ebwtFw, const Ebwt* ebwtBw, const Read& read, const Scoring& pens,
	AlignmentCacheIface& cache, SeedResults& sr, SeedSearchMetrics& met, PerReadMetrics& prm)
{
	ebwtFw_ = ebwtFw; ebwtBw_ = ebwtBw; sc_ = &pens; read_ = &read;
	bwops_ = bwedits_ = 0;
	uint64_t possearches = 0, seedsearches = 0, ooms = 0;
	if(ebwtBw == NULL) { fprintf(stderr, "bt2gpu shim: searchAllSeeds needs the mirror index\n"); throw 1; }
	// which policy (Seed::mmSeeds, aligner_seed.cpp:365-448): one exact seed, or the 2 / 3 multi-zone seeds of -N 1 / -N 2
	int mms = -1;
	if(seeds.size() == 1 && seeds[0].type == SEED_TYPE_EXACT) mms = 0;
	else if(seeds.size() == 2 && seeds[0].type == SEED_TYPE_LEFT_TO_RIGHT && seeds[1].type == SEED_TYPE_RIGHT_TO_LEFT) mms = 1;
	else if(seeds.size() == 3 && seeds[0].type == SEED_TYPE_LEFT_TO_RIGHT && seeds[1].type == SEED_TYPE_RIGHT_TO_LEFT &&
	        seeds[2].type == SEED_TYPE_INSIDE_OUT) mms = 2;
	else { fprintf(stderr, "bt2gpu shim: seed policy outside Seed::mmSeeds(0..2) is not on the CUDA path\n"); throw 1; }
	if(mms == 0) {
	bt2shim::SeedsCall sd;
	sd.fw = ebwtFw; sd.bw = ebwtBw; sd.sc = &pens; sd.rd = &read;
	for(int fwi = 0; fwi < 2; fwi++) {
		const bool fw = (fwi == 0);
		for(size_t i = 0; i < sr.numOffs(); i++) {
			EList<InstantiatedSeed>& iss = sr.instantiatedSeeds(fw, i);
			if(iss.empty()) continue;              // seed window with an N, or a cross-read cache hit
			possearches++;
			seedsearches++;
			bt2gpu_seed_job j;
			j.read = 0; j.off = (uint16_t)iss[0].seedoff; j.len = (uint8_t)sr.seqs(fw)[i].length(); j.fw = fw ? 1 : 0;
			sd.jobs.push_back(j);
		}
	}
	bt2shim::seeds(sd);
	size_t k = 0;
	for(int fwi = 0; fwi < 2; fwi++) {
		const bool fw = (fwi == 0);
		for(size_t i = 0; i < sr.numOffs(); i++) {
			if(sr.instantiatedSeeds(fw, i).empty()) continue;
			const bt2gpu_seed_hit& h = sd.out[k++];
			bwops_ += h.bwops;
			const BTDnaString& seq = sr.seqs(fw)[i];
			SeedSearchCache srcache(seq, sr.quals(fw)[i]);
			if(h.botf > h.topf) srcache.addOnTheFly(seq, h.topf, h.botf, h.topb, h.botb);   // reportHit, exact: rf = seq
			if(srcache.beginAlign(cache) == -1) { ooms++; continue; }
			if(!srcache.addAllCached()) { ooms++; continue; }
			srcache.finishAlign();
			if(srcache.qvValid()) sr.add(srcache.getQv(), cache.current(), (uint32_t)i, fw);
		}
	}
	} else {
	// -N 1 / -N 2: one job per seed offset; the device instantiates the offset's seed types as Seed::instantiate does
	// and returns their hits in the order of the reference's reportHit calls (the types of an offset in lock step,
	// mismatch branches depth first), each with the reference string it stands for
	bt2shim::SeedsMmCall sd;
	sd.fw = ebwtFw; sd.bw = ebwtBw; sd.sc = &pens; sd.rd = &read; sd.mms = mms;
	for(int fwi = 0; fwi < 2; fwi++) {
		const bool fw = (fwi == 0);
		for(size_t i = 0; i < sr.numOffs(); i++) {
			EList<InstantiatedSeed>& iss = sr.instantiatedSeeds(fw, i);
			if(iss.empty()) continue;
			possearches++;
			seedsearches += iss.size();
			if(sr.seqs(fw)[i].length() > 32) { fprintf(stderr, "bt2gpu shim: seeds longer than 32 are not cacheable by the reference either\n"); throw 1; }
			bt2gpu_seed_job j;
			j.read = 0; j.off = (uint16_t)iss[0].seedoff; j.len = (uint8_t)sr.seqs(fw)[i].length(); j.fw = fw ? 1 : 0;
			sd.jobs.push_back(j);
		}
	}
	bt2shim::seedsMm(sd);
	size_t k = 0;
	BTDnaString rf;
	for(int fwi = 0; fwi < 2; fwi++) {
		const bool fw = (fwi == 0);
		for(size_t i = 0; i < sr.numOffs(); i++) {
			if(sr.instantiatedSeeds(fw, i).empty()) continue;
			const bt2gpu_mmseed_result& r = sd.res[k++];
			bwops_ += r.bwops; bwedits_ += r.bwedits;
			const BTDnaString& seq = sr.seqs(fw)[i];
			const size_t len = seq.length();
			SeedSearchCache srcache(seq, sr.quals(fw)[i]);
			for(uint32_t h = 0; h < r.nhits; h++) {
				const bt2gpu_mmseed_hit& hit = sd.hits[r.hit_off + h];
				rf.resize(len);
				for(size_t p = 0; p < len; p++) rf.set((int)((hit.key >> (2 * (len - 1 - p))) & 3), p);
				srcache.addOnTheFly(rf, hit.topf, hit.botf, hit.topb, hit.botb);       // reportHit: rf = seed with its edits applied
			}
			if(srcache.beginAlign(cache) == -1) { ooms++; continue; }
			if(!srcache.addAllCached()) { ooms++; continue; }
			srcache.finishAlign();
			if(srcache.qvValid()) sr.add(srcache.getQv(), cache.current(), (uint32_t)i, fw);
		}
	}
	}
	prm.nSeedRanges = sr.numRanges();
	prm.nSeedElts = sr.numElts();
	prm.nSeedRangesFw = sr.numRangesFw();
	prm.nSeedRangesRc = sr.numRangesRc();
	prm.nSeedEltsFw = sr.numEltsFw();
	prm.nSeedEltsRc = sr.numEltsRc();
	prm.seedMedian = (uint64_t)(sr.medianHitsPerSeed() + 0.5);
	prm.seedMean = (uint64_t)sr.averageHitsPerSeed();
	prm.nSdFmops += bwops_;
	met.seedsearch += seedsearches;
	met.nrange += sr.numRanges();
	met.nelt += sr.numElts();
	met.possearch += possearches;
	met.ooms += ooms;
	met.bwops += bwops_;
	met.bweds += bwedits_;
}
