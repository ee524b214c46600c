// ref_seed_dump.cpp — TEST INFRASTRUCTURE (our code; links the UNMODIFIED reference objects).
//
// Drives the reference's own seed-side hot path on a FASTQ file and dumps what it computes,
// for use as golden vectors (tests/golden/fm_*.txt) and as the CPU baseline of bench.py:
//   SeedAligner::exactSweep, ::oneMmSearch, ::instantiateSeeds + ::searchAllSeeds (exact
//   seeds, -N 0), SwDriver::extend and Ebwt::getOffset.
// Built by oracle/Makefile with -fno-access-control (SwDriver::extend and
// SeedAligner::bwops_ are private in the reference).
//
// Output (one record per line):
//   R <i> <len> <seq> <qual>
//   SW <nelt> <mineFw> <mineRc> <exFw.top> <exFw.bot> <exRc.top> <exRc.bot> <bwops>
//   M1 <ret> <nmm1> <exFw.top> <exFw.bot> <exRc.top> <exRc.bot> <bwops>
//   H <fw> <top> <bot> <score> <pos> <chr> <qchr>                 (one per 1-mismatch hit)
//   SD <L> <per> <nseeds> <bwops>
//   S <fw> <offidx> <off> <nranges> {<topf> <botf> <topb> <botb>}*
//   X <fw> <off> <len> <topf> <botf> <topb> <botb> <nlex> <nrex> <fmops>
//   L <row> <off>
#include <iostream>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>
#include <thread>
#include <chrono>
#include <atomic>
#include <cmath>
#include "bt2_idx.h"
#include "aligner_seed.h"
#include "aligner_sw_driver.h"
#include "aligner_cache.h"
#include "scoring.h"
#include "simple_func.h"
#include "read.h"

using namespace std;

// globals the reference defines in bt2_search.cpp (not linked here)
int gVerbose = 0;
bool gReportOverhangs = false;
int gDefaultSeedLen = 22;

struct Fq { string name, seq, qual; };

static vector<Fq> readFastq(const string& fn, size_t limit) {
	vector<Fq> v;
	ifstream in(fn.c_str());
	string l1, l2, l3, l4;
	while(v.size() < limit && getline(in, l1) && getline(in, l2) && getline(in, l3) && getline(in, l4)) {
		Fq f; f.name = l1.substr(1); f.seq = l2; f.qual = l4;
		v.push_back(f);
	}
	return v;
}

static Ebwt* loadEbwt(const string& base, bool fw) {
	Ebwt* e = new Ebwt(base, 0, fw ? -1 : 1, fw, -1, 0, false, false, false, true, true, true, true,
	                   false, false, false, false);
	e->loadIntoMemory(0, fw ? -1 : 1, fw, true, fw, true, false);
	return e;
}

int main(int argc, char** argv) {
	string idx, reads, hitsOut;   // --hits-out FILE (bench mode): one byte per read, 1 = an exact seed of the read hit the index
	size_t limit = (size_t)-1, maxloc = 16;
	int seedLen = 22, nthreads = 1, reps = 1;
	double ivC = 1.0, ivL = 1.15;     // default end-to-end interval S,1,1.15 (--sensitive)
	int seedMms = 0;
	bool local = false, bench = false;
	for(int i = 1; i < argc; i++) {
		string a = argv[i];
		if(a == "--index") idx = argv[++i];
		else if(a == "--reads") reads = argv[++i];
		else if(a == "--limit") limit = (size_t)atol(argv[++i]);
		else if(a == "-L") seedLen = atoi(argv[++i]);
		else if(a == "--interval") { ivC = atof(argv[++i]); ivL = atof(argv[++i]); }
		else if(a == "--local") local = true;
		else if(a == "--maxloc") maxloc = (size_t)atol(argv[++i]);
		else if(a == "--bench") bench = true;
		else if(a == "--threads") nthreads = atoi(argv[++i]);
		else if(a == "--reps") reps = atoi(argv[++i]);
		else if(a == "--hits-out") hitsOut = argv[++i];
		else if(a == "--seedmms") seedMms = atoi(argv[++i]);      // -N: 1 / 2 = multi-zone seeds (Seed::oneMmSeeds / twoMmSeeds)
		else { cerr << "unknown arg " << a << endl; return 2; }
	}
	Ebwt* ebwtFw = loadEbwt(idx, true);
	Ebwt* ebwtBw = loadEbwt(idx + ".rev", false);
	SimpleFunc scoreMin, nCeil, msIval;
	if(local) scoreMin.init(SIMPLE_FUNC_LOG, 20.0, 8.0);
	else scoreMin.init(SIMPLE_FUNC_LINEAR, DEFAULT_MIN_CONST, DEFAULT_MIN_LINEAR);
	nCeil.init(SIMPLE_FUNC_LINEAR, 0.0f, std::numeric_limits<double>::max(), 0.0, 0.15);
	msIval.init(SIMPLE_FUNC_SQRT, ivC, ivL);
	Scoring sc(local ? DEFAULT_MATCH_BONUS_LOCAL : DEFAULT_MATCH_BONUS, DEFAULT_MM_PENALTY_TYPE,
	           DEFAULT_MM_PENALTY_MAX, DEFAULT_MM_PENALTY_MIN, scoreMin, nCeil,
	           DEFAULT_N_PENALTY_TYPE, DEFAULT_N_PENALTY, DEFAULT_N_CAT_PAIR,
	           DEFAULT_READ_GAP_CONST, DEFAULT_REF_GAP_CONST, DEFAULT_READ_GAP_LINEAR,
	           DEFAULT_REF_GAP_LINEAR, 4);
	vector<Fq> fq = readFastq(reads, limit);

	if(bench) {
		// Threaded timing of the same calls (no output): one SeedAligner per thread as
		// bt2_search.cpp:3128 does.  Prints one JSON line with reads/s and LF steps/s.
		atomic<size_t> next(0);
		atomic<uint64_t> totOps(0), totSeeds(0), totLoc(0);
		const size_t total = fq.size() * (size_t)reps;
		vector<uint8_t> hitflag(fq.size(), 0);
		auto t0 = chrono::steady_clock::now();
		vector<thread> th;
		for(int t = 0; t < nthreads; t++) th.emplace_back([&]() {
			SeedAligner al; SeedResults shs; SeedSearchMetrics sdm; PerReadMetrics prm;
			AlignmentCache scCurrent(20 * 1024 * 1024, false);
			AlignmentCacheIface ca(&scCurrent, NULL, NULL);
			EList<Seed> seeds; Constraint gc = Constraint::penaltyFuncBased(scoreMin);
			Read rd; uint64_t ops = 0, nsd = 0, sink = 0, nloc = 0;
			SwDriver sd(4 * 1024 * 1024); EList<SATuple, 16> satups;
			for(;;) {
				size_t k = next.fetch_add(64);
				if(k >= total) break;
				for(size_t q = k; q < min(total, k + 64); q++) {
					const Fq& f = fq[q % fq.size()];
					rd.init(f.name.c_str(), f.seq.c_str(), f.qual.c_str());
					size_t len = rd.length();
					shs.clear(); shs.nextRead(rd); ca.nextRead();
					size_t mineFw = 0, mineRc = 0;
					al.bwops_ = 0;
					al.exactSweep(*ebwtFw, rd, sc, false, false, 2, mineFw, mineRc, true, shs, sdm);
					ops += al.bwops_;
					shs.clearExactE2eHits();
					int interval = msIval.f<int>((double)len); if(interval < 1) interval = 1;
					seeds.clear(); Seed::mmSeeds(0, seedLen, seeds, gc);
					pair<int,int> instFw, instRc;
					pair<int,int> inst = al.instantiateSeeds(seeds, 0, interval, rd, sc, false, false, ca, shs, sdm, instFw, instRc);
					if(inst.first + inst.second > 0) {
						al.searchAllSeeds(seeds, ebwtFw, ebwtBw, rd, sc, ca, shs, sdm, prm);
						ops += al.bwops_; nsd += inst.first;
						// extend every hit, then resolve the first row of the narrowest range
						size_t sl = min<size_t>((size_t)seedLen, len);
						TIndexOffU bestTop = 0, bestSz = 0;
						for(int fwi = 0; fwi < 2; fwi++) for(size_t i = 0; i < shs.numOffs(); i++) {
							const QVal& qv = shs.hitsAtOffIdx(fwi == 0, i);
							if(!qv.valid()) continue;
							satups.clear(); size_t nr = 0, ne = 0;
							ca.queryQval(qv, satups, nr, ne);
							for(size_t k = 0; k < satups.size(); k++) {
								TIndexOffU sz = (TIndexOffU)satups[k].offs.size();
								size_t nlex = 0, nrex = 0; PerReadMetrics prm2;
								sd.extend(rd, *ebwtFw, ebwtBw, satups[k].topf, satups[k].topf + sz, satups[k].topb,
								          satups[k].topb + sz, fwi == 0, (size_t)interval * i, sl, prm2, nlex, nrex);
								ops += prm2.nSdFmops;
								if(bestSz == 0 || sz < bestSz) { bestSz = sz; bestTop = satups[k].topf; }
							}
						}
						if(bestSz > 0) { sink += ebwtFw->getOffset(bestTop); nloc++; hitflag[q % fq.size()] = 1; }
					}
				}
			}
			totOps += ops; totSeeds += nsd; totLoc += nloc + (sink == 0xdeadbeefdeadull);
		});
		for(auto& t : th) t.join();
		double sec = chrono::duration<double>(chrono::steady_clock::now() - t0).count();
		if(!hitsOut.empty()) { ofstream ho(hitsOut.c_str(), ios::binary); ho.write((const char*)hitflag.data(), (streamsize)hitflag.size()); }
		printf("{\"reads\": %zu, \"seconds\": %.6f, \"reads_per_s\": %.1f, \"lf_steps\": %llu, \"lf_steps_per_s\": %.1f, \"seeds\": %llu, \"located\": %llu, \"threads\": %d}\n",
		       total, sec, total / sec, (unsigned long long)totOps.load(), totOps.load() / sec,
		       (unsigned long long)totSeeds.load(), (unsigned long long)totLoc.load(), nthreads);
		return 0;
	}

	SeedAligner al;
	SwDriver sd(20 * 1024 * 1024);
	SeedResults shs;
	SeedSearchMetrics sdm;
	PerReadMetrics prm;
	AlignmentCache scCurrent(20 * 1024 * 1024, false);
	AlignmentCacheIface ca(&scCurrent, NULL, NULL);
	EList<Seed> seeds;
	Constraint gc = Constraint::penaltyFuncBased(scoreMin);
	EList<SATuple, 16> satups;
	for(size_t r = 0; r < fq.size(); r++) {
		Read rd;
		rd.init(fq[r].name.c_str(), fq[r].seq.c_str(), fq[r].qual.c_str());
		size_t len = rd.length();
		cout << "R " << r << ' ' << len << ' ' << fq[r].seq << ' ' << fq[r].qual << '\n';
		// --- exactSweep (bt2_search.cpp:3454-3482) ---
		shs.clear(); shs.nextRead(rd); ca.nextRead();
		size_t mineFw = 0, mineRc = 0;
		al.bwops_ = 0;
		size_t nelt = al.exactSweep(*ebwtFw, rd, sc, false, false, 2, mineFw, mineRc, true, shs, sdm);
		EEHit ef = shs.exactFwEEHit(), er = shs.exactRcEEHit();
		cout << "SW " << nelt << ' ' << mineFw << ' ' << mineRc << ' ' << ef.top << ' ' << ef.bot << ' '
		     << er.top << ' ' << er.bot << ' ' << al.bwops_ << '\n';
		shs.clearExactE2eHits();
		// --- oneMmSearch (bt2_search.cpp:3649-3668) ---
		{
			TAlScore minsc = local ? (TAlScore)max<double>(scoreMin.f<double>((double)len), 0.0)
			                       : (TAlScore)min<double>(scoreMin.f<double>((double)len), 0.0);
			al.bwops_ = 0;
			bool ret = al.oneMmSearch(ebwtFw, ebwtBw, rd, sc, minsc, false, false, local, true, true, shs, sdm);
			ef = shs.exactFwEEHit(); er = shs.exactRcEEHit();
			const EList<EEHit>& mm = shs.mm1EEHits();
			cout << "M1 " << (ret ? 1 : 0) << ' ' << mm.size() << ' ' << ef.top << ' ' << ef.bot << ' '
			     << er.top << ' ' << er.bot << ' ' << al.bwops_ << ' ' << minsc << '\n';
			for(size_t i = 0; i < mm.size(); i++) {
				cout << "H " << (mm[i].fw ? 1 : 0) << ' ' << mm[i].top << ' ' << mm[i].bot << ' ' << mm[i].score
				     << ' ' << mm[i].e1.pos << ' ' << (char)mm[i].e1.chr << ' ' << (char)mm[i].e1.qchr << '\n';
			}
			shs.clearExactE2eHits(); shs.clear1mmE2eHits();
		}
		// --- instantiateSeeds + searchAllSeeds, exact seeds (bt2_search.cpp:3862-3906) ---
		int interval = msIval.f<int>((double)len);
		if(interval < 1) interval = 1;
		ca.nextRead(); shs.clearSeeds();
		seeds.clear();
		Seed::mmSeeds(seedMms, seedLen, seeds, gc);
		pair<int,int> instFw, instRc;
		pair<int,int> inst = al.instantiateSeeds(seeds, 0, interval, rd, sc, false, false, ca, shs, sdm, instFw, instRc);
		if(inst.first + inst.second == 0) { cout << "SD " << seedLen << ' ' << interval << " 0 0\n"; continue; }
		al.searchAllSeeds(seeds, ebwtFw, ebwtBw, rd, sc, ca, shs, sdm, prm);
		cout << "SD " << seedLen << ' ' << interval << ' ' << shs.numOffs() << ' ' << al.bwops_ << '\n';
		size_t sl = min<size_t>((size_t)seedLen, len);
		for(int fwi = 0; fwi < 2; fwi++) {
			bool fw = (fwi == 0);
			for(size_t i = 0; i < shs.numOffs(); i++) {
				const QVal& qv = shs.hitsAtOffIdx(fw, i);
				size_t off = (size_t)interval * i;
				cout << "S " << (fw ? 1 : 0) << ' ' << i << ' ' << off;
				satups.clear();
				size_t nrange = 0, neltq = 0;
				if(qv.valid()) ca.queryQval(qv, satups, nrange, neltq);
				cout << ' ' << satups.size();
				for(size_t k = 0; k < satups.size(); k++) {
					TIndexOffU sz = (TIndexOffU)satups[k].offs.size();
					cout << ' ' << satups[k].topf << ' ' << satups[k].topf + sz << ' ' << satups[k].topb << ' ' << satups[k].topb + sz;
				}
				cout << '\n';
				if(seedMms > 0) {
					// the reference strings the ranges stand for (SAKey: 2 bits per base, first base most significant):
					// the seed with its mismatches applied, in the cache's order = the order of the reportHit calls
					cout << "K " << (fw ? 1 : 0) << ' ' << i << ' ' << satups.size();
					for(size_t k = 0; k < satups.size(); k++) cout << ' ' << satups[k].key.seq << ' ' << satups[k].key.len;
					cout << '\n';
				}
				for(size_t k = 0; k < satups.size(); k++) {
					TIndexOffU sz = (TIndexOffU)satups[k].offs.size();
					size_t nlex = 0, nrex = 0;
					PerReadMetrics prm2;
					sd.extend(rd, *ebwtFw, ebwtBw, satups[k].topf, satups[k].topf + sz, satups[k].topb, satups[k].topb + sz,
					          fw, off, sl, prm2, nlex, nrex);
					cout << "X " << (fw ? 1 : 0) << ' ' << off << ' ' << sl << ' ' << satups[k].topf << ' ' << satups[k].topf + sz
					     << ' ' << satups[k].topb << ' ' << satups[k].topb + sz << ' ' << nlex << ' ' << nrex << ' ' << prm2.nSdFmops << '\n';
					for(TIndexOffU e = 0; e < sz && e < maxloc; e++) {
						TIndexOffU row = satups[k].topf + e;
						cout << "L " << row << ' ' << ebwtFw->getOffset(row) << '\n';
					}
				}
			}
		}
	}
	return 0;
}
