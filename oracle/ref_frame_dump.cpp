// ref_frame_dump.cpp — TEST INFRASTRUCTURE (our code; links the UNMODIFIED reference objects).
//
// Feeds seeded pseudo-random inputs to the reference's own window framing and paired-end
// geometry and prints input -> output, one record per line, as golden vectors
// (tests/golden/frame_vectors.txt) for oracle/frame_oracle.c and the device framer:
//   S  DynProgFramer::frameSeedExtensionRect   (dp_framer.cpp:81-129)
//   M  PairedEndPolicy::otherMate (pe.cpp:161-353) followed, when a window exists, by
//      DynProgFramer::frameFindMateRect (dp_framer.h:155-197)  -- the call pair of
//      SwDriver::extendSeedsPaired (aligner_sw_driver.cpp:1994-2023)
//   C  PairedEndPolicy::peClassifyPair         (pe.cpp:37-136)
//   G  Scoring::maxReadGaps / maxRefGaps       (scoring.cpp:42-105)
//   J  Ebwt::joinedToTextOff                   (bt2_idx.cpp:54-124), when --index is given
// Usage: ref_frame_dump [--n N] [--seed S] [--index BASE]
#include <iostream>
#include <string>
#include <cstdint>
#include <cstdlib>
#include "bt2_idx.h"
#include "dp_framer.h"
#include "pe.h"
#include "scoring.h"
#include "simple_func.h"
#include "aligner_seed_policy.h"

using namespace std;

int gVerbose = 0;
bool gReportOverhangs = false;
int gDefaultSeedLen = 22;

static uint64_t s_state;
static uint64_t rnd() {              // splitmix64
	uint64_t z = (s_state += 0x9e3779b97f4a7c15ull);
	z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
	z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
	return z ^ (z >> 31);
}
static int64_t rng(int64_t lo, int64_t hi) { return lo + (int64_t)(rnd() % (uint64_t)(hi - lo + 1)); }

static void printRect(bool ok, const DPRect& r) {
	cout << " -> " << (ok ? 1 : 0) << ' ' << r.refl << ' ' << r.refr << ' ' << r.refl_pretrim << ' ' << r.refr_pretrim
	     << ' ' << r.triml << ' ' << r.trimr << ' ' << r.corel << ' ' << r.corer << ' ' << r.maxgap << '\n';
}

int main(int argc, char** argv) {
	size_t n = 2000; uint64_t seed = 20261104; string idx;
	for(int i = 1; i < argc; i++) {
		string a = argv[i];
		if(a == "--n") n = (size_t)atol(argv[++i]);
		else if(a == "--seed") seed = (uint64_t)atoll(argv[++i]);
		else if(a == "--index") idx = argv[++i];
		else { cerr << "unknown arg " << a << endl; return 2; }
	}
	s_state = seed;
	// ---- S: seed-extension rectangles, hits near both reference ends included ----
	for(size_t i = 0; i < n; i++) {
		const int64_t reflen = (i % 3 == 0) ? rng(30, 400) : rng(1000, 5000000);
		const size_t rdlen = (size_t)rng(20, 300);
		const int64_t off = (i % 2) ? rng(-(int64_t)rdlen - 40, reflen + 40) : rng(0, reflen);
		const size_t rdgap = (size_t)rng(0, 40), rfgap = (size_t)rng(0, 40), maxhalf = (size_t)rng(1, 30);
		int64_t maxns = rng(0, 45); if(i % 17 == 0) maxns = (int64_t)rdlen;
		const bool trim = (i % 5) != 0;
		DynProgFramer fr(trim); DPRect r;
		bool ok = fr.frameSeedExtensionRect(off, rdlen, reflen, rdgap, rfgap, maxns, maxhalf, r);
		cout << "S " << off << ' ' << rdlen << ' ' << reflen << ' ' << rdgap << ' ' << rfgap << ' ' << maxns << ' ' << maxhalf
		     << ' ' << (trim ? 1 : 0);
		printRect(ok, r);
	}
	// ---- M: otherMate + frameFindMateRect ----
	for(size_t i = 0; i < n; i++) {
		const int pol = (i % 4 == 0) ? (int)rng(1, 4) : PE_POLICY_FR;
		const size_t minfrag = (i % 3 == 0) ? (size_t)rng(0, 300) : 0;
		const size_t maxfrag = minfrag + (size_t)rng(0, 800) + ((i % 3) ? 200 : 0);
		const bool local = rnd() & 1, flip = (i % 7 == 0), dove = (i % 5 == 0), cont = (i % 11) != 0, olap = (i % 6) != 0;
		const bool expand = (i % 9) != 0;
		PairedEndPolicy pe(pol, maxfrag, minfrag, local, flip, dove, cont, olap, expand);
		const bool is1 = rnd() & 1, fw = rnd() & 1;
		const size_t len1 = (size_t)rng(20, 300), len2 = (size_t)rng(20, 300);
		const int64_t reflen = (i % 3 == 0) ? rng(300, 3000) : rng(10000, 5000000);
		const int64_t off = rng(-50, reflen + 50);
		const size_t orows = is1 ? len2 : len1;
		const size_t rdgap = (size_t)rng(0, 40), rfgap = (size_t)rng(0, 40), maxhalf = (size_t)rng(1, 30);
		const int64_t maxalcols = (i % 13 == 0) ? -1 : (int64_t)(orows + rdgap);
		int64_t maxns = rng(0, 45); if(i % 17 == 0) maxns = (int64_t)orows;
		const bool trim = (i % 5) != 0;
		bool oleft = false, ofw = false; int64_t oll = 0, olr = 0, orl = 0, orr = 0;
		bool found = pe.otherMate(is1, fw, off, maxalcols, (size_t)reflen, len1, len2, oleft, oll, olr, orl, orr, ofw);
		cout << "M " << pol << ' ' << maxfrag << ' ' << minfrag << ' ' << local << flip << dove << cont << olap << expand
		     << ' ' << is1 << ' ' << fw << ' ' << off << ' ' << maxalcols << ' ' << reflen << ' ' << len1 << ' ' << len2
		     << " | " << (found ? 1 : 0) << ' ' << oleft << ' ' << oll << ' ' << olr << ' ' << orl << ' ' << orr << ' ' << ofw
		     << " | " << orows << ' ' << rdgap << ' ' << rfgap << ' ' << maxns << ' ' << maxhalf << ' ' << (trim ? 1 : 0);
		DPRect r;
		bool ok = false;
		if(found) {
			DynProgFramer fr(trim);
			ok = fr.frameFindMateRect(!oleft, oll, olr, orl, orr, orows, reflen, rdgap, rfgap, maxns, maxhalf, r);
		} else {
			r.refl_pretrim = r.refr_pretrim = 0; r.maxgap = 0;
		}
		printRect(ok, r);
	}
	// ---- C: peClassifyPair ----
	for(size_t i = 0; i < n; i++) {
		const int pol = (i % 4 == 0) ? (int)rng(1, 4) : PE_POLICY_FR;
		const size_t minfrag = (i % 3 == 0) ? (size_t)rng(0, 300) : 0;
		const size_t maxfrag = minfrag + (size_t)rng(0, 800);
		const bool flip = (i % 7 == 0), dove = (i % 5 == 0), cont = (i % 11) != 0, olap = (i % 6) != 0, expand = (i % 9) != 0;
		PairedEndPolicy pe(pol, maxfrag, minfrag, false, flip, dove, cont, olap, expand);
		const size_t len1 = (size_t)rng(20, 300), len2 = (size_t)rng(20, 300);
		const int64_t off1 = rng(-20, 2000);
		const int64_t off2 = off1 + rng(-700, 700);
		const bool fw1 = rnd() & 1, fw2 = (i % 3 == 0) ? (bool)(rnd() & 1) : !fw1;
		int cl = pe.peClassifyPair(off1, len1, fw1, off2, len2, fw2);
		cout << "C " << pol << ' ' << maxfrag << ' ' << minfrag << ' ' << 0 << flip << dove << cont << olap << expand
		     << ' ' << off1 << ' ' << len1 << ' ' << fw1 << ' ' << off2 << ' ' << len2 << ' ' << fw2 << " -> " << cl << '\n';
	}
	// ---- G: gap budgets, end-to-end and local defaults ----
	for(int local = 0; local < 2; local++) {
		SimpleFunc scoreMin, nCeil;
		if(local) scoreMin.init(SIMPLE_FUNC_LOG, 20.0, 8.0);
		else scoreMin.init(SIMPLE_FUNC_LINEAR, DEFAULT_MIN_CONST, DEFAULT_MIN_LINEAR);
		nCeil.init(SIMPLE_FUNC_LINEAR, 0.0f, std::numeric_limits<double>::max(), 0.0, 0.15);
		Scoring sc(local ? DEFAULT_MATCH_BONUS_LOCAL : DEFAULT_MATCH_BONUS, DEFAULT_MM_PENALTY_TYPE,
		           DEFAULT_MM_PENALTY_MAX, DEFAULT_MM_PENALTY_MIN, scoreMin, nCeil,
		           DEFAULT_N_PENALTY_TYPE, DEFAULT_N_PENALTY, DEFAULT_N_CAT_PAIR,
		           DEFAULT_READ_GAP_CONST, DEFAULT_REF_GAP_CONST, DEFAULT_READ_GAP_LINEAR,
		           DEFAULT_REF_GAP_LINEAR, 4);
		for(size_t i = 0; i < n / 4; i++) {
			const size_t rdlen = (size_t)rng(20, 400);
			const int64_t perfect = (int64_t)rdlen * sc.match(30);
			const int64_t minsc = perfect - rng(0, 250);
			cout << "G " << local << ' ' << minsc << ' ' << rdlen << ' ' << sc.match(30) << ' ' << sc.readGapOpen() << ' '
			     << sc.readGapExtend() << ' ' << sc.refGapOpen() << ' ' << sc.refGapExtend()
			     << " -> " << sc.maxReadGaps(minsc, rdlen) << ' ' << sc.maxRefGaps(minsc, rdlen) << '\n';
		}
	}
	// ---- J: joined offset -> (text id, offset, length) ----
	if(!idx.empty()) {
		for(int fwi = 1; fwi >= 1; fwi--) {   // the reference only ever asks the forward index (aligner_sw_driver.cpp:940,1639)
			const bool fw = fwi != 0;
			Ebwt* e = new Ebwt(fw ? idx : idx + ".rev", 0, fw ? -1 : 1, fw, -1, 0, false, false, false, true, true, true, true,
			                   false, false, false, false);
			e->loadIntoMemory(0, fw ? -1 : 1, fw, true, fw, true, false);
			const TIndexOffU len = e->_eh._len;
			for(size_t i = 0; i < n; i++) {
				const TIndexOffU qlen = (TIndexOffU)rng(1, 40);
				TIndexOffU off;
				if(i % 3 == 0 && e->_nFrag > 1) {       // close to a fragment boundary
					const TIndexOffU f = (TIndexOffU)rng(1, (int64_t)e->_nFrag - 1);
					const int64_t o = (int64_t)e->rstarts()[f * 3] + rng(-45, 5);
					off = (TIndexOffU)(o < 0 ? 0 : o);
				} else off = (TIndexOffU)rng(0, (int64_t)len - 1);
				const bool reject = (i % 4) != 0;
				TIndexOffU tidx = 0, textoff = 0, tlen = 0; bool straddled = false;
				e->joinedToTextOff(qlen, off, tidx, textoff, tlen, reject, straddled);
				cout << "J " << (fw ? 1 : 0) << ' ' << qlen << ' ' << off << ' ' << (reject ? 1 : 0) << " -> ";
				if(tidx == (TIndexOffU)OFF_MASK) cout << "4294967295 0 0 " << straddled << '\n';
				else cout << tidx << ' ' << textoff << ' ' << tlen << ' ' << straddled << '\n';
			}
			delete e;
		}
	}
	return 0;
}
