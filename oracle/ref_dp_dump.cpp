// oracle/ref_dp_dump.cpp — TEST INFRASTRUCTURE (our code, links the UNMODIFIED
// reference objects built by oracle/Makefile).
//
// Replays dynamic-programming problems through the reference's own
// SwAligner (aligner_sw.h:241-374: initRead / initRef / align / nextAlignment)
// and dumps everything the B200 path must reproduce bit-exactly:
//   * align(): aligned flag, best score, which width succeeded (u8 / i16),
//     the sorted candidate list btncand_ (aligner_sw.cpp:722-724)
//   * every nextAlignment(): score, reference offset, soft trims, ns, gaps,
//     refns, the edit list (after invertEdits for '-' strand), and a probe of
//     the RandomSource after the call (re-seed protocol aligner_sw.cpp:794,877,879,932)
//
// Inputs
//   --log FILE --index BASE   problems logged by `bowtie2-align-server-s --log-dp`
//                             (format bt2_dp.cpp:566-631); the window is re-fetched
//                             from BASE.{3,4}.bt2 exactly like the server does
//                             (initRef overload aligner_sw.cpp:155-271, incl. the
//                             extra right-hand column :174-176)
//   --problems FILE           self-contained problems (the "P" lines this tool
//                             prints); window given explicitly incl. extra column,
//                             run through the raw-buffer overload like bt2_dp.cpp:724
// Scoring options mirror bt2_dp.cpp:96-130: --local --ma --mp MX,MN --np --rdg C,L
// --rfg C,L --gbar --no-sse8 --nceil C,L  --maxaln K
//
// Output: one "P" line, one "R" line, zero or more "A" lines and one "E" line per problem.

#include <iostream>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>
#include <cstring>
#include <cstdlib>
#include <limits>
#include <thread>
#include <chrono>
#define private public
#define protected public
#include "aligner_sw.h"
#undef private
#undef protected
#include "reference.h"
#include "scoring.h"
#include "simple_func.h"
#include "aligner_seed_policy.h"
#include "mask.h"
#include "alphabet.h"
#include "tokenize.h"

using namespace std;

int gVerbose = 0;
int gQuiet = 0;

struct Prob {
	TRefId refidx; TRefOff reflen; TAlScore minsc; bool fw; DPRect rect;
	string window; // mask chars, includes the extra column (may be empty => fetch from index)
};

static void parseRect(istream& is, DPRect& r) {
	char comma;
	is >> r.refl >> comma >> r.refr >> comma >> r.refl_pretrim >> comma >> r.refr_pretrim
	   >> comma >> r.triml >> comma >> r.trimr >> comma >> r.corel >> comma >> r.corer
	   >> comma >> r.maxgap;
}

static string rectStr(const DPRect& r) {
	ostringstream os;
	os << r.refl << ',' << r.refr << ',' << r.refl_pretrim << ',' << r.refr_pretrim << ','
	   << r.triml << ',' << r.trimr << ',' << r.corel << ',' << r.corer << ',' << r.maxgap;
	return os.str();
}

static pair<int,int> parsePair(const char* s) {
	pair<int,int> p(0, 0);
	const char* c = strchr(s, ',');
	p.first = atoi(s);
	p.second = c ? atoi(c + 1) : p.first;
	return p;
}

int main(int argc, char** argv) {
	string logfn, idx, probfn, ungfn;
	bool local = false, enable8 = true;
	int ma = -1, mpmax = DEFAULT_MM_PENALTY_MAX, mpmin = DEFAULT_MM_PENALTY_MIN;
	int np = DEFAULT_N_PENALTY, gbar = 4;
	int rdgc = DEFAULT_READ_GAP_CONST, rdgl = DEFAULT_READ_GAP_LINEAR;
	int rfgc = DEFAULT_REF_GAP_CONST, rfgl = DEFAULT_REF_GAP_LINEAR;
	double nceilC = 0.0, nceilL = 0.15;
	size_t maxaln = 8, cminlen = 2000, cpow2 = 4, upto = (size_t)-1;
	bool bench = false; int nthreads = 1;
	bool printMet = false;
	for(int i = 1; i < argc; i++) {
		string a = argv[i];
		if(a == "--log") logfn = argv[++i];
		else if(a == "--index") idx = argv[++i];
		else if(a == "--problems") probfn = argv[++i];
		else if(a == "--ungapped") ungfn = argv[++i];
		else if(a == "--local") local = true;
		else if(a == "--no-sse8") enable8 = false;
		else if(a == "--ma") ma = atoi(argv[++i]);
		else if(a == "--mp") { pair<int,int> p = parsePair(argv[++i]); mpmax = p.first; mpmin = p.second; }
		else if(a == "--np") np = atoi(argv[++i]);
		else if(a == "--rdg") { pair<int,int> p = parsePair(argv[++i]); rdgc = p.first; rdgl = p.second; }
		else if(a == "--rfg") { pair<int,int> p = parsePair(argv[++i]); rfgc = p.first; rfgl = p.second; }
		else if(a == "--gbar") gbar = atoi(argv[++i]);
		else if(a == "--nceil") { const char* s = argv[++i]; nceilC = atof(s); const char* c = strchr(s, ','); nceilL = c ? atof(c+1) : 0.0; }
		else if(a == "--maxaln") maxaln = (size_t)atol(argv[++i]);
		else if(a == "--cp-min") cminlen = (size_t)atol(argv[++i]);
		else if(a == "--upto") upto = (size_t)atol(argv[++i]);
		else if(a == "--met") printMet = true;                  // SSEMetrics of the replay, one MET line per counter set, at the end
		else if(a == "--bench") bench = true;
		else if(a == "--threads") nthreads = atoi(argv[++i]);
		else { cerr << "unknown arg " << a << endl; return 2; }
	}
	if(ma < 0) ma = local ? DEFAULT_MATCH_BONUS_LOCAL : DEFAULT_MATCH_BONUS;
	SimpleFunc scoreMin, nCeil;
	scoreMin.init(SIMPLE_FUNC_LINEAR, DEFAULT_MIN_CONST, DEFAULT_MIN_LINEAR);
	nCeil.init(SIMPLE_FUNC_LINEAR, 0.0f, std::numeric_limits<double>::max(), nceilC, nceilL);
	Scoring sc(ma, DEFAULT_MM_PENALTY_TYPE, mpmax, mpmin, scoreMin, nCeil,
	           DEFAULT_N_PENALTY_TYPE, np, DEFAULT_N_CAT_PAIR,
	           rdgc, rfgc, rdgl, rfgl, gbar);
	BitPairReference* refs = NULL;
	if(!idx.empty()) {
		refs = new BitPairReference(idx, false, false, NULL, NULL, false, false, false, false, false, false);
	}

	if(!ungfn.empty()) {
		// --ungapped FILE --index BASE: SwAligner::ungappedAlign (aligner_sw.cpp:286-494) on lines
		//   U <rd as passed (oriented)> <qual> <+/-> <refidx> <off> <ohang 0/1> <minsc>
		// prints the line back with reflen and the window the reference scored against (codes as ACGTN,
		// overhang = N), then  UR ret [score ns refns refoff trim5 trim3 nedits edits]
		if(refs == NULL) { cerr << "need --index" << endl; return 2; }
		ifstream uin(ungfn.c_str());
		string ul;
		SwAligner usw(NULL);
		BTDnaString rd; BTString qu;
		while(getline(uin, ul)) {
			EList<string> b; tokenize(ul, '\t', b);
			if(b.size() < 8 || b[0] != "U") continue;
			rd.install(b[1].c_str(), true); qu = b[2];
			bool fw = (b[3] == "+");
			TRefId refidx = (TRefId)atol(b[4].c_str());
			TRefOff off = (TRefOff)atoll(b[5].c_str());
			bool ohang = atoi(b[6].c_str()) != 0;
			TAlScore minsc = (TAlScore)atoll(b[7].c_str());
			size_t reflen = refs->approxLen(refidx);
			Coord coord(refidx, off, fw);
			SwResult res;
			int ret = usw.ungappedAlign(rd, qu, coord, *refs, reflen, sc, ohang, minsc, res);
			cout << "U\t" << b[1] << '\t' << b[2] << '\t' << b[3] << '\t' << refidx << '\t' << off << '\t'
			     << (ohang ? 1 : 0) << '\t' << minsc << '\t' << reflen << '\t';
			for(size_t i = 0; i < rd.length(); i++) {
				TRefOff p = off + (TRefOff)i;
				cout << ((p < 0 || (size_t)p >= reflen) ? 'N' : "ACGTN"[refs->getBase(refidx, (size_t)p)]);
			}
			cout << '\n' << "UR\t" << ret;
			if(ret == 1) {
				const AlnRes& r = res.alres;
				cout << '\t' << r.score().score() << '\t' << r.score().ns() << '\t' << r.refNs() << '\t' << r.refoff()
				     << '\t' << r.trimmed5p(true) << '\t' << r.trimmed3p(true) << '\t' << r.ned().size() << '\t';
				for(size_t i = 0; i < r.ned().size(); i++) {
					const Edit& e = r.ned()[i];
					if(i) cout << ';';
					cout << e.pos << ':' << (char)e.chr << ':' << (char)e.qchr << ':' << (int)e.type;
				}
			}
			cout << '\n';
		}
		return 0;
	}

	if(bench) {
		// --bench: time the reference's own SwAligner (initRead/initRef/align/nextAlignment*)
		// over all problems of a --problems file with `nthreads` worker threads (one
		// SwAligner each, as bt2_search.cpp:3128-3130 does); prints one JSON line.
		struct BP { string seq, qual; Prob p; };
		vector<BP> all;
		{
			ifstream bin(probfn.c_str());
			string bl;
			while(getline(bin, bl)) {
				EList<string> b; tokenize(bl, '\t', b);
				if(b.size() < 9 || b[0] != "P") continue;
				BP x; x.seq = b[1]; x.qual = b[2];
				x.p.fw = (b[3] == "+"); x.p.refidx = (TRefId)atol(b[4].c_str());
				x.p.reflen = (TRefOff)atoll(b[5].c_str()); x.p.minsc = (TAlScore)atoll(b[6].c_str());
				istringstream is(b[7]); parseRect(is, x.p.rect); x.p.window = b[8];
				all.push_back(x);
			}
		}
		vector<uint64_t> cells(nthreads, 0), alns(nthreads, 0);
		auto t0 = chrono::steady_clock::now();
		vector<thread> th;
		for(int t = 0; t < nthreads; t++) th.emplace_back([&, t]() {
			SwAligner swl(NULL);
			BTDnaString sq, sqrc; BTString ql, qlrc;
			vector<char> rfb;
			for(size_t i = t; i < all.size(); i += nthreads) {
				BP& x = all[i];
				sq.install(x.seq.c_str(), true); ql = x.qual;
				sqrc = sq; sqrc.reverseComp(); qlrc = ql; qlrc.reverse();
				swl.initRead(sq, sqrc, ql, qlrc, 0, sq.length(), sc);
				size_t n = x.p.window.length();
				rfb.resize(n + 16);
				for(size_t k = 0; k < n; k++) { int m = asc2dnamask[(int)x.p.window[k]]; if(m == 15) m = 16; rfb[k] = (char)m; }
				swl.initRef(x.p.fw, x.p.refidx, x.p.rect, &rfb[0], 0, n - 1, x.p.reflen, sc, x.p.minsc,
				            enable8, cminlen, cpow2, false, true);
				TAlScore best = std::numeric_limits<TAlScore>::min();
				bool aligned = swl.align(best);
				cells[t] += (uint64_t)sq.length() * (n - 1);
				if(aligned) {
					RandomSource rnd((uint32_t)(i * 7919u + 17u));
					SwResult res; size_t na = 0;
					while(na < maxaln && !swl.done()) { res.reset(); if(!swl.nextAlignment(res, x.p.minsc, rnd)) break; na++; }
					alns[t] += na;
				}
			}
		});
		for(auto& x : th) x.join();
		double sec = chrono::duration<double>(chrono::steady_clock::now() - t0).count();
		uint64_t c = 0, a = 0; for(int t = 0; t < nthreads; t++) { c += cells[t]; a += alns[t]; }
		cout << "{\"problems\": " << all.size() << ", \"cells\": " << c << ", \"alignments\": " << a
		     << ", \"seconds\": " << sec << ", \"threads\": " << nthreads << "}" << endl;
		return 0;
	}
	const string& fn = logfn.empty() ? probfn : logfn;
	bool isLog = !logfn.empty();
	ifstream in(fn.c_str());
	if(!in.good()) { cerr << "cannot open " << fn << endl; return 2; }
	SwAligner sw(NULL);
	string ln;
	size_t pid = 0, nread = 0;
	BTDnaString seq, seqrc; BTString qual, qualrc;
	while(getline(in, ln)) {
		if(ln.empty()) continue;
		EList<string> buf;
		tokenize(ln, '\t', buf);
		if(buf.size() < 3) continue;
		vector<Prob> probs;
		size_t first = 2;
		string sseq, squal;
		if(isLog) {
			sseq = buf[0]; squal = buf[1];
			for(size_t i = 2; i < buf.size(); i++) {
				Prob p; istringstream is(buf[i]); char comma, tmp;
				is >> p.refidx >> comma >> p.reflen >> comma >> p.minsc >> comma >> tmp >> comma;
				p.fw = (tmp == '+');
				parseRect(is, p.rect);
				probs.push_back(p);
			}
		} else {
			// P seq qual +/- refidx reflen minsc rect window
			if(buf[0] != "P") continue;
			sseq = buf[1]; squal = buf[2];
			Prob p;
			p.fw = (buf[3] == "+");
			p.refidx = (TRefId)atol(buf[4].c_str());
			p.reflen = (TRefOff)atoll(buf[5].c_str());
			p.minsc = (TAlScore)atoll(buf[6].c_str());
			istringstream is(buf[7]);
			parseRect(is, p.rect);
			p.window = buf[8];
			probs.push_back(p);
			(void)first;
		}
		seq.install(sseq.c_str(), true);
		qual = squal;
		seqrc = seq; seqrc.reverseComp();
		qualrc = qual; qualrc.reverse();
		sw.initRead(seq, seqrc, qual, qualrc, 0, seq.length(), sc);
		nread++;
		for(size_t j = 0; j < probs.size(); j++) {
			Prob& p = probs[j];
			RandomSource rnd((uint32_t)(pid * 7919u + 17u));
			vector<char> rfbuf;
			if(p.window.empty()) {
				if(refs == NULL) { cerr << "need --index for log input" << endl; return 2; }
				size_t nsUpto = 0;
				sw.initRef(p.fw, p.refidx, p.rect, *refs, p.reflen, sc, p.minsc, enable8,
				           cminlen, cpow2, false, true, 0, nsUpto);
			} else {
				size_t n = p.window.length(); // includes the extra column
				rfbuf.resize(n + 16);
				for(size_t i = 0; i < n; i++) {
					int m = asc2dnamask[(int)p.window[i]];
					if(m == 15) m = 16;
					rfbuf[i] = (char)m;
				}
				sw.initRef(p.fw, p.refidx, p.rect, &rfbuf[0], 0, n - 1, p.reflen, sc, p.minsc,
				           enable8, cminlen, cpow2, false, true);
			}
			// P line (self-contained problem, window incl. extra column)
			cout << "P\t" << sseq << '\t' << squal << '\t' << (p.fw ? '+' : '-') << '\t'
			     << p.refidx << '\t' << p.reflen << '\t' << p.minsc << '\t' << rectStr(p.rect) << '\t';
			for(TRefOff i = sw.rfi_; i <= sw.rff_; i++) cout << mask2dna[(int)sw.rf_[i]];
			cout << '\n';
			TAlScore best = std::numeric_limits<TAlScore>::min();
			bool aligned = sw.align(best);
			cout << "R\t" << (aligned ? 1 : 0) << '\t' << best << '\t' << (sw.sse8succ_ ? 1 : 0)
			     << '\t' << (sw.sse16succ_ ? 1 : 0) << '\t' << sw.btncand_.size() << '\t';
			for(size_t i = 0; i < sw.btncand_.size(); i++) {
				if(i) cout << ';';
				cout << sw.btncand_[i].row << ',' << sw.btncand_[i].col << ',' << sw.btncand_[i].score;
			}
			cout << '\n';
			size_t naln = 0;
			if(aligned) {
				SwResult res;
				while(naln < maxaln && !sw.done()) {
					res.reset();
					bool ok = sw.nextAlignment(res, p.minsc, rnd);
					RandomSource probe = rnd;
					if(!ok) break;
					const AlnRes& r = res.alres;
					cout << "A\t" << r.score().score() << '\t' << r.refoff() << '\t'
					     << r.trimmed5p(true) << '\t' << r.trimmed3p(true) << '\t'
					     << r.score().ns() << '\t' << r.score().gaps() << '\t' << r.refNs() << '\t'
					     << r.ned().size() << '\t';
					for(size_t i = 0; i < r.ned().size(); i++) {
						const Edit& e = r.ned()[i];
						if(i) cout << ';';
						cout << e.pos << ':' << (char)e.chr << ':' << (char)e.qchr << ':' << (int)e.type;
					}
					cout << '\t' << sw.cural_ << '\t' << probe.nextU32() << '\n';
					naln++;
				}
			}
			RandomSource probe = rnd;
			cout << "E\t" << naln << '\t' << sw.cural_ << '\t' << probe.nextU32() << '\n';
			pid++;
			if(pid >= upto) break;
		}
		if(pid >= upto) break;
	}
	if(printMet) {
		const SSEMetrics *ms[4] = { &sw.sseU8ExtendMet_, &sw.sseU8MateMet_, &sw.sseI16ExtendMet_, &sw.sseI16MateMet_ };
		const char *nm[4] = { "u8ext", "u8mate", "i16ext", "i16mate" };
		for(int k = 0; k < 4; k++) {
			const SSEMetrics& m = *ms[k];
			cout << "MET\t" << nm[k] << "\tdp=" << m.dp << "\tdpsat=" << m.dpsat << "\tdpfail=" << m.dpfail << "\tdpsucc=" << m.dpsucc
			     << "\tcol=" << m.col << "\tcell=" << m.cell << "\tinner=" << m.inner << "\tfixup=" << m.fixup << "\tgathsol=" << m.gathsol
			     << "\tbt=" << m.bt << "\tbtfail=" << m.btfail << "\tbtsucc=" << m.btsucc << "\tbtcell=" << m.btcell
			     << "\tcorerej=" << m.corerej << "\tnrej=" << m.nrej << '\n';
		}
	}
	return 0;
}
