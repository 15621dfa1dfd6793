// TEST INFRASTRUCTURE ONLY -- not part of the product, never linked or imported by cnf2freq_b200/.
//
// CPU restatement (plain C++, scalar, single-threaded unless OpenMP is asked for) of the hot path of
// cnettel/cnF2freq that the CUDA engine replaces: emission (trackpossible), forward/backward sweeps
// (realanalyze/adjustprobs/initfwbw), posterior enumeration (the marker x state x shift x flag2 loop of
// doit with ignoreflag2) and the transfer into the EM sufficient statistics (moveinfprobs/movehaplos).
//
// Parity: PINNED.  The reference ships no tests or golden vectors, so this file is
// pinned against the reference's own code compiled here (oracle/_ref/cnf2ref, recipe oracle/build_ref.sh):
// tests/test_oracle_vs_ref.py compares alpha-/alpha+/beta, log-factors and the statistics with dumps the
// reference binary wrote (committed under tests/golden/).
//
// Settings restated: settings.h at HEAD -- NUMGEN 3 (64 states, 128 orientations, 8 shift modes),
// SELFING false, RELSKEWSTATES false, HAPLOTYPING true, DOINFPROBS true, DOIMPOSSIBLE false,
// DOREMAPDISTANCES false; CORRECTIONINFERENCE is false during doit (cnF2freq.cpp:8083-8085).
// Compiled with -DCNF2_ORACLE_SELFING=1 (a second library, tests/oracle_lib.py) it restates the SELFING = true
// instantiation instead (row g1: 256 states = 64 x 4 values of the two selfing bits, 192 of them valid, settings.h:7-46),
// pinned on tests/golden/selfing_*.npz, which oracle/_ref/cnf2ref_selfing wrote.
//
// Every function names the reference lines it follows (paths relative to /root/reference).

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace
{
#ifndef CNF2_ORACLE_SELFING
#define CNF2_ORACLE_SELFING 0
#endif
constexpr bool SELFING = CNF2_ORACLE_SELFING;  // settings.h:14
constexpr int TYPEBITS = 6, SELFBITS = 2, SELFMASK = 3;  // settings.h:7, 20, 38
constexpr int NONSELFNUMTYPES = 1 << TYPEBITS;           // settings.h:39
constexpr int NUMTYPES = SELFING ? 256 : 64;             // settings.h:27
constexpr int VALIDSELFNUMTYPES = NUMTYPES - SELFING * (NUMTYPES >> 2);  // settings.h:46
constexpr int NUMPATHS = 128;  // settings.h:32
constexpr int NUMSHIFTS = 8;   // settings.h:35
constexpr int TURNBITS = 7;    // settings.h:40
// settings.h:29 -- a float constant, used as a double
const double MINFACTOR = (double)(float)-1e15;
constexpr int U_HAPLOS = 1, U_GENOS = 2, U_HOMOZYGOUS = 4, U_GENOSPROBE = 8;  // cnF2freq.cpp:792-795
constexpr int UNKNOWN = 0;     // cnF2freq.cpp:225
constexpr int SEXMARKER = 9;   // cnF2freq.cpp:226

struct Problem
{
	int K = 0;  // individual ids 0..K-1, id 0 = "no individual"
	int M = 0;  // markers (all chromosomes)
	int C = 0;  // chromosomes
	const int32_t* par = nullptr;          // [K][2], 0 = none            (individ::pars :868)
	const uint8_t* empty = nullptr;        // [K]                         (:864)
	std::vector<uint8_t> founder;          // [K] sticky, see fixtrees    (:866, :3176)
	const int32_t* gen = nullptr;          // [K]                         (:859)
	const int32_t* descendants = nullptr;  // [K]                         (:862)
	const double* markerpos = nullptr;     // [M]                         (:233)
	const uint32_t* chromstarts = nullptr; // [C+1]                       (:239)
	double genrec[3] = { 0, 0, 0 };        //                             (:238)
	const uint8_t* alleles = nullptr;      // [K][M][2]                   (markerdata :876)
	const double* sure = nullptr;          // [K][M][2]                   (markersure :878)
	const double* haploweight = nullptr;   // [K][M]                      (:887)
};

// Thread-local accumulators of the reference (haplos/infprobs, cnF2freq.cpp:379-380), kept per individual id.
struct Scratch
{
	std::map<int, std::array<double, 2>> haplos;
	std::map<int, std::array<std::map<int, double>, 2>> infprobs;
};

struct Env
{
	const Problem* p;
	Scratch* sc;
	int shift;  // shiftflagmode (cnF2freq.cpp:336)
};

// cnF2freq.cpp:303-316 with zeropropagate == 0 (the only mode the E-step's kept results depend on)
inline bool markermiss(int& a, int b)
{
	if (a == UNKNOWN)
	{
		a = b;
		return false;
	}
	if (b == UNKNOWN && a != SEXMARKER) return false;
	return a != b;
}

// cnF2freq.cpp:321-329
inline int upflagit(int flag, int parnum, int genwidth)
{
	if (flag < 0) return flag;
	flag >>= parnum * (genwidth - 1);
	flag &= (1 << (genwidth - 1)) - 1;
	return flag;
}

double track(const Env& e, int update, int ind, int inval, double second, int m, unsigned flag, int flag99,
             int localshift, double updateval, int genwidth);

// recursetrackpossible, cnF2freq.cpp:955-1058 (DOIMPOSSIBLE false)
double up(const Env& e, int update, int ind, int par, int val, double second, int m, int upflag, int upflag2,
          int upshift, double updateval, int genwidth)
{
	int p = e.p->par[ind * 2 + par];
	if (!p) return 1 + second;  // :1043-1046
	int f = upflagit(upflag, par, genwidth);
	int f2 = upflagit(upflag2, par, genwidth);       // NUMGEN - NUMFLAG2GEN = 0      :985
	int sh = upflagit(upshift, par, genwidth >> 1);  // NUMGEN - NUMSHIFTGEN = 1      :986
	return track(e, update, p, val, second, m, f, f2, sh, updateval, genwidth >> 1);
}

// trackpossible<update, 0>, cnF2freq.cpp:1075-1359
double track(const Env& e, int update, int ind, int inval, double second, int m, unsigned flag, int flag99,
             int localshift, double updateval, int genwidth)
{
	const Problem& P = *e.p;
	const bool attopnow = !(update & U_HOMOZYGOUS) && (genwidth == 1 || P.founder[ind]);  // :1120
	const uint8_t* const A0 = P.alleles + ((size_t)ind * P.M + m) * 2;
	const double* const S0 = P.sure + ((size_t)ind * P.M + m) * 2;
	const double hw = P.haploweight[(size_t)ind * P.M + m];
	// :1122-1127: bits above TYPEBITS + 1 of the flag are the selfing bits; they only act in the root generation
	const int selfval = (flag >> (TYPEBITS + 1)) & SELFMASK;
	const bool selfingNOW = SELFING && genwidth == 4 && selfval;
	uint8_t selfmarker[2];
	double selfsure[2];
	const uint8_t* const A = selfingNOW ? selfmarker : A0;  // :1138-1139
	const double* const S = selfingNOW ? selfsure : S0;

	int upflag2 = -1;
	const int upflag = flag >> 1;
	const int upshift = localshift >> 1;
	int f2s = 0, f2end = 2;
	if (flag99 != -1 && genwidth > 0)  // :1141
	{
		upflag2 = flag99 >> 1;
		f2s = flag99 & 1;
		f2end = (flag99 & 1) + 1;
	}
	const int firstpar = flag & 1;  // :1156
	double ok = 0;

	for (int f2 = f2s; f2 < f2end; f2++)
	{
		int f2n = f2 & 1;
		if (selfingNOW)  // :1169-1190
		{
			const int selfindex = (selfval >> 1) ^ f2n;
			selfmarker[0] = selfmarker[1] = UNKNOWN;
			selfsure[0] = selfsure[1] = 0;
			int first = A0[0];
			if (!markermiss(first, A0[1]))
			{
				selfmarker[selfindex] = (uint8_t)first;
				selfsure[selfindex] = 1 - (1 - S0[0]) * (1 - S0[1]);
			}
			else
			{
				selfmarker[selfindex] = A0[1];
				if (S0[0] == 0) return 0;
				selfsure[selfindex] = 1 - (S0[0]) * (1 - S0[1]);
			}
		}
		const bool allthesame = A[0] == A[1];
		int val = inval;
		double base;
		const int realf2n = f2n;
		double main2 = 0;
		if (markermiss(val, A[f2n]))  // :1198
		{
			base = S[f2n];
			if (S[f2n] && second) main2 = (1.0 - S[f2n]) * second;
		}
		else
		{
			double eff2 = (inval == UNKNOWN && val != UNKNOWN) ? 1 : second;  // :1205
			base = 1.0 - S[f2n];
			double effsure = (A[f2n] == UNKNOWN ? 1 : S[f2n]);
			main2 = effsure * eff2;
		}
		// `update & (GENOS || GENOSPROBE)` in the reference is `update & 1`           :1213
		if (attopnow || (update & 1))
		{
			base += main2;
			main2 = 0;
		}
		else
		{
			if (main2) main2 /= base;
		}
		bool doupd = true;
		f2n ^= ((firstpar ^ localshift) & 1);  // :1227
		if (!genwidth)                        // :1229 (zeropropagate == 0)
		{
			base *= 0.5;
			doupd = false;
		}
		else if ((allthesame && S[0] == S[1]) || selfingNOW)  // :1235 with CORRECTIONINFERENCE false
		{
			base *= (f2n ? 1.0 : 0.0);
			doupd = false;
		}
		else
		{
			base *= fabs((f2n ? 1.0 : 0.0) - hw);  // :1245
		}

		if (!base || attopnow)
		{
		}
		else
		{
			// first branch is evaluated lazily in the reference (:1279-1289, :1336-1340)
			if (!(update & U_GENOS))  // :1291 (zeropropagate == 0)
			{
				double sec2 = 0;
				int secmark = A[!realf2n];
				if (!(update & U_HOMOZYGOUS))
				{
					if (S[!realf2n])
					{
						base *= (1 - S[!realf2n]);
						sec2 = S[!realf2n] / (1 - S[!realf2n]);
					}
				}
				else
				{
					if (val != secmark)
					{
						if (secmark != UNKNOWN) base *= S[!realf2n];
						secmark = val;
						sec2 = 0;
					}
					else
					{
						base *= (1 - S[!realf2n]);
						sec2 = 0;
					}
				}
				base *= up(e, update & ~U_HOMOZYGOUS, ind, !firstpar, secmark, sec2, m, upflag, upflag2, upshift,
				           updateval, genwidth);
			}
			if (base) base *= up(e, update & ~U_HOMOZYGOUS, ind, firstpar, val, main2, m, upflag, upflag2, upshift,
			                     updateval, genwidth);
		}

		if (base)
		{
			ok += base;
			if ((update & U_HAPLOS) && doupd) e.sc->haplos[ind][f2n] += updateval;   // :1347-1350
			if (update & U_GENOS) e.sc->infprobs[ind][realf2n][val] += updateval;    // :1351-1354
		}
	}
	return ok;
}

// ---------------------------------------------------------------------------------------------------
// Focus-tree tables, fixtrees (cnF2freq.cpp:3099-3187)
struct Tree
{
	int root = 0;
	int flag2ignore = 0, shiftignore = 0;
	int ordered[TURNBITS] = { 0, 0, 0, 0, 0, 0, 0 };  // reltreeordered: non-empty occupant of each slot, 0 = none
	std::map<int, int> relmap, relmapshift;           // distinct non-empty relative -> slot mask / shift mask
	std::vector<int> reltree;                         // all distinct tree members incl. empty ones
};

Tree fixtrees(Problem& P, int ind)
{
	Tree t;
	t.root = ind;
	t.reltree.push_back(ind);
	t.ordered[0] = ind;
	t.relmap[ind] = 1;
	t.relmapshift[ind] = 1;
	int flag2ignore = 1, shiftignore = 0;
	bool anylev1 = false;
	for (int lev1 = 0; lev1 < 2; lev1++)
	{
		int l1 = P.par[ind * 2 + lev1];
		if (!l1) continue;
		int flag2index = 1 + lev1 * 3;  // :3124
		int shiftval = 2 << lev1;       // :3125
		if (!P.empty[l1])
		{
			flag2ignore |= 1 << flag2index;
			t.relmap[l1] |= 1 << flag2index;
			t.relmapshift[l1] |= shiftval;
			t.ordered[flag2index] = l1;
		}
		bool anypars = false;
		t.reltree.push_back(l1);
		for (int lev2 = 0; lev2 < 2; lev2++)
		{
			int l2 = P.par[l1 * 2 + lev2];
			if (!l2) continue;
			if (!P.empty[l2])
			{
				flag2ignore |= 1 << (flag2index + lev2 + 1);
				t.relmap[l2] |= 1 << (flag2index + lev2 + 1);
				t.relmapshift[l2] |= 0;
				t.ordered[flag2index + lev2 + 1] = l2;
				anypars = true;
			}
			t.reltree.push_back(l2);
		}
		if (anypars) shiftignore |= shiftval;
		if (anypars || !P.empty[l1]) anylev1 = true;
	}
	if (anylev1) shiftignore |= 1;
	else P.founder[ind] = 1;  // sticky side effect, :3176
	t.flag2ignore = flag2ignore ^ (NUMPATHS - 1);
	t.shiftignore = shiftignore ^ (NUMSHIFTS - 1);
	std::sort(t.reltree.begin(), t.reltree.end());
	t.reltree.resize(std::unique(t.reltree.begin(), t.reltree.end()) - t.reltree.begin());
	return t;
}

// ignoreflag2, cnF2freq.cpp:3462-3496
bool ignoreflag2(const Problem& P, const Tree& t, int flag2, int g, int shift, int marker)
{
	if (flag2 & t.flag2ignore) return true;
	auto j = t.relmapshift.begin();
	for (auto i = t.relmap.begin(); i != t.relmap.end(); ++i, ++j)
	{
		int currfilter = i->second;
		int filtered = (flag2 ^ (g * 2)) & currfilter;
		if (filtered && filtered != currfilter) return true;  // :3486
		const uint8_t* A = P.alleles + ((size_t)i->first * P.M + marker) * 2;
		const double* S = P.sure + ((size_t)i->first * P.M + marker) * 2;
		if (!SELFING && A[0] == A[1] && S[0] == S[1] && !(((bool)filtered) ^ ((bool)(shift & j->second)))) return true;  // :3488-3489
	}
	return false;
}

// ---------------------------------------------------------------------------------------------------
// Sweeps.  Storage mirrors fwbw[shift][marker][pad][state] / fwbwfactors (cnF2freq.cpp:392-393):
// pad 0 = alpha before the emission at the marker, pad 2 = after emission+normalisation, pad 1 = beta.
struct Sweeps
{
	int c0 = 0, Mc = 0;
	std::vector<double> v;    // [NUMSHIFTS][Mc][3][NUMTYPES]
	std::vector<double> fac;  // [NUMSHIFTS][Mc][3]
	bool done[NUMSHIFTS] = { false };
	double* at(int s, int m, int pad) { return &v[(((size_t)s * Mc + (m - c0)) * 3 + pad) * NUMTYPES]; }
	double& f(int s, int m, int pad) { return fac[((size_t)s * Mc + (m - c0)) * 3 + pad]; }
};

// adjustprobs, cnF2freq.cpp:1579-1670
void adjustprobs(const Env& e, int root, double* probs, int marker, double& factor, int flag99)
{
	double sum = 0;
	for (int i = 0; i < NUMTYPES; i++)
	{
		double val = probs[i];
		if (val < 1e-300)
		{
			probs[i] = 0;
			continue;
		}
		// calltrackpossible<false,false>, :1380-1385
		double realok = track(e, 0, root, UNKNOWN, 0, marker, i * 2, flag99, e.shift, 0.0, 4);
		val *= realok;
		sum += val;
		probs[i] = val;
	}
	if (sum <= 0) factor = MINFACTOR;
	else
	{
		for (int i = 0; i < NUMTYPES; i++) probs[i] /= sum;
		factor += log(sum);
	}
}

// transition part of realanalyze, cnF2freq.cpp:2270-2367 (dense XOR-convolution, as the reference does it)
void transition(const Problem& P, double* probs, double dist, int selfgen)
{
	if (!(dist > 0)) return;
	double recprob[2 + SELFING][2], other[2][2][2];
	for (int gen = 0; gen < 2 + SELFING; gen++)
		for (int k = 0; k < 2; k++)
			recprob[gen][k] = 0.5 * (1.0 - exp((gen == 2 ? selfgen : 1) * P.genrec[gen] * dist));  // :2286, getactrec :777-780
	for (int gen = 0; gen < 2; gen++)
		for (int mm = 0; mm < 2; mm++)
			for (int k = 0; k < 2; k++) other[gen][mm][k] = mm ? 1.0 - recprob[gen][k] : recprob[gen][k];
	static const int TYPESEXES[6] = { 0, 0, 1, 1, 0, 1 };  // settings.h:21
	static const int TYPEGENS[6] = { 1, 0, 0, 1, 0, 0 };   // settings.h:23
	double recombprec[NONSELFNUMTYPES];
	for (int i = 0; i < NONSELFNUMTYPES; i++) recombprec[i] = 1;
	double selfprec[5][5] = { { 0 } };  // :2316-2326 (only filled under SELFING)
	if (SELFING)
	{
		const double rs = recprob[2 * SELFING][0];  // the selfing generations' rate, recprob[2][0]
		selfprec[0][1] = selfprec[0][2] = rs;
		selfprec[0][0] = 1 - 2 * rs;
		selfprec[1][0] = selfgen ? selfprec[0][1] * 2.0 / ((1 << selfgen) - 1) : 1;
		selfprec[1][2] = selfprec[1][0] * selfprec[0][1];
		selfprec[2][0] = selfprec[1][0];
		selfprec[2][1] = selfprec[1][2];
		selfprec[2][2] = selfprec[1][1] = 1 - selfprec[1][0] - selfprec[1][2];
	}
	for (int t = 0; t < 6; t++)
		for (int i = 0; i < NONSELFNUMTYPES; i++) recombprec[i] *= other[TYPEGENS[t]][!((i >> t) & 1)][TYPESEXES[t]];
	double probs2[NUMTYPES] = { 0 };
	for (int from = 0; from < VALIDSELFNUMTYPES; from++)  // :2352-2364 (the double-bit selfing value is never entered or left)
	{
		double fromval = probs[from];
		if (fromval <= 0) continue;
		for (int to = 0; to < VALIDSELFNUMTYPES; to++)
		{
			const int xored = from ^ to;
			probs2[to] += fromval * recombprec[xored & (NONSELFNUMTYPES - 1)] *
			              (SELFING ? selfprec[(from >> TYPEBITS) & SELFMASK][(to >> TYPEBITS) & SELFMASK] : 1);
		}
	}
	std::copy(probs2, probs2 + NUMTYPES, probs);
}

// initfwbw + realanalyze<STORE|FORWARD|1> / <STORE|BACKWARD|1>, cnF2freq.cpp:2074-2120, 2145-2418
void runsweeps(const Env& e, int root, Sweeps& sw, int c0, int c1)
{
	const Problem& P = *e.p;
	const int s = e.shift;
	if (sw.done[s]) return;
	sw.done[s] = true;
	const int E = c1 - 1;
	double probs[NUMTYPES];
	const int selfgen = P.gen[root] - 2;  // :2094
	double selfingfactors[4] = { 1, 1, 1, 1 };
	if (SELFING)  // :2092-2099
	{
		selfingfactors[0] = 1.0 / (1 << selfgen);
		selfingfactors[1] = (1 - selfingfactors[0]) * 0.5;
		selfingfactors[2] = (1 - selfingfactors[0]) * 0.5;
		selfingfactors[3] = 0;
	}
	for (int i = 0; i < NUMTYPES; i++) probs[i] = (1.0 / NUMTYPES) * (SELFING ? selfingfactors[(i >> TYPEBITS) & SELFMASK] : 1.0);  // EVENGEN :2100-2104
	double factor = 0;
	for (int m = c0; m <= E; m++)
	{
		std::copy(probs, probs + NUMTYPES, sw.at(s, m, 0));
		sw.f(s, m, 0) = factor;
		adjustprobs(e, root, probs, m, factor, -1);
		std::copy(probs, probs + NUMTYPES, sw.at(s, m, 2));
		sw.f(s, m, 2) = factor;
		if (m < E) transition(P, probs, P.markerpos[m + 1] - P.markerpos[m], selfgen);
	}
	for (int i = 0; i < NUMTYPES; i++) probs[i] = 1.0;  // :2111-2114
	factor = 0;
	std::copy(probs, probs + NUMTYPES, sw.at(s, E, 1));
	sw.f(s, E, 1) = factor;
	for (int j = E - 1; j >= c0; j--)
	{
		adjustprobs(e, root, probs, j + 1, factor, -1);
		transition(P, probs, P.markerpos[j + 1] - P.markerpos[j], selfgen);
		std::copy(probs, probs + NUMTYPES, sw.at(s, j, 1));
		sw.f(s, j, 1) = factor;
	}
}

// doanalyze + quickanalyze with NONESTOP (cnF2freq.cpp:1936-2032, 2122-2131): total log-likelihood of a shift mode.
// The reference's one extra transition towards markerposes[E+1] (:1984, an out-of-range read at the last
// chromosome) preserves the sum up to rounding and is left out.
double loglik_nonestop(Sweeps& sw, int s, int E, double minfactor_d)
{
	const float minfactor = (float)minfactor_d;
	double factor = sw.f(s, E, 2);
	if (factor < minfactor) return factor;
	double sum = 0;
	const double* a = sw.at(s, E, 2);
	const double* b = sw.at(s, E, 1);
	for (int k = 0; k < NUMTYPES; k++) sum += a[k] * b[k];
	factor += sw.f(s, E, 1);
	if (sum <= 0) factor = MINFACTOR;
	else
	{
		sum = 1 / sum;
		factor -= log(sum);
	}
	return factor;
}

// doanalyze + quickanalyze with classicstop(marker, g) and a fixed orientation flag2 (:1963-2018, :2229-2266)
double logpost_term(const Env& e, int root, Sweeps& sw, int marker, int g, int flag2, double minfactor_d)
{
	const float minfactor = (float)minfactor_d;
	const int s = e.shift;
	double factor = sw.f(s, marker, 0);
	if (factor < minfactor) return factor;
	// realanalyze<4>: filter to state g, adjustprobs with flag2, stop (:2229-2238, 2266)
	double probs[NUMTYPES];
	const double* a = sw.at(s, marker, 0);
	for (int i = 0; i < NUMTYPES; i++) probs[i] = a[i] * (i == g);
	double inner = 0;
	adjustprobs(e, root, probs, marker, inner, flag2);
	factor += inner;
	if (factor < minfactor) return factor;
	const double* b = sw.at(s, marker, 1);
	double sum = 0;
	for (int k = 0; k < NUMTYPES; k++)
	{
		probs[k] *= b[k];
		sum += probs[k];
	}
	factor += sw.f(s, marker, 1);
	if (sum <= 0) factor = MINFACTOR;
	else
	{
		sum = 1 / sum;
		factor -= log(sum);
	}
	return factor;
}

// ---------------------------------------------------------------------------------------------------
// One chromosome of one analysed individual: the body of the OpenMP loop in doit, cnF2freq.cpp:5302-5907
struct Outputs
{
	double* haplobase;   // [K][M]
	double* haplocount;  // [K][M]
	double* infprobs;    // [K][M][2][2]  (side, allele-1)
	double* homozyg;     // [K][M][2]
	double* loglik;      // [n_dous][C][2]  (max_s L_s, logsumexp_s L_s) -- the pair printed at :5399-5401
	long long* counters; // [2] terms tried (after ignoreflag2), terms kept (val > -200); may be null
	double* negshift;    // [K][M] phase-inversion scores (:3640-3715); may be null (no turn scoring)
	const int32_t* children;  // [K] individ::children as doit rebuilds it (:5246-5259); needed with negshift
};


// doanalyze<aroundturner>(classicstop(-1000 - marker, -1), flag2 = -1) for shift mode s (cnF2freq.cpp:1936-2018 with
// genotype -1 -> pad 2; realanalyze<4 | 2> :2229-2266 applies the turner and stops; aroundturner :498-554): alpha+ of
// shift s at the marker, permuted by the turn mask, against beta of shift s ^ flagmodeshift.
double turn_term(const Env& e0, int root, Sweeps& sw, int c0, int c1, int marker, int turn, double minfactor_d)
{
	const float minfactor = (float)minfactor_d;
	const int s = e0.shift;
	const int mask = turn & 54;                                                   // :506
	const int fms = (turn >> 6) | ((turn & 1) ? 2 : 0) | ((turn & 8) ? 4 : 0);   // :507-508
	double factor = sw.f(s, marker, 2);
	double probs[NUMTYPES];
	const double* a = sw.at(s, marker, 2);
	if (factor < minfactor) return factor;                                       // :1966
	for (int i = 0; i < NUMTYPES; i++) probs[i ^ mask] = a[i];                    // turner, :531-545
	const int s2 = s ^ fms;                                                       // :548
	Env e2 = e0;
	e2.shift = s2;
	runsweeps(e2, root, sw, c0, c1);                                              // initfwbw(..., 2) for the new shift mode, :1988
	if (factor < minfactor) return factor;
	const double* b = sw.at(s2, marker, 1);
	double sum = 0;
	for (int k = 0; k < NUMTYPES; k++)
	{
		probs[k] *= b[k];
		sum += probs[k];
	}
	factor += sw.f(s2, marker, 1);
	if (sum <= 0) factor = MINFACTOR;
	else
	{
		sum = 1 / sum;
		factor -= log(sum);
	}
	return factor;
}

// The turn scoring of one marker (cnF2freq.cpp:5668-5757 with DOTOULBAR 0: only the 8 turns of at most one bit) and
// updatenegshifts (:3640-3715).
void score_turns(Problem& P, const Env& e, const Tree& t, int root, Sweeps& sw, int c0, int c1, int marker, int shiftend,
                 double factor, const Outputs& out)
{
	constexpr int NUMTURNS = 128;
	const int M = P.M;
	for (int g = 0; g < NUMTURNS; g++)
	{
		if (g & (t.flag2ignore >> 1)) continue;
		if (__builtin_popcount(g & 127) > 1) continue;  // !DOTOULBAR, :5690-5698
		bool valid = false;
		double rawvals[NUMSHIFTS];
		for (int s = 0; s < NUMSHIFTS; s++) rawvals[s] = -1;
		for (int s = 0; s < shiftend; s++)
		{
			if (s & t.shiftignore) continue;
			Env es = e;
			es.shift = s;
			const double rawer = turn_term(es, root, sw, c0, c1, marker, g, -50000 + factor) - factor;
			valid = true;
			const double expval = exp(rawer);
			if (!std::isfinite(expval)) rawvals[s] = expval > 1e300 ? 1e300 : 0;
			else rawvals[s] = expval;
		}
		if (!valid) continue;
		// updatenegshifts, :3640-3715
		double val = 0;
		for (int s = 0; s < shiftend; s++)
		{
			if (s & t.shiftignore) continue;
			if (rawvals[s] < 0) continue;
			val += rawvals[s];
		}
		if (!std::isfinite(val) || val < 1e-174) val = 1e-174;
		int g2 = g;
		if (!g) g2 = (1 << 15) - 1;
		auto add = [&](int ind, int bit, int div) {
			double v = log(val) * (1.0 - ((g >> bit) & 1) * 2) * ((g2 >> bit) & 1);
			if (div) v = v / div;
#pragma omp atomic update
			out.negshift[(size_t)ind * M + marker] += v;
		};
		add(root, 6, 0);
		const int p0 = P.par[root * 2], p1 = P.par[root * 2 + 1];
		if (p0) add(p0, 0, 0);
		if (p1) add(p1, 3, 0);
		if (P.gen[root] >= 2)
		{
			if (p0 && P.par[p0 * 2]) add(P.par[p0 * 2], 1, out.children[p0]);
			if (p0 && P.par[p0 * 2 + 1]) add(P.par[p0 * 2 + 1], 2, out.children[p0]);
			if (p1 && P.par[p1 * 2]) add(P.par[p1 * 2], 4, out.children[p1]);
			if (p1 && P.par[p1 * 2 + 1]) add(P.par[p1 * 2 + 1], 5, out.children[p1]);
		}
	}
}

void estep_individual(Problem& P, int jidx, int root, int chrom, const Outputs& out, Sweeps* keep)
{
	const int c0 = P.chromstarts[chrom], c1 = P.chromstarts[chrom + 1];
	const int M = P.M;
	Scratch sc;
	Env e{ &P, &sc, 0 };
	Tree t = fixtrees(P, root);  // :5329
	int shiftend = NUMSHIFTS;
	if (P.gen[root] < 2) shiftend = std::min(2, shiftend);  // :5359

	Sweeps local;
	Sweeps& sw = keep ? *keep : local;
	sw.c0 = c0;
	sw.Mc = c1 - c0;
	sw.v.assign((size_t)NUMSHIFTS * sw.Mc * 3 * NUMTYPES, 0.0);
	sw.fac.assign((size_t)NUMSHIFTS * sw.Mc * 3, 0.0);
	for (bool& d : sw.done) d = false;

	// :5373-5382
	double factor = -1e15;
	double factors[NUMSHIFTS];
	for (int s = 0; s < shiftend; s++)
	{
		if (s & t.shiftignore) factors[s] = -1e30;
		else
		{
			e.shift = s;
			runsweeps(e, root, sw, c0, c1);
			factors[s] = loglik_nonestop(sw, s, c1 - 1, -40000 + factor);
		}
		factor = std::max(factor, factors[s]);
	}
	// :5385-5400
	double realfactor = 0;
	for (int s = 0; s < shiftend; s++)
	{
		if (s & t.shiftignore) continue;
		realfactor += exp(factors[s] - factor);
	}
	out.loglik[((size_t)jidx * P.C + chrom) * 2 + 0] = factor;
	factor += log(realfactor);
	out.loglik[((size_t)jidx * P.C + chrom) * 2 + 1] = factor;
	if (std::isnan(factor) || factor < MINFACTOR) return;  // :5403

	const double unusualstate = -200;  // :5414
	const double desc = P.descendants[root];
	for (int marker = c0; marker < c1; marker++)  // :5406
	{
		for (int g = 0; g < NUMTYPES; g++)
		{
			for (int s = 0; s < shiftend; s++)
			{
				if (s & t.shiftignore) continue;
				if (factor - factors[s] > 40) continue;  // :5421
				e.shift = s;
				for (int flag2 = 0; flag2 < NUMPATHS; flag2++)
				{
					if (ignoreflag2(P, t, flag2, g, s, marker)) continue;  // :5464
					if (out.counters) out.counters[0]++;
					double val = logpost_term(e, root, sw, marker, g, flag2, unusualstate + factor) - factor;  // :5499
					if (!(std::isfinite(val) && val > unusualstate)) continue;
					if (out.counters) out.counters[1]++;
					val = exp(val);
					double sidevals[2][2] = { { 0, 0 }, { 0, 0 } };
					double sidevalsums[2] = { 0, 0 };
					double hz[2] = { 0, 0 };
					for (int side = 0; side < 2; side++)  // :5519-5529
						for (int a = 1; a <= 2; a++)
						{
							double sv = track(e, U_GENOSPROBE, root, a, 0, marker, g * 2 + side, flag2 ^ side, s, 0.0, 4);
							sidevals[side][a - 1] += sv;
							sidevalsums[side] += sv;
						}
					for (int a = 1; a <= 2; a++)  // :5532-5537
						hz[a - 1] += track(e, U_HOMOZYGOUS, root, a, 0, marker, g * 2, flag2, s, 0.0, 4);
					// updatehaplo, :1561-1575
					double ok = track(e, 0, root, UNKNOWN, 0, marker, g * 2, flag2, s, val, 4);
					if (ok) track(e, U_HAPLOS, root, UNKNOWN, 0, marker, g * 2, flag2, s, val, 4);
					for (int side = 0; side < 2; side++)  // :5560-5568
						for (int a = 1; a <= 2; a++)
						{
							double updateval = val * sidevals[side][a - 1] / sidevalsums[side];
							track(e, U_GENOS, root, a, 0, marker, g * 2 + side, flag2 ^ side, s, updateval, 4);
						}
					for (int a = 1; a <= 2; a++)  // :5571-5575
						out.homozyg[((size_t)root * M + marker) * 2 + a - 1] += val * hz[a - 1] / sidevalsums[0];
				}
			}
		}
		if (out.negshift) score_turns(P, e, t, root, sw, c0, c1, marker, shiftend, factor, out);  // :5668-5871
		// :5875-5904
		double sum = 0;
		for (auto& kv : sc.infprobs[root][0]) sum += kv.second;
		sum = 1 / sum;
		for (int a = 0; a < 2; a++) out.homozyg[((size_t)root * M + marker) * 2 + a] *= sum;
		for (int rel : t.reltree)
		{
			// moveinfprobs, :3577-3597
			double norm = sum * 2;
			for (int k = 0; k < TURNBITS; k++)
				if (t.ordered[k] == rel) norm /= 2;
			norm *= desc;
			auto ip = sc.infprobs.find(rel);
			if (ip != sc.infprobs.end())
			{
				for (int side = 0; side < 2; side++)
				{
					for (auto& kv : ip->second[side])
					{
						if (kv.first >= 1 && kv.first <= 2)
						{
							const double add = kv.second * norm;
#pragma omp atomic update
							out.infprobs[(((size_t)rel * M + marker) * 2 + side) * 2 + kv.first - 1] += add;
						}
					}
					ip->second[side].clear();
				}
			}
			// movehaplos, :3599-3616
			auto hp = sc.haplos.find(rel);
			if (hp != sc.haplos.end() && (hp->second[0] || hp->second[1]))
			{
				if (fabs(P.haploweight[(size_t)rel * M + marker] - 0.5) < 0.5 - 1e-12)
				{
					const float maxdiff = 0.000005;  // :228
					double b1 = hp->second[0] + exp(-400) * maxdiff * maxdiff * 0.5;
					double b2 = hp->second[1] + exp(-400) * maxdiff * maxdiff * 0.5;
					const double add = b1 / (b1 + b2) * desc;
#pragma omp atomic update
					out.haplobase[(size_t)rel * M + marker] += add;
#pragma omp atomic update
					out.haplocount[(size_t)rel * M + marker] += desc;
				}
				hp->second[0] = 0;
				hp->second[1] = 0;
			}
		}
	}
}
} // namespace

extern "C"
{
// One E-step over all chromosomes and all analysed individuals, in the reference's order
// (chromosome outer, individual inner, cnF2freq.cpp:5283-5295).  Statistics arrays must be zeroed by the
// caller (the reference zeroes haplobase/haplocount at :5222-5234 and the M-step clears infprobs/homozyg).
// founder is in/out: fixtrees' sticky side effect (:3176) is applied in analysis order.
// fwbw_dous_index >= 0 additionally copies that individual's sweeps for chromosome fwbw_chrom into
// fwbw [8][Mc][3][NUMTYPES] / fwbwfactors [8][Mc][3] / fwbwmask [8].
// negshift [K][M] (zeroed by the caller; null = no turn scoring) accumulates the phase-inversion scores; children [K]
// is individ::children as doit rebuilds it from the analysed individuals (:5246-5259).
int cnf2_oracle_estep(int K, int M, int C, int n_dous, const int32_t* par, const uint8_t* empty, uint8_t* founder,
                      const int32_t* gen, const int32_t* descendants, const int32_t* dous, const double* markerpos,
                      const uint32_t* chromstarts, const double* genrec, const uint8_t* alleles, const double* sure,
                      const double* haploweight, double* haplobase, double* haplocount, double* infprobs,
                      double* homozyg, double* loglik, long long* counters, int fwbw_dous_index, int fwbw_chrom,
                      double* fwbw, double* fwbwfactors, int32_t* fwbwmask, double* negshift, const int32_t* children)
{
	Problem P;
	P.K = K;
	P.M = M;
	P.C = C;
	P.par = par;
	P.empty = empty;
	P.founder.assign(founder, founder + K);
	P.gen = gen;
	P.descendants = descendants;
	P.markerpos = markerpos;
	P.chromstarts = chromstarts;
	for (int i = 0; i < 3; i++) P.genrec[i] = genrec[i];
	P.alleles = alleles;
	P.sure = sure;
	P.haploweight = haploweight;
	Outputs out{ haplobase, haplocount, infprobs, homozyg, loglik, counters, negshift, children };
	// CNF2_ORACLE_THREADS > 1 (opt-in, for sampled checks at full size): units run in parallel; rows shared between units
	// are then summed in a nondeterministic order (atomic adds), rows only one unit adds to are unaffected.  The sticky
	// founder flags must already be steady (cnf2_oracle_steady_founder) in that mode.
	int nthreads = 1;
	if (const char* t = getenv("CNF2_ORACLE_THREADS")) nthreads = std::max(1, atoi(t));
	for (int c = 0; c < C; c++)
	{
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads) if (nthreads > 1)
		for (int j = 0; j < n_dous; j++)
		{
			Sweeps keep;
			bool want = (j == fwbw_dous_index && c == fwbw_chrom);
			estep_individual(P, j, dous[j], c, out, want ? &keep : nullptr);
			if (want)
			{
				std::copy(keep.v.begin(), keep.v.end(), fwbw);
				std::copy(keep.fac.begin(), keep.fac.end(), fwbwfactors);
				for (int s = 0; s < NUMSHIFTS; s++) fwbwmask[s] = keep.done[s] ? 3 : 0;
			}
		}
	}
	std::copy(P.founder.begin(), P.founder.end(), founder);
	return 0;
}

// Steady state of the sticky founder flag: what fixtrees (:3170-3177) leaves after every analysed
// individual has been visited once.
void cnf2_oracle_steady_founder(int K, int n_dous, const int32_t* par, const uint8_t* empty, const int32_t* dous,
                                uint8_t* founder)
{
	for (int j = 0; j < n_dous; j++)
	{
		int ind = dous[j];
		bool anylev1 = false;
		for (int lev1 = 0; lev1 < 2; lev1++)
		{
			int l1 = par[ind * 2 + lev1];
			if (!l1) continue;
			bool anypars = false;
			for (int lev2 = 0; lev2 < 2; lev2++)
			{
				int l2 = par[l1 * 2 + lev2];
				if (l2 && !empty[l2]) anypars = true;
			}
			if (anypars || !empty[l1]) anylev1 = true;
		}
		if (!anylev1) founder[ind] = 1;
	}
}
}
