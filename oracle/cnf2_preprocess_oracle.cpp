// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// CPU restatement of the reference's pre-processing, postmarkerdata (cnF2freq.cpp:3191-3412), row f3 of SURVEY.md
// section 8: fixkid :1469-1487, the descendant / child counts :3228-3277, fixparents :1392-1467 with its emission scan
// (calltrackpossible<false,false> :1380-1385 -> trackpossible :1075-1359, zeropropagate 0, CORRECTIONINFERENCE true), the
// corrections :3283-3351 (dosureval :3082-3097), addvariance :1489-1558 (trackpossible with NO_EQUIVALENCE) and lockhaplos
// :3045-3079.  Everything runs in the reference's single-threaded order: individuals by id, markers inside, every effect
// (a cleared genotype, :1432-1437) visible to whatever comes later -- which is also what pins the engine's parallel
// formulation (cnf2_preprocess).  Pinned against the reference binary: tests/test_preprocess_oracle.py compares the state
// this leaves with the reference's own state after postmarkerdata (iteration-0 dumps in tests/golden/host/, priors and
// "before iteration 1" parameters in tests/golden/mstep_*.npz).
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <map>
#include <utility>
#include <vector>

namespace
{
#ifndef CNF2_ORACLE_SELFING
#define CNF2_ORACLE_SELFING 0
#endif
// -DCNF2_ORACLE_SELFING=1: the SELFING = true instantiation (settings.h:14; row g1), see cnf2_oracle.cpp
constexpr bool SELFING = CNF2_ORACLE_SELFING;
constexpr int TYPEBITS = 6, SELFMASK = 3;
constexpr int UNKNOWN = 0, SEXMARKER = 9, NUMTYPES = SELFING ? 256 : 64, NUMPATHS = 128;

struct Pre
{
	int K, M;
	const int32_t* par;
	uint8_t* al;
	double* su;
	double* hw;
	std::vector<uint8_t> founder;  // individ::founder as it stands at this moment
};

bool markermiss(int zp, int& a, int b)  // :303-316 (zeropropagate 0 or -1)
{
	if (a == UNKNOWN)
	{
		if (zp == 0) a = b;
		return false;
	}
	if (b == UNKNOWN && a != SEXMARKER) return false;
	return a != b;
}

int slice(int flag, int par, int w)  // upflagit, :321-329
{
	if (flag < 0) return flag;
	return (flag >> (par * (w - 1))) & ((1 << (w - 1)) - 1);
}

// trackpossible<update = 0, zp> (:1075-1359) with CORRECTIONINFERENCE true
double track(const Pre& P, int zp, int ind, int inval, double secondval, int m, unsigned flag, int flag99, int localshift, int genwidth)
{
	const bool rootgen = genwidth == 4;
	const bool attopnow = genwidth == 1 || P.founder[ind];
	const int A0[2] = { P.al[((size_t)ind * P.M + m) * 2], P.al[((size_t)ind * P.M + m) * 2 + 1] };
	const double S0[2] = { P.su[((size_t)ind * P.M + m) * 2], P.su[((size_t)ind * P.M + m) * 2 + 1] };
	int A[2] = { A0[0], A0[1] };
	double S[2] = { S0[0], S0[1] };
	const int selfval = (int)(flag >> (TYPEBITS + 1)) & SELFMASK;  // :1122
	const bool selfingNOW = SELFING && rootgen && selfval;          // :1126
	const double hw = P.hw[(size_t)ind * P.M + m];
	int upflag2 = -1;
	const int upflag = (int)(flag >> 1), upshift = localshift >> 1;
	int f2s = 0, f2end = 2;
	if (flag99 != -1)
	{
		upflag2 = flag99 >> 1;
		f2s = flag99 & 1;
		f2end = f2s + 1;
	}
	const int firstpar = flag & 1;
	double ok = 0;
	for (int flag2 = f2s; flag2 < f2end; flag2++)
	{
		int f2n = flag2 & 1;
		const int realf2n = f2n;
		if (selfingNOW)  // :1169-1190
		{
			const int selfindex = (selfval >> 1) ^ f2n;
			A[0] = A[1] = UNKNOWN;
			S[0] = S[1] = 0;
			int first = A0[0];
			if (!markermiss(0, first, A0[1]))
			{
				A[selfindex] = first;
				S[selfindex] = 1 - (1 - S0[0]) * (1 - S0[1]);
			}
			else
			{
				A[selfindex] = A0[1];
				if (S0[0] == 0) return 0;
				S[selfindex] = 1 - (S0[0]) * (1 - S0[1]);
			}
		}
		const bool allthesame = A[0] == A[1];
		int markerval = inval;
		double baseval, mainsecondval = 0;
		if (markermiss(zp, markerval, A[f2n]))
		{
			baseval = S[f2n];
			if (S[f2n] && secondval) mainsecondval = (1.0 - S[f2n]) * secondval;
		}
		else
		{
			const double eff = (inval == UNKNOWN && markerval != UNKNOWN) ? 1 : secondval;
			baseval = 1.0 - S[f2n];
			mainsecondval = (A[f2n] == UNKNOWN ? 1 : S[f2n]) * eff;
		}
		if (attopnow)
		{
			baseval += mainsecondval;
			mainsecondval = 0;
		}
		else if (mainsecondval) mainsecondval /= baseval;
		f2n ^= (firstpar ^ localshift) & 1;
		if (zp || !genwidth) baseval *= 0.5;
		else if (allthesame || selfingNOW) baseval *= f2n ? 1.0 : 0.0;  // CORRECTIONINFERENCE, :1235
		else baseval *= std::fabs((f2n ? 1.0 : 0.0) - hw);
		if (baseval && !attopnow)
		{
			auto up = [&](int par, int val, double second) -> double {  // recursetrackpossible, :955-1058
				const int who = P.par[(size_t)ind * 2 + par];
				if (!who) return 1 + second;
				return track(P, zp, who, val, second, m, (unsigned)slice(upflag, par, genwidth), slice(upflag2, par, genwidth),
				             slice(upshift, par, genwidth >> 1), genwidth >> 1);
			};
			if (!zp || rootgen)
			{
				double sec2 = 0;
				if (S[!realf2n])
				{
					baseval *= 1 - S[!realf2n];
					sec2 = S[!realf2n] / (1 - S[!realf2n]);
				}
				baseval *= up(!firstpar, A[!realf2n], sec2);
			}
			if (baseval) baseval *= up(firstpar, markerval, mainsecondval);
		}
		if (baseval) ok += baseval;
	}
	return ok;
}

int g_cleared = 0;  // genotypes cleared by the last run (:1432-1437)

double dosureval(int what, double prob)  // :3082-3097
{
	if (prob == 0) return 0;
	double toret = std::log(prob);
	toret /= what;
	toret *= 4;
	toret = std::exp(toret);
	toret /= (1 + toret);
	return toret;
}
}  // namespace

extern "C" int cnf2_oracle_preprocess(int K, int M, int C, const uint32_t* chromstarts, const int32_t* par, const uint8_t* exists,
                                      const uint8_t* empty, uint8_t* al, double* su, double* hw, int32_t* descendants, int32_t* children,
                                      uint8_t* founder, int32_t* n_corrections)
{
	Pre P{ K, M, par, al, su, hw, std::vector<uint8_t>(K, 0) };
	g_cleared = 0;
	auto md = [&](int i, int g, int s) -> uint8_t& { return al[((size_t)i * M + g) * 2 + s]; };
	auto ms = [&](int i, int g, int s) -> double& { return su[((size_t)i * M + g) * 2 + s]; };
	typedef std::map<int, std::pair<int, double>> Evidence;  // individ::markervals[marker]
	std::vector<Evidence> vals((size_t)K * M);
	int any, anyrem, total = 0;
	do
	{
		for (int i = 1; i < K; i++) children[i] = 0;
		for (int i = 1; i < K; i++)  // fixkid
		{
			if (!exists[i]) continue;
			for (int g = 0; g < M; g++)
			{
				if (md(i, g, 0) != UNKNOWN || md(i, g, 1) != UNKNOWN) continue;
				for (int p = 0; p < 2; p++)
				{
					const int q = par[(size_t)i * 2 + p];
					if (!q) continue;
					if (md(q, g, 0) == UNKNOWN || md(q, g, 0) != md(q, g, 1)) continue;
					md(i, g, p) = md(q, g, 0);
					ms(i, g, p) = 0.5;
				}
			}
		}
		{
			std::vector<int> upsent(K, 0);
			bool anychange;
			do
			{
				anychange = false;
				for (int i = 1; i < K; i++)
				{
					if (!exists[i]) continue;
					int nowdesc = descendants[i];
					if (!nowdesc) nowdesc = 1;
					nowdesc -= upsent[i];
					if (nowdesc > 0)
					{
						for (int p = 0; p < 2; p++)
							if (const int q = par[(size_t)i * 2 + p]) descendants[q] += nowdesc;
						upsent[i] += nowdesc;
						anychange = true;
					}
				}
			} while (anychange);
			for (int i = 1; i < K; i++)
				if (exists[i] && descendants[i] == 0) descendants[i] = 1;
		}
		for (int i = 1; i < K; i++)  // fixparents
		{
			if (!exists[i]) continue;
			for (int p = 0; p < 2; p++)
				if (const int q = par[(size_t)i * 2 + p]) children[q]++;
			for (int g = 0; g < M; g++)
			{
				double okvals[2] = { 0, 0 };
				for (int st = 0; st < NUMTYPES; st++)
				{
					for (int flag2 = 0; flag2 < NUMPATHS; flag2++)
					{
						if (okvals[flag2 & 1]) continue;
						const double ok = track(P, 0, i, UNKNOWN, 0, g, (unsigned)(st * 2), flag2, 0, 4);
						if (!ok) continue;
						okvals[flag2 & 1] += ok;
						if (okvals[0] && okvals[1]) break;
					}
					if (okvals[0] && okvals[1]) break;
				}
				if (!okvals[0] && !okvals[1])
				{
					g_cleared++;
					md(i, g, 0) = md(i, g, 1) = UNKNOWN;
					ms(i, g, 0) = ms(i, g, 1) = 0;
				}
				if ((bool)okvals[0] ^ (bool)okvals[1])
					for (int flag2 = 0; flag2 < 2; flag2++)
					{
						if (!okvals[flag2]) continue;
						for (int k = 0; k < 2; k++)
						{
							const int q = par[(size_t)i * 2 + k];
							if (!q) continue;
							const int u = (k ^ flag2) & 1;
							if (md(i, g, u) == UNKNOWN) continue;
							double old1 = 1;
							int old2 = 0;
							Evidence& ev = vals[(size_t)q * M + g];
							auto it = ev.find(md(i, g, u));
							if (it != ev.end()) old1 = it->second.second, old2 = it->second.first;
							double probit = ms(i, g, 0) + ms(i, g, 1);
							probit /= (1 - probit);
							ev[md(i, g, u)] = std::make_pair(old2 + 1, old1 * probit);
						}
					}
			}
		}
		any = anyrem = 0;
		for (int i = 1; i < K; i++)  // corrections
		{
			if (!exists[i]) continue;
			for (int g = 0; g < M; g++)
			{
				Evidence& ev = vals[(size_t)i * M + g];
				const int known = (md(i, g, 0) != UNKNOWN) + (md(i, g, 1) != UNKNOWN);
				if (known == 2) continue;
				ev.erase(UNKNOWN);
				if (md(i, g, 0) != UNKNOWN) ev.insert(std::make_pair((int)md(i, g, 0), std::make_pair(children[i], ms(i, g, 0))));
				if (md(i, g, 1) != UNKNOWN) ev.insert(std::make_pair((int)md(i, g, 1), std::make_pair(children[i], ms(i, g, 1))));
				if (ev.size() == 2)
				{
					auto a = ev.begin(), b = ++ev.begin();
					const int knowncount = a->second.first + b->second.first;
					md(i, g, 0) = (uint8_t)a->first;
					md(i, g, 1) = (uint8_t)b->first;
					ms(i, g, 0) = dosureval(knowncount, a->second.second);
					ms(i, g, 1) = dosureval(knowncount, b->second.second);
					any++;
				}
				else if (ev.size() == 1 && known == 0)
				{
					any++;
					md(i, g, 0) = (uint8_t)ev.begin()->first;
					md(i, g, 1) = UNKNOWN;
					ms(i, g, 0) = dosureval(ev.begin()->second.first, ev.begin()->second.second);
					ms(i, g, 1) = 0;
				}
			}
			for (int g = 0; g < M; g++)
			{
				if (md(i, g, 0) == SEXMARKER) std::swap(md(i, g, 0), md(i, g, 1));
				vals[(size_t)i * M + g].clear();
			}
		}
		total += any;
	} while (any > anyrem);
	if (n_corrections) *n_corrections = total;

	for (int i = 1; i < K; i++)  // fixtrees + addvariance + lockhaplos, individuals in id order
	{
		if (!exists[i]) continue;
		int flag2ignore = 1;
		bool anylev1 = false;
		for (int lev1 = 0; lev1 < 2; lev1++)
		{
			const int l1 = par[(size_t)i * 2 + lev1];
			if (!l1) continue;
			const int idx = 1 + lev1 * 3;
			if (!empty[l1]) flag2ignore |= 1 << idx;
			bool anypars = false;
			for (int lev2 = 0; lev2 < 2; lev2++)
			{
				const int l2 = par[(size_t)l1 * 2 + lev2];
				if (!l2 || empty[l2]) continue;
				flag2ignore |= 1 << (idx + lev2 + 1);
				anypars = true;
			}
			if (anypars || !empty[l1]) anylev1 = true;
		}
		if (!anylev1) P.founder[i] = 1;  // sticky (:3176)
		flag2ignore ^= NUMPATHS - 1;
		std::vector<double> variances(M, 0.0);
		for (int g = 0; g < M; g++)
		{
			double sum = 0, sqsum = 0;
			for (int shift = 0; shift < 2; shift++)
				for (int majori = 0; majori < 2; majori++)
					for (int majorflag2 = 0; majorflag2 < 2; majorflag2++)
					{
						double fullok = 0, ok = 0;
						for (int subi = 0, st = majori; subi < NUMTYPES * 2; subi += 2, st += 2)
							for (int flag2 = majorflag2; flag2 < NUMPATHS; flag2 += 2)
							{
								if (flag2 & flag2ignore) continue;
								for (int allele = 0; allele < 2; allele++)
								{
									const double term = track(P, -1, i, md(i, g, allele), ms(i, g, allele), g, (unsigned)st, flag2, shift, 4);
									ok += term * (allele ? 1 : -1);
									fullok += term;
								}
							}
						ok = std::fabs(ok);
						sum += fullok;
						sqsum += ok * ok;
					}
			if (sum) variances[g] = sqsum;
		}
		for (int c = 0; c < C; c++)
		{
			long bestpos = -1;
			double bestvar = 0;
			for (uint32_t j = chromstarts[c]; j != chromstarts[c + 1]; j++)
				if (variances[j] > bestvar) bestpos = j, bestvar = variances[j];
			if (bestpos == -1) continue;
			hw[(size_t)i * M + bestpos] = hw[(size_t)i * M + bestpos] <= 0.5 ? 0 : 1;
		}
	}
	for (int i = 1; i < K; i++) founder[i] = P.founder[i];
	return 0;
}

extern "C" int cnf2_oracle_preprocess_cleared() { return g_cleared; }
