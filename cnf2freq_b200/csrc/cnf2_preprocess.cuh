// Pre-processing on the device (row f3 of SURVEY.md section 8): the two emission scans of postmarkerdata
// (cnF2freq.cpp:3191-3412) -- fixparents' search for an admissible (state, orientation) per root orientation
// (:1392-1467) and addvariance's signed sums over every (flag, orientation) pair (:1489-1558).  Everything else of
// postmarkerdata is sequential bookkeeping over individuals and stays host logic in cnf2_preprocess (cnf2_engine.cu).
//
// Both scans run over EVERY individual, analysed or not, so the emission here walks the pedigree by individual id
// (parents [K][2]) over the full [K][M] parameter arrays instead of a focus-tree unit.  ptrack is trackpossible
// (:1075-1359) with update = 0 for the two zeropropagate modes the scans use: 0 (binding, fixparents through
// calltrackpossible :1380-1385) and NO_EQUIVALENCE = -1 (addvariance), and with CORRECTIONINFERENCE true (:8083): the
// homozygous rule applies whatever the sureness (:1235).  Products and sums are written with the rounding intrinsics so
// that no multiply-add is fused: addvariance's result decides by `>` which marker's phase is locked (:3058), and the
// reference binary (x86-64, no FMA contraction) is the yardstick.
// SELF: the SELFING = true instantiation (row g1): both scans run over 256 states, and the root's emission sees the selfing
// bits of the flag (:1115-1127, :1169-1190) exactly as track<UPD, GW, true> does in the E-step.
#pragma once
#include "cnf2_device.cuh"

namespace cnf2
{
struct PreArgs
{
	const uchar2* alleles;   // [K][M]
	const double2* sure;     // [K][M]
	const double* hw;        // [K][M]
	const int2* par;         // [K] (0 = none)
	const uint8_t* steady;   // [K] what fixtrees leaves in individ::founder (:3170-3177) once it has seen the individual
	int32_t K, M;
};

__device__ __forceinline__ double pmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double padd(double a, double b) { return __dadd_rn(a, b); }

// markermiss<zeropropagate> (:303-316) for zeropropagate 0 (binds an unknown value) and -1 (does not)
template <int ZP>
__device__ __forceinline__ bool pmiss(int& a, int b)
{
	if (a == UNKNOWN)
	{
		if (ZP == 0) a = b;
		return false;
	}
	if (b == UNKNOWN && a != SEXMARKER) return false;
	return a != b;
}

// founders_upto: individ::founder is false for everybody while fixparents runs and becomes steady[j] when the
// addvariance loop reaches individual j (its fixtrees call, :3377) -- single-threaded order, individuals by id
template <int ZP, int GW, bool SELF = false>
__device__ double ptrack(const PreArgs& a, int ind, int m, int inval, double second, unsigned flag, int flag99, int localshift, int founders_upto)
{
	const bool rootgen = GW == 4;
	const bool founder = ind <= founders_upto && a.steady[ind];
	const bool attopnow = GW == 1 || founder;  // :1120
	const size_t cell = (size_t)ind * a.M + m;
	uchar2 A = a.alleles[cell];
	double2 S = a.sure[cell];
	const uchar2 A0 = A;
	const double2 S0 = S;
	const int selfval = (SELF && GW == 4) ? (int)((flag >> (TYPEBITS + 1)) & SELFMASK) : 0;  // :1122
	const bool selfingNOW = selfval != 0;                                                       // :1126
	const double hw = a.hw[cell];
	const int2 pars = a.par[ind];
	int upflag2 = -1;
	const unsigned upflag = flag >> 1;
	const int upshift = localshift >> 1;
	int f2s = 0, f2end = 2;
	if (flag99 != -1)
	{
		upflag2 = flag99 >> 1;
		f2s = flag99 & 1;
		f2end = f2s + 1;
	}
	const int firstpar = flag & 1;
	bool allthesame = A.x == A.y;
	double ok = 0;
	for (int f2 = f2s; f2 < f2end; f2++)
	{
		int f2n = f2;
		const int realf2n = f2;
		if constexpr (SELF && GW == 4)
		{
			if (selfingNOW)  // :1169-1190
			{
				const int selfindex = (selfval >> 1) ^ f2;
				int first = A0.x;
				int mk;
				double su;
				if (!pmiss<0>(first, A0.y))
				{
					mk = first;
					su = padd(1.0, -pmul(1.0 - S0.x, 1.0 - S0.y));
				}
				else
				{
					mk = A0.y;
					if (S0.x == 0) return 0;  // :1186
					su = padd(1.0, -pmul(S0.x, 1.0 - S0.y));
				}
				A = selfindex ? make_uchar2(UNKNOWN, (unsigned char)mk) : make_uchar2((unsigned char)mk, UNKNOWN);
				S = selfindex ? make_double2(0.0, su) : make_double2(su, 0.0);
				allthesame = A.x == A.y;
			}
		}
		const int Af = f2 ? A.y : A.x, Ao = f2 ? A.x : A.y;
		const double Sf = f2 ? S.y : S.x, So = f2 ? S.x : S.y;
		int val = inval;
		double base, main2 = 0;
		if (pmiss<ZP>(val, Af))  // :1198-1203
		{
			base = Sf;
			if (Sf != 0 && second != 0) main2 = pmul(1.0 - Sf, second);
		}
		else  // :1204-1210
		{
			const double eff2 = (inval == UNKNOWN && val != UNKNOWN) ? 1.0 : second;
			base = 1.0 - Sf;
			main2 = pmul(Af == UNKNOWN ? 1.0 : Sf, eff2);
		}
		if (attopnow)  // :1213-1217 (update = 0)
		{
			base = padd(base, main2);
			main2 = 0;
		}
		else if (main2 != 0) main2 = main2 / base;
		f2n ^= (firstpar ^ localshift) & 1;  // :1227
		if (ZP != 0) base = pmul(base, 0.5);  // :1229-1233 (genwidth is never 0 here)
		else if (allthesame || selfingNOW) base = pmul(base, f2n ? 1.0 : 0.0);  // :1235 with CORRECTIONINFERENCE true
		else base = pmul(base, fabs((f2n ? 1.0 : 0.0) - hw));    // :1245

		if constexpr (GW > 1)
		{
			if (base != 0 && !attopnow)  // :1271-1341
			{
				auto up = [&](int p, int v, double sec) -> double {  // recursetrackpossible, :955-1058
					const int who = p ? pars.y : pars.x;
					if (!who) return padd(1.0, sec);  // :1043-1046
					const unsigned f = (upflag >> (p * (GW - 1))) & ((1u << (GW - 1)) - 1u);
					const int f99 = upflag2 < 0 ? upflag2 : (upflag2 >> (p * (GW - 1))) & ((1 << (GW - 1)) - 1);
					const int sh = (upshift >> (p * ((GW >> 1) - 1))) & ((1 << ((GW >> 1) - 1)) - 1);  // NUMGEN - NUMSHIFTGEN = 1, :986
					return ptrack<ZP, (GW >> 1)>(a, who, m, v, sec, f, f99, sh, founders_upto);
				};
				if (ZP == 0 || rootgen)  // :1291 (update & GENOS is 0)
				{
					double sec2 = 0;
					if (So != 0)  // :1298-1302
					{
						base = pmul(base, 1.0 - So);
						sec2 = So / (1.0 - So);
					}
					base = pmul(base, up(!firstpar, Ao, sec2));
				}
				if (base != 0) base = pmul(base, up(firstpar, val, main2));
			}
		}
		if (base != 0) ok = padd(ok, base);
	}
	return ok;
}

// fixparents' scan (:1413-1430) for every (individual, marker): bit f of okmask = some (state i, orientation flag2 with
// flag2 & 1 = f) has a non-zero emission at shift mode 0.  One thread per cell; the loops end as the reference's do.
template <bool SELF>
__global__ void __launch_bounds__(128) pre_okmask_kernel(PreArgs a, const uint8_t* __restrict__ exists, uint8_t* __restrict__ okmask)
{
	constexpr int NS = SELF ? NUMTYPES_SELFING : NUMTYPES;
	const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= (size_t)a.K * a.M) return;
	const int ind = (int)(idx / a.M), m = (int)(idx % a.M);
	if (!exists[ind])
	{
		okmask[idx] = 0;
		return;
	}
	bool ok0 = false, ok1 = false;
	for (int i = 0; i < NS && !(ok0 && ok1); i++)
		for (int flag2 = 0; flag2 < NUMPATHS && !(ok0 && ok1); flag2++)
		{
			if ((flag2 & 1) ? ok1 : ok0) continue;
			const double v = ptrack<0, 4, SELF>(a, ind, m, UNKNOWN, 0.0, (unsigned)(i * 2), flag2, 0, -1);
			if (!(v != 0)) continue;  // `if (!ok) continue`: a NaN counts as found, as it does there
			if (flag2 & 1) ok1 = true;
			else ok0 = true;
		}
	okmask[idx] = (uint8_t)((ok0 ? 1 : 0) | (ok1 ? 2 : 0));
}

// addvariance (:1489-1558) for every (individual, marker): one thread per (cell, shift mode, majori, majorflag2) walks its
// 64 x 64 (flag, orientation) pairs in the reference's order with the reference's running sums; the eight partial results
// are combined in that order too.  flag2ignore: what fixtrees returns for the individual (absent / empty slots).
// list: null = every cell; else the n_list cells (ind * M + marker) to evaluate, results in list order.
template <bool SELF>
__global__ void __launch_bounds__(128) pre_variance_kernel(PreArgs a, const uint8_t* __restrict__ exists, const uint8_t* __restrict__ flag2ignore,
                                                           double* __restrict__ part_sq, double* __restrict__ part_sum,
                                                           const unsigned long long* __restrict__ list, size_t n_list)
{
	constexpr int NS = SELF ? NUMTYPES_SELFING : NUMTYPES;
	const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (tid >= (list ? n_list : (size_t)a.K * a.M) * 8) return;
	const size_t idx = list ? (size_t)list[tid >> 3] : tid >> 3;
	const int combo = (int)(tid & 7), shift = combo >> 2, majori = (combo >> 1) & 1, majorflag2 = combo & 1;
	const int ind = (int)(idx / a.M), m = (int)(idx % a.M);
	double fullok = 0, ok = 0;
	if (exists[ind])
	{
		const size_t cell = (size_t)ind * a.M + m;
		const uchar2 A = a.alleles[cell];
		const double2 S = a.sure[cell];
		const int ign = flag2ignore[ind];
		for (int subi = 0, i = majori; subi < NS * 2; subi += 2, i += 2)
			for (int flag2 = majorflag2; flag2 < NUMPATHS; flag2 += 2)
			{
				if (flag2 & ign) continue;
				for (int allele = 0; allele < 2; allele++)
				{
					const double term = ptrack<-1, 4, SELF>(a, ind, m, allele ? A.y : A.x, allele ? S.y : S.x, (unsigned)i, flag2, shift, ind);
					ok = padd(ok, pmul(term, allele ? 1.0 : -1.0));
					fullok = padd(fullok, term);
				}
			}
		ok = fabs(ok);
	}
	part_sq[tid] = pmul(ok, ok);
	part_sum[tid] = fullok;
}

// ---- addvariance, factorised (SELFING = false) ----
// In addvariance's mode (NO_EQUIVALENCE) the phase weight is a plain 1/2 (:1229-1233), so a term is a product of one factor per
// tree node and the 64 x 64 (flag, orientation) pairs of a (majori, majorflag2) class sum subtree by subtree:
//   sum = root(majori, majorflag2) * U(parent !firstpar; other allele) * U(parent firstpar; matched allele),
//   U(q) = 4 * sum_{f2 of q} sum_{firstpar' of q} node(q, f2) * [sum_{f2 of that grandparent} leaf] * |orientations of the other one|
// (a node below the root follows one parent only, :1291; bits nobody reads count by their multiplicity; flag2ignore restricts an
// absent or empty slot to orientation 0, :1524).  ~500 node evaluations per (individual, marker) instead of 65,536 emission walks.
// The result differs from the literal sums by rounding, and lockhaplos decides by `>` (:3058): so this value only RANKS the
// markers -- cnf2_preprocess re-evaluates the few that can hold the maximum with the literal kernel and decides on those.
struct PNode
{
	double base, main2;
};

// :1198-1210 for zeropropagate = NO_EQUIVALENCE (nothing is bound, so the value handed on is the incoming one)
__device__ __forceinline__ PNode pnode(uchar2 A, double2 S, int v, double sec, int f2)
{
	const int Af = f2 ? A.y : A.x;
	const double Sf = f2 ? S.y : S.x;
	const bool miss = v != UNKNOWN && !(Af == UNKNOWN && v != SEXMARKER) && v != Af;
	PNode r;
	if (miss)
	{
		r.base = Sf;
		r.main2 = (Sf != 0 && sec != 0) ? (1.0 - Sf) * sec : 0.0;
	}
	else
	{
		r.base = 1.0 - Sf;
		r.main2 = (Af == UNKNOWN ? 1.0 : Sf) * sec;
	}
	return r;
}

// sum over the three flag bits and the admissible three orientation bits of the subtree of parent q (tree slot sb = 1 or 4)
__device__ double pfast_subtree(const PreArgs& a, int q, int m, int v, double sec, int sb, int ign, int root)
{
	auto cnt = [&](int slot) { return ((ign >> slot) & 1) ? 1.0 : 2.0; };
	if (!q) return (1.0 + sec) * 8.0 * cnt(sb) * cnt(sb + 1) * cnt(sb + 2);  // :1043-1046
	const size_t cell = (size_t)q * a.M + m;
	const uchar2 A = a.alleles[cell];
	const double2 S = a.sure[cell];
	const bool founder = q <= root && a.steady[q];
	const int2 pars = a.par[q];
	double sum = 0;
	for (int f2 = 0; f2 < 2; f2++)
	{
		if (f2 && ((ign >> sb) & 1)) break;
		const PNode n = pnode(A, S, v, sec, f2);
		if (founder)  // at the top (:1120): everything in one factor, no recursion; both grandparents' bits are free
		{
			sum += (n.base + n.main2) * 0.5 * 2.0 * cnt(sb + 1) * cnt(sb + 2);
			continue;
		}
		const double base = n.base * 0.5;
		if (!(base != 0)) continue;
		const double m2 = n.main2 != 0 ? n.main2 / n.base : 0.0;
		for (int fp = 0; fp < 2; fp++)
		{
			const int who = fp ? pars.y : pars.x;
			const int gs = sb + 1 + fp;
			double g = 0;
			if (!who) g = (1.0 + m2) * cnt(gs);
			else
			{
				const size_t gc = (size_t)who * a.M + m;
				const uchar2 GA = a.alleles[gc];
				const double2 GS = a.sure[gc];
				for (int g2 = 0; g2 < 2; g2++)
				{
					if (g2 && ((ign >> gs) & 1)) break;
					const PNode l = pnode(GA, GS, v, m2, g2);
					g += (l.base + l.main2) * 0.5;  // genwidth 1: at the top
				}
			}
			sum += base * g * cnt(sb + 1 + (1 - fp));
		}
	}
	return sum * 4.0;  // the grandparents' flag bits
}

// out[cell] = the variance addvariance would store, up to rounding (0 where it returns early, :1553); negative sign = at
// rounding-noise level relative to the sum (the ranking is then meaningless).  sig[cell] = a hash of everything the value depends on
// at this marker: cells with equal signatures have bit-identical literal variances.
__global__ void __launch_bounds__(128) pre_variance_fast_kernel(PreArgs a, const uint8_t* __restrict__ exists, const uint8_t* __restrict__ flag2ignore,
                                                                double* __restrict__ out, unsigned long long* __restrict__ sig)
{
	const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= (size_t)a.K * a.M) return;
	const int ind = (int)(idx / a.M), m = (int)(idx % a.M);
	if (!exists[ind])
	{
		out[idx] = 0;
		sig[idx] = 0;
		return;
	}
	const uchar2 A = a.alleles[idx];
	const double2 S = a.sure[idx];
	const int ign = flag2ignore[ind];
	const int2 pars = a.par[ind];
	const bool founder = a.steady[ind];
	auto cnt = [&](int slot) { return ((ign >> slot) & 1) ? 1.0 : 2.0; };
	double sum = 0, sqsum = 0;
	for (int majori = 0; majori < 2; majori++)
		for (int f2 = 0; f2 < 2; f2++)
		{
			if (f2 && (ign & 1)) continue;
			double T[2];
			for (int allele = 0; allele < 2; allele++)
			{
				const int inval = allele ? A.y : A.x;
				const double second = allele ? S.y : S.x;
				const PNode n = pnode(A, S, inval, second, f2);
				double t;
				if (founder) t = (n.base + n.main2) * 0.5 * 64.0 * cnt(1) * cnt(2) * cnt(3) * cnt(4) * cnt(5) * cnt(6);
				else
				{
					t = n.base * 0.5;
					if (t != 0)
					{
						const double m2 = n.main2 != 0 ? n.main2 / n.base : 0.0;
						const int Ao = f2 ? A.x : A.y;
						const double So = f2 ? S.x : S.y;
						double sec2 = 0;
						if (So != 0)
						{
							t *= 1.0 - So;
							sec2 = So / (1.0 - So);
						}
						const int firstpar = majori;
						t *= pfast_subtree(a, firstpar ? pars.x : pars.y, m, Ao, sec2, firstpar ? 1 : 4, ign, ind);
						t *= pfast_subtree(a, firstpar ? pars.y : pars.x, m, inval, m2, firstpar ? 4 : 1, ign, ind);
					}
				}
				T[allele] = t;
			}
			const double ok = fabs(T[1] - T[0]);
			sum += T[0] + T[1];
			sqsum += ok * ok;
		}
	// the two shift modes of :1511 give the same sums (the shift only enters through the phase weight, which is 1/2 here)
	sum *= 2.0, sqsum *= 2.0;
	const double var = sum != 0 ? sqsum : 0.0;
	out[idx] = (var != 0 && var <= 1e-20 * sum * sum) ? -var : var;
	// signature: genotype and sureness of the seven tree members at this marker (FNV-1a over their bits)
	unsigned long long h = 1469598103934665603ull;
	auto mix = [&](unsigned long long x) {
		h ^= x;
		h *= 1099511628211ull;
	};
	auto feed = [&](int who) {
		if (!who)
		{
			mix(0x9e3779b97f4a7c15ull);
			return;
		}
		const size_t c = (size_t)who * a.M + m;
		const uchar2 X = a.alleles[c];
		const double2 Y = a.sure[c];
		mix(((unsigned long long)X.x << 8) | X.y);
		mix((unsigned long long)__double_as_longlong(Y.x));
		mix((unsigned long long)__double_as_longlong(Y.y));
	};
	feed(ind);
	for (int p = 0; p < 2; p++)
	{
		const int q = p ? pars.y : pars.x;
		feed(q);
		const int2 gp = q ? a.par[q] : make_int2(0, 0);
		feed(gp.x);
		feed(gp.y);
	}
	sig[idx] = h;
}
}  // namespace cnf2
