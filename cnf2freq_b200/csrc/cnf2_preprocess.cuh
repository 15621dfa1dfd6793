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
template <bool SELF>
__global__ void __launch_bounds__(128) pre_variance_kernel(PreArgs a, const uint8_t* __restrict__ exists, const uint8_t* __restrict__ flag2ignore,
                                                           double* __restrict__ part_sq, double* __restrict__ part_sum)
{
	constexpr int NS = SELF ? NUMTYPES_SELFING : NUMTYPES;
	const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (tid >= (size_t)a.K * a.M * 8) return;
	const size_t idx = tid >> 3;
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
}  // namespace cnf2
