// Device-side data model and the emission routine of the cnF2freq hot path, for sm_100a.
//
// State space (settings.h:19-42 at HEAD): 64 inheritance states g (6 meiosis bits), 128 orientations
// flag2 (7 focus-tree slots), 8 shift modes.  Slots: 0 root, 1 parent 0, 2/3 its parents, 4 parent 1,
// 5/6 its parents (fixtrees, cnF2freq.cpp:3124, 3146).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace cnf2
{
constexpr int NUMTYPES = 64;   // settings.h:27
// SELFING = true (settings.h:7, 14, 27, 38, 46; row g1): two selfing bits above the six meiosis bits, 256 states of which the
// 192 with selfing value 0..2 are valid; only the literal kernel estep_kernel<true> is instantiated for it
constexpr int TYPEBITS = 6, SELFMASK = 3, NUMTYPES_SELFING = 256;
constexpr int NUMPATHS = 128;  // settings.h:32
constexpr int NUMSHIFTS = 8;   // settings.h:35
constexpr int NSLOTS = 7;      // TURNBITS, settings.h:40
constexpr int TILE = 32;       // markers staged in shared memory at a time
constexpr int NTHREADS = NUMSHIFTS * 32;  // one warp per shift mode, two states per lane

constexpr int U_HAPLOS = 1, U_GENOS = 2, U_HOMOZYGOUS = 4, U_GENOSPROBE = 8;  // cnF2freq.cpp:792-795
constexpr int UNKNOWN = 0;    // cnF2freq.cpp:225
constexpr int SEXMARKER = 9;  // cnF2freq.cpp:226

// Per-thread accumulator columns in shared memory (the reference's thread-local haplos / infprobs,
// cnF2freq.cpp:379-380, keyed by focus-tree slot instead of individual id):
//   [0,14)   haplos[slot][f2n]
//   [14,42)  infprobs[slot][realf2n][allele-1]
//   [42,44)  homozyg[allele-1]
constexpr int ACC_HAPLOS = 0, ACC_INF = 14, ACC_HZ = 42, NACC = 44;
constexpr int STATS_PER_CELL = 9;  // haplobase, haplocount, infprobs[2][2], homozyg[2], negshift

// One analysed individual = one focus tree (what fixtrees leaves in reltree / reltreeordered / relmap /
// relmapshift, cnF2freq.cpp:3099-3187), built on the host by cnf2_set_pedigree.
struct Unit
{
	int32_t row[NSLOTS];         // slot -> row of the compact parameter arrays; -1 = null parent pointer
	int32_t relrow[NSLOTS];      // distinct tree members (reltree, empty ones included) -> statistics row
	uint8_t slotrel[NSLOTS];     // slot -> index into relrow
	uint8_t founder[NSLOTS];     // individ::founder of the slot's occupant
	uint8_t relfilter[NSLOTS];   // relmap[rel]: slots the (non-empty) relative occupies; 0 = not in relmap
	uint8_t relshift[NSLOTS];    // relmapshift[rel]
	uint8_t relhalvings[NSLOTS]; // number of reltreeordered slots equal to rel (moveinfprobs, :3584-3588)
	uint8_t nrel;
	uint8_t flag2ignore;
	uint8_t shiftignore;
	uint8_t shiftend;            // NUMSHIFTS, or 2 when gen < 2 (:5359)
	int32_t out;                 // row of the loglik output
	double descendants;
	// dense enumeration of the surviving (g, flag2) pairs through x = flag2 ^ (g << 1) (see cnf2_fast_kernel.cuh)
	uint8_t gfixmask;            // state bits pinned by x: bit j set when orientation bit j+1 is in flag2ignore
	uint8_t nfree;               // number of free state bits
	uint8_t vismask;             // slots trackpossible visits from a non-founder root (pointer chain + founder flags)
	uint8_t pad0;
	uint8_t expand[NUMTYPES];    // j -> j deposited into the free state bits
	int32_t selfgen;             // gen - 2: generations of selfing behind the listed pedigree (cnF2freq.cpp:2094, 2280); SELFING only
	// batched kernel (cnf2_batch_kernel.cuh): x = flag2 ^ (g << 1) values passing rule 2 of ignoreflag2
	// (bit x of xmask2), and the local configurations (3 meiosis bits | 3 orientation bits << 3) of the two
	// parents' subtrees that some surviving (x, g) pair uses (bit cfg of cfgmask[p])
	uint32_t xmask2[4];
	unsigned long long cfgmask[2];
	// turn scoring (updatenegshifts, cnF2freq.cpp:3640-3715): weight of the slot's negshift contribution -- 1 for the
	// root and the two parents (:3665-3671), 1 / children of the parent for its parents and only when gen >= 2
	// (:3673-3686), 0 when the pointer chain to the slot is broken
	double nsscale[NSLOTS];
};

// Shared-memory view handed to the emission routine.
struct Ctx
{
	const uchar2* A;         // [NSLOTS][TILE] allele pair
	const double2* S;        // [NSLOTS][TILE] markersure pair
	const double* HW;        // [NSLOTS][TILE] haploweight
	const int32_t* row;      // [NSLOTS]
	const uint8_t* founder;  // [NSLOTS]
	double* acc;             // this thread's accumulator column; stride NTHREADS
};

// upflagit, cnF2freq.cpp:321-329
template <int GW>
__device__ __forceinline__ int upflagit(int flag, int parnum)
{
	if (flag < 0) return flag;
	return (flag >> (parnum * (GW - 1))) & ((1 << (GW - 1)) - 1);
}

template <int UPD, int GW, bool SELF = false>
__device__ double track(const Ctx& c, int slot, int inval, double second, int t, unsigned flag, int flag99,
                        int localshift, double updateval);

// recursetrackpossible, cnF2freq.cpp:955-1058 (DOIMPOSSIBLE false).  GW is the CALLER's genwidth.
template <int UPD, int GW>
__device__ __forceinline__ double up(const Ctx& c, int slot, int par, int val, double second, int t, int upflag,
                                     int upflag2, int upshift, double updateval)
{
	const int child = (GW == 4) ? 1 + 3 * par : slot + 1 + par;
	if (c.row[child] < 0) return 1.0 + second;  // :1043-1046
	const int f = upflagit<GW>(upflag, par);
	const int f2 = upflagit<GW>(upflag2, par);          // NUMGEN - NUMFLAG2GEN = 0, :985
	const int sh = upflagit<(GW >> 1)>(upshift, par);   // NUMGEN - NUMSHIFTGEN = 1, :986
	return track<UPD, (GW >> 1)>(c, child, val, second, t, (unsigned)f, f2, sh, updateval);
}

// trackpossible<update, zeropropagate = 0>, cnF2freq.cpp:1075-1359: probability of the observed alleles
// of the focus tree below `slot` given inheritance flag, orientation flag99 (-1 = both) and shift mode;
// in the update modes it also scatters updateval into the accumulator columns.
// SELF: the SELFING = true instantiation; the selfing bits of the flag (above TYPEBITS + 1) act in the root generation only
// (:1115-1127): a state with selfing value 1 / 2 sees the root's genotype as one known allele -- the agreed one, or the second
// one of a heterozygote -- on the side the value and the orientation pick, and an unknown one on the other (:1169-1190).
template <int UPD, int GW, bool SELF>
__device__ double track(const Ctx& c, int slot, int inval, double second, int t, unsigned flag, int flag99,
                        int localshift, double updateval)
{
	const bool attopnow = !(UPD & U_HOMOZYGOUS) && (GW == 1 || c.founder[slot]);  // :1120
	uchar2 a = c.A[slot * TILE + t];
	double2 s = c.S[slot * TILE + t];
	const double hw = c.HW[slot * TILE + t];
	const uchar2 a0 = a;
	const double2 s0 = s;
	const int selfval = (SELF && GW == 4) ? (int)((flag >> (TYPEBITS + 1)) & SELFMASK) : 0;  // :1122
	const bool selfingNOW = selfval != 0;                                                       // :1126

	int upflag2 = -1;
	const int upflag = flag >> 1;
	const int upshift = localshift >> 1;
	int f2s = 0, f2end = 2;
	if (flag99 != -1)  // :1141
	{
		upflag2 = flag99 >> 1;
		f2s = flag99 & 1;
		f2end = f2s + 1;
	}
	const int firstpar = flag & 1;  // :1156
	bool allthesame = a.x == a.y;
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
				int first = a0.x;
				bool miss0;  // markermiss<false>(first, second), :303-316
				if (first == UNKNOWN)
				{
					first = a0.y;
					miss0 = false;
				}
				else if (a0.y == UNKNOWN && first != SEXMARKER) miss0 = false;
				else miss0 = first != a0.y;
				int mk;
				double su;
				if (!miss0)
				{
					mk = first;
					su = 1.0 - (1.0 - s0.x) * (1.0 - s0.y);
				}
				else
				{
					mk = a0.y;
					if (s0.x == 0) return 0;  // :1186
					su = 1.0 - s0.x * (1.0 - s0.y);
				}
				a = selfindex ? make_uchar2(UNKNOWN, (unsigned char)mk) : make_uchar2((unsigned char)mk, UNKNOWN);
				s = selfindex ? make_double2(0.0, su) : make_double2(su, 0.0);
				allthesame = a.x == a.y;
			}
		}
		const int Af = f2 ? a.y : a.x, Ao = f2 ? a.x : a.y;
		const double Sf = f2 ? s.y : s.x, So = f2 ? s.x : s.y;
		int val = inval;
		double base, main2 = 0;
		bool miss;  // markermiss<0>, :303-316
		if (val == UNKNOWN)
		{
			val = Af;
			miss = false;
		}
		else if (Af == UNKNOWN && val != SEXMARKER) miss = false;
		else miss = val != Af;

		if (miss)  // :1198-1203
		{
			base = Sf;
			if (Sf != 0 && second != 0) main2 = (1.0 - Sf) * second;
		}
		else  // :1204-1210
		{
			const double eff2 = (inval == UNKNOWN && val != UNKNOWN) ? 1.0 : second;
			base = 1.0 - Sf;
			main2 = (Af == UNKNOWN ? 1.0 : Sf) * eff2;
		}
		if (attopnow || (UPD & 1))  // `update & (GENOS || GENOSPROBE)` is `update & 1`, :1213
		{
			base += main2;
			main2 = 0;
		}
		else if (main2 != 0) main2 /= base;

		bool doupd = true;
		f2n ^= (firstpar ^ localshift) & 1;  // :1227
		if ((allthesame && s.x == s.y) || selfingNOW)  // :1235 (CORRECTIONINFERENCE false inside doit)
		{
			base *= f2n ? 1.0 : 0.0;
			doupd = false;
		}
		else base *= fabs((f2n ? 1.0 : 0.0) - hw);  // :1245

		if constexpr (GW > 1)
		{
			if (base != 0 && !attopnow)  // :1271-1341
			{
				if (!(UPD & U_GENOS))  // :1291
				{
					double sec2 = 0;
					int secmark = Ao;
					if (!(UPD & U_HOMOZYGOUS))
					{
						if (So != 0)  // :1298-1302
						{
							base *= 1.0 - So;
							sec2 = So / (1.0 - So);
						}
					}
					else if (val != secmark)  // :1306-1314
					{
						if (secmark != UNKNOWN) base *= So;
						secmark = val;
					}
					else base *= 1.0 - So;  // :1315-1319
					base *= up<(UPD & ~U_HOMOZYGOUS), GW>(c, slot, !firstpar, secmark, sec2, t, upflag, upflag2, upshift,
					                                      updateval);
				}
				if (base != 0)
					base *= up<(UPD & ~U_HOMOZYGOUS), GW>(c, slot, firstpar, val, main2, t, upflag, upflag2, upshift,
					                                      updateval);
			}
		}

		if (base != 0)  // :1343-1355
		{
			ok += base;
			if ((UPD & U_HAPLOS) && doupd) c.acc[(ACC_HAPLOS + slot * 2 + f2n) * NTHREADS] += updateval;
			if ((UPD & U_GENOS) && (val == 1 || val == 2))
				c.acc[(ACC_INF + (slot * 2 + realf2n) * 2 + (val - 1)) * NTHREADS] += updateval;
		}
	}
	return ok;
}
}  // namespace cnf2

namespace cnf2
{
// The root-level part of trackpossible<UPD, 0> (cnF2freq.cpp:1166-1320) for a root that is not "at the top"
// (genwidth 4, not a founder): everything up to, but not including, the two recursive calls.  base already
// holds the factor of the other allele (:1298-1320); (val, main2) is what recursetrackpossible hands to
// parent `firstpar`, (secmark, sec2) what it hands to the other parent.  incoming second is 0 at the root.
struct RootLocal
{
	double base, main2, sec2;
	int val, secmark;
};

template <int UPD>
__device__ __forceinline__ RootLocal root_local(uchar2 a, double2 s, double hw, int inval, int firstpar, int f2,
                                                int localshift)
{
	RootLocal r;
	const int Af = f2 ? a.y : a.x, Ao = f2 ? a.x : a.y;
	const double Sf = f2 ? s.y : s.x, So = f2 ? s.x : s.y;
	int val = inval;
	bool miss;
	if (val == UNKNOWN)
	{
		val = Af;
		miss = false;
	}
	else if (Af == UNKNOWN && val != SEXMARKER) miss = false;
	else miss = val != Af;
	double base, main2 = 0;
	if (miss) base = Sf;  // incoming second is 0: no main2 (:1200-1201)
	else
	{
		const double eff2 = (inval == UNKNOWN && val != UNKNOWN) ? 1.0 : 0.0;
		base = 1.0 - Sf;
		main2 = (Af == UNKNOWN ? 1.0 : Sf) * eff2;
	}
	if (UPD & 1)
	{
		base += main2;
		main2 = 0;
	}
	else if (main2 != 0) main2 /= base;
	const int f2n = (f2 ^ firstpar ^ localshift) & 1;
	if (a.x == a.y && s.x == s.y) base *= f2n ? 1.0 : 0.0;
	else base *= fabs((f2n ? 1.0 : 0.0) - hw);
	double sec2 = 0;
	int secmark = Ao;
	if (!(UPD & U_GENOS))
	{
		if (!(UPD & U_HOMOZYGOUS))
		{
			if (So != 0)
			{
				base *= 1.0 - So;
				sec2 = So / (1.0 - So);
			}
		}
		else if (val != secmark)
		{
			if (secmark != UNKNOWN) base *= So;
			secmark = val;
		}
		else base *= 1.0 - So;
	}
	r.base = base;
	r.main2 = main2;
	r.sec2 = sec2;
	r.val = val;
	r.secmark = secmark;
	return r;
}
}  // namespace cnf2

namespace cnf2
{
// Per (slot, marker) constants of the emission routine, computed once when a tile is staged so that the
// per-call path of the table-driven kernel has no division: for allele position j of the individual
//   S    markersure                                  (cnF2freq.cpp:878)
//   RM   (1 - S) / S      : main2 / base after a mismatch, per unit of incoming `second`   (:1200-1201, 1220)
//   RN   effS / (1 - S)   : main2 / base after a match, effS = 1 for an unknown allele     (:1205-1209, 1220)
//   SEC  S / (1 - S) or 0 : `secondval` handed up with the other allele                    (:1298-1302)
//   W    phase weight for effective phase bit j: j ? 1 : 0 when homozygous with equal sureness (:1235-1238),
//        |j - haploweight| otherwise (:1245)
struct NodeTab
{
	double S[2], RM[2], RN[2], SEC[2], W[2];
};

__device__ __forceinline__ NodeTab make_nodetab(uchar2 a, double2 s, double hw)
{
	NodeTab n;
	n.S[0] = s.x;
	n.S[1] = s.y;
	n.RM[0] = s.x != 0 ? (1.0 - s.x) / s.x : 0.0;
	n.RM[1] = s.y != 0 ? (1.0 - s.y) / s.y : 0.0;
	n.RN[0] = (a.x == UNKNOWN ? 1.0 : s.x) / (1.0 - s.x);
	n.RN[1] = (a.y == UNKNOWN ? 1.0 : s.y) / (1.0 - s.y);
	n.SEC[0] = s.x != 0 ? s.x / (1.0 - s.x) : 0.0;
	n.SEC[1] = s.y != 0 ? s.y / (1.0 - s.y) : 0.0;
	const bool hom = a.x == a.y && s.x == s.y;
	n.W[0] = hom ? 0.0 : fabs(0.0 - hw);
	n.W[1] = hom ? 1.0 : fabs(1.0 - hw);
	return n;
}

// Shared-memory view of the table-driven kernel.
struct FCtx
{
	const uchar2* A;       // [NSLOTS][TILE]
	const NodeTab* N;      // [NSLOTS][TILE]
	const int32_t* row;    // [NSLOTS]
	const uint8_t* founder;
};

// trackpossible<0, 0> (cnF2freq.cpp:1075-1359) below the root, on the node constants: the value
// recursetrackpossible returns for the subtree rooted at `slot` (genwidth GW = 2 parents, 1 grandparents).
template <int GW>
__device__ __forceinline__ double ftrack(const FCtx& c, int slot, int inval, double second, int t, int flag, int flag99,
                                         int localshift)
{
	const bool attop = GW == 1 || c.founder[slot];
	const uchar2 a = c.A[slot * TILE + t];
	const NodeTab& n = c.N[slot * TILE + t];
	const int firstpar = flag & 1;
	int f2s = 0, f2end = 2;
	if (flag99 != -1)
	{
		f2s = flag99 & 1;
		f2end = f2s + 1;
	}
	double ok = 0;
	for (int f2 = f2s; f2 < f2end; f2++)
	{
		const int Af = f2 ? a.y : a.x;
		const double Sf = n.S[f2];
		int val = inval;
		bool miss;
		if (val == UNKNOWN)
		{
			val = Af;
			miss = false;
		}
		else if (Af == UNKNOWN && val != SEXMARKER) miss = false;
		else miss = val != Af;
		double base, up2;  // up2: the `second` handed to parent firstpar (main2 / base)
		if (miss)
		{
			const bool on = Sf != 0 && second != 0;
			base = attop ? (on ? Sf + (1.0 - Sf) * second : Sf) : Sf;
			up2 = on ? n.RM[f2] * second : 0.0;
		}
		else
		{
			const double eff2 = (inval == UNKNOWN && val != UNKNOWN) ? 1.0 : second;
			base = 1.0 - Sf;
			if (attop) base += (Af == UNKNOWN ? 1.0 : Sf) * eff2;
			up2 = n.RN[f2] * eff2;
		}
		base *= n.W[(f2 ^ firstpar ^ localshift) & 1];
		if constexpr (GW > 1)
		{
			if (!attop && base != 0)
			{
				const int o = f2 ^ 1;
				const int c0 = slot + 1, upflag = flag >> 1, upflag2 = flag99 < 0 ? -1 : flag99 >> 1;
				if (n.S[o] != 0) base *= 1.0 - n.S[o];
				// other allele to parent !firstpar, matched allele to parent firstpar (:1322-1340)
				const int po = firstpar ^ 1;
				const int cho = c0 + po, chf = c0 + firstpar;
				base *= c.row[cho] < 0 ? 1.0 + n.SEC[o]
				                       : ftrack<1>(c, cho, o ? a.y : a.x, n.SEC[o], t, (upflag >> po) & 1,
				                                   upflag2 < 0 ? -1 : (upflag2 >> po) & 1, 0);
				if (base != 0)
					base *= c.row[chf] < 0 ? 1.0 + up2
					                       : ftrack<1>(c, chf, val, up2, t, (upflag >> firstpar) & 1,
					                                   upflag2 < 0 ? -1 : (upflag2 >> firstpar) & 1, 0);
			}
		}
		ok += base;
	}
	return ok;
}

// Root-level factor and messages (see RootLocal) from the node constants; kind: 0 plain emission,
// 1 GENOSPROBE (inval = allele), 2 HOMOZYGOUS (inval = allele).  The incoming `second` is 0 at the root, so
// only the plain emission hands a non-zero `second` (RN) to parent firstpar.
__device__ __forceinline__ RootLocal froot(uchar2 a, const NodeTab& n, int kind, int inval, int firstpar, int f2,
                                           int localshift)
{
	RootLocal r;
	const int Af = f2 ? a.y : a.x, Ao = f2 ? a.x : a.y;
	const double Sf = n.S[f2], So = n.S[f2 ^ 1];
	int val = inval;
	bool miss;
	if (val == UNKNOWN)
	{
		val = Af;
		miss = false;
	}
	else if (Af == UNKNOWN && val != SEXMARKER) miss = false;
	else miss = val != Af;
	double base = miss ? Sf : 1.0 - Sf;
	r.main2 = (!miss && inval == UNKNOWN && val != UNKNOWN) ? n.RN[f2] : 0.0;
	base *= n.W[(f2 ^ firstpar ^ localshift) & 1];
	r.sec2 = 0;
	r.secmark = Ao;
	if (kind != 2)
	{
		if (So != 0) base *= 1.0 - So;
		r.sec2 = n.SEC[f2 ^ 1];
	}
	else if (val != Ao)
	{
		if (Ao != UNKNOWN) base *= So;
		r.secmark = val;
	}
	else base *= 1.0 - So;
	r.base = base;
	r.val = val;
	return r;
}
}  // namespace cnf2
