// The batched E-step kernel (the product path).
//
// Work decomposition: one persistent CTA per SM, one (analysed individual, chromosome) unit at a time.
//   * sweeps: warp = shift mode, lane l holds states l and l + 32 (cnF2freq.cpp:2145-2418).  The emission with
//     all orientations summed (flag2 = -1) factorises as  sum_f R[f][s0] * Q1[f][s2][g>>3] * Q0[f][s1][g&7]
//     where the 64 + 4 table values per marker do not depend on the shift mode except through the single
//     shift bit of the root / the parent (they flip the phase-weight index, :1227).  The tables are built once
//     per marker by the whole CTA, used by all 8 shift modes and by both sweeps (the forward sweep stashes
//     them).  The vectors are kept scaled by exact powers of two instead of being renormalised (:1656-1669),
//     so no log or division is on the sweep path; the log-factors the reference carries are k * ln 2.
//   * posterior (cnF2freq.cpp:5406-5583): the backward sweep runs over a batch of TB markers and leaves
//     W_s[g] = alpha_s[g] * beta_s[g] * exp(F + B - factor) for the 8 shift modes in shared memory.  Then one
//     warp per marker enumerates the surviving (x = flag2 ^ (g << 1), g) pairs densely; each lane handles one
//     pair for ALL 8 shift modes at once: the shift mode only selects which phase weight the root and the
//     two parents contribute (W[x_k ^ s_k]), everything else -- the orientation-fixed subtree values and the
//     GENOSPROBE / HOMOZYGOUS ratios of :5517-5538, tabulated per (marker, parent, local configuration,
//     root orientation) -- is common to the 8 terms, and so are the scatter targets of the GENOS calls.
//   * the accumulator columns are private per lane, reduced by the marker's own warp and transferred into the
//     EM sufficient statistics (:5875-5904) without a block-wide barrier per marker.
//
// ignoreflag2 (:3462-3496): rule 1 pins state bits (dense enumeration), rule 2 is in Unit::xmask2, rule 3 only
// drops terms whose emission is identically zero (the homozygous phase weight f2n ? 1 : 0, :1235-1238); it is
// applied for the relatives whose rule does not depend on the shift mode (relmapshift 0) to shorten the
// enumeration, and left to the zero weights otherwise.
#pragma once
#include "cnf2_estep_kernel.cuh"

namespace cnf2
{
constexpr int TB = 4;    // markers per backward batch (one warp each in the posterior phase)
constexpr int ACCS = TB * 32;  // accumulator columns: one per lane of the posterior warps
constexpr int TF = 32;   // markers per forward tile
constexpr int QS = 68;   // sweep-emission table per marker: Q[p][s_p][g_p][f] (64) + R[s0][f] (4)
constexpr int RC = 16;   // root constants per marker
constexpr int TABW = 5;  // doubles per posterior table entry
constexpr double LN2 = 0.693147180559945309417232121458;
constexpr double EXP_M200_B = 1.3838965267367376e-87;  // exp(-200): unusualstate, cnF2freq.cpp:5414

struct BwdSmem
{
	double acc[NACC * ACCS];
	double Wt[TB * NUMSHIFTS * NUMTYPES];
	double tab[TB * 2 * 2 * 64 * TABW];  // [marker][parent][f][cfg][TABW]
	double qh[TB * QS];
	NodeTab nt[NSLOTS * TB];
	double HW[NSLOTS * TB];
	double tot[TB * NACC];
	double rootc[TB * RC];
	uchar2 A[NSLOTS * TB];
	unsigned char xl[TB * NUMPATHS];
	unsigned char el[TB * 128];
	int nX[TB], nE[TB], homslot[TB];
};

struct FwdSmem
{
	double qh[TF * QS];
	NodeTab nt[NSLOTS * TF];
	uchar2 A[NSLOTS * TF];
};

struct __align__(16) BatchSmem
{
	union
	{
		BwdSmem b;
		FwdSmem f;
	};
	Unit unit;
	double Ls[NUMSHIFTS];
	int32_t unit_index;
};

struct BatchArgs
{
	EstepArgs e;
	int32_t* kexp;           // [n_ctas][Mc][NUMSHIFTS] power-of-two exponent of the stored alpha
	double* qstash;          // [n_ctas][Mc][QS] sweep-emission tables (forward pass -> backward pass)
	int32_t* literal_units;  // units the literal kernel has to do
	unsigned int* literal_count;
};

// markermiss<0> (cnF2freq.cpp:303-316) for an incoming value v against the individual's allele A
__device__ __forceinline__ bool bmiss(int v, int A)
{
	return v != UNKNOWN && !(A == UNKNOWN && v != SEXMARKER) && v != A;
}

// trackpossible at a node that is "at the top" (genwidth 1 or founder, :1120), without the phase weight:
// base + main2 of :1198-1217 for orientation o, incoming (v, sec)
__device__ __forceinline__ double gval(uchar2 a, const NodeTab& n, int o, int v, double sec)
{
	const int A = o ? a.y : a.x;
	const double S = n.S[o];
	if (bmiss(v, A)) return (S != 0 && sec != 0) ? S + (1.0 - S) * sec : S;
	const double eff2 = (v == UNKNOWN && A != UNKNOWN) ? 1.0 : sec;
	return (1.0 - S) + (A == UNKNOWN ? 1.0 : S) * eff2;
}

// the same for a node that recurses (:1198-1220): base, the value and the `second` handed to parent firstpar
struct PMsg
{
	double nb, up2;
	int v2;
};
__device__ __forceinline__ PMsg pmsg(uchar2 a, const NodeTab& n, int o, int v, double sec)
{
	PMsg r;
	const int A = o ? a.y : a.x;
	const double S = n.S[o];
	if (bmiss(v, A))
	{
		r.nb = S;
		r.up2 = (S != 0 && sec != 0) ? n.RM[o] * sec : 0.0;
		r.v2 = v;
	}
	else
	{
		const double eff2 = (v == UNKNOWN && A != UNKNOWN) ? 1.0 : sec;
		r.nb = 1.0 - S;
		r.up2 = n.RN[o] * eff2;
		r.v2 = v == UNKNOWN ? A : v;
	}
	return r;
}

// Scale (p0, p1) of the whole warp by the power of two that brings the largest entry into [1, 2).
__device__ __forceinline__ void rescale(double& p0, double& p1, int& k)
{
	const int e = max(__double2hiint(p0), __double2hiint(p1)) >> 20;  // non-negative doubles order like their high words
	const int w = __reduce_max_sync(0xffffffffu, e);
	if (w == 0) return;  // all zero
	const double sc = __hiloint2double((2046 - w) << 20, 0);
	p0 *= sc;
	p1 *= sc;
	k += w - 1023;
}

// emission of the lane's two states from the marker's sweep table (all orientations summed)
__device__ __forceinline__ void sweep_emission(const double* __restrict__ q, int lane, int shift, double& e0, double& e1)
{
	const int s0 = shift & 1, s1 = (shift >> 1) & 1, s2 = shift >> 2;
	const double2 q0 = *reinterpret_cast<const double2*>(q + ((s1 * 8) + (lane & 7)) * 2);
	const double2 qa = *reinterpret_cast<const double2*>(q + (((2 + s2) * 8) + (lane >> 3)) * 2);
	const double2 qb = *reinterpret_cast<const double2*>(q + (((2 + s2) * 8) + (lane >> 3) + 4) * 2);
	const double2 R = *reinterpret_cast<const double2*>(q + 64 + s0 * 2);
	const double t0 = R.x * q0.x, t1 = R.y * q0.y;
	e0 = t0 * qa.x + t1 * qa.y;
	e1 = t0 * qb.x + t1 * qb.y;
}

// Stages markers [m0, m0 + n) of the unit's 7 slots: alleles, node constants, optionally the haploweight.
// Returns (to the calling thread) whether it saw a markersure outside [0,1) or a haploweight outside [0,1].
template <int STRIDE>
__device__ __forceinline__ int stage_markers(const Unit& u, const EstepArgs& a, int m0, int n, int tid, uchar2* A, NodeTab* N,
                                             double* HW)
{
	int bad = 0;
	for (int i = tid; i < NSLOTS * STRIDE; i += NTHREADS)
	{
		const int slot = i / STRIDE, t = i % STRIDE;
		const int row = u.row[slot];
		uchar2 al = make_uchar2(0, 0);
		double2 su = make_double2(0, 0);
		double h = 0;
		if (row >= 0 && t < n)
		{
			const size_t o = (size_t)row * a.ld + m0 + t;
			al = a.alleles[o];
			su = a.sure[o];
			h = a.hw[o];
			if (!(su.x >= 0 && su.x < 1 && su.y >= 0 && su.y < 1 && h >= 0 && h <= 1)) bad = 1;
		}
		A[i] = al;
		N[i] = make_nodetab(al, su, h);
		if (HW) HW[i] = h;
	}
	return bad;
}

// value of a grandparent slot k (or 1 + sec when the pointer is null, :1043-1046) with its orientation summed
template <int STRIDE>
__device__ __forceinline__ double ghat(const Unit& u, const uchar2* A, const NodeTab* N, int k, int t, int v, double sec, int mk)
{
	if (u.row[k] < 0) return 1.0 + sec;
	const uchar2 a = A[k * STRIDE + t];
	const NodeTab& n = N[k * STRIDE + t];
	return gval(a, n, 0, v, sec) * n.W[mk] + gval(a, n, 1, v, sec) * n.W[1 ^ mk];
}

// One family (parent p, root orientation f, parent meiosis bit m) of the sweep-emission table of marker t:
// the 8 entries over the grandparents' meiosis bits and the parent's shift bit.
template <int STRIDE>
__device__ __forceinline__ void sweep_family(const Unit& u, const uchar2* A, const NodeTab* N, int t, int fam, double* __restrict__ q)
{
	const int p = fam >> 2, f = (fam >> 1) & 1, m = fam & 1;
	const uchar2 a0 = A[t];
	const NodeTab& n0 = N[t];
	int v;
	double sec;
	if (p == 0)
	{
		v = f ? a0.y : a0.x;
		sec = v != UNKNOWN ? n0.RN[f] : 0.0;  // main2 / base of the root's matched allele (:1205-1220)
	}
	else
	{
		v = f ? a0.x : a0.y;
		sec = n0.SEC[f ^ 1];  // :1298-1302
	}
	if (fam < 4)  // R[s0 = fam >> 1][f = fam & 1]: root factor without the subtrees
	{
		const int s0 = fam >> 1, ff = fam & 1;
		double rb = 1.0 - n0.S[ff];
		if (n0.S[ff ^ 1] != 0) rb *= 1.0 - n0.S[ff ^ 1];
		q[64 + s0 * 2 + ff] = rb * n0.W[ff ^ s0];
	}
	const int sp = 1 + 3 * p;
	double out[2][4];  // [s_p][ma | mb << 1]
	if (u.row[sp] < 0)
	{
#pragma unroll
		for (int j = 0; j < 4; j++) out[0][j] = out[1][j] = 1.0 + sec;
	}
	else
	{
		const uchar2 ap = A[sp * STRIDE + t];
		const NodeTab& np = N[sp * STRIDE + t];
		if (u.founder[sp])
		{
			const double g0 = gval(ap, np, 0, v, sec), g1 = gval(ap, np, 1, v, sec);
#pragma unroll
			for (int s = 0; s < 2; s++)
			{
				const double val = g0 * np.W[m ^ s] + g1 * np.W[1 ^ m ^ s];
#pragma unroll
				for (int j = 0; j < 4; j++) out[s][j] = val;
			}
		}
		else
		{
			const int kf = sp + 1 + m, ko = sp + 1 + (m ^ 1);
			double Gf[2][2], Go[2][2], nb[2];
#pragma unroll
			for (int o = 0; o < 2; o++)
			{
				const PMsg pm = pmsg(ap, np, o, v, sec);
				nb[o] = pm.nb;
				if (np.S[o ^ 1] != 0) nb[o] *= 1.0 - np.S[o ^ 1];
				const int vo = o ? ap.x : ap.y;
#pragma unroll
				for (int mk = 0; mk < 2; mk++)
				{
					Gf[o][mk] = ghat<STRIDE>(u, A, N, kf, t, pm.v2, pm.up2, mk);
					Go[o][mk] = ghat<STRIDE>(u, A, N, ko, t, vo, np.SEC[o ^ 1], mk);
				}
			}
#pragma unroll
			for (int j = 0; j < 4; j++)
			{
				const int ma = j & 1, mb = j >> 1;
				const int mkf = m ? mb : ma, mko = m ? ma : mb;
				const double t0 = nb[0] * Go[0][mko] * Gf[0][mkf], t1 = nb[1] * Go[1][mko] * Gf[1][mkf];
#pragma unroll
				for (int s = 0; s < 2; s++) out[s][j] = t0 * np.W[m ^ s] + t1 * np.W[1 ^ m ^ s];
			}
		}
	}
#pragma unroll
	for (int s = 0; s < 2; s++)
#pragma unroll
		for (int j = 0; j < 4; j++) q[(((p * 2 + s) * 8) + (m | (j << 1))) * 2 + f] = out[s][j];
}

// value of grandparent slot k at a fixed orientation o, phase-weight index w (or 1 + sec for a null pointer)
__device__ __forceinline__ double gfix(const Unit& u, const uchar2* A, const NodeTab* N, int k, int t, int v, double sec, int o, int w)
{
	if (u.row[k] < 0) return 1.0 + sec;
	const NodeTab& n = N[k * TB + t];
	return gval(A[k * TB + t], n, o, v, sec) * n.W[w];
}

// Root constants of marker t for the posterior phase (see RC): rbase[f], W[w], mm[al][f], hfd[al][f].
__device__ __forceinline__ void root_constants(const uchar2 a0, const NodeTab& n0, int f, double* __restrict__ rc)
{
	const int Af = f ? a0.y : a0.x, Ao = f ? a0.x : a0.y;
	const double Sf = n0.S[f], So = n0.S[f ^ 1];
	const double d = So != 0 ? 1.0 - So : 1.0;  // :1298-1302
	rc[f] = (1.0 - Sf) * d;                      // plain emission, root level, without the phase weight
	rc[2 + f] = n0.W[f];
#pragma unroll
	for (int al = 0; al < 2; al++)
	{
		const int v = al + 1;
		rc[4 + al * 2 + f] = bmiss(v, Af) ? Sf : 1.0 - Sf;                                  // GENOSPROBE root factor (:1198-1217)
		const double hf = (v != Ao) ? (Ao != UNKNOWN ? So : 1.0) : 1.0 - So;                // HOMOZYGOUS, :1306-1319
		rc[8 + al * 2 + f] = hf / d;
	}
}

// One entry of the posterior tables of marker t: parent p's subtree at local configuration cfg, for both root
// orientations.  Q(v, sec) is the value recursetrackpossible returns for the subtree WITHOUT the parent's own
// phase weight (which carries the shift mode).  Entry: [0] Q for the plain emission's message,
// parent 0: [1..2] rho0[al] = share of allele al in the side-0 GENOSPROBE pair, [3..4] rho0[al] * hfd[al][f];
// parent 1: [1..2] rho1[al] (side 1), [3..4] Q(al, 0) / Q(plain).
__device__ __forceinline__ void post_entry(const Unit& u, const uchar2* A, const NodeTab* N, const double* __restrict__ rc, int t, int p,
                                           int cfg, double* __restrict__ tab)
{
	const int m = cfg & 1, ma = (cfg >> 1) & 1, mb = (cfg >> 2) & 1, o = (cfg >> 3) & 1, oa = (cfg >> 4) & 1, ob = (cfg >> 5) & 1;
	const int sp = 1 + 3 * p;
	const uchar2 a0 = A[t];
	const NodeTab& n0 = N[t];
	const bool absent = u.row[sp] < 0, attop = !absent && u.founder[sp];
	const uchar2 ap = A[sp * TB + t];
	const NodeTab& np = N[sp * TB + t];
	const int kf = sp + 1 + m, of = m ? ob : oa, wf = of ^ (m ? mb : ma);
	double pre = 1.0;
	if (!absent && !attop)
	{
		const int ko = sp + 1 + (m ^ 1), oo = m ? oa : ob, wo = oo ^ (m ? ma : mb);
		if (np.S[o ^ 1] != 0) pre = 1.0 - np.S[o ^ 1];
		pre *= gfix(u, A, N, ko, t, o ? ap.x : ap.y, np.SEC[o ^ 1], oo, wo);
	}
	auto Q = [&](int v, double sec) -> double {
		if (absent) return 1.0 + sec;
		if (attop) return gval(ap, np, o, v, sec);
		const PMsg pm = pmsg(ap, np, o, v, sec);
		return pm.nb * pre * gfix(u, A, N, kf, t, pm.v2, pm.up2, of, wf);
	};
	const double P1 = Q(1, 0.0), P2 = Q(2, 0.0);
#pragma unroll
	for (int f = 0; f < 2; f++)
	{
		double* e = tab + ((size_t)((p * 2 + f) * 64 + cfg)) * TABW;
		if (p == 0)
		{
			const int v = f ? a0.y : a0.x;
			e[0] = Q(v, v != UNKNOWN ? n0.RN[f] : 0.0);
			const double A1 = rc[4 + f] * P1, A2 = rc[6 + f] * P2;
			const double sum = A1 + A2, inv = sum > 0 ? 1.0 / sum : 0.0;
			e[1] = A1 * inv;
			e[2] = A2 * inv;
			e[3] = A1 * inv * rc[8 + f];
			e[4] = A2 * inv * rc[10 + f];
		}
		else
		{
			const int v = f ? a0.x : a0.y;
			const double E = Q(v, n0.SEC[f ^ 1]);
			e[0] = E;
			const double A1 = rc[4 + (f ^ 1)] * P1, A2 = rc[6 + (f ^ 1)] * P2;  // side 1 probes the root at orientation f ^ 1
			const double sum = A1 + A2, inv = sum > 0 ? 1.0 / sum : 0.0, invE = E > 0 ? 1.0 / E : 0.0;
			e[1] = A1 * inv;
			e[2] = A2 * inv;
			e[3] = P1 * invE;
			e[4] = P2 * invE;
		}
	}
}

// Posterior of one marker by one warp: dense enumeration of the (x, g) pairs, 8 shift modes per pair.
__device__ __forceinline__ void batch_posterior(BwdSmem& b, const Unit& u, int t, int lane, double* __restrict__ acc)
{
	const double* rc = b.rootc + t * RC;
	const double* Wt = b.Wt + t * (NUMSHIFTS * NUMTYPES);
	const double* tab = b.tab + (size_t)t * (2 * 2 * 64 * TABW);
	const unsigned char* xl = b.xl + t * NUMPATHS;
	const int nX = b.nX[t], nfree = u.nfree, gfix_ = u.gfixmask;
	const int T = nX << nfree;
	const int vis = u.vismask, hmask = vis & ~b.homslot[t];
	double wp0[2] = { 1.0, 1.0 }, wp1[2] = { 1.0, 1.0 };
	if (u.row[1] >= 0) wp0[0] = b.nt[1 * TB + t].W[0], wp0[1] = b.nt[1 * TB + t].W[1];
	if (u.row[4] >= 0) wp1[0] = b.nt[4 * TB + t].W[0], wp1[1] = b.nt[4 * TB + t].W[1];
	for (int i = lane; i < T; i += 32)
	{
		const int x = xl[i >> nfree];
		const int g = ((x >> 1) & gfix_) | u.expand[i & ((1 << nfree) - 1)];
		const int flag2 = x ^ (g << 1);
		const int f = x & 1, x1 = (x >> 1) & 1, x4 = (x >> 4) & 1;
		const int cfg0 = (g & 7) | (((flag2 >> 1) & 7) << 3), cfg1 = ((g >> 3) & 7) | (((flag2 >> 4) & 7) << 3);
		const double* T0 = tab + ((0 * 2 + f) * 64 + cfg0) * TABW;
		const double* T1 = tab + ((1 * 2 + f) * 64 + cfg1) * TABW;
		const double ep = rc[f] * T0[0] * T1[0];
		if (!(ep > 0)) continue;
		// the 8 shift modes: s = s0 | s1 << 1 | s2 << 2 flips the phase-weight index of root / parent 0 / parent 1
		const double r0 = ep * rc[2 + f], r1 = ep * rc[2 + (f ^ 1)];              // s0 = 0, 1
		const double c00 = wp0[x1] * wp1[x4], c10 = wp0[x1 ^ 1] * wp1[x4];       // [s1][s2]
		const double c01 = wp0[x1] * wp1[x4 ^ 1], c11 = wp0[x1 ^ 1] * wp1[x4 ^ 1];
		double ts[8];
		ts[0] = r0 * c00 * Wt[0 * NUMTYPES + g];
		ts[1] = r1 * c00 * Wt[1 * NUMTYPES + g];
		ts[2] = r0 * c10 * Wt[2 * NUMTYPES + g];
		ts[3] = r1 * c10 * Wt[3 * NUMTYPES + g];
		ts[4] = r0 * c01 * Wt[4 * NUMTYPES + g];
		ts[5] = r1 * c01 * Wt[5 * NUMTYPES + g];
		ts[6] = r0 * c11 * Wt[6 * NUMTYPES + g];
		ts[7] = r1 * c11 * Wt[7 * NUMTYPES + g];
#pragma unroll
		for (int s = 0; s < 8; s++) ts[s] = ts[s] > EXP_M200_B ? ts[s] : 0.0;  // unusualstate, :5414, 5502
		const double p01 = ts[0] + ts[1], p23 = ts[2] + ts[3], p45 = ts[4] + ts[5], p67 = ts[6] + ts[7];
		const double V = (p01 + p23) + (p45 + p67);
		if (!(V > 0)) continue;
		// updatehaplo (:1561-1575): visited, non-homozygous slots at their effective phase bit x_k ^ s_k
		if (hmask & 1)
		{
			acc[(ACC_HAPLOS + f) * ACCS] += (ts[0] + ts[2]) + (ts[4] + ts[6]);        // s0 = 0: bit f
			acc[(ACC_HAPLOS + (f ^ 1)) * ACCS] += (ts[1] + ts[3]) + (ts[5] + ts[7]);  // s0 = 1
		}
		if (hmask & 2)
		{
			acc[(ACC_HAPLOS + 2 + x1) * ACCS] += p01 + p45;
			acc[(ACC_HAPLOS + 2 + (x1 ^ 1)) * ACCS] += p23 + p67;
		}
		if (hmask & 16)
		{
			acc[(ACC_HAPLOS + 8 + x4) * ACCS] += p01 + p23;
			acc[(ACC_HAPLOS + 8 + (x4 ^ 1)) * ACCS] += p45 + p67;
		}
#pragma unroll
		for (int k = 2; k < NSLOTS; k++)
			if (k != 4 && ((hmask >> k) & 1)) acc[(ACC_HAPLOS + k * 2 + ((x >> k) & 1)) * ACCS] += V;
		// the four GENOS calls (:5560-5568): root, parent `side`, that parent's parent picked by its meiosis bit
#pragma unroll
		for (int side = 0; side < 2; side++)
		{
			const double* Ts = side ? T1 : T0;
			const double u0 = V * Ts[1], u1 = V * Ts[2];
			const int sp = 1 + 3 * side, gs = sp + 1 + ((g >> (3 * side)) & 1);
			double* a0 = acc + (ACC_INF + (f ^ side) * 2) * ACCS;
			a0[0] += u0;
			a0[ACCS] += u1;
			if ((vis >> sp) & 1)
			{
				double* ap = acc + (ACC_INF + (sp * 2 + ((flag2 >> sp) & 1)) * 2) * ACCS;
				ap[0] += u0;
				ap[ACCS] += u1;
				if ((vis >> gs) & 1)
				{
					double* ag = acc + (ACC_INF + (gs * 2 + ((flag2 >> gs) & 1)) * 2) * ACCS;
					ag[0] += u0;
					ag[ACCS] += u1;
				}
			}
		}
		acc[(ACC_HZ + 0) * ACCS] += V * T0[3] * T1[3];  // :5571-5575
		acc[(ACC_HZ + 1) * ACCS] += V * T0[4] * T1[4];
	}
}

// Per-marker lists of the posterior phase, built by the marker's warp: valid x values, needed table entries,
// root constants; then the table entries themselves.
__device__ __forceinline__ void batch_prepare(BwdSmem& b, const Unit& u, int t, int lane)
{
	const unsigned FULL = 0xffffffffu;
	int homslot = 0, hom = 0;
	for (int k = 0; k < NSLOTS; k++)
	{
		if (u.row[k] < 0) continue;
		const uchar2 al = b.A[k * TB + t];
		const NodeTab& n = b.nt[k * TB + t];
		if (al.x == al.y && n.S[0] == n.S[1])  // :1235, :3488
		{
			homslot |= 1 << k;
			hom |= 1 << u.slotrel[k];
		}
	}
	int gpmust = 0;  // rule 3 (:3488-3493) of the relatives whose rule does not involve the shift mode
	for (int i = 0; i < u.nrel; i++)
		if (((hom >> i) & 1) && u.relshift[i] == 0) gpmust |= u.relfilter[i];
	int nX = 0, nE = 0;
#pragma unroll
	for (int r4 = 0; r4 < 4; r4++)
	{
		const int x = r4 * 32 + lane;
		const bool ok = ((u.xmask2[r4] >> lane) & 1) && (x & gpmust) == gpmust;
		const unsigned mask = __ballot_sync(FULL, ok);
		if (ok) b.xl[t * NUMPATHS + nX + __popc(mask & ((1u << lane) - 1u))] = (unsigned char)x;
		nX += __popc(mask);
	}
#pragma unroll
	for (int r4 = 0; r4 < 4; r4++)
	{
		const int c = r4 * 32 + lane, p = c >> 6, cfg = c & 63;
		const int sub = (gpmust >> (1 + 3 * p)) & 7, xc = (cfg & 7) ^ (cfg >> 3);
		const bool ok = ((u.cfgmask[p] >> cfg) & 1ull) && (xc & sub) == sub;
		const unsigned mask = __ballot_sync(FULL, ok);
		if (ok) b.el[t * 128 + nE + __popc(mask & ((1u << lane) - 1u))] = (unsigned char)c;
		nE += __popc(mask);
	}
	if (lane < 2) root_constants(b.A[t], b.nt[t], lane, b.rootc + t * RC);
	if (lane == 0)
	{
		b.nX[t] = nX;
		b.nE[t] = nE;
		b.homslot[t] = homslot;
	}
	__syncwarp();
	for (int i = lane; i < nE; i += 32)
	{
		const int c = b.el[t * 128 + i];
		post_entry(u, b.A, b.nt, b.rootc + t * RC, t, c >> 6, c & 63, b.tab + (size_t)t * (2 * 2 * 64 * TABW));
	}
	__syncwarp();
}

__global__ void __launch_bounds__(NTHREADS, 2) batch_estep_kernel(BatchArgs ba)
{
	extern __shared__ __align__(16) unsigned char smem_raw[];
	BatchSmem& sm = *reinterpret_cast<BatchSmem*>(smem_raw);
	BwdSmem& b = sm.b;
	const EstepArgs& a = ba.e;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, shift = warp;
	const int Mc = a.Mc, E = Mc - 1;
	double* const alpha = a.alpha + (size_t)blockIdx.x * Mc * (NUMSHIFTS * NUMTYPES);
	int32_t* const kexp = ba.kexp + (size_t)blockIdx.x * Mc * NUMSHIFTS;
	double* const qs = ba.qstash + (size_t)blockIdx.x * Mc * QS;
	constexpr int FWD_DOUBLES = (int)((sizeof(FwdSmem) + 7) / 8);

	for (int i = tid; i < NACC * ACCS; i += NTHREADS) b.acc[i] = 0;

	for (;;)
	{
		__syncthreads();
		if (tid == 0) sm.unit_index = (int)atomicAdd(a.counter, 1u);
		__syncthreads();
		const int ui = sm.unit_index;
		if (ui >= a.n_units) break;
		{
			const int* src = reinterpret_cast<const int*>(a.units + ui);
			int* dst = reinterpret_cast<int*>(&sm.unit);
			for (int i = tid; i < (int)(sizeof(Unit) / 4); i += NTHREADS) dst[i] = src[i];
		}
		__syncthreads();
		const Unit& u = sm.unit;
		bool literal = u.founder[0] != 0;  // a root "at the top" never recurses (:1120): literal kernel
		const bool active = shift < u.shiftend && !(shift & u.shiftignore);

		// ---- forward sweep (initfwbw + realanalyze<STORE|FORWARD>, cnF2freq.cpp:2086-2107) ----
		double p0 = 1.0 / NUMTYPES, p1 = 1.0 / NUMTYPES;  // EVENGEN, :2100-2104
		int ka = 0;
		for (int m0 = 0; m0 < Mc && !literal; m0 += TF)
		{
			const int n = min(TF, Mc - m0);
			__syncthreads();
			const int bad = stage_markers<TF>(u, a, m0, n, tid, sm.f.A, sm.f.nt, nullptr);
			if (__syncthreads_or(bad))
			{
				literal = true;
				break;
			}
			if ((tid >> 3) < n) sweep_family<TF>(u, sm.f.A, sm.f.nt, tid >> 3, tid & 7, sm.f.qh + (tid >> 3) * QS);
			__syncthreads();
			for (int i = tid; i < n * QS; i += NTHREADS) qs[(size_t)m0 * QS + i] = sm.f.qh[i];
			if (!active) continue;
			for (int t = 0; t < n; t++)
			{
				const int m = m0 + t;
				double* dst = alpha + ((size_t)m * NUMSHIFTS + shift) * NUMTYPES;
				dst[lane] = p0;
				dst[lane + 32] = p1;
				if (lane == 0) kexp[(size_t)m * NUMSHIFTS + shift] = ka;
				double e0, e1;
				sweep_emission(sm.f.qh + t * QS, lane, shift, e0, e1);
				p0 *= e0;
				p1 *= e1;
				rescale(p0, p1, ka);
				if (m < E) transition(p0, p1, a.rec0[m], a.rec1[m]);
			}
		}
		__syncthreads();
		for (int i = tid; i < min(FWD_DOUBLES, NACC * ACCS); i += NTHREADS) b.acc[i] = 0;  // the forward tile aliases the accumulators
		if (literal)
		{
			if (tid == 0) ba.literal_units[atomicAdd(ba.literal_count, 1u)] = ui;
			continue;
		}
		double L = 0;
		if (active)
		{
			// doanalyze / quickanalyze with NONESTOP (:1984-2018): log-factor of alpha+ at the last marker
			const double sum = warp_sum(p0 + p1);
			L = sum > 0 ? (double)ka * LN2 + log(sum) : MINFACTOR_D;
			if (lane == 0) sm.Ls[shift] = L;
		}
		__syncthreads();
		double factor = -1e15, realfactor = 0;  // :5373-5403
		for (int s = 0; s < u.shiftend; s++) factor = fmax(factor, (s & u.shiftignore) ? -1e30 : sm.Ls[s]);
		for (int s = 0; s < u.shiftend; s++)
			if (!(s & u.shiftignore)) realfactor += exp(sm.Ls[s] - factor);
		const double maxfactor = factor;
		factor += log(realfactor);
		if (tid == 0)
		{
			a.loglik[(size_t)u.out * 2 + 0] = maxfactor;
			a.loglik[(size_t)u.out * 2 + 1] = factor;
		}
		if (isnan(factor) || factor < MINFACTOR_D) continue;  // :5403
		const bool contributing = active && !(factor - L > 40);  // :5421
		const double minfactor = (double)(float)(-200.0 + factor);  // unusualstate + factor as a float, :1936
		// exp(k ln 2 - factor) = 2^(k + kf) * exp(frem) with -factor = kf ln 2 + frem: one exp per unit
		const int kf = (int)rint(-factor / LN2);
		const double efrem = exp(-factor - (double)kf * LN2);

		// ---- backward sweep (:2109-2116) in batches of TB markers, each followed by its posterior phase ----
		double b0 = 1.0, b1 = 1.0;  // :2111-2114
		int kb = 0;
		for (int m0 = ((Mc - 1) / TB) * TB; m0 >= 0; m0 -= TB)
		{
			const int n = min(TB, Mc - m0);
			__syncthreads();
			stage_markers<TB>(u, a, m0, n, tid, b.A, b.nt, b.HW);
			for (int i = tid; i < n * QS; i += NTHREADS) b.qh[i] = qs[(size_t)m0 * QS + i];
			double a0r[TB], a1r[TB];
			int kar[TB];
#pragma unroll
			for (int t = 0; t < TB; t++)
			{
				a0r[t] = a1r[t] = 0;
				kar[t] = 0;
				if (contributing && t < n)
				{
					const double* src = alpha + ((size_t)(m0 + t) * NUMSHIFTS + shift) * NUMTYPES;
					a0r[t] = src[lane];
					a1r[t] = src[lane + 32];
					kar[t] = kexp[(size_t)(m0 + t) * NUMSHIFTS + shift];
				}
			}
			__syncthreads();
			if (warp < n) batch_prepare(b, u, warp, lane);
#pragma unroll
			for (int t = TB - 1; t >= 0; t--)
			{
				if (t >= n) continue;
				const int m = m0 + t;
				double w0 = 0, w1 = 0;
				if (contributing)
				{
					// val = exp(F- + log(alpha e) + B + log beta - factor) (:1963-2018, :5499-5508); the factors are k ln 2 here
					const int ke = kar[t] + kb + kf;
					if (!((double)kar[t] * LN2 + 5.0 < minfactor) && ke > -860 && ke < 860)  // |F + B - factor| < 600 in the literal kernel
					{
						const double sc = __hiloint2double((1023 + ke) << 20, 0) * efrem;
						w0 = a0r[t] * b0 * sc;
						w1 = a1r[t] * b1 * sc;
					}
				}
				b.Wt[(t * NUMSHIFTS + shift) * NUMTYPES + lane] = w0;
				b.Wt[(t * NUMSHIFTS + shift) * NUMTYPES + lane + 32] = w1;
				if (active && m > 0)
				{
					double e0, e1;
					sweep_emission(b.qh + t * QS, lane, shift, e0, e1);
					b0 *= e0;
					b1 *= e1;
					rescale(b0, b1, kb);
					transition(b0, b1, a.rec0[m - 1], a.rec1[m - 1]);
				}
			}
			__syncthreads();
			if (warp < n)
			{
				const int t = warp;
				batch_posterior(b, u, t, lane, b.acc + tid);
				__syncwarp();
				// the warp reduces its own 32 accumulator columns; the reader clears what it read
				for (int k = lane; k < NACC; k += 32)
				{
					double* col = b.acc + k * ACCS + warp * 32;
					double v = 0;
#pragma unroll 8
					for (int j = 0; j < 32; j++)
					{
						double* p = col + ((j + lane) & 31);
						v += *p;
						*p = 0;
					}
					b.tot[warp * NACC + k] = v;
				}
				__syncwarp();
				// transfer into the EM sufficient statistics (:5875-5904, moveinfprobs :3577-3597, movehaplos :3599-3616)
				if (lane < u.nrel) transfer_stats(a, u, b.tot + warp * NACC, b.HW, lane, t, m0 + t, TB);
			}
		}
	}
}
}  // namespace cnf2
