// Emission tables of the E-step kernel (shared device helpers).
//
// With all orientations summed (flag2 = -1, the sweeps) trackpossible (cnF2freq.cpp:1075-1359) factorises as
//   e(g; shift) = sum_f R[f][s0] * Q1[f][s2][g >> 3] * Q0[f][s1][g & 7]
// where f is the root's orientation and Q_p the value recursetrackpossible (:955-1058) returns for parent p's
// subtree given what the root hands it.  The 64 + 4 table values per marker depend on the shift mode only through
// the single shift bit of the root / the parent (it flips the phase-weight index, :1227), so one table serves all
// 8 shift modes and both sweeps.  With a fixed orientation (the posterior loop, :5406-5583) a subtree value
// depends on the subtree's local configuration (3 meiosis bits | 3 orientation bits << 3); those tables are
// built per marker for the configurations in use (warp_tables_from_factors in cnf2_warp_kernel.cuh).
//
// Validity: markersure in [0,1) and haploweight in [0,1] (then every factor is finite and non-negative and the
// `if (base)` gates of :1271-1355 only skip zero products); other units go to the literal kernel.
#pragma once
#include "cnf2_estep_kernel.cuh"

namespace cnf2
{
constexpr int QS = 68;   // sweep-emission table per marker: Q[p][s_p][g_p][f] (64) + R[s0][f] (4)
constexpr int RC = 16;   // root constants per marker
constexpr double LN2 = 0.693147180559945309417232121458;
// Reciprocal for the per-marker tables: rcp.approx (>= 20 bits) + two Newton steps, ~1 ulp for normal x > 0.
// x = 0 gives NaN instead of inf; every caller either guards x != 0 or only multiplies a value that is then NaN in
// the reference too (0 * inf).  The parity tolerance is 1e-9 relative; an IEEE division costs ~4x the instructions
// on the dependent chain of a latency-bound warp.
__device__ __forceinline__ double frcp(double x)
{
	double r;
	asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
	r = fma(fma(-x, r, 1.0), r, r);
	r = fma(fma(-x, r, 1.0), r, r);
	return r;
}
// log(y) for an ordinary positive y (no zero, subnormal, inf or NaN: the solver only takes logs of (1 - x) / x with x in
// [eps, 1 - eps]).  The usual reduction y = 2^e m, m in [sqrt(1/2), sqrt(2)), log m = 2 atanh((m - 1) / (m + 1)) as an
// odd series in u = 2 (m - 1) / (m + 1) with the division's residual added back; <= 1.2 ulp (checked against long
// double over 2^-23..2^23 and around 1).  The library log costs three times the instructions in this loop: its
// coefficients are 64-bit immediates (two moves each on sm_100) and it carries the special-case branches.  The series
// coefficients 1 / ((2k + 1) 4^k) sit in constant memory and are read as instruction operands.
__constant__ double MS_LC[10] = { 1.0 / 12,     1.0 / 80,     1.0 / 448,     1.0 / 2304,    1.0 / 11264,
	                              1.0 / 53248,  1.0 / 245760, 1.0 / 1114112, 1.0 / 4980736, 1.0 / 22020096 };
__constant__ double MS_LN2[2] = { 6.93147180369123816490e-01, 1.90821492927058770002e-10 };  // hi (21 trailing zero bits), lo
__device__ __forceinline__ double ms_log(double y)
{
	int hi = __double2hiint(y);
	const int lo = __double2loint(y);
	int e = (hi >> 20) - 1023;
	hi = (hi & 0x000fffff) | 0x3ff00000;
	if (hi >= 0x3ff6a09f) hi -= 0x00100000, e += 1;
	const double m = __hiloint2double(hi, lo);
	const double a = m - 1.0, r = frcp(m + 1.0);
	const double h = a * r, u = h + h;
	double res = a - u;
	res += res;
	res = fma(a, -u, res);  // 2 a - u (2 + a): what the rounded quotient leaves
	const double c = r * res;
	const double sq = u * u;
	double p = MS_LC[9];
#pragma unroll
	for (int k = 8; k >= 0; k--) p = fma(p, sq, MS_LC[k]);
	const double t = fma(u * sq, p, c);
	const double ed = __hiloint2double(0x43300000, e ^ 0x80000000) - 4503601774854144.0;  // (double)e
	return fma(ed, MS_LN2[0], u) + fma(ed, MS_LN2[1], t);
}
// NodeTab (cnf2_device.cuh) with two reciprocals per allele position instead of three divisions
__device__ __forceinline__ NodeTab make_nodetab_fast(uchar2 a, double2 s, double hw)
{
	NodeTab n;
	n.S[0] = s.x;
	n.S[1] = s.y;
	const double ix = s.x != 0 ? frcp(s.x) : 0.0, iy = s.y != 0 ? frcp(s.y) : 0.0;
	const double jx = frcp(1.0 - s.x), jy = frcp(1.0 - s.y);
	n.RM[0] = (1.0 - s.x) * ix;
	n.RM[1] = (1.0 - s.y) * iy;
	n.RN[0] = (a.x == UNKNOWN ? 1.0 : s.x) * jx;
	n.RN[1] = (a.y == UNKNOWN ? 1.0 : s.y) * jy;
	n.SEC[0] = s.x * jx;
	n.SEC[1] = s.y * jy;
	const bool hom = a.x == a.y && s.x == s.y;
	n.W[0] = hom ? 0.0 : fabs(0.0 - hw);
	n.W[1] = hom ? 1.0 : fabs(1.0 - hw);
	return n;
}

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
				// selects, not indexed reads: a runtime index would put the four 2 x 2 arrays into local memory
				const double go0 = mko ? Go[0][1] : Go[0][0], go1 = mko ? Go[1][1] : Go[1][0];
				const double gf0 = mkf ? Gf[0][1] : Gf[0][0], gf1 = mkf ? Gf[1][1] : Gf[1][0];
				const double t0 = nb[0] * go0 * gf0, t1 = nb[1] * go1 * gf1;
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
template <int STRIDE>
__device__ __forceinline__ double gfix(const Unit& u, const uchar2* A, const NodeTab* N, int k, int t, int v, double sec, int o, int w)
{
	if (u.row[k] < 0) return 1.0 + sec;
	const NodeTab& n = N[k * STRIDE + t];
	return gval(A[k * STRIDE + t], n, o, v, sec) * n.W[w];
}

// Root constants of marker t for the posterior phase (see RC): rbase[f], W[w], mm[al][f], hfd[al][f].
__device__ __forceinline__ void root_constants(const uchar2 a0, const NodeTab& n0, int f, double* __restrict__ rc)
{
	const int Af = f ? a0.y : a0.x, Ao = f ? a0.x : a0.y;
	const double Sf = n0.S[f], So = n0.S[f ^ 1];
	const double d = So != 0 ? 1.0 - So : 1.0;  // :1298-1302
	const double invd = frcp(d);
	rc[f] = (1.0 - Sf) * d;                      // plain emission, root level, without the phase weight
	rc[2 + f] = n0.W[f];
#pragma unroll
	for (int al = 0; al < 2; al++)
	{
		const int v = al + 1;
		rc[4 + al * 2 + f] = bmiss(v, Af) ? Sf : 1.0 - Sf;                                  // GENOSPROBE root factor (:1198-1217)
		const double hf = (v != Ao) ? (Ao != UNKNOWN ? So : 1.0) : 1.0 - So;                // HOMOZYGOUS, :1306-1319
		rc[8 + al * 2 + f] = hf * invd;
	}
}
}  // namespace cnf2
