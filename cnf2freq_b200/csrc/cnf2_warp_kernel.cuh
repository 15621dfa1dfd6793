// The warp-per-unit E-step kernel (the product path).
//
// One warp = one (analysed individual, chromosome) unit from start to end; CTAs are single warps, several per
// SM, fed from an atomic queue.  No block-wide barrier exists anywhere: the warp runs its forward sweep, its
// backward sweep and the posterior of every marker on its own.
//
//   * sweeps (cnF2freq.cpp:2145-2418): ALL 8 shift modes live in one warp: lane = shift * 4 + (state bits 4-5),
//     16 states (bits 0-3) per lane in registers.  The transition R(r)^{(x)6} is four in-register butterfly
//     stages and two shfl.xor stages on 16 independent values per lane; the emission comes from the marker's
//     shift-independent table (see cnf2_tables.cuh); the vectors are scaled by exact powers of two per
//     shift mode instead of being renormalised.
//   * posterior (:5406-5583): W_s[g] = alpha_s[g] beta_s[g] exp(F + B - factor) is transposed through shared
//     memory; one lane owns one (parent subtree configuration, root orientation) entry and sums over the
//     compatible entries of the other parent's side for all 8 shift modes (warp_posterior_pairs), scatters once into
//     an accumulator column, the warp reduces the columns and transfers them into the EM sufficient statistics
//     (:5875-5904).
//   * the orientation-fixed subtree tables are direct-indexed by local configuration; one lane builds one
//     (entry, root orientation) pair.
#pragma once
#include "cnf2_tables.cuh"

namespace cnf2
{
// tuning knobs (compile-time; the defaults are the measured best, profiles/r2_summary.md)
#ifndef CNF2_WB
#define CNF2_WB 4
#endif
#ifndef CNF2_WARPS_PER_SM
#define CNF2_WARPS_PER_SM 12
#endif
#ifndef CNF2_PAIR_UNROLL
#define CNF2_PAIR_UNROLL 2
#endif
constexpr int WB = CNF2_WB;   // markers staged at a time
constexpr int PAIR_UNROLL = CNF2_PAIR_UNROLL;
constexpr int T0W = 3, T1W = 5;
constexpr int NCOL = 8;       // accumulator columns; lanes scatter in 32 / NCOL passes
constexpr int NACCP = NACC + 2; // column stride of the accumulators (doubles): 16-byte aligned pairs, columns on distinct banks

// Swizzled layout of W_s[g] (8 shift modes x 64 states) in shared memory: the 8 shift modes of a state are one
// 64-byte row (four 16-byte chunks (s1, s2), each holding s0 = 0, 1); rows and chunks are permuted by XORs of the
// state's bits so that the sweep lanes' 64-bit stores and the posterior lanes' 128-bit loads are conflict-free.
// The address is GF(2)-linear in g:  ws_addr(g, s) = ws_base(g) ^ s,  ws_base(g0 ^ g1) = ws_base(g0) ^ ws_base(g1).
__host__ __device__ constexpr int ws_base(int g)
{
	const int b3 = (g >> 3) & 1, b4 = (g >> 4) & 1, b5 = (g >> 5) & 1;
	return ((g ^ b3 ^ b5) << 3) | ((((g >> 1) & 3) ^ (2 * b4 + b5)) << 1);
}

// One output of the transfer step (relative r, quantity q: 0 haplobase, 1 haplocount, 2..5 infprobs, 6..7 homozyg),
// fixed for a (unit, chromosome): where it goes and what it sums.
struct TrMeta
{
	double* ptr0;   // address of the output at marker 0 of the chromosome; marker m is `stride` doubles further per marker
	double scale;   // q < 2: descendants; q 2..5: 2 / 2^halvings * descendants (moveinfprobs, :3584-3594); else 1
	uint8_t mask;   // tree slots the relative occupies (non-null ones)
	uint8_t first;  // lowest of them: its haploweight is the relative's
	uint8_t q;
	uint8_t stride;
	uint8_t pad[4];
};
constexpr int NTR = NSLOTS * 6 + 2;
constexpr int CLS = 2 * 16 + 2 * 16 * 3;  // class sums: parent 0 [f][class] E; parent 1 [f][class] (E, E * Q(1,0)/Q, E * Q(2,0)/Q)

struct __align__(16) WarpSmem
{
	TrMeta tr[NTR];
	double acc[NCOL * NACCP];      // [column][accumulator]
	double Ws[NUMSHIFTS * NUMTYPES]; // ws_base(g) ^ s
	double qh[WB * QS];
	NodeTab nt[NSLOTS * WB];
	double HW[NSLOTS * WB];
	double tot[NACC];
	double rootc[RC];
	uchar2 A[NSLOTS * WB];
	unsigned char el[128];         // list of the needed configurations: parent << 6 | cfg; parent 0 first, phase bit 0 first
	unsigned char vmk[256];        // per (entry, root orientation): bit xo = rule 2 admits the other parent's x = xo
	unsigned short ainfo[128];     // per entry: ws_base of its meiosis bits at the parent's place | x << 10
	int listtag;
	short nE, nE0, nX0, nX1;        // nX0 / nX1: entries of parent 0 / 1 with phase bit x & 1 = 0
	short ntr, rootmask;            // (sizes kept tight: 12 single-warp CTAs per SM need <= 18,432 B each)
	double* nsptr[NSLOTS];         // turn scoring: negshift cell of the slot's occupant at marker 0 (null: no contribution)
	unsigned char homslot[WB], gpmust[WB];
	Unit unit;
	// followed by the posterior tables of the current marker, sized by the host for the pedigree (WarpArgs::cap0/cap1):
	//   tab0[cap0][f][T0W]: Q(plain), rho0[1], rho0[2];  tab1[cap1][f][T1W]: Q(plain), rho1[1], rho1[2], Q(al,0)/Q(plain)
	// entry k of a table is the k-th needed configuration of that parent at this marker (its position in el);
	// (units that rule 2 does not constrain use the same space for their factors and class sums: warp_posterior_factored)
};

// Posterior of the current marker, pairwise form.  A surviving term factorises as
//   val = rbase[f] * E0[f][cfg0] * E1[f][cfg1] * sum_s Wr[f ^ s0] Wp0[x1 ^ s1] Wp1[x4 ^ s2] W_s[g0 | g1 << 3]
// and every scatter target of :5556-5575 depends on the term only through (f, cfg0) [root as side 0, parent 0's
// subtree] or (f, cfg1) [root as side 1, parent 1's subtree]; homozyg is a product of one factor of each side.
// One lane owns one (side, local configuration, root orientation) entry, runs over the compatible entries of
// the OTHER side (rule 2 of ignoreflag2 couples the sides through Unit::xmask2, precomputed as vmk) summing in
// registers, and scatters its sums once.  The sums are kept raw -- per (own shift bit, root shift bit), weighted
// only by the other side's emission and phase weight -- and combined with the lane's own weights after the loop.
// The unusualstate cut of single terms (val > e^-200, :5414) is not applied here (it removes at most e^-200 of
// the posterior mass); whole shift modes below it are cut where W is formed.
__device__ __forceinline__ void warp_posterior_pairs(WarpSmem& w, const Unit& u, int homslot, int lane, const double* __restrict__ tab0,
                                                     const double* __restrict__ tab1)
{
	const double* rc = w.rootc;
	double* const acc = w.acc + (lane & (NCOL - 1)) * NACCP;
	const int vis = u.vismask, hmask = vis & ~homslot;
	const int nE = w.nE, nE0 = w.nE0;
	const double2* const Ws2 = reinterpret_cast<const double2*>(w.Ws);
	for (int ibase = 0; ibase < 2 * nE; ibase += 32)
	{
		const int i = ibase + lane;
		const bool valid = i < 2 * nE;
		const int ie = valid ? i >> 1 : 0;
		const int c = w.el[ie], f = i & 1;
		const int p = c >> 6, cfg = c & 63;
		const int gown = cfg & 7, xown = gown ^ (cfg >> 3);
		// phase weights: own parent at its shift bit 0 / 1, and the table of the other parent
		const double own0 = rc[12 + p * 2 + (xown & 1)], own1 = rc[12 + p * 2 + ((xown & 1) ^ 1)];
		const double wo0 = rc[12 + (p ^ 1) * 2], wo1 = rc[12 + (p ^ 1) * 2 + 1];
		const double rA = rc[2 + f], rB = rc[2 + (f ^ 1)];  // root at s0 = 0, 1
		const int aown = w.ainfo[ie] & 1023;
		const int vm = w.vmk[valid ? i : 0];
		// chunk (s1, s2) of a row is at ^ (s1 + 2 s2) << 1; the own parent's shift bit is s1 for parent 0, s2 for parent 1
		const int kown = p ? 4 : 2, koth = p ? 2 : 4;
		const double* Tp = p ? tab0 + f * T0W : tab1 + f * T1W;  // the other parent's table, this root orientation
		const int tw = 2 * (p ? T0W : T1W);
		const int jlo = p ? 0 : nE0, jhi = valid ? (p ? nE0 : nE) : jlo;
		const int jmid = min(jhi, jlo + (p ? w.nX0 : w.nX1));  // the other parent's entries with phase bit 0 come first
		const double k00 = own0 * rA, k01 = own0 * rB, k10 = own1 * rA, k11 = own1 * rB;
		double Rx0 = 0, Ry0 = 0, Rx1 = 0, Ry1 = 0, Sx1 = 0, Sx2 = 0;
		auto run = [&](int ja, int jb, double qa, double qb) {  // qa / qb: the other parent's weight at its shift bit 0 / 1
			const double* T = Tp + (ja - jlo) * tw;
#pragma unroll PAIR_UNROLL
			for (int j = ja; j < jb; j++, T += tw)
			{
				const int info = w.ainfo[j];
				const int a0 = aown ^ (info & 1023);
				const double E = ((vm >> (info >> 10)) & 1) ? T[0] : 0.0;  // rule 2 (:3483-3486)
				const double Ea = E * qa, Eb = E * qb;
				const double2 c00 = Ws2[a0 >> 1], cT = Ws2[(a0 ^ koth) >> 1], cO = Ws2[(a0 ^ kown) >> 1], c11 = Ws2[(a0 ^ 6) >> 1];
				const double tx0 = Ea * c00.x + Eb * cT.x, ty0 = Ea * c00.y + Eb * cT.y;  // own shift bit 0; root shift bit 0 / 1
				const double tx1 = Ea * cO.x + Eb * c11.x, ty1 = Ea * cO.y + Eb * c11.y;  // own shift bit 1
				Rx0 += tx0, Ry0 += ty0, Rx1 += tx1, Ry1 += ty1;
				if (!p)
				{
					const double tot = k00 * tx0 + k01 * ty0 + k10 * tx1 + k11 * ty1;
					Sx1 += tot * T[3];
					Sx2 += tot * T[4];
				}
			}
		};
		run(jlo, jmid, wo0, wo1);
		run(jmid, jhi, wo1, wo0);
		const int slot = ie - (p ? nE0 : 0);
		const double* To = p ? tab1 + (slot * 2 + f) * T1W : tab0 + (slot * 2 + f) * T0W;
		const double F = rc[f] * To[0];
		const double Sz0 = F * rA * (own0 * Rx0 + own1 * Rx1), Sz1 = F * rB * (own0 * Ry0 + own1 * Ry1);  // root shift bit 0 / 1
		const double So0 = F * own0 * (rA * Rx0 + rB * Ry0), So1 = F * own1 * (rA * Rx1 + rB * Ry1);    // own shift bit 0 / 1
		const double S = So0 + So1;
		const bool live = valid && S > 0;
		const int sp = 1 + 3 * p, o_ = cfg >> 3, xb = xown & 1;
		const int gs = sp + 1 + (gown & 1);
		const double u0 = S * To[1], u1 = S * To[2];
		// scatter targets: pairs of adjacent accumulators (128-bit read-modify-write) and two single ones
		int pk[6], sk[2];
		double2 pv[6];
		bool pe[6], se[2];
		pk[0] = ACC_HAPLOS, pv[0] = f ? make_double2(Sz1, Sz0) : make_double2(Sz0, Sz1), pe[0] = !p && (hmask & 1);  // root: s0 = 0 -> bit f
		pk[1] = ACC_HAPLOS + sp * 2, pv[1] = xb ? make_double2(So1, So0) : make_double2(So0, So1), pe[1] = (hmask >> sp) & 1;
		sk[0] = ACC_HAPLOS + (sp + 1) * 2 + ((xown >> 1) & 1), se[0] = (hmask >> (sp + 1)) & 1;
		sk[1] = ACC_HAPLOS + (sp + 2) * 2 + ((xown >> 2) & 1), se[1] = (hmask >> (sp + 2)) & 1;
		// the GENOS calls of side p (:5560-5568): root at orientation f ^ side, the parent, its parent picked by the meiosis bit
		const double2 uu = make_double2(u0, u1);
		pk[2] = ACC_INF + (f ^ p) * 2, pv[2] = uu, pe[2] = true;
		pk[3] = ACC_INF + (sp * 2 + (o_ & 1)) * 2, pv[3] = uu, pe[3] = (vis >> sp) & 1;
		pk[4] = ACC_INF + (gs * 2 + ((o_ >> (gs - sp)) & 1)) * 2, pv[4] = uu, pe[4] = pe[3] && ((vis >> gs) & 1);
		// homozyg (:5571-5575): rho0[al] * hfd[al][f] * sum over the other side of val * Q1(al, 0) / Q1(plain)
		pk[5] = ACC_HZ, pv[5] = make_double2(F * Sx1 * To[1] * rc[8 + f], F * Sx2 * To[2] * rc[10 + f]), pe[5] = !p;
		// NCOL lanes at a time add into their own accumulator column; the targets of one entry are distinct, so
		// all are read before any is written
#pragma unroll 1
		for (int pass = 0; pass < 32 / NCOL; pass++)
		{
			if (live && (lane / NCOL) == pass)
			{
				double2 cur[6];
				double cs[2];
#pragma unroll
				for (int j = 0; j < 6; j++)
					if (pe[j]) cur[j] = *reinterpret_cast<const double2*>(acc + pk[j]);
#pragma unroll
				for (int j = 0; j < 2; j++)
					if (se[j]) cs[j] = acc[sk[j]];
#pragma unroll
				for (int j = 0; j < 6; j++)
					if (pe[j]) *reinterpret_cast<double2*>(acc + pk[j]) = make_double2(cur[j].x + pv[j].x, cur[j].y + pv[j].y);
#pragma unroll
				for (int j = 0; j < 2; j++)
					if (se[j]) acc[sk[j]] = cs[j] + S;
			}
			__syncwarp();
		}
	}
}

// The posterior tables of the pairwise form, from FACTORS rather than one evaluation per (entry, root orientation): the value
// of a subtree configuration is H(o, m) * X(first grandparent) * Y(other grandparent) (see warp_posterior_factored), so the
// 32 lanes evaluate the 16 (side, f, o, m) rows once -- lane = row * 2 + h holds H_E, H_1, H_2 and, for grandparent
// orientation h, the first grandparent's three values and the other grandparent's one, 4 gval calls instead of 4 per
// entry -- and every entry's lane fetches its factors by shuffle, applies the grandparents' phase weights and multiplies:
// nb * [c * gfix(other)] * gfix(first) of the per-entry form, with the same zeros.
__device__ __forceinline__ void warp_tables_from_factors(const WarpSmem& w, const Unit& u, int t, int lane, double* __restrict__ tab0,
                                                         double* __restrict__ tab1)
{
	const unsigned FULL = 0xffffffffu;
	const uchar2* A = w.A;
	const NodeTab* N = w.nt;
	const double* rc = w.rootc;
	double H[3], GX[3], GY = 1.0;
	{
		const int h = lane & 1, m = (lane >> 1) & 1, o = (lane >> 2) & 1, f = (lane >> 3) & 1, p = lane >> 4;
		const int sp = 1 + 3 * p;
		const uchar2 a0 = A[t];
		const NodeTab& n0 = N[t];
		const int vroot = p == 0 ? (f ? a0.y : a0.x) : (f ? a0.x : a0.y);
		const double secroot = p == 0 ? (vroot != UNKNOWN ? n0.RN[f] : 0.0) : n0.SEC[f ^ 1];
		const bool absent = u.row[sp] < 0, attop = !absent && u.founder[sp];
		const uchar2 ap = A[sp * WB + t];
		const NodeTab& np = N[sp * WB + t];
		const int kf = sp + 1 + m, ko = sp + 1 + (m ^ 1);
#pragma unroll
		for (int k = 0; k < 3; k++)
		{
			const int v = k == 0 ? vroot : k;
			const double sec = k == 0 ? secroot : 0.0;
			if (absent) H[k] = 1.0 + sec, GX[k] = 1.0;
			else if (attop) H[k] = gval(ap, np, o, v, sec), GX[k] = 1.0;
			else
			{
				const PMsg pm = pmsg(ap, np, o, v, sec);
				H[k] = pm.nb * (np.S[o ^ 1] != 0 ? 1.0 - np.S[o ^ 1] : 1.0);
				GX[k] = u.row[kf] < 0 ? 1.0 + pm.up2 : gval(A[kf * WB + t], N[kf * WB + t], h, pm.v2, pm.up2);
			}
		}
		if (!absent && !attop)
		{
			const int vo = o ? ap.x : ap.y;
			const double so = np.SEC[o ^ 1];
			GY = u.row[ko] < 0 ? 1.0 + so : gval(A[ko * WB + t], N[ko * WB + t], h, vo, so);
		}
	}
	const int nE = w.nE, nE0 = w.nE0;
	for (int ibase = 0; ibase < 2 * nE; ibase += 32)
	{
		const int i = ibase + lane;
		const bool valid = i < 2 * nE;
		const int c = w.el[valid ? i >> 1 : 0], f = i & 1;
		const int p = c >> 6, cfg = c & 63, sp = 1 + 3 * p;
		const int m = cfg & 1, ma = (cfg >> 1) & 1, mb = (cfg >> 2) & 1, o = (cfg >> 3) & 1, oa = (cfg >> 4) & 1, ob = (cfg >> 5) & 1;
		const int of = m ? ob : oa, oo = m ? oa : ob, mF = m ? mb : ma, mO = m ? ma : mb;
		const int kf = sp + 1 + m, ko = sp + 1 + (m ^ 1);
		const int rowl = ((((p * 2 + f) * 2 + o) * 2 + m) << 1);
		const double HE = __shfl_sync(FULL, H[0], rowl), H1 = __shfl_sync(FULL, H[1], rowl), H2 = __shfl_sync(FULL, H[2], rowl);
		double XE = __shfl_sync(FULL, GX[0], rowl | of), X1 = __shfl_sync(FULL, GX[1], rowl | of), X2 = __shfl_sync(FULL, GX[2], rowl | of);
		double Y = __shfl_sync(FULL, GY, rowl | oo);
		if (!valid) continue;
		const bool leaf = u.row[sp] < 0 || u.founder[sp];
		if (!leaf)
		{
			if (u.row[kf] >= 0)
			{
				const double wF = N[kf * WB + t].W[of ^ mF];
				XE *= wF, X1 *= wF, X2 *= wF;
			}
			if (u.row[ko] >= 0) Y *= N[ko * WB + t].W[oo ^ mO];
		}
		const double E = HE * Y * XE, P1 = H1 * Y * X1, P2 = H2 * Y * X2;
		const int slot = (i >> 1) - (p ? nE0 : 0);
		if (p == 0)
		{
			double* e = tab0 + (slot * 2 + f) * T0W;
			e[0] = E;
			const double A1 = rc[4 + f] * P1, A2 = rc[6 + f] * P2;
			const double sum = A1 + A2, inv = sum > 0 ? frcp(sum) : 0.0;
			e[1] = A1 * inv;
			e[2] = A2 * inv;
		}
		else
		{
			double* e = tab1 + (slot * 2 + f) * T1W;
			e[0] = E;
			const double A1 = rc[4 + (f ^ 1)] * P1, A2 = rc[6 + (f ^ 1)] * P2;  // side 1 probes the root at orientation f ^ 1
			const double sum = A1 + A2, inv = sum > 0 ? frcp(sum) : 0.0, invE = E > 0 ? frcp(E) : 0.0;
			e[1] = A1 * inv;
			e[2] = A2 * inv;
			e[3] = P1 * invE;
			e[4] = P2 * invE;
		}
	}
}

// Posterior of the current marker for a unit that rule 2 of ignoreflag2 does not constrain (no individual on both sides:
// Unit::xmask2 all ones), fully factorised.  Per side the emission of a subtree configuration (m, ma, mb | o, oa, ob) is
//   E = H(o, m) * X(o, m; first grandparent's orientation and phase index) * Y(o, m; other grandparent's)
// (warp_tables_from_factors: nb * [c * gfix(other)] * gfix(first); "first" = the grandparent the parent's meiosis bit m points to), the
// configurations in use are a product set over the three tree slots (rule 1 pins the orientation of an absent / empty slot,
// rule 3 the x bit of a homozygous grandparent), and every scatter target of :5556-5575 groups the configurations by a few
// of these bits.  So nothing is built per configuration: 16 (o, m) rows of factors per side and root orientation
// (FT: 8 gval evaluations each), sums over the grandparents' orientations where a target does not tell them apart, and one
// lane per (side, root orientation, phase bit, meiosis bits) CLASS -- 64 classes, two rounds -- that meets the other
// side's class sums through W (AR / AH, as in the pairwise form) and scatters its ~20 results.  For three-generation
// outbred pedigrees this replaces 256 table entries + 256 pair-loop items per marker.
// per (side, f, o, m): H_E, H_1, H_2; X_k[mF][q] (k = E, 1, 2) and Y[mO][q]: the grandparents' values with their phase weight
// W[q ^ meiosis bit] applied and zero where the slot's admissible set excludes (orientation q, meiosis bit); GX_1[q], GX_2[q] raw
constexpr int FT = 23, FT_X = 3, FT_Y = 15, FT_G = 19;
constexpr int FT_DOUBLES = 16 * FT, CLS_AT = FT_DOUBLES;  // class sums (CLS doubles) sit behind the factors

__device__ __forceinline__ void warp_posterior_factored(WarpSmem& w, const Unit& u, int homslot, int gpmust, int t, int lane,
                                                        double* __restrict__ ft)
{
	const unsigned FULL = 0xffffffffu;
	const double* rc = w.rootc;
	double* const cls = ft + CLS_AT;
	const uchar2* A = w.A;
	const NodeTab* N = w.nt;
	// ---- factors: lane = (side, f, o, m) * 2 + h; h = the orientation (of the first resp. other grandparent) this lane does ----
	{
		const int h = lane & 1, m = (lane >> 1) & 1, o = (lane >> 2) & 1, f = (lane >> 3) & 1, p = lane >> 4;
		const int sp = 1 + 3 * p;
		const uchar2 a0 = A[t];
		const NodeTab& n0 = N[t];
		const int vroot = p == 0 ? (f ? a0.y : a0.x) : (f ? a0.x : a0.y);
		const double secroot = p == 0 ? (vroot != UNKNOWN ? n0.RN[f] : 0.0) : n0.SEC[f ^ 1];
		const bool absent = u.row[sp] < 0, attop = !absent && u.founder[sp];
		const uchar2 ap = A[sp * WB + t];
		const NodeTab& np = N[sp * WB + t];
		const int kf = sp + 1 + m, ko = sp + 1 + (m ^ 1);
		double H[3], GX[3], GY = 1.0;
#pragma unroll
		for (int k = 0; k < 3; k++)
		{
			const int v = k == 0 ? vroot : k;
			const double sec = k == 0 ? secroot : 0.0;
			if (absent) H[k] = 1.0 + sec, GX[k] = 1.0;
			else if (attop) H[k] = gval(ap, np, o, v, sec), GX[k] = 1.0;
			else
			{
				const PMsg pm = pmsg(ap, np, o, v, sec);
				H[k] = pm.nb * (np.S[o ^ 1] != 0 ? 1.0 - np.S[o ^ 1] : 1.0);
				GX[k] = u.row[kf] < 0 ? 1.0 + pm.up2 : gval(A[kf * WB + t], N[kf * WB + t], h, pm.v2, pm.up2);
			}
		}
		if (!absent && !attop)
		{
			const int vo = o ? ap.x : ap.y;
			const double so = np.SEC[o ^ 1];
			GY = u.row[ko] < 0 ? 1.0 + so : gval(A[ko * WB + t], N[ko * WB + t], h, vo, so);
		}
		double* row = ft + (lane >> 1) * FT;
		if (h == 0) row[0] = H[0], row[1] = H[1], row[2] = H[2];
		const bool leaf = absent || attop;  // the subtree ends at the parent: no grandparent is visited
		const int ign = u.flag2ignore;
#pragma unroll
		for (int mb = 0; mb < 2; mb++)  // the grandparent's meiosis bit: its x bit (the phase index of its weight) is h ^ mb
		{
			const bool okF = !(((ign >> kf) & 1) && h) && !(((gpmust >> kf) & 1) && !(h ^ mb));
			const bool okO = !(((ign >> ko) & 1) && h) && !(((gpmust >> ko) & 1) && !(h ^ mb));
			const double wF = (leaf || u.row[kf] < 0) ? 1.0 : N[kf * WB + t].W[h ^ mb];
			const double wO = (leaf || u.row[ko] < 0) ? 1.0 : N[ko * WB + t].W[h ^ mb];
#pragma unroll
			for (int k = 0; k < 3; k++) row[FT_X + (mb * 2 + h) * 3 + k] = okF ? GX[k] * wF : 0.0;
			row[FT_Y + mb * 2 + h] = okO ? GY * wO : 0.0;
		}
		row[FT_G + h] = GX[1], row[FT_G + 2 + h] = GX[2];
	}
	__syncwarp();
	// ---- per class (side, f, phase bit, meiosis bits): the factors with the phase weights and the slots' admissible sets ----
	struct ClassView
	{
		double HE, H1, H2, XE[2], X1[2], X2[2], YO[2];  // X / Y: zero where the configuration is not in use
		int o, m, mF, mO, kF, kO;
		bool any;
	};
	auto view = [&](int p, int f, int ph, int g) -> ClassView {
		ClassView c;
		const int sp = 1 + 3 * p;
		c.m = g & 1;
		c.o = ph ^ c.m;
		const int ma = (g >> 1) & 1, mb = (g >> 2) & 1;
		c.mF = c.m ? mb : ma, c.mO = c.m ? ma : mb;
		c.kF = sp + 1 + c.m, c.kO = sp + 1 + (c.m ^ 1);
		c.any = !(((u.flag2ignore >> sp) & 1) && c.o);
		const double* row = ft + (((p * 2 + f) * 2 + c.o) * 2 + c.m) * FT;
		c.HE = row[0], c.H1 = row[1], c.H2 = row[2];
#pragma unroll
		for (int q = 0; q < 2; q++)  // q: the grandparent's orientation
		{
			const double* x = row + FT_X + (c.mF * 2 + q) * 3;
			c.XE[q] = x[0], c.X1[q] = x[1], c.X2[q] = x[2];
			c.YO[q] = row[FT_Y + c.mO * 2 + q];
		}
		return c;
	};
	// class sums of the emissions (and, for parent 1, of the two probing values: homozyg) per (root orientation, phase bit, meiosis bits)
#pragma unroll
	for (int pass = 0; pass < 2; pass++)
	{
		const int i = pass * 32 + lane, p = i >> 5, f = (i >> 4) & 1, k = i & 15;
		const ClassView c = view(p, f, k >> 3, k & 7);
		const double SY = c.any ? c.YO[0] + c.YO[1] : 0.0;
		const double sE = c.HE * (c.XE[0] + c.XE[1]) * SY;
		if (p)
		{
			double* o = cls + 32 + (f * 16 + k) * 3;
			// the probing values only where the configuration's own emission is not zero: a term without posterior mass adds
			// nothing to homozyg (:5571-5575), whatever its probing value
			const double e0 = c.HE * c.XE[0], e1 = c.HE * c.XE[1];
			o[0] = sE;
			o[1] = c.H1 * ((e0 != 0 ? c.X1[0] : 0.0) + (e1 != 0 ? c.X1[1] : 0.0)) * SY;
			o[2] = c.H2 * ((e0 != 0 ? c.X2[0] : 0.0) + (e1 != 0 ? c.X2[1] : 0.0)) * SY;
		}
		else cls[f * 16 + k] = sE;
	}
	__syncwarp();
	// ---- the other side's class sums through W: lane = side * 16 + meiosis bits * 2 + f (as in the pairwise form) ----
	const double2* const Ws2 = reinterpret_cast<const double2*>(w.Ws);
	double AR[4] = { 0, 0, 0, 0 }, AH1[4] = { 0, 0, 0, 0 }, AH2[4] = { 0, 0, 0, 0 };
	{
		const int p = lane >> 4, g = (lane >> 1) & 7, f = lane & 1;
		const int aown = ws_base(g << (p ? 3 : 0));
		const int kown = p ? 4 : 2, koth = p ? 2 : 4;
		const double wo0 = rc[12 + (p ^ 1) * 2], wo1 = rc[12 + (p ^ 1) * 2 + 1];
		const double* C = p ? cls + f * 16 : cls + 32 + f * 48;
		const int cw = p ? 1 : 3;
		// per meiosis-bits value of the other side: its two phase classes first meet the other parent's phase weights (class h
		// carries weight wo_h at shift bit 0, wo_(1-h) at shift bit 1), then the W block once
#pragma unroll 2
		for (int go = 0; go < 8; go++)
		{
			const int a0 = aown ^ (p ? ws_base(go) : ws_base(go << 3));
			const double* C0 = C + go * cw;
			const double* C1 = C + (8 + go) * cw;
			const double2 c00 = Ws2[a0 >> 1], cT = Ws2[(a0 ^ koth) >> 1], cO = Ws2[(a0 ^ kown) >> 1], c11 = Ws2[(a0 ^ 6) >> 1];
			{
				const double Ea = C0[0] * wo0 + C1[0] * wo1, Eb = C0[0] * wo1 + C1[0] * wo0;
				AR[0] += Ea * c00.x + Eb * cT.x, AR[1] += Ea * c00.y + Eb * cT.y, AR[2] += Ea * cO.x + Eb * c11.x, AR[3] += Ea * cO.y + Eb * c11.y;
			}
			if (!p)
			{
				const double Ha = C0[1] * wo0 + C1[1] * wo1, Hb = C0[1] * wo1 + C1[1] * wo0;
				AH1[0] += Ha * c00.x + Hb * cT.x, AH1[1] += Ha * c00.y + Hb * cT.y, AH1[2] += Ha * cO.x + Hb * c11.x, AH1[3] += Ha * cO.y + Hb * c11.y;
				const double Ga = C0[2] * wo0 + C1[2] * wo1, Gb = C0[2] * wo1 + C1[2] * wo0;
				AH2[0] += Ga * c00.x + Gb * cT.x, AH2[1] += Ga * c00.y + Gb * cT.y, AH2[2] += Ga * cO.x + Gb * c11.x, AH2[3] += Ga * cO.y + Gb * c11.y;
			}
		}
	}
	// ---- the classes' own results, scattered into the accumulator columns ----
	double* const acc = w.acc + (lane & (NCOL - 1)) * NACCP;
	const int vis = u.vismask, hmask = vis & ~homslot;
#pragma unroll 1
	for (int round = 0; round < 2; round++)
	{
		const int i = round * 32 + lane, p = i >> 5, f = (i >> 4) & 1, ph = (i >> 3) & 1, g = i & 7;
		const int sp = 1 + 3 * p;
		const ClassView c = view(p, f, ph, g);
		const int src = p * 16 + g * 2 + f;
		const double Rx0 = __shfl_sync(FULL, AR[0], src), Ry0 = __shfl_sync(FULL, AR[1], src);
		const double Rx1 = __shfl_sync(FULL, AR[2], src), Ry1 = __shfl_sync(FULL, AR[3], src);
		const double h10 = __shfl_sync(FULL, AH1[0], src), h11 = __shfl_sync(FULL, AH1[1], src);
		const double h12 = __shfl_sync(FULL, AH1[2], src), h13 = __shfl_sync(FULL, AH1[3], src);
		const double h20 = __shfl_sync(FULL, AH2[0], src), h21 = __shfl_sync(FULL, AH2[1], src);
		const double h22 = __shfl_sync(FULL, AH2[2], src), h23 = __shfl_sync(FULL, AH2[3], src);
		const double own0 = rc[12 + p * 2 + ph], own1 = rc[12 + p * 2 + (ph ^ 1)];
		const double rA = rc[2 + f], rB = rc[2 + (f ^ 1)];
		const double z0 = rA * (own0 * Rx0 + own1 * Rx1), z1 = rB * (own0 * Ry0 + own1 * Ry1);  // root shift bit 0 / 1
		const double s0 = own0 * (rA * Rx0 + rB * Ry0), s1 = own1 * (rA * Rx1 + rB * Ry1);    // own shift bit 0 / 1
		const double tS = s0 + s1;
		const double base = c.any ? rc[f] * c.HE : 0.0;
		const double SX = c.XE[0] + c.XE[1], SY = c.YO[0] + c.YO[1];
		const double M = base * SX * SY;
		const bool live = M * tS > 0;
		// probing ratios of the GENOS calls (:5517-5538): A_al / (A_1 + A_2) per orientation of the first grandparent
		const int fq = p ? f ^ 1 : f;
		double U1[2], U2[2], hz1 = 0, hz2 = 0;
		const double hzw1 = own0 * (rA * h10 + rB * h11) + own1 * (rA * h12 + rB * h13);
		const double hzw2 = own0 * (rA * h20 + rB * h21) + own1 * (rA * h22 + rB * h23);
		const double* row = ft + (((p * 2 + f) * 2 + c.o) * 2 + c.m) * FT;
#pragma unroll
		for (int q = 0; q < 2; q++)
		{
			const double A1 = rc[4 + fq] * c.H1 * row[FT_G + q], A2 = rc[6 + fq] * c.H2 * row[FT_G + 2 + q];
			const double sum = A1 + A2, inv = sum > 0 ? frcp(sum) : 0.0;
			const double mq = base * c.XE[q] * SY;  // the class's mass with the first grandparent at orientation q
			U1[q] = mq * tS * (A1 * inv), U2[q] = mq * tS * (A2 * inv);
			hz1 += mq * (A1 * inv), hz2 += mq * (A2 * inv);
		}
		double2 pv[9];
		int pk[9];
		bool pe[9];
		pk[0] = ACC_HAPLOS, pv[0] = f ? make_double2(M * z1, M * z0) : make_double2(M * z0, M * z1), pe[0] = !p && (hmask & 1);
		pk[1] = ACC_HAPLOS + sp * 2, pv[1] = ph ? make_double2(M * s1, M * s0) : make_double2(M * s0, M * s1), pe[1] = (hmask >> sp) & 1;
		{
			// the grandparents' haplos: index = the x bit of the slot = orientation ^ meiosis bit
			const double f0 = base * tS * c.XE[0] * SY, f1 = base * tS * c.XE[1] * SY, o0 = base * tS * SX * c.YO[0], o1 = base * tS * SX * c.YO[1];
			pk[2] = ACC_HAPLOS + c.kF * 2, pv[2] = c.mF ? make_double2(f1, f0) : make_double2(f0, f1), pe[2] = (hmask >> c.kF) & 1;
			pk[3] = ACC_HAPLOS + c.kO * 2, pv[3] = c.mO ? make_double2(o1, o0) : make_double2(o0, o1), pe[3] = (hmask >> c.kO) & 1;
		}
		const double2 uu = make_double2(U1[0] + U1[1], U2[0] + U2[1]);
		pk[4] = ACC_INF + (f ^ p) * 2, pv[4] = uu, pe[4] = true;
		pk[5] = ACC_INF + (sp * 2 + c.o) * 2, pv[5] = uu, pe[5] = (vis >> sp) & 1;
		pk[6] = ACC_INF + (c.kF * 2 + 0) * 2, pv[6] = make_double2(U1[0], U2[0]), pe[6] = pe[5] && ((vis >> c.kF) & 1);
		pk[7] = ACC_INF + (c.kF * 2 + 1) * 2, pv[7] = make_double2(U1[1], U2[1]), pe[7] = pe[6];
		pk[8] = ACC_HZ, pv[8] = make_double2(hz1 * hzw1 * rc[8 + f], hz2 * hzw2 * rc[10 + f]), pe[8] = !p;
#pragma unroll 1
		for (int pass = 0; pass < 32 / NCOL; pass++)
		{
			if (live && (lane / NCOL) == pass)
			{
#pragma unroll
				for (int j = 0; j < 9; j++)
					if (pe[j])
					{
						double2* a2 = reinterpret_cast<double2*>(acc + pk[j]);
						const double2 cur = *a2;
						*a2 = make_double2(cur.x + pv[j].x, cur.y + pv[j].y);
					}
			}
			__syncwarp();
		}
	}
}

// Transfer of one marker's reduced accumulators into the EM sufficient statistics (cnF2freq.cpp:5875-5904,
// moveinfprobs :3577-3597, movehaplos :3599-3616; same arithmetic as transfer_stats), one lane per output.
// transfer_setup fills the per-(unit, chromosome) description of every output once; transfer_out is the
// per-marker part.
__device__ __forceinline__ void transfer_setup(WarpSmem& w, const EstepArgs& a, const Unit& u, int lane)
{
	const int rootrel = u.slotrel[0], nout = u.nrel * 6 + 2;
	for (int i = lane; i < nout; i += 32)
	{
		const int n6 = u.nrel * 6;
		const int r = i < n6 ? i / 6 : rootrel, q = i < n6 ? i % 6 : 6 + i - n6;
		int mask = 0;
		for (int slot = NSLOTS - 1; slot >= 0; slot--)
			if (u.row[slot] >= 0 && u.slotrel[slot] == r) mask |= 1 << slot;
		const size_t Mc = a.Mc;
		const int srow = u.relrow[r];
		const bool sh = srow < a.n_shared;
		double* base = sh ? a.stats_shared : a.stats_local;
		const size_t nr = sh ? a.n_shared : a.n_local, cell = (size_t)(sh ? srow : srow - a.n_shared) * Mc, plane = nr * Mc;
		TrMeta t;
		t.mask = (uint8_t)mask;
		t.first = (uint8_t)(mask ? __ffs(mask) - 1 : 0);
		t.q = (uint8_t)q;
		if (q < 2)
		{
			t.ptr0 = base + q * plane + cell;
			t.stride = 1;
			t.scale = u.descendants;
		}
		else if (q < 6)
		{
			t.ptr0 = base + 2 * plane + cell * 4 + (q - 2);
			t.stride = 4;
			double sc = 2;
			for (int k = 0; k < u.relhalvings[r]; k++) sc /= 2;
			t.scale = sc * u.descendants;
		}
		else
		{
			t.ptr0 = base + 6 * plane + cell * 2 + (q - 6);
			t.stride = 2;
			t.scale = 1.0;
		}
		w.tr[i] = t;
	}
	if (lane == 0)
	{
		int rm = 0;
		for (int slot = 0; slot < NSLOTS; slot++)
			if (u.row[slot] >= 0 && u.slotrel[slot] == rootrel) rm |= 1 << slot;
		w.rootmask = (short)rm;
		w.ntr = (short)nout;
	}
	if (lane < NSLOTS)
	{
		double* p = nullptr;
		if (u.row[lane] >= 0 && u.nsscale[lane] != 0)
		{
			const int srow = u.relrow[u.slotrel[lane]];
			const bool sh = srow < a.n_shared;
			p = stats_cell(sh ? a.stats_shared : a.stats_local, sh ? a.n_shared : a.n_local, a.Mc, sh ? srow : srow - a.n_shared, 0).negshift;
		}
		w.nsptr[lane] = p;
	}
}
__device__ __forceinline__ void transfer_out(const WarpSmem& w, const double* __restrict__ tot, int t, int m, int lane)
{
	double rootsum = 0;
	for (int rm = w.rootmask; rm; rm &= rm - 1)
	{
		const int slot = __ffs(rm) - 1;
		rootsum += tot[ACC_INF + slot * 4 + 0] + tot[ACC_INF + slot * 4 + 1];
	}
	const double inv = frcp(rootsum);
	for (int i = lane; i < w.ntr; i += 32)
	{
		const TrMeta mt = w.tr[i];
		const int q = mt.q;
		double* const dst = mt.ptr0 + (size_t)m * mt.stride;
		if (q >= 6)
		{
			atomicAdd(dst, tot[ACC_HZ + q - 6] * inv);
			continue;
		}
		if (q >= 2)
		{
			double ip = 0;
			for (int sm = mt.mask; sm; sm &= sm - 1) ip += tot[ACC_INF + (__ffs(sm) - 1) * 4 + (q - 2)];
			if (ip != 0) atomicAdd(dst, ip * (inv * mt.scale));
			continue;
		}
		double h0 = 0, h1 = 0;
		for (int sm = mt.mask; sm; sm &= sm - 1)
		{
			const int slot = __ffs(sm) - 1;
			h0 += tot[ACC_HAPLOS + slot * 2 + 0];
			h1 += tot[ACC_HAPLOS + slot * 2 + 1];
		}
		const double hw = w.HW[mt.first * WB + t];  // the relative's own haploweight; any of its slots holds it
		if ((h0 != 0 || h1 != 0) && fabs(hw - 0.5) < 0.5 - 1e-12)
		{
			const double eps = 1.9151695967140057e-174 * (double)(0.000005f) * (double)(0.000005f) * 0.5;  // exp(-400) * maxdiff^2 / 2, :228
			const double x1 = h0 + eps, x2 = h1 + eps;
			atomicAdd(dst, q ? mt.scale : x1 * frcp(x1 + x2) * mt.scale);
		}
	}
}

// ---- sweeps: lane = shift * 4 + gh, state g = r | gh << 4, r = 0..15 in registers ----

// Transition (cnF2freq.cpp:2270-2367): bits 0-3 inside the lane, bits 4-5 across lanes 1 and 2 apart.
// TYPEGENS = {1,0,0,1,0,0} (settings.h:23): genrec[1] (r1) for the child meioses, bits 0 and 3.
__device__ __forceinline__ void transition16(double (&p)[16], double r0, double r1)
{
	if (!(r0 > 0) && !(r1 > 0)) return;  // same position: identity (:2273)
#pragma unroll
	for (int b = 0; b < 4; b++)
	{
		const double r = (b == 0 || b == 3) ? r1 : r0;
#pragma unroll
		for (int i = 0; i < 16; i++)
			if (!(i & (1 << b)))
			{
				const double d = r * (p[i | (1 << b)] - p[i]);
				p[i] += d;
				p[i | (1 << b)] -= d;
			}
	}
#pragma unroll
	for (int b = 1; b <= 2; b <<= 1)
#pragma unroll
		for (int i = 0; i < 16; i++)
		{
			const double q = __shfl_xor_sync(0xffffffffu, p[i], b);
			p[i] += r0 * (q - p[i]);
		}
}

// p *= emission (all orientations summed) from the marker's sweep table q (layout: sweep_family)
__device__ __forceinline__ void emit16(double (&p)[16], const double* __restrict__ q, int lane)
{
	const int s = lane >> 2, gh = lane & 3;
	const int s0 = s & 1, s1 = (s >> 1) & 1, s2 = s >> 2;
	const double2 R = *reinterpret_cast<const double2*>(q + 64 + s0 * 2);
	const double2* Q1 = reinterpret_cast<const double2*>(q + (((2 + s2) * 8) + gh * 2) * 2);
	const double2* Q0 = reinterpret_cast<const double2*>(q + (s1 * 8) * 2);
	const double2 qa = Q1[0], qb = Q1[1];
	const double c0a = R.x * qa.x, c1a = R.y * qa.y, c0b = R.x * qb.x, c1b = R.y * qb.y;
#pragma unroll
	for (int j = 0; j < 8; j++)
	{
		const double2 q0 = Q0[j];
		p[j] *= c0a * q0.x + c1a * q0.y;
		p[j + 8] *= c0b * q0.x + c1b * q0.y;
	}
}
// fp64 tensor-core step of the turn scoring: C[8][8] += A[8][4] B[4][8] (mma.sync m8n8k4).  Lane T supplies
// A[T / 4][T % 4] and B[T % 4][T / 4] and holds C[T / 4][2 (T % 4)], C[T / 4][2 (T % 4) + 1].
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
{
	asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// One marker of the backward pass, fused: W = alpha- beta exp(F- + B - factor) into shared memory for the posterior
// (:1963-2018, :5499-5508), and the turn scoring (cnF2freq.cpp:5668-5757 with DOTOULBAR 0: the turns of at most one
// individual; aroundturner :498-554; quickanalyze with classicstop(marker, -1), :1960-2018; updatenegshifts :3640-3715):
//   rawvals[turn][s] = exp(F+_s + B_s' + log sum_k alpha+_s[k ^ mask] beta_s'[k] - factor),  s' = s ^ flagmodeshift
// with alpha+ = alpha- (x) emission, streamed state by state so that neither W nor alpha+ nor the emission values are
// ever held as 16-value register arrays.
//   * the root's turn and the two parents' turns flip a SHIFT bit (s' = s ^ 1, 2, 4): together with the unturned sum they
//     are four entries per row of the 8 x 8 matrix C[s][s'] = sum_g alpha+_s[g] beta_s'[g], a contraction over the 64
//     states.  In the sweeps' layout (lane = shift * 4 + state bits 4-5, registers = state bits 0-3) a lane's alpha+ and
//     beta values of one register index ARE the A and B fragments of mma.sync.m8n8k4.f64: 16 tensor-core steps, no
//     shuffle, and the sum over the four lanes of a shift mode comes with it.
//   * the grandparents' turns flip a state bit: bits 1, 2 inside the lane (register renaming), bits 4, 5 one and two lanes
//     apart -- alpha+ goes through shared memory (xbuf, the posterior tables' space, dead at this point) once.
// an: alpha- (power of two kan per shift mode), b: beta (kb), sc: the scale of W (0 for a shift mode that does not
// contribute), 2^kf * efrem = exp(-factor).
__device__ __forceinline__ void warp_w_and_turns(WarpSmem& w, const Unit& u, const double (&an)[16], const double (&b)[16], double sc, int kan,
                                                 int kb, int kf, double efrem, const double* __restrict__ q, double2* __restrict__ xbuf, int m,
                                                 int lane, bool turns)
{
	const unsigned FULL = 0xffffffffu;
	const int s = lane >> 2, gh = lane & 3;
	const int s0 = s & 1, s1 = (s >> 1) & 1, s2 = s >> 2;
	const double2 R = *reinterpret_cast<const double2*>(q + 64 + s0 * 2);
	const double2* Q1 = reinterpret_cast<const double2*>(q + (((2 + s2) * 8) + gh * 2) * 2);
	const double2* Q0 = reinterpret_cast<const double2*>(q + (s1 * 8) * 2);
	const double2 qa = Q1[0], qb = Q1[1];
	const double c0a = R.x * qa.x, c1a = R.y * qa.y, c0b = R.x * qb.x, c1b = R.y * qb.y;
	const int wbase = ws_base(gh << 4) ^ s;
	double C0 = 0, C1 = 0, D2 = 0, D3 = 0;
#pragma unroll
	for (int j = 0; j < 8; j++)
	{
		w.Ws[wbase ^ ws_base(j)] = an[j] * (b[j] * sc);
		w.Ws[wbase ^ ws_base(j + 8)] = an[j + 8] * (b[j + 8] * sc);
		if (turns)
		{
			const double2 q0 = Q0[j];
			const double alo = an[j] * (c0a * q0.x + c1a * q0.y), ahi = an[j + 8] * (c0b * q0.x + c1b * q0.y);  // alpha+ of states j, j + 8
			dmma884(C0, C1, alo, b[j]);
			dmma884(C0, C1, ahi, b[j + 8]);
			D2 = fma(alo, b[j ^ 2], fma(ahi, b[(j ^ 2) + 8], D2));  // slot 2: state bit 1
			D3 = fma(alo, b[j ^ 4], fma(ahi, b[(j ^ 4) + 8], D3));  // slot 3: state bit 2
			xbuf[j * 32 + lane] = make_double2(alo, ahi);
		}
	}
	if (!turns) return;
	__syncwarp();
	double D5 = 0, D6 = 0;
#pragma unroll
	for (int j = 0; j < 8; j++)
	{
		const double2 x1 = xbuf[j * 32 + (lane ^ 1)], x2 = xbuf[j * 32 + (lane ^ 2)];
		D5 = fma(x1.x, b[j], fma(x1.y, b[j + 8], D5));  // slot 5: state bit 4
		D6 = fma(x2.x, b[j], fma(x2.y, b[j + 8], D6));  // slot 6: state bit 5
	}
	auto pow2 = [&](int kbx) -> double {
		const int ke = kan + kbx + kf;
		// out of range below: the term underflows as exp() does in the reference; above (never in practice): its 1e300 cap (:5724)
		return ke < -1000 ? 0.0 : ke > 1000 ? 1e300 : __hiloint2double((1023 + ke) << 20, 0) * efrem;
	};
	// C0, C1 = C[s][2 gh], C[s][2 gh + 1]: column c is the unturned sum for c = s, the root's turn for c = s ^ 1, slot 1's for
	// c = s ^ 2, slot 4's for c = s ^ 4.  jj = gh ^ (s >> 1) says which pair this lane holds; the pair moves to lane position jj.
	{
		const int kA = __shfl_sync(FULL, kb, gh * 8), kB = __shfl_sync(FULL, kb, gh * 8 + 4);
		C0 *= pow2(kA), C1 *= pow2(kB);
	}
	const int jj = gh ^ (s >> 1);
	double pA = s0 ? C1 : C0, pB = s0 ? C0 : C1;  // pA: column with (c ^ s) even = 2 jj, pB: column 2 jj + 1
	if (jj) pB = 0.0;                             // c ^ s = 3, 5, 7: not a turn that is scored here
	if (jj == 3) pA = 0.0;                        // c ^ s = 6
	{
		const int src = (lane & ~3) | (gh ^ (s >> 1));  // position jj = gh of the receiving lane
		pA = __shfl_sync(FULL, pA, src);
		pB = __shfl_sync(FULL, pB, src);
	}
	// the four in-lane / in-shift sums over the four lanes of the shift mode: gh = 0 ends with D2, 1 with D5, 2 with D3, 3 with D6
	double pC;
	{
		const bool hi1 = gh & 1, hi2 = gh & 2;
		const double k0 = hi1 ? D5 : D2, t0 = hi1 ? D2 : D5, k1 = hi1 ? D6 : D3, t1 = hi1 ? D3 : D6;
		const double v0 = k0 + __shfl_xor_sync(FULL, t0, 1), v1 = k1 + __shfl_xor_sync(FULL, t1, 1);
		pC = (hi2 ? v1 : v0) + __shfl_xor_sync(FULL, hi2 ? v0 : v1, 2);
		pC *= pow2(kb);
	}
	// sum over the shift modes (those not in use carry alpha = 0); every group of four lanes ends with all totals:
	//   gh 0: (none, root, slot 2)   gh 1: (slot 1, -, slot 5)   gh 2: (slot 4, -, slot 3)   gh 3: (-, -, slot 6)
#pragma unroll
	for (int o = 4; o < 32; o <<= 1)
	{
		pA += __shfl_xor_sync(FULL, pA, o);
		pB += __shfl_xor_sync(FULL, pB, o);
		pC += __shfl_xor_sync(FULL, pC, o);
	}
	// one logarithm per lane: shift group 0 takes the first value of its position, group 1 the second, group 2 the third
	double v1 = s == 0 ? pA : s == 1 ? pB : pC;
	if (!isfinite(v1) || v1 < 1e-174) v1 = 1e-174;  // :3655
	const double lv = ms_log(v1);
	const double l0 = __shfl_sync(FULL, lv, 0);
	int slot = -1;  // the tree slot whose turn this lane holds
	if (s == 0) slot = gh == 1 ? 1 : gh == 2 ? 4 : -1;
	else if (s == 1) slot = gh == 0 ? 0 : -1;
	else if (s == 2) slot = gh == 0 ? 2 : gh == 1 ? 5 : gh == 2 ? 3 : 6;
	if (slot >= 0)
	{
		double* const p = w.nsptr[slot];
		if (p)
		{
			const bool scored = slot == 0 || !((u.flag2ignore >> slot) & 1);  // `g & (flag2ignore >> 1)`, :5684
#ifndef CNF2_X_NONSATOMIC
			atomicAdd(p + m, (l0 - (scored ? lv : 0.0)) * u.nsscale[slot]);
#endif
		}
	}
}

// Scale the 64 entries of each shift mode (4 lanes x 16) by the power of two that brings the largest into [1, 2).
__device__ __forceinline__ void rescale16(double (&p)[16], int& k)
{
	int e = 0;
#pragma unroll
	for (int i = 0; i < 16; i++) e = max(e, __double2hiint(p[i]));  // non-negative doubles order like their high words
	e = max(e, __shfl_xor_sync(0xffffffffu, e, 1));
	e = max(e, __shfl_xor_sync(0xffffffffu, e, 2));
	const int w = e >> 20;
	if (w == 0) return;  // all zero
	const double sc = __hiloint2double((2046 - w) << 20, 0);
#pragma unroll
	for (int i = 0; i < 16; i++) p[i] *= sc;
	k += w - 1023;
}

// scratch layout of one marker's alpha: chunk j (two consecutive states) of lane l at double2 index j * 32 + l
__device__ __forceinline__ void store_alpha(double2* __restrict__ dst, const double (&p)[16], int lane)
{
#pragma unroll
	for (int j = 0; j < 8; j++) dst[j * 32 + lane] = make_double2(p[2 * j], p[2 * j + 1]);
}
__device__ __forceinline__ void load_alpha(const double2* __restrict__ src, double (&p)[16], int lane)
{
#pragma unroll
	for (int j = 0; j < 8; j++)
	{
		const double2 v = src[j * 32 + lane];
		p[2 * j] = v.x;
		p[2 * j + 1] = v.y;
	}
}

// Raw marker data of one (slot, marker) pair, held in registers between its load and its use.
struct RawItem
{
	uchar2 al;
	double2 su;
	double h;
};
// lane i < NSLOTS * WB loads (slot = i / WB, t = i % WB) of markers [m0, m0 + n)
__device__ __forceinline__ RawItem raw_load(const Unit& u, const EstepArgs& a, int m0, int n, int lane)
{
	RawItem r;
	r.al = make_uchar2(0, 0);
	r.su = make_double2(0, 0);
	r.h = 0;
	const int slot = lane / WB, t = lane % WB;
	if (slot < NSLOTS && t < n && m0 >= 0)
	{
		const int row = u.row[slot];
		if (row >= 0)
		{
			const size_t o = (size_t)row * a.ld + m0 + t;
			r.al = a.alleles[o];
			r.su = a.sure[o];
			r.h = a.hw[o];
		}
	}
	return r;
}
// writes the staged item; returns whether it is outside the regular range (markersure in [0,1), haploweight in [0,1])
__device__ __forceinline__ bool raw_commit(WarpSmem& w, const RawItem& r, int lane)
{
	const int slot = lane / WB, t = lane % WB;
	if (slot >= NSLOTS) return false;
	const int i = slot * WB + t;
	w.A[i] = r.al;
	w.nt[i] = make_nodetab_fast(r.al, r.su, r.h);
	w.HW[i] = r.h;
	return !(r.su.x >= 0 && r.su.x < 1 && r.su.y >= 0 && r.su.y < 1 && r.h >= 0 && r.h <= 1);
}

// Lists of the posterior phase for a marker whose shift-independent rule-3 mask is gpmust (see batch_prepare):
// the needed configurations of parent 0, then of parent 1, each with the phase bit x & 1 = 0 ones first; per entry
// the W row offset of its meiosis bits and its x; per (entry, root orientation) the rule-2 mask over the other side's x.
__device__ __forceinline__ void warp_lists(WarpSmem& w, const Unit& u, int gpmust, int lane)
{
	const unsigned FULL = 0xffffffffu;
	int nE = 0, nE0 = 0, nX[2] = { 0, 0 };
#pragma unroll
	for (int r8 = 0; r8 < 8; r8++)
	{
		// candidates in meiosis-bits-major order, so that the entries of one (phase bit, meiosis bits) class are consecutive
		const int p = r8 >> 2, xb = (r8 >> 1) & 1, half = r8 & 1, gc = half * 4 + (lane >> 3), cfg = gc | (lane & 7) << 3;
		const int sub = (gpmust >> (1 + 3 * p)) & 7, xc = (cfg & 7) ^ (cfg >> 3);
		const bool ok = ((u.cfgmask[p] >> cfg) & 1ull) && (xc & sub) == sub && (xc & 1) == xb;
		const unsigned mask = __ballot_sync(FULL, ok);
		if (ok)
		{
			const int pos = nE + __popc(mask & ((1u << lane) - 1u));
			w.el[pos] = (unsigned char)(p << 6 | cfg);
			w.ainfo[pos] = (unsigned short)(ws_base((cfg & 7) << (p ? 3 : 0)) | xc << 10);
		}
		nE += __popc(mask);
		if (!xb) nX[p] += __popc(mask);
		if (r8 == 3) nE0 = nE;  // entries of parent 0 come first
	}
	__syncwarp();
	for (int i = lane; i < 2 * nE; i += 32)
	{
		const int c = w.el[i >> 1], f = i & 1, p = c >> 6, xown = (c & 7) ^ ((c & 63) >> 3);
		int vm = 0;
#pragma unroll
		for (int xo = 0; xo < 8; xo++)
		{
			const int x = f | ((p ? xo : xown) << 1) | ((p ? xown : xo) << 4);
			vm |= ((u.xmask2[x >> 5] >> (x & 31)) & 1) << xo;
		}
		w.vmk[i] = (unsigned char)vm;
	}
	if (lane == 0)
	{
		w.nE = (short)nE;
		w.nE0 = (short)nE0;
		w.nX0 = (short)nX[0];
		w.nX1 = (short)nX[1];
		w.listtag = gpmust;
	}
	__syncwarp();
}

struct WarpArgs
{
	EstepArgs e;             // e.alpha: alpha ring [n_warps][seg][512] doubles in store_alpha's layout (alpha before each marker's emission)
	int32_t* kexp;           // [n_warps][seg][NUMSHIFTS] power-of-two exponent of the stored alpha
	double* qstash;          // [n_warps][seg][QS] sweep-emission tables (forward pass -> backward pass)
	double* ckalpha;         // [n_warps][nseg][512] checkpoints: alpha before the first marker of segments 1.. (slot 0 unused)
	int32_t* ckexp;          // [n_warps][nseg][NUMSHIFTS]
	int32_t seg, nseg;       // markers per segment (a multiple of WB), segments per chromosome
	int32_t* literal_units;  // units the literal kernel has to do
	unsigned int* literal_count;
	int32_t cap0, cap1;      // capacity (entries) of the posterior tables of parent 0 / parent 1
	int32_t cls;             // CLS when the class sums are in use (some unit has an all-ones rule-2 mask), else 0
	int32_t turns;           // score the phase-inversion turns (HAPLOTYPING && !early && !full, :5668)
};
constexpr size_t XBUF_BYTES = 8 * 32 * sizeof(double2);  // warp_w_and_turns' exchange buffer lives in the posterior tables' space
static_assert((FT_DOUBLES + CLS) * sizeof(double) <= XBUF_BYTES, "the factorised posterior's tables fit the minimum table space");
__host__ __device__ inline size_t warp_smem_bytes(int cap0, int cap1, int cls)
{
	// (the factorised posterior of uncoupled units keeps its factors and class sums, FT_DOUBLES + CLS doubles, in the same space)
	const size_t tables = ((size_t)cap0 * 2 * T0W + (size_t)cap1 * 2 * T1W + cls) * sizeof(double);
	return sizeof(WarpSmem) + (tables > XBUF_BYTES ? tables : XBUF_BYTES);
}

// AGG: built with the factorised posterior for units without rule-2 coupling (picked by the host when the pedigree has any)
//
// Memory of the sweeps (row a5): the forward pass keeps alpha only at the start of every segment of `seg` markers (a
// checkpoint: 4 KB per segment) and, for the chromosome's LAST segment, for every marker in a per-warp ring of `seg`
// entries.  The backward pass walks the segments from the last to the first; before it enters a segment it re-runs the
// forward sweep over it from the checkpoint, filling the ring.  With seg >= Mc that is the plain store-everything
// schedule (one segment, nothing recomputed); the host picks seg so that the ring fits its budget whatever Mc is.
template <bool AGG>
__global__ void __launch_bounds__(32, CNF2_WARPS_PER_SM) warp_estep_kernel(WarpArgs wa)
{
	extern __shared__ __align__(16) unsigned char smem_raw[];
	WarpSmem& w = *reinterpret_cast<WarpSmem*>(smem_raw);
	double* const tab0 = reinterpret_cast<double*>(smem_raw + sizeof(WarpSmem));
	double* const tab1 = tab0 + (size_t)wa.cap0 * 2 * T0W;
	double2* const xbuf = reinterpret_cast<double2*>(tab0);
	const EstepArgs& a = wa.e;
	const unsigned FULL = 0xffffffffu;
	const int lane = threadIdx.x, s = lane >> 2;
	const int Mc = a.Mc, E = Mc - 1, SEG = wa.seg, nseg = wa.nseg;
	constexpr int AV = NUMSHIFTS * NUMTYPES / 2;  // double2 per stored alpha
	double2* const alpha = reinterpret_cast<double2*>(a.alpha) + (size_t)blockIdx.x * SEG * AV;
	int32_t* const kexp = wa.kexp + (size_t)blockIdx.x * SEG * NUMSHIFTS;
	double* const qs = wa.qstash + (size_t)blockIdx.x * SEG * QS;
	double2* const ckalpha = reinterpret_cast<double2*>(wa.ckalpha) + (size_t)blockIdx.x * nseg * AV;
	int32_t* const ckexp = wa.ckexp + (size_t)blockIdx.x * nseg * NUMSHIFTS;

	for (int i = lane; i < NCOL * NACCP; i += 32) w.acc[i] = 0;

	for (;;)
	{
		int ui = 0;
		if (lane == 0) ui = (int)atomicAdd(a.counter, 1u);
		ui = __shfl_sync(FULL, ui, 0);
		if (ui >= a.n_units) break;
		__syncwarp();
		{
			const int* src = reinterpret_cast<const int*>(a.units + ui);
			int* dst = reinterpret_cast<int*>(&w.unit);
			for (int i = lane; i < (int)(sizeof(Unit) / 4); i += 32) dst[i] = src[i];
			if (lane == 0) w.listtag = -1;
		}
		__syncwarp();
		const Unit& u = w.unit;
		transfer_setup(w, a, u, lane);
		const bool agg = AGG && wa.cls != 0 && (u.xmask2[0] & u.xmask2[1] & u.xmask2[2] & u.xmask2[3]) == 0xffffffffu;
		bool literal = u.founder[0] != 0;  // a root "at the top" never recurses (:1120): literal kernel
		const bool active = s < u.shiftend && !(s & u.shiftignore);

		// the forward sweep over markers [mb, me) from (p, ka) (realanalyze<STORE|FORWARD>, cnF2freq.cpp:2086-2107); stash: alpha
		// before each marker, its exponents and the marker's sweep table go to ring entry (marker - mb); check: parameters outside
		// the regular range end the sweep (the unit is the literal kernel's)
		double p[16];  // the forward vector; in the backward pass: alpha before the current marker's emission
		int ka = 0;
		auto forward = [&](const int mb, const int me, const bool stash, const bool check) -> bool {
			RawItem raw = raw_load(u, a, mb, min(WB, me - mb), lane);
			for (int m0 = mb; m0 < me; m0 += WB)
			{
				const int n = min(WB, me - m0);
				__syncwarp();
				const bool bad = raw_commit(w, raw, lane);
				if (m0 + WB < me) raw = raw_load(u, a, m0 + WB, min(WB, me - m0 - WB), lane);
				if (check && __any_sync(FULL, bad)) return false;
				__syncwarp();
				if ((lane >> 3) < n) sweep_family<WB>(u, w.A, w.nt, lane >> 3, lane & 7, w.qh + (lane >> 3) * QS);
				__syncwarp();
				if (stash)
					for (int i = lane; i < n * QS; i += 32) qs[(size_t)(m0 - mb) * QS + i] = w.qh[i];
				for (int t = 0; t < n; t++)
				{
					const int m = m0 + t;
					double r0 = 0, r1 = 0;  // asked for ahead of the stores below, used after the emission
					if (m < E) r0 = a.rec0[m], r1 = a.rec1[m];
					if (stash)
					{
						store_alpha(alpha + (size_t)(m - mb) * AV, p, lane);
						if ((lane & 3) == 0) kexp[(size_t)(m - mb) * NUMSHIFTS + s] = ka;
					}
					emit16(p, w.qh + t * QS, lane);
					rescale16(p, ka);
					if (m < E) transition16(p, r0, r1);
				}
			}
			return true;
		};

		// ---- forward pass (initfwbw, :2086-2107): checkpoints at the segment starts, the last segment into the ring ----
#pragma unroll
		for (int i = 0; i < 16; i++) p[i] = active ? 1.0 / NUMTYPES : 0.0;  // EVENGEN, :2100-2104
		for (int sg = 0; sg < nseg && !literal; sg++)
		{
			if (sg > 0)
			{
				store_alpha(ckalpha + (size_t)sg * AV, p, lane);
				if ((lane & 3) == 0) ckexp[sg * NUMSHIFTS + s] = ka;
			}
			if (!forward(sg * SEG, min(Mc, (sg + 1) * SEG), sg == nseg - 1, true)) literal = true;
		}
		if (literal)
		{
			if (lane == 0) wa.literal_units[atomicAdd(wa.literal_count, 1u)] = ui;
			continue;
		}
		// doanalyze / quickanalyze with NONESTOP (:1984-2018): log-factor of alpha+ at the last marker, per shift mode
		double L;
		{
			double sum = 0;
#pragma unroll
			for (int i = 0; i < 16; i++) sum += p[i];
			sum += __shfl_xor_sync(FULL, sum, 1);
			sum += __shfl_xor_sync(FULL, sum, 2);
			L = sum > 0 ? (double)ka * LN2 + log(sum) : MINFACTOR_D;
		}
		double factor = -1e15, realfactor = 0;  // :5373-5403
		{
			const unsigned actmask = __ballot_sync(FULL, active);
#pragma unroll
			for (int j = 0; j < NUMSHIFTS; j++)
			{
				const double Lj = __shfl_sync(FULL, L, j * 4);
				if ((actmask >> (j * 4)) & 1) factor = fmax(factor, Lj);
			}
#pragma unroll
			for (int j = 0; j < NUMSHIFTS; j++)
			{
				const double Lj = __shfl_sync(FULL, L, j * 4);
				if ((actmask >> (j * 4)) & 1) realfactor += exp(Lj - factor);
			}
		}
		const double maxfactor = factor;
		if (!(maxfactor > MINFACTOR_D))
		{
			// impossible under every shift mode: the reference goes on with factor = MINFACTOR + log(n) and scatters the
			// NaNs of its 0/0 terms (:5403 compares against a float constant); the literal kernel reproduces that
			if (lane == 0) wa.literal_units[atomicAdd(wa.literal_count, 1u)] = ui;
			continue;
		}
		factor += log(realfactor);
		if (lane == 0)
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

		// ---- backward sweep (:2109-2116), the posterior of each marker right behind it ----
		double b[16];
#pragma unroll
		for (int i = 0; i < 16; i++) b[i] = 1.0;  // :2111-2114
		int kb = 0;
		for (int sg = nseg - 1; sg >= 0; sg--)
		{
			const int mb = sg * SEG, me = min(Mc, mb + SEG);
			if (sg != nseg - 1)
			{
				// the segment's alphas are not in the ring: forward again from its checkpoint
				if (sg > 0)
				{
					load_alpha(ckalpha + (size_t)sg * AV, p, lane);
					ka = ckexp[sg * NUMSHIFTS + s];
				}
				else
				{
#pragma unroll
					for (int i = 0; i < 16; i++) p[i] = active ? 1.0 / NUMTYPES : 0.0;
					ka = 0;
				}
				__syncwarp();
				forward(mb, me, true, false);
			}
			const int mlast = mb + ((me - 1 - mb) / WB) * WB;
			RawItem raw = raw_load(u, a, mlast, me - mlast, lane);
			load_alpha(alpha + (size_t)(me - 1 - mb) * AV, p, lane);
			int kan = kexp[(size_t)(me - 1 - mb) * NUMSHIFTS + s];
			for (int m0 = mlast; m0 >= mb; m0 -= WB)
			{
				const int n = min(WB, me - m0);
				__syncwarp();
				raw_commit(w, raw, lane);
				for (int i = lane; i < n * QS; i += 32) w.qh[i] = qs[(size_t)(m0 - mb) * QS + i];
				// the stash and the alphas of the units in flight are far larger than L2: ask for the next block's lines now
				if (m0 >= mb + WB && lane * 16 < WB * QS) asm volatile("prefetch.global.L2 [%0];" ::"l"(qs + (size_t)(m0 - mb - WB) * QS + lane * 16));
				// homozygous with equal sureness (:1235, :3488), per (slot, marker) of the batch
				const bool homi = (lane / WB) < NSLOTS && u.row[(lane / WB) % NSLOTS] >= 0 && raw.al.x == raw.al.y && raw.su.x == raw.su.y;
				const unsigned homb = __ballot_sync(FULL, homi);
				if (m0 >= mb + WB) raw = raw_load(u, a, m0 - WB, WB, lane);
				if (lane < WB)
				{
					int hs = 0, hom = 0, gp = 0;
					for (int k = 0; k < NSLOTS; k++)
						if ((homb >> (k * WB + lane)) & 1)
						{
							hs |= 1 << k;
							hom |= 1 << u.slotrel[k];
						}
					for (int i = 0; i < u.nrel; i++)  // rule 3 (:3488-3493) of the relatives whose rule does not involve the shift mode
						if (((hom >> i) & 1) && u.relshift[i] == 0) gp |= u.relfilter[i];
					w.homslot[lane] = hs;
					w.gpmust[lane] = gp;
				}
				__syncwarp();
				for (int t = n - 1; t >= 0; t--)
				{
					const int m = m0 + t;
					// W = alpha beta exp(F- + B - factor) (:1963-2018, :5499-5508); the factors are k ln 2 here
					{
						double sc = 0;
						const int ke = kan + kb + kf;
						if (contributing && !((double)kan * LN2 + 5.0 < minfactor) && ke > -860 && ke < 860)
							sc = __hiloint2double((1023 + ke) << 20, 0) * efrem;
#ifdef CNF2_X_NOTURNS
						warp_w_and_turns(w, u, p, b, sc, kan, kb, kf, efrem, w.qh + t * QS, xbuf, m, lane, false);
#else
						warp_w_and_turns(w, u, p, b, sc, kan, kb, kf, efrem, w.qh + t * QS, xbuf, m, lane, wa.turns != 0);
#endif
					}
					if (m > mb + WB) asm volatile("prefetch.global.L2 [%0];" ::"l"(alpha + (size_t)(m - 1 - WB - mb) * AV + lane * 8));
					const int gp = w.gpmust[t];
					if (!agg && gp != w.listtag) warp_lists(w, u, gp, lane);
					if (lane < 2) root_constants(w.A[t], w.nt[t], lane, w.rootc);
					else if (lane < 4)
					{
						const int sp = lane == 2 ? 1 : 4;
						const bool there = u.row[sp] >= 0;
						w.rootc[12 + (lane - 2) * 2 + 0] = there ? w.nt[sp * WB + t].W[0] : 1.0;
						w.rootc[12 + (lane - 2) * 2 + 1] = there ? w.nt[sp * WB + t].W[1] : 1.0;
					}
					__syncwarp();
#ifndef CNF2_X_NOTAB
if (!agg) warp_tables_from_factors(w, u, t, lane, tab0, tab1);
#endif
					__syncwarp();
#ifndef CNF2_X_NOPAIRS
					if (agg) warp_posterior_factored(w, u, w.homslot[t], gp, t, lane, tab0);
					else warp_posterior_pairs(w, u, w.homslot[t], lane, tab0, tab1);
#endif
					__syncwarp();
#ifndef CNF2_X_NOREDUCE
					// the warp reduces its accumulator columns; the reader clears what it read
					for (int k = lane; k < NACC; k += 32)
					{
						double* col = w.acc + k;
						double v0 = 0, v1 = 0;
#pragma unroll
						for (int j = 0; j < NCOL; j += 2)
						{
							v0 += col[j * NACCP], v1 += col[(j + 1) * NACCP];
							col[j * NACCP] = 0, col[(j + 1) * NACCP] = 0;
						}
						w.tot[k] = v0 + v1;
					}
#endif
					__syncwarp();
					double r0 = 0, r1 = 0;  // for the step to marker m - 1, asked for ahead of the transfer
					if (m > 0) r0 = a.rec0[m - 1], r1 = a.rec1[m - 1];
#ifndef CNF2_X_NOTRANSFER
					// transfer into the EM sufficient statistics (:5875-5904, moveinfprobs :3577-3597, movehaplos :3599-3616)
					transfer_out(w, w.tot, t, m, lane);
#endif
					// alpha before the next marker to the left: asked for here, behind the posterior (whose loops need the registers),
					// and in flight while beta takes its step
					if (m > mb)
					{
						load_alpha(alpha + (size_t)(m - 1 - mb) * AV, p, lane);
						kan = kexp[(size_t)(m - 1 - mb) * NUMSHIFTS + s];
					}
					if (m > 0)
					{
						emit16(b, w.qh + t * QS, lane);
						rescale16(b, kb);
						transition16(b, r0, r1);
					}
					__syncwarp();
				}
			}
		}
	}
}
}  // namespace cnf2
