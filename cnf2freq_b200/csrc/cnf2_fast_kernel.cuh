// The table-driven E-step kernel (the product path).  Same work decomposition as the literal kernel in
// cnf2_estep_kernel.cuh -- one persistent CTA per unit at a time, one warp per shift mode, two states per
// lane -- but the emission routine is never evaluated per term:
//
//  * With a fixed orientation flag2, trackpossible (cnF2freq.cpp:1075-1359) is a product of a root-level
//    factor and the two values recursetrackpossible returns for the parents' subtrees.  A subtree value
//    depends only on what the root hands it (inval, second) and on the subtree's own three meiosis bits and
//    three orientation bits ("local configuration").  Per (marker, shift) the warp tabulates those values
//    for the 8 + 10 distinct root messages of the posterior loop (plain emission, the four GENOSPROBE calls,
//    the two HOMOZYGOUS calls, :5512-5538) and for the local configurations that actually occur, by calling
//    the literal subtree routine track<0,2>; each surviving term is then 7 three-factor products.
//  * With flag2 = -1 (the sweeps) the emission is sum_{root orientation} root factor x Q1[g>>3] x Q0[g&7]
//    with 32 subtree values per (marker, shift): one literal call per lane.  The forward sweep stores them
//    for the backward sweep.
//  * ignoreflag2 (:3462-3496) depends on (g, flag2) only through x = flag2 ^ (g << 1) (rules 2 and 3) and
//    through flag2 & flag2ignore (rule 1), which pins the state bits of absent/empty slots to the bits of x.
//    The warp lists the valid x once per marker (ballot) and enumerates the free state bits densely, so
//    every lane always holds a surviving term.
//  * Scatter: HAPLOS targets are bits of x ^ shiftmask; GENOS targets are the chain root -> parent(side) ->
//    grandparent(side, meiosis bit).  For markers whose markersure lie in [0,1) and haploweights in [0,1] a
//    subtree value is zero in update mode exactly when it is zero in plain mode (all terms are non-negative
//    and affine in `second`), so the `if (base)` gates of :1343-1355 are implied by val > 0.  Units with any
//    marker outside that range, and units whose root is "at the top" (founder), are handed to the literal
//    kernel through a device-side list.
#pragma once
#include "cnf2_estep_kernel.cuh"

namespace cnf2
{
constexpr int NMSG0 = 8, NMSG1 = 10, NKIND = 7;
constexpr double EXP_M200 = 1.3838965267367376e-87;  // exp(-200): unusualstate, cnF2freq.cpp:5414

struct FastWarp
{
	double P0[NMSG0 * NUMTYPES];  // parent-0 subtree value [f2root*4 + j][local cfg]
	double P1[NMSG1 * NUMTYPES];  // parent-1 subtree value [f2root*5 + j][local cfg]
	double beta[NUMTYPES];
	double rbase[2 * NKIND];      // root-level factor [f2root][kind]
	double msec0[NMSG0], msec1[NMSG1];
	int minv0[NMSG0], minv1[NMSG1];
	unsigned need[4];             // local configurations in use: parent 0 bits 0..63, parent 1 bits 64..127
	unsigned char xlist[NUMPATHS];
	unsigned char cfglist[2 * NUMTYPES];
};

struct __align__(16) FastSmem
{
	double acc[NACC * NTHREADS];
	FastWarp w[NUMSHIFTS];
	NodeTab N[NSLOTS * TILE];
	double2 S[NSLOTS * TILE];
	double HW[NSLOTS * TILE];
	double tot[NACC];
	double Ls[NUMSHIFTS];
	uchar2 A[NSLOTS * TILE];
	Unit unit;
	int32_t unit_index;
};

struct FastArgs
{
	EstepArgs e;
	double* qscratch;        // [n_ctas][Mc][NUMSHIFTS][32] subtree values of the sweeps
	int32_t* literal_units;  // units the literal kernel has to do
	unsigned int* literal_count;
};

// value recursetrackpossible returns for parent p's subtree (cnF2freq.cpp:955-1058)
__device__ __forceinline__ double subtree(const FCtx& c, int p, int inval, double second, int t, int fp3, int f23, int shift)
{
	const int sp = 1 + 3 * p;
	if (c.row[sp] < 0) return 1.0 + second;
	return ftrack<2>(c, sp, inval, second, t, fp3, f23, (shift >> (1 + p)) & 1);
}

// emission with all orientations summed for the two states of this lane, from the 32 subtree values
// (lane = f2root*16 + parent*8 + meiosis bits) and the root factors
__device__ __forceinline__ void combine_q(double q, double rb, int lane, double& e0, double& e1)
{
	const unsigned F = 0xffffffffu;
	const double R0 = __shfl_sync(F, rb, 0), R1 = __shfl_sync(F, rb, 16);
	const double q0a = __shfl_sync(F, q, lane & 7), q0b = __shfl_sync(F, q, 16 | (lane & 7));
	const double q1a = __shfl_sync(F, q, 8 | (lane >> 3)), q1b = __shfl_sync(F, q, 24 | (lane >> 3));
	const double q1c = __shfl_sync(F, q, 12 | (lane >> 3)), q1d = __shfl_sync(F, q, 28 | (lane >> 3));
	e0 = R0 * q1a * q0a + R1 * q1b * q0b;
	e1 = R0 * q1c * q0a + R1 * q1d * q0b;
}

// the lane's item of the sweep emission: root factor of its root orientation and one subtree value
__device__ __forceinline__ void sweep_item(const FCtx& c, int t, int lane, int shift, double& q, double& rb)
{
	const int f = lane >> 4, p = (lane >> 3) & 1, fp3 = lane & 7;
	const RootLocal r = froot(c.A[t], c.N[t], 0, UNKNOWN, 0, f, shift & 1);
	rb = r.base;
	q = p ? subtree(c, 1, r.secmark, r.sec2, t, fp3, -1, shift) : subtree(c, 0, r.val, r.main2, t, fp3, -1, shift);
}

__device__ __forceinline__ void apply_emission(double e0, double e1, double& p0, double& p1, double& F)
{
	p0 = (p0 < 1e-300) ? 0.0 : p0 * e0;
	p1 = (p1 < 1e-300) ? 0.0 : p1 * e1;
	const double sum = warp_sum(p0 + p1);
	if (sum <= 0) F = MINFACTOR_D;
	else
	{
		p0 /= sum;
		p1 /= sum;
		F += log(sum);
	}
}

// Tile load with the regularity test of the fast path; returns true (to every thread) when some visited
// slot has a markersure outside [0,1) or a haploweight outside [0,1] in this tile.
__device__ __forceinline__ bool load_tile_checked(FastSmem& sm, const EstepArgs& a, int m0, int tid)
{
	int bad = 0;
	for (int i = tid; i < NSLOTS * TILE; i += NTHREADS)
	{
		const int slot = i / TILE, t = i % TILE;
		const int row = sm.unit.row[slot];
		const int m = m0 + t;
		uchar2 al = make_uchar2(0, 0);
		double2 su = make_double2(0, 0);
		double h = 0;
		if (row >= 0 && m < a.Mc)
		{
			const size_t o = (size_t)row * a.ld + m;
			al = a.alleles[o];
			su = a.sure[o];
			h = a.hw[o];
			if (!(su.x >= 0 && su.x < 1 && su.y >= 0 && su.y < 1 && h >= 0 && h <= 1)) bad = 1;
		}
		sm.A[i] = al;
		sm.S[i] = su;
		sm.HW[i] = h;
		sm.N[i] = make_nodetab(al, su, h);
	}
	return __syncthreads_or(bad) != 0;
}

// Posterior of one (marker, shift): tables, dense term enumeration, scatter into the accumulator columns.
__device__ __forceinline__ void fast_posterior(FastSmem& sm, FastWarp& w, const FCtx& c, double* acc, const Unit& u, int t,
                                               int lane, int shift, const double* __restrict__ alpha_m, double scale)
{
	const unsigned FULL = 0xffffffffu;
	// ---- per-marker masks: relatives / slots that are homozygous with equal sureness (:1235, :3488) ----
	int hom = 0, homslot = 0;
	for (int k = 0; k < NSLOTS; k++)
	{
		if (u.row[k] < 0) continue;
		const uchar2 al = sm.A[k * TILE + t];
		const double2 su = sm.S[k * TILE + t];
		if (al.x == al.y && su.x == su.y)
		{
			homslot |= 1 << k;
			hom |= 1 << u.slotrel[k];
		}
	}
	// ---- root-level factors and the messages handed to the parents' subtrees ----
	if (lane < 2 * NKIND)
	{
		const int f = lane / NKIND, kind = lane % NKIND;
		const uchar2 a0 = sm.A[t];
		const NodeTab& n0 = sm.N[t];
		const int al = 1 + ((kind - 1) & 1);  // kinds 1,3,5 probe allele 1; kinds 2,4,6 allele 2
		RootLocal r;
		if (kind == 0) r = froot(a0, n0, 0, UNKNOWN, 0, f, shift & 1);
		else if (kind <= 2) r = froot(a0, n0, 1, al, 0, f, shift & 1);
		else if (kind <= 4) r = froot(a0, n0, 1, al, 1, f ^ 1, shift & 1);
		else r = froot(a0, n0, 2, al, 0, f, shift & 1);
		w.rbase[lane] = r.base;
		if (kind == 0)
		{
			w.minv0[f * 4] = r.val, w.msec0[f * 4] = r.main2;
			w.minv1[f * 5] = r.secmark, w.msec1[f * 5] = r.sec2;
		}
		else if (kind <= 2) w.minv0[f * 4 + kind] = r.val, w.msec0[f * 4 + kind] = r.main2;
		else if (kind <= 4)
		{
			w.minv1[f * 5 + kind - 2] = r.val, w.msec1[f * 5 + kind - 2] = r.main2;
			if (kind == 3) w.minv0[f * 4 + 3] = r.secmark, w.msec0[f * 4 + 3] = r.sec2;
		}
		else w.minv1[f * 5 + kind - 2] = r.secmark, w.msec1[f * 5 + kind - 2] = r.sec2;
	}
	if (lane < 4) w.need[lane] = 0;
	// ---- valid x = flag2 ^ (g << 1): rules 2 and 3 of ignoreflag2 (:3483-3493) ----
	int nX = 0;
#pragma unroll
	for (int r4 = 0; r4 < NUMPATHS / 32; r4++)
	{
		const int x = r4 * 32 + lane;
		bool ok = true;
		for (int i = 0; i < u.nrel; i++)
		{
			const int filt = u.relfilter[i];
			if (!filt) continue;
			const int filtered = x & filt;
			if (filtered && filtered != filt) ok = false;
			if (((hom >> i) & 1) && !((filtered != 0) ^ ((shift & u.relshift[i]) != 0))) ok = false;
		}
		const unsigned mask = __ballot_sync(FULL, ok);
		if (ok) w.xlist[nX + __popc(mask & ((1u << lane) - 1u))] = (unsigned char)x;
		nX += __popc(mask);
	}
	__syncwarp();
	const int nfree = u.nfree, gfix = u.gfixmask;
	const int T = nX << nfree;
	// ---- which local configurations occur ----
	for (int i = lane; i < T; i += 32)
	{
		const int x = w.xlist[i >> nfree];
		const int g = ((x >> 1) & gfix) | u.expand[i & ((1 << nfree) - 1)];
		const int flag2 = x ^ (g << 1);
		const int cfg0 = (g & 7) | (((flag2 >> 1) & 7) << 3), cfg1 = ((g >> 3) & 7) | (((flag2 >> 4) & 7) << 3);
		atomicOr(&w.need[cfg0 >> 5], 1u << (cfg0 & 31));
		atomicOr(&w.need[2 + (cfg1 >> 5)], 1u << (cfg1 & 31));
	}
	__syncwarp();
	int n0 = 0, n1 = 0;
#pragma unroll
	for (int r4 = 0; r4 < 4; r4++)
	{
		const unsigned word = w.need[r4];
		const bool on = (word >> lane) & 1;
		int& n = (r4 < 2) ? n0 : n1;
		if (on) w.cfglist[(r4 < 2 ? 0 : NUMTYPES) + n + __popc(word & ((1u << lane) - 1u))] = (unsigned char)((r4 & 1) * 32 + lane);
		n += __popc(word);
	}
	__syncwarp();
	// ---- subtree tables ----
	const int items0 = n0 * NMSG0, items = items0 + n1 * NMSG1;
	for (int i = lane; i < items; i += 32)
	{
		if (i < items0)
		{
			const int cfg = w.cfglist[i >> 3], msg = i & 7;
			w.P0[msg * NUMTYPES + cfg] = subtree(c, 0, w.minv0[msg], w.msec0[msg], t, cfg & 7, cfg >> 3, shift);
		}
		else
		{
			const int i1 = i - items0;
			const int cfg = w.cfglist[NUMTYPES + i1 / NMSG1], msg = i1 % NMSG1;
			w.P1[msg * NUMTYPES + cfg] = subtree(c, 1, w.minv1[msg], w.msec1[msg], t, cfg & 7, cfg >> 3, shift);
		}
	}
	__syncwarp();
	// ---- terms ----
	const int smask = (shift & 1) | ((shift & 2) ? 2 : 0) | ((shift & 4) ? 16 : 0);  // shift bit of slots 0, 1, 4
	const int vis = u.vismask, hmask = vis & ~homslot;
	for (int i = lane; i < T; i += 32)
	{
		const int x = w.xlist[i >> nfree];
		const int g = ((x >> 1) & gfix) | u.expand[i & ((1 << nfree) - 1)];
		const int flag2 = x ^ (g << 1);
		const double am = alpha_m[g];
		if (am < 1e-300) continue;
		const int f = flag2 & 1;
		const int cfg0 = (g & 7) | (((flag2 >> 1) & 7) << 3), cfg1 = ((g >> 3) & 7) | (((flag2 >> 4) & 7) << 3);
		const double* rb = w.rbase + f * NKIND;
		const double* P0 = w.P0 + (f * 4) * NUMTYPES + cfg0;
		const double* P1 = w.P1 + (f * 5) * NUMTYPES + cfg1;
		const double pe0 = P0[0], pe1 = P1[0];
		const double e = rb[0] * pe1 * pe0;
		const double val = am * e * w.beta[g] * scale;
		if (!(val > EXP_M200)) continue;
		const double pa0 = P0[1 * NUMTYPES], pa1 = P0[2 * NUMTYPES], po0 = P0[3 * NUMTYPES];
		const double sv00 = rb[1] * pe1 * pa0, sv01 = rb[2] * pe1 * pa1;                                  // side 0 (:5519-5529)
		const double sv10 = rb[3] * po0 * P1[1 * NUMTYPES], sv11 = rb[4] * po0 * P1[2 * NUMTYPES];        // side 1
		const double hz0 = rb[5] * P1[3 * NUMTYPES] * pa0, hz1 = rb[6] * P1[4 * NUMTYPES] * pa1;          // :5532-5537
		const double r0 = val / (sv00 + sv01), r1 = val / (sv10 + sv11);
		// updatehaplo (:1561-1575): every visited, non-homozygous slot, at its effective phase bit
		const int y = x ^ smask;
#pragma unroll
		for (int k = 0; k < NSLOTS; k++)
			if ((hmask >> k) & 1) acc[(ACC_HAPLOS + k * 2 + ((y >> k) & 1)) * NTHREADS] += val;
		// the four GENOS calls (:5560-5568): root, parent `side`, that parent's parent picked by its meiosis bit
#pragma unroll
		for (int side = 0; side < 2; side++)
		{
			const double u0 = (side ? sv10 : sv00) * (side ? r1 : r0), u1 = (side ? sv11 : sv01) * (side ? r1 : r0);
			const int sp = 1 + 3 * side, gs = sp + 1 + ((g >> (3 * side)) & 1);
			double* a0 = acc + (ACC_INF + (f ^ side) * 2) * NTHREADS;
			a0[0] += u0;
			a0[NTHREADS] += u1;
			if ((vis >> sp) & 1)
			{
				double* ap = acc + (ACC_INF + (sp * 2 + ((flag2 >> sp) & 1)) * 2) * NTHREADS;
				ap[0] += u0;
				ap[NTHREADS] += u1;
				if ((vis >> gs) & 1)
				{
					double* ag = acc + (ACC_INF + (gs * 2 + ((flag2 >> gs) & 1)) * 2) * NTHREADS;
					ag[0] += u0;
					ag[NTHREADS] += u1;
				}
			}
		}
		acc[(ACC_HZ + 0) * NTHREADS] += hz0 * r0;  // :5571-5575
		acc[(ACC_HZ + 1) * NTHREADS] += hz1 * r0;
	}
	__syncwarp();
}

__global__ void __launch_bounds__(NTHREADS, 1) fast_estep_kernel(FastArgs fa)
{
	extern __shared__ __align__(16) unsigned char smem_raw[];
	FastSmem& sm = *reinterpret_cast<FastSmem*>(smem_raw);
	const EstepArgs& a = fa.e;
	const int tid = threadIdx.x, lane = tid & 31, shift = tid >> 5;
	const int Mc = a.Mc, E = Mc - 1;
	const int ntiles = (Mc + TILE - 1) / TILE;
	double* const alpha = a.alpha + (size_t)blockIdx.x * Mc * (NUMSHIFTS * NUMTYPES);
	double* const alphafac = a.alphafac + (size_t)blockIdx.x * Mc * NUMSHIFTS;
	double* const qs = fa.qscratch + (size_t)blockIdx.x * Mc * (NUMSHIFTS * 32);
	FCtx c{ sm.A, sm.N, sm.unit.row, sm.unit.founder };
	double* const acc = sm.acc + tid;
	FastWarp& w = sm.w[shift];

	for (int i = tid; i < NACC * NTHREADS; i += NTHREADS) sm.acc[i] = 0;

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

		// ---- forward sweep ----
		double p0 = 1.0 / NUMTYPES, p1 = 1.0 / NUMTYPES, F = 0, L = 0;
		for (int tile = 0; tile < ntiles && !literal; tile++)
		{
			__syncthreads();
			if (load_tile_checked(sm, a, tile * TILE, tid))
			{
				literal = true;
				break;
			}
			if (!active) continue;
			const int tend = min(TILE, Mc - tile * TILE);
			for (int t = 0; t < tend; t++)
			{
				const int m = tile * TILE + t;
				double* dst = alpha + ((size_t)m * NUMSHIFTS + shift) * NUMTYPES;
				dst[lane] = p0;
				dst[lane + 32] = p1;
				if (lane == 0) alphafac[(size_t)m * NUMSHIFTS + shift] = F;
				double q, rb, e0, e1;
				sweep_item(c, t, lane, shift, q, rb);
				qs[((size_t)m * NUMSHIFTS + shift) * 32 + lane] = q;
				combine_q(q, rb, lane, e0, e1);
				apply_emission(e0, e1, p0, p1, F);
				if (m < E) transition(p0, p1, a.rec0[m], a.rec1[m]);
			}
		}
		if (literal)
		{
			if (tid == 0) fa.literal_units[atomicAdd(fa.literal_count, 1u)] = ui;
			continue;
		}
		if (active)
		{
			L = F;
			const double sum = warp_sum(p0 + p1);
			if (sum <= 0) L = MINFACTOR_D;
			else L -= log(1.0 / sum);
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
		const float minfactor = (float)(-200.0 + factor);

		// ---- backward sweep fused with the posterior ----
		double b0 = 1.0, b1 = 1.0, B = 0;
		for (int tile = ntiles - 1; tile >= 0; tile--)
		{
			__syncthreads();
			load_tile_checked(sm, a, tile * TILE, tid);
			const int tend = min(TILE, Mc - tile * TILE);
			for (int t = tend - 1; t >= 0; t--)
			{
				const int m = tile * TILE + t;
				if (contributing)
				{
					const double Fm = alphafac[(size_t)m * NUMSHIFTS + shift];
					const double lscale = Fm + B - factor;
					if (!(Fm < minfactor) && lscale > -600.0 && lscale < 600.0)
					{
						w.beta[lane] = b0;
						w.beta[lane + 32] = b1;
						__syncwarp();
						fast_posterior(sm, w, c, acc, u, t, lane, shift, alpha + ((size_t)m * NUMSHIFTS + shift) * NUMTYPES, exp(lscale));
					}
				}
				__syncthreads();
				for (int k = shift; k < NACC; k += NUMSHIFTS)
				{
					double v = 0;
#pragma unroll
					for (int j = 0; j < NTHREADS / 32; j++)
					{
						double* p = &sm.acc[k * NTHREADS + j * 32 + lane];
						v += *p;
						*p = 0;
					}
					v = warp_sum(v);
					if (lane == 0) sm.tot[k] = v;
				}
				__syncthreads();
				if (tid < u.nrel) transfer_stats(a, u, sm.tot, sm.HW, tid, t, m);
				if (active && m > 0)
				{
					const double q = qs[((size_t)m * NUMSHIFTS + shift) * 32 + lane];
					const RootLocal r = froot(sm.A[t], sm.N[t], 0, UNKNOWN, 0, lane >> 4, shift & 1);
					double e0, e1;
					combine_q(q, r.base, lane, e0, e1);
					apply_emission(e0, e1, b0, b1, B);
					transition(b0, b1, a.rec0[m - 1], a.rec1[m - 1]);
				}
			}
		}
	}
}
}  // namespace cnf2
