// The E-step kernel: one persistent CTA per resident slot, each looping over (analysed individual,
// chromosome) units.  One warp per shift mode, two inheritance states per lane (bits 0-4 of the state
// across lanes, bit 5 inside the thread), so the Kronecker transition R(r)^{(x)6} (cnF2freq.cpp:2270-2367)
// is five __shfl_xor butterfly stages plus one in-register stage.
//
// estep_kernel<true> is the SELFING = true instantiation of the reference (settings.h:14; row g1 of the scope table): 256
// states = 64 x the value of two selfing bits, eight per lane (state = lane + 32 k: k & 1 is meiosis bit 5, k >> 1 the
// selfing value), the transition a Kronecker product of the six 2 x 2 matrices with the 3 x 3 selfing matrix selfprec
// (cnF2freq.cpp:2316-2326, 2352-2364), the prior of :2092-2104, the root's emission of :1169-1190, rule 3 of ignoreflag2 off
// (:3489).  It is the only kernel built for that setting: every unit goes through it.
#pragma once
#include "cnf2_device.cuh"

namespace cnf2
{
struct EstepArgs
{
	const uchar2* alleles;   // [n_rows][ld]
	const double2* sure;     // [n_rows][ld]
	const double* hw;        // [n_rows][ld]
	const double* rec0;      // [Mc] recombination fraction towards the next marker, genrec[0] (0 = same position)
	const double* rec1;      // [Mc] same with genrec[1]
	const double* dist;      // [Mc] SELFING only: distance to the next marker (cM), for the selfing generations' rate (:2286)
	double genrec2;          // SELFING only: genrec[2] (:7943)
	const Unit* units;
	int32_t n_units;
	int32_t Mc;              // markers of this chromosome
	int32_t ld;              // row stride of the parameter arrays (markers)
	unsigned int* counter;   // work queue head
	double* alpha;           // scratch [n_ctas][Mc][NUMSHIFTS][NUMTYPES]: alpha before the emission at each marker
	double* alphafac;        // scratch [n_ctas][Mc][NUMSHIFTS]: its log-factor
	double* stats_shared;    // statistics block of the rows shared between ranks (layout: StatsBlock)
	double* stats_local;     // statistics block of the other rows
	int32_t n_shared;        // rows [0, n_shared) live in stats_shared
	int32_t n_local;
	double* loglik;          // [n_units][2]
	const int32_t* unit_list;      // optional indirection: work item i is unit unit_list[i] ...
	const unsigned int* list_count;  // ... for i < *list_count (device-side count written by the fast kernel)
};

// Statistics of nr rows x Mc markers, one array per statistic so that each can be copied straight into
// the caller's per-individual vectors: haplobase [nr][Mc], haplocount [nr][Mc], infprobs [nr][Mc][2][2],
// homozyg [nr][Mc][2], negshift [nr][Mc]  (9 doubles per cell in total).
struct StatsCell
{
	double *haplobase, *haplocount, *infprobs, *homozyg, *negshift;
};
__host__ __device__ __forceinline__ StatsCell stats_cell(double* base, size_t nr, size_t Mc, size_t idx, size_t m)
{
	const size_t cell = idx * Mc + m, plane = nr * Mc;
	return StatsCell{ base + cell, base + plane + cell, base + 2 * plane + cell * 4, base + 6 * plane + cell * 2, base + 8 * plane + cell };
}

template <int NS>  // NS: states (64, or 256 under SELFING)
struct __align__(16) SmemT
{
	double acc[NACC * NTHREADS];
	double2 S[NSLOTS * TILE];
	double HW[NSLOTS * TILE];
	double tot[NACC];
	double Ls[NUMSHIFTS];
	double beta[NUMSHIFTS * NS];
	double aplus[NUMSHIFTS * NS];           // turn scoring: alpha after the emission at the current marker
	double Bs[NUMSHIFTS], Fp[NUMSHIFTS];    // log-factors of beta / alpha+
	double raw[8 * NUMSHIFTS];              // rawvals[turn][shift], -1 = not scored (:5678-5681)
	unsigned short queue[NUMSHIFTS * 64];
	uchar2 A[NSLOTS * TILE];
	Unit unit;
	int32_t unit_index;
};
using Smem = SmemT<NUMTYPES>;
using SmemSelfing = SmemT<NUMTYPES_SELFING>;

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}

// Transition part of realanalyze (cnF2freq.cpp:2270-2367): p'[to] = sum_from p[from] * prod_t (bit_t(from^to) ? r_t : 1-r_t).
// TYPEGENS = {1,0,0,1,0,0} (settings.h:23) picks genrec[1] for the two child meioses (bits 0 and 3).
__device__ __forceinline__ void transition(double& p0, double& p1, double r0, double r1)
{
	if (!(r0 > 0) && !(r1 > 0)) return;  // same position: identity (:2273)
#pragma unroll
	for (int b = 0; b < 5; b++)
	{
		const double r = (b == 0 || b == 3) ? r1 : r0;
		const double q0 = __shfl_xor_sync(0xffffffffu, p0, 1 << b);
		const double q1 = __shfl_xor_sync(0xffffffffu, p1, 1 << b);
		p0 = (1.0 - r) * p0 + r * q0;
		p1 = (1.0 - r) * p1 + r * q1;
	}
	const double n0 = (1.0 - r0) * p0 + r0 * p1;
	const double n1 = (1.0 - r0) * p1 + r0 * p0;
	p0 = n0;
	p1 = n1;
}

// The same under SELFING (:2276-2364): probs2[to] = sum_from probs[from] recombprec[(from ^ to) & 63] selfprec[sv(from)][sv(to)]
// over the 192 valid states -- the 64-state product per selfing value, then the 3 x 3 mix of the selfing values; the
// fourth value is never entered and is zeroed by any transition over a distance (probs2 starts at 0).
__device__ __forceinline__ void transition_selfing(double (&p)[8], double r0, double r1, double dist, double genrec2, int selfgen)
{
	if (!(r0 > 0) && !(r1 > 0)) return;
#pragma unroll
	for (int sv = 0; sv < 3; sv++) transition(p[2 * sv], p[2 * sv + 1], r0, r1);
	const double rs = 0.5 * (1.0 - exp((double)selfgen * genrec2 * dist));  // recprob[2][0], :2286
	const double s01 = rs, s00 = 1.0 - 2.0 * rs;                   // selfprec, :2319-2325
	const double s10 = selfgen ? s01 * 2.0 / (double)((1 << selfgen) - 1) : 1.0;
	const double s12 = s10 * s01, s11 = 1.0 - s10 - s12;
#pragma unroll
	for (int h = 0; h < 2; h++)
	{
		const double a = p[h], b = p[2 + h], c = p[4 + h];
		p[h] = a * s00 + b * s10 + c * s10;
		p[2 + h] = a * s01 + b * s11 + c * s12;
		p[4 + h] = a * s01 + b * s12 + c * s11;
		p[6 + h] = 0;
	}
}

constexpr double MINFACTOR_D = (double)(float)-1e15;  // settings.h:29, a float constant

// adjustprobs (cnF2freq.cpp:1579-1670) for one warp = one shift mode, all orientations summed (flag2 = -1).
template <bool SELF, int NP>  // NP states per lane: state k of the lane is lane + 32 k
__device__ __forceinline__ void emit_step(const Ctx& c, int t, int lane, int shift, double (&p)[NP], double& F)
{
	double part = 0;
#pragma unroll
	for (int k = 0; k < NP; k++)
	{
		if (p[k] < 1e-300) p[k] = 0;
		else p[k] *= track<0, 4, SELF>(c, 0, UNKNOWN, 0.0, t, (unsigned)((lane + 32 * k) * 2), -1, shift, 0.0);
		part += p[k];
	}
	const double sum = warp_sum(part);
	if (sum <= 0) F = MINFACTOR_D;
	else
	{
#pragma unroll
		for (int k = 0; k < NP; k++) p[k] /= sum;
		F += log(sum);
	}
}

template <int NS>
__device__ __forceinline__ void load_tile(SmemT<NS>& sm, const EstepArgs& a, int m0, int tid)
{
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
		}
		sm.A[i] = al;
		sm.S[i] = su;
		sm.HW[i] = h;
	}
}

// ignoreflag2, cnF2freq.cpp:3462-3496, rules 2 and 3 (rule 1 is applied by enumerating only the
// orientations outside flag2ignore).  hom bit i = relative i is "homozygous with equal sureness" here.
template <bool SELF>
__device__ __forceinline__ bool ignored(const Unit& u, int hom, int flag2, int g, int shift)
{
	const int x = flag2 ^ (g << 1);
	for (int i = 0; i < u.nrel; i++)
	{
		const int filt = u.relfilter[i];
		if (!filt) continue;
		const int filtered = x & filt;
		if (filtered && filtered != filt) return true;
		if (!SELF && ((hom >> i) & 1) && !((filtered != 0) ^ ((shift & u.relshift[i]) != 0))) return true;  // rule 3: off under SELFING (:3489)
	}
	return false;
}

// Transfer of one marker's reduced accumulators into the EM sufficient statistics, for relative r of the
// unit (cnF2freq.cpp:5875-5904, moveinfprobs :3577-3597, movehaplos :3599-3616).
__device__ __forceinline__ void transfer_stats(const EstepArgs& a, const Unit& u, const double* tot, const double* HW,
                                               int r, int t, int m, int hwstride = TILE)
{
	const int Mc = a.Mc;
	const int rootrel = u.slotrel[0];
	double rootsum = 0, h0 = 0, h1 = 0, ip[4] = { 0, 0, 0, 0 };
	for (int slot = 0; slot < NSLOTS; slot++)
	{
		if (u.row[slot] < 0) continue;
		const int rel = u.slotrel[slot];
		if (rel == rootrel) rootsum += tot[ACC_INF + slot * 4 + 0] + tot[ACC_INF + slot * 4 + 1];
		if (rel == r)
		{
			h0 += tot[ACC_HAPLOS + slot * 2 + 0];
			h1 += tot[ACC_HAPLOS + slot * 2 + 1];
			for (int q = 0; q < 4; q++) ip[q] += tot[ACC_INF + slot * 4 + q];
		}
	}
	const double sum = 1.0 / rootsum;
	const int srow = u.relrow[r];
	const StatsCell cell = srow < a.n_shared ? stats_cell(a.stats_shared, a.n_shared, Mc, srow, m)
	                                         : stats_cell(a.stats_local, a.n_local, Mc, srow - a.n_shared, m);
	double norm = sum * 2;
	for (int k = 0; k < u.relhalvings[r]; k++) norm /= 2;
	norm *= u.descendants;
	for (int q = 0; q < 4; q++)
		if (ip[q] != 0) atomicAdd(cell.infprobs + q, ip[q] * norm);
	if (h0 != 0 || h1 != 0)
	{
		// the relative's own haploweight at this marker; any of its slots holds it
		int slot = 0;
		while (u.slotrel[slot] != r || u.row[slot] < 0) slot++;
		const double w = HW[slot * hwstride + t];
		if (fabs(w - 0.5) < 0.5 - 1e-12)
		{
			const double eps = exp(-400.0) * (double)(0.000005f) * (double)(0.000005f) * 0.5;  // maxdiff, :228
			const double x1 = h0 + eps, x2 = h1 + eps;
			atomicAdd(cell.haplobase, x1 / (x1 + x2) * u.descendants);
			atomicAdd(cell.haplocount, u.descendants);
		}
	}
	if (r == rootrel)
	{
		atomicAdd(cell.homozyg + 0, tot[ACC_HZ + 0] * sum);
		atomicAdd(cell.homozyg + 1, tot[ACC_HZ + 1] * sum);
	}
}

// One surviving (state g, shift, orientation flag2) term of the posterior loop, cnF2freq.cpp:5464-5577:
// its weight val = exp(doanalyze(classicstop(marker, g), flag2) - factor) and the scatter of val into the
// accumulator columns (haplos via updatehaplo, infprobs via the four GENOS calls, homozyg).
template <bool SELF>
__device__ __forceinline__ void posterior_term(const Ctx& c, int t, int g, int flag2, int shift, double am, double bm,
                                               double Fm, double B, double factor, float minfactor)
{
	// doanalyze with classicstop(marker, g) and orientation flag2 (:1963-2018, 2229-2266)
	double lp = Fm;
	double e = 0;
	if (!(lp < minfactor))
	{
		double v = 0;
		if (!(am < 1e-300))
		{
			e = track<0, 4, SELF>(c, 0, UNKNOWN, 0.0, t, (unsigned)(g * 2), flag2, shift, 0.0);
			v = am * e;
		}
		lp += (v <= 0) ? MINFACTOR_D : log(v);
		if (!(lp < minfactor))
		{
			lp += B;
			if (bm <= 0) lp = MINFACTOR_D;
			else lp -= log(1.0 / bm);
		}
	}
	double val = lp - factor;
	if (!(isfinite(val) && val > -200.0)) return;  // unusualstate, :5414, 5502
	val = exp(val);
	double sv[2][2], svs[2], hz[2];
#pragma unroll
	for (int side = 0; side < 2; side++)  // :5519-5529
	{
		sv[side][0] = track<U_GENOSPROBE, 4, SELF>(c, 0, 1, 0.0, t, (unsigned)(g * 2 + side), flag2 ^ side, shift, 0.0);
		sv[side][1] = track<U_GENOSPROBE, 4, SELF>(c, 0, 2, 0.0, t, (unsigned)(g * 2 + side), flag2 ^ side, shift, 0.0);
		svs[side] = sv[side][0] + sv[side][1];
	}
#pragma unroll
	for (int al = 0; al < 2; al++)  // :5532-5537
		hz[al] = track<U_HOMOZYGOUS, 4, SELF>(c, 0, al + 1, 0.0, t, (unsigned)(g * 2), flag2, shift, 0.0);
	// updatehaplo, :1561-1575 (its first, update-free call is the emission above)
	if (e != 0) track<U_HAPLOS, 4, SELF>(c, 0, UNKNOWN, 0.0, t, (unsigned)(g * 2), flag2, shift, val);
#pragma unroll
	for (int side = 0; side < 2; side++)  // :5560-5568
#pragma unroll
		for (int al = 0; al < 2; al++)
			track<U_GENOS, 4, SELF>(c, 0, al + 1, 0.0, t, (unsigned)(g * 2 + side), flag2 ^ side, shift, val * sv[side][al] / svs[side]);
#pragma unroll
	for (int al = 0; al < 2; al++)  // :5571-5575
		c.acc[(ACC_HZ + al) * NTHREADS] += val * hz[al] / svs[0];
}

template <bool SELF, int NP>
__device__ __forceinline__ void step_over(double (&p)[NP], const EstepArgs& a, int m, int selfgen)
{
	if constexpr (SELF) transition_selfing(p, a.rec0[m], a.rec1[m], a.dist[m], a.genrec2, selfgen);
	else transition(p[0], p[1], a.rec0[m], a.rec1[m]);
}

template <bool SELF>
__global__ void __launch_bounds__(NTHREADS, 1) estep_kernel(EstepArgs a)
{
	constexpr int NS = SELF ? NUMTYPES_SELFING : NUMTYPES;  // states
	constexpr int NP = NS / 32;                              // per lane: state k of the lane is lane + 32 k
	extern __shared__ __align__(16) unsigned char smem_raw[];
	SmemT<NS>& sm = *reinterpret_cast<SmemT<NS>*>(smem_raw);
	const int tid = threadIdx.x, lane = tid & 31, shift = tid >> 5;
	double* const alpha = a.alpha + (size_t)blockIdx.x * a.Mc * (NUMSHIFTS * NS);
	double* const alphafac = a.alphafac + (size_t)blockIdx.x * a.Mc * NUMSHIFTS;
	Ctx c{ sm.A, sm.S, sm.HW, sm.unit.row, sm.unit.founder, sm.acc + tid };
	const int Mc = a.Mc, E = Mc - 1;
	const int ntiles = (Mc + TILE - 1) / TILE;

	for (int i = tid; i < NACC * NTHREADS; i += NTHREADS) sm.acc[i] = 0;

	for (;;)
	{
		__syncthreads();
		if (tid == 0) sm.unit_index = (int)atomicAdd(a.counter, 1u);
		__syncthreads();
		int ui = sm.unit_index;
		if (ui >= (a.unit_list ? (int)*a.list_count : a.n_units)) break;
		if (a.unit_list) ui = a.unit_list[ui];
		{
			const int* src = reinterpret_cast<const int*>(a.units + ui);
			int* dst = reinterpret_cast<int*>(&sm.unit);
			for (int i = tid; i < (int)(sizeof(Unit) / 4); i += NTHREADS) dst[i] = src[i];
		}
		__syncthreads();
		const Unit& u = sm.unit;
		const bool active = shift < u.shiftend && !(shift & u.shiftignore);

		// ---- forward sweep: initfwbw + realanalyze<STORE|FORWARD> (cnF2freq.cpp:2086-2107) ----
		double p[NP], F = 0;
		const int selfgen = SELF ? u.selfgen : 0;
#pragma unroll
		for (int k = 0; k < NP; k++)
		{
			p[k] = 1.0 / NS;  // EVENGEN, :2100-2104
			if constexpr (SELF)  // selfingfactors, :2092-2099
			{
				const double f0 = 1.0 / (double)(1 << selfgen);
				p[k] *= (k >> 1) == 0 ? f0 : (k >> 1) == 3 ? 0.0 : (1.0 - f0) * 0.5;
			}
		}
		double L = 0;
		for (int tile = 0; tile < ntiles; tile++)
		{
			__syncthreads();
			load_tile(sm, a, tile * TILE, tid);
			__syncthreads();
			if (!active) continue;
			const int tend = min(TILE, Mc - tile * TILE);
			for (int t = 0; t < tend; t++)
			{
				const int m = tile * TILE + t;
				double* dst = alpha + ((size_t)m * NUMSHIFTS + shift) * NS;
#pragma unroll
				for (int k = 0; k < NP; k++) dst[lane + 32 * k] = p[k];
				if (lane == 0) alphafac[(size_t)m * NUMSHIFTS + shift] = F;
				emit_step<SELF>(c, t, lane, shift, p, F);
				if (m < E) step_over<SELF>(p, a, m, selfgen);
			}
		}
		if (active)
		{
			// doanalyze/quickanalyze with NONESTOP (:1984-2018): alpha+ at the last marker against beta = 1
			L = F;
			double part = 0;
#pragma unroll
			for (int k = 0; k < NP; k++) part += p[k];
			const double sum = warp_sum(part);
			if (sum <= 0) L = MINFACTOR_D;
			else L -= log(1.0 / sum);
			if (lane == 0) sm.Ls[shift] = L;
		}
		__syncthreads();
		// :5373-5403
		double factor = -1e15, realfactor = 0;
		for (int s = 0; s < u.shiftend; s++)
		{
			const double f = (s & u.shiftignore) ? -1e30 : sm.Ls[s];
			factor = fmax(factor, f);
		}
		for (int s = 0; s < u.shiftend; s++)
		{
			if (s & u.shiftignore) continue;
			realfactor += exp(sm.Ls[s] - factor);
		}
		const double maxfactor = factor;
		factor += log(realfactor);
		if (tid == 0)
		{
			a.loglik[(size_t)u.out * 2 + 0] = maxfactor;
			a.loglik[(size_t)u.out * 2 + 1] = factor;
		}
		if (isnan(factor) || factor < MINFACTOR_D) continue;  // :5403
		const bool contributing = active && !(factor - L > 40);  // :5421
		const float minfactor = (float)(-200.0 + factor);         // unusualstate + factor as a float, :1936
		const int allowed = (NUMPATHS - 1) & ~u.flag2ignore;

		// ---- backward sweep fused with the posterior enumeration (cnF2freq.cpp:2109-2116, 5406-5583) ----
		double b[NP], B = 0;  // :2111-2114
#pragma unroll
		for (int k = 0; k < NP; k++) b[k] = 1.0;
		for (int tile = ntiles - 1; tile >= 0; tile--)
		{
			__syncthreads();
			load_tile(sm, a, tile * TILE, tid);
			__syncthreads();
			const int tend = min(TILE, Mc - tile * TILE);
			for (int t = tend - 1; t >= 0; t--)
			{
				const int m = tile * TILE + t;
				const double* src = alpha + ((size_t)m * NUMSHIFTS + shift) * NS;
				const double Fm = alphafac[(size_t)m * NUMSHIFTS + shift];
				// beta of this shift mode, readable by every lane of the warp (and, for the turns, by the other warps)
#pragma unroll
				for (int k = 0; k < NP; k++) sm.beta[shift * NS + lane + 32 * k] = b[k];
				// ---- turn scoring (cnF2freq.cpp:5668-5757, DOTOULBAR 0) and updatenegshifts (:3640-3715) ----
				{
					double ap[NP], Fp = 0;
#pragma unroll
					for (int k = 0; k < NP; k++) ap[k] = 0;
					if (active)
					{
#pragma unroll
						for (int k = 0; k < NP; k++) ap[k] = src[lane + 32 * k];
						Fp = Fm;
						emit_step<SELF>(c, t, lane, shift, ap, Fp);  // alpha+ and its factor: pad 2 of fwbw (:2248-2251)
					}
#pragma unroll
					for (int k = 0; k < NP; k++) sm.aplus[shift * NS + lane + 32 * k] = ap[k];
					if (lane == 0) sm.Bs[shift] = B, sm.Fp[shift] = Fp;
					__syncthreads();
					const float minfturn = (float)(-50000.0 + factor);
					for (int tn = 0; tn < 8; tn++)
					{
						const int g = tn == 0 ? 0 : tn == 7 ? 64 : 1 << (tn - 1);  // turn of slot tn (7: the root)
						double rv = -1;
						if (active && !(g & (u.flag2ignore >> 1)))
						{
							const int mask = g & 54, fms = (g >> 6) | ((g & 1) ? 2 : 0) | ((g & 8) ? 4 : 0);  // aroundturner, :506-508
							const int s2 = shift ^ fms;
							double fac = sm.Fp[shift];
							if (!(fac < minfturn))
							{
								double part = 0;  // the turn permutes the meiosis bits only (:506, 536: i ^ turn)
#pragma unroll
								for (int k = 0; k < NP; k++)
									part += sm.aplus[shift * NS + ((lane + 32 * k) ^ mask)] * sm.beta[s2 * NS + lane + 32 * k];
								const double sum = warp_sum(part);
								fac += sm.Bs[s2];
								if (sum <= 0) fac = MINFACTOR_D;
								else fac -= log(1.0 / sum);
							}
							rv = exp(fac - factor);
							if (!isfinite(rv)) rv = rv > 1e300 ? 1e300 : 0;
						}
						if (lane == 0) sm.raw[tn * NUMSHIFTS + shift] = rv;
					}
					__syncthreads();
					if (tid < NSLOTS && u.row[tid] >= 0 && u.nsscale[tid] != 0)
					{
						const int slot = tid, tn = slot == 0 ? 7 : slot;
						double l[2];
						bool valid[2];
						for (int q = 0; q < 2; q++)
						{
							const int tq = q ? tn : 0;
							double val = 0;
							valid[q] = false;
							for (int s = 0; s < NUMSHIFTS; s++)
							{
								const double r = sm.raw[tq * NUMSHIFTS + s];
								if (r < 0) continue;
								val += r;
								valid[q] = true;
							}
							if (!isfinite(val) || val < 1e-174) val = 1e-174;
							l[q] = log(val);
						}
						const int srow = u.relrow[u.slotrel[slot]];
						const StatsCell cell = srow < a.n_shared ? stats_cell(a.stats_shared, a.n_shared, Mc, srow, m)
						                                         : stats_cell(a.stats_local, a.n_local, Mc, srow - a.n_shared, m);
						double add = 0;
						if (valid[0]) add += l[0] * u.nsscale[slot];
						if (valid[1]) add -= l[1] * u.nsscale[slot];
						atomicAdd(cell.negshift, add);
					}
				}
				if (contributing)
				{
					int hom = 0;
					for (int i = 0; i < u.nrel; i++)
					{
						if (!u.relfilter[i]) continue;
						// any slot of the relative holds the same individual's data
						const int slot = __ffs(u.relfilter[i]) - 1;
						const uchar2 al = sm.A[slot * TILE + t];
						const double2 su = sm.S[slot * TILE + t];
						if (al.x == al.y && su.x == su.y) hom |= 1 << i;
					}
					// The surviving (g, flag2) pairs are sparse and differ from lane to lane, so they are first
					// compacted into a per-warp queue (ballot + prefix count) and then handed out one per lane.
					unsigned short* q = sm.queue + shift * 64;  // entries: flag2 << 8 | g
					int cnt = 0;
					for (int flag2 = allowed;; flag2 = (flag2 - 1) & allowed)
					{
#pragma unroll
						for (int gi = 0; gi < NP; gi++)
						{
							const int g = lane + 32 * gi;
							// a state without forward or backward mass ends in posterior_term before it scatters anything (v = 0 resp.
							// bm <= 0 put the term at MINFACTOR, 1e15 below an ordinary factor): not queued.  Under SELFING that is a
							// quarter of the states at least.  Not so for a unit impossible under every shift mode, whose factor sits at
							// MINFACTOR itself: there such terms pass the cut and the reference scatters them; they stay queued
							const bool massless = factor > -1e14 && (src[g] < 1e-300 || !(sm.beta[shift * NS + g] > 0));
							const bool ok = !massless && !ignored<SELF>(u, hom, flag2, g, shift);
							const unsigned mask = __ballot_sync(0xffffffffu, ok);
							if (ok) q[cnt + __popc(mask & ((1u << lane) - 1u))] = (unsigned short)((flag2 << 8) | g);
							cnt += __popc(mask);
							__syncwarp();
							if (cnt >= 32)
							{
								cnt -= 32;
								const int entry = q[cnt + lane];
								__syncwarp();
								posterior_term<SELF>(c, t, entry & 255, entry >> 8, shift, src[entry & 255], sm.beta[shift * NS + (entry & 255)],
								                     Fm, B, factor, minfactor);
							}
						}
						if (flag2 == 0) break;
					}
					if (lane < cnt)
					{
						const int entry = q[lane];
						posterior_term<SELF>(c, t, entry & 255, entry >> 8, shift, src[entry & 255], sm.beta[shift * NS + (entry & 255)], Fm,
						                     B, factor, minfactor);
					}
					__syncwarp();
				}
				__syncthreads();
				// block-level reduction of the accumulator columns; the reader clears what it read
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
				// transfer into the EM sufficient statistics (:5875-5904, moveinfprobs :3577-3597, movehaplos :3599-3616)
				if (tid < u.nrel) transfer_stats(a, u, sm.tot, sm.HW, tid, t, m);
				// beta for the previous marker: adjustprobs at m, then the transition (:2390-2396)
				if (m > 0)  // every shift mode: a turn pairs alpha+ of one mode with beta of another (initfwbw(..., 2), :1988)
				{
					emit_step<SELF>(c, t, lane, shift, b, B);
					step_over<SELF>(b, a, m - 1, selfgen);
				}
			}
		}
	}
}
}  // namespace cnf2
