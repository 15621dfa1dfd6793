// The M-step on the device (row a6 / f1 of SURVEY.md section 8): one pass of the reference's block at
// cnF2freq.cpp:6356-6367 for one chromosome's fresh statistics, plus the re-run of updatehaploweights that the
// reference makes on the chromosomes it finished earlier in the same iteration (the block sits inside the
// chromosome loop and walks all markers, :6239-6367).
//
//   mstep_infprobs_kernel     processinfprobs (:4179-4323): one thread per (row, marker), both sides
//   mstep_relskew_kernel      relskewhmm (:4325-4467): one thread per (row, chromosome), sequential two-state sweep
//   mstep_haploweight_kernel  updatehaploweights (:4533-4734): one thread per (row, marker)
//
// The step solver (cappedgd :4040-4177, caplogitchange :4006-4038, 15-point Gauss-Legendre as
// boost::math::quadrature::gauss<double, 15>) is the same arithmetic in the same order as the reference; the only
// difference is the device's log().  lastinved is -1 (no phase inversion in flight, :6290), so breakathalf is false.
#pragma once
#include <type_traits>
#include "cnf2_tables.cuh"

namespace cnf2
{
constexpr float MS_MAXDIFF = 0.000005f;  // const float maxdiff, :228
// `1 - maxdiff` (:4667) and `maxdiff / (children + 1)` (:4242, :4703) are float arithmetic in the reference
constexpr double MS_ONE_MINUS_MAXDIFF = (double)(1.0f - MS_MAXDIFF);
__device__ __forceinline__ double ms_epsilon(int children) { return (double)(MS_MAXDIFF / (float)(children + 1)); }

struct MstepRow  // per statistics row
{
	int32_t children;     // individ::children as rebuilt by doit (:5246-5259)
	int32_t descendants;  // :872
	uint8_t empty, hasprior, pad[2];
};

// The EM control state of the reference (scalefactor / entropyfactor :3573-3574, the hitnnn history :6373-6390), kept on
// the device so that a pass needs no host round trip.
struct EmCtl
{
	double scalefactor, entropyfactor;
	double sf_pass;               // scalefactor of the pass in flight: 0 while phase inversions are applied (:6339-6341)
	double hit_local_d;           // local hits as a double: the all-reduce slot between ranks
	unsigned long long hit[2];    // hits of the pass: [0] rows only this rank holds, [1] shared rows (every rank counts those)
	long long hitnnn_last;        // global hitnnn of the last finished pass
	int32_t oldhitnnn, oldhitnnn2;
	int32_t n_dous_total, any;
	// measurement (bench.py's M-step roofline): gradient evaluations the solves made since the last reset --
	// [0] processinfprobs (MODE 0), [1] updatehaploweights (MODE 1); 15 per integral, 1 per midpoint and start
	unsigned long long evals[2];
	// solves that ended on integral-free steps (see ms_cappedgd_warp): [0] confirmed by the closing integral, [1] sent back
	unsigned long long closing[2];
};
struct MsCount
{
	unsigned evals = 0, confirmed = 0, rejected = 0;
};
__device__ __forceinline__ void ms_count_flush(EmCtl* ctl, int mode, MsCount c)
{
	const unsigned FULL = 0xffffffffu;
	c.evals = __reduce_add_sync(FULL, c.evals);
	c.confirmed = __reduce_add_sync(FULL, c.confirmed);
	c.rejected = __reduce_add_sync(FULL, c.rejected);
	if ((threadIdx.x & 31) == 0)
	{
		if (c.evals) atomicAdd(ctl->evals + mode, (unsigned long long)c.evals);
		if (c.confirmed) atomicAdd(ctl->closing + 0, (unsigned long long)c.confirmed);
		if (c.rejected) atomicAdd(ctl->closing + 1, (unsigned long long)c.rejected);
	}
}

struct MstepArgs
{
	// parameters of the chromosome, [n_rows][Mc] blocks (in/out)
	uchar2* alleles;
	double2* sure;
	double* hw;
	const double* relhaplo;
	const uchar2* prioralleles;
	const double2* priorsure;
	const MstepRow* rows;
	// statistics: the chromosome's E-step block (fresh) or the carried haplobase/haplocount of an earlier pass
	double *stats_shared, *stats_local;
	double *carry_base, *carry_count;  // [n_rows][Mc]
	double *ratio, *fw;                // scratch [n_rows][Mc], [n_rows][Mc][2]
	uint8_t* anyinfo;                  // [n_rows]
	EmCtl* ctl;
	int32_t n_rows, n_shared, Mc;
	int32_t fresh;  // 1: statistics come from the E-step block; 0: from the carry arrays
};

// One launch serves every chromosome of a pass (the fresh one and the ones carried from earlier in the iteration, which
// the reference's block walks again, :6239-6367): blockIdx.y picks the chromosome.  `g` holds the genome-wide base
// pointers (chromosome c's block starts n_rows * c0 cells in; anyinfo at n_rows * index).
constexpr int MS_MAXLIST = 24;
struct MstepPass
{
	MstepArgs g;
	int32_t n, fresh_index;  // chromosomes in this launch; which of them is the fresh one (-1: none)
	int32_t c0[MS_MAXLIST], Mc[MS_MAXLIST], cidx[MS_MAXLIST];
};
__device__ __forceinline__ MstepArgs ms_view(const MstepPass& p, int k)
{
	MstepArgs a = p.g;
	const size_t off = (size_t)a.n_rows * p.c0[k];
	a.alleles += off, a.sure += off, a.hw += off, a.relhaplo += off, a.prioralleles += off, a.priorsure += off;
	a.carry_base += off, a.carry_count += off, a.ratio += off, a.fw += 2 * off;
	a.anyinfo += (size_t)a.n_rows * p.cidx[k];
	a.Mc = p.Mc[k];
	a.fresh = k == p.fresh_index;
	return a;
}

__device__ __forceinline__ double ms_caplogit(double intended, double orig, double epsilon, unsigned long long& hit)
{
	const double nnn = 3;
	const double limn = (nnn - 1.0) * orig * (-1 + orig);
	const double limd1 = -1 - (nnn - 1.0) * orig;
	const double limd2 = (nnn - 1.0) * orig - nnn;
	intended = fmin(intended, 1.0 - epsilon);
	intended = fmax(intended, epsilon);
	const double diff = intended - orig;
	if (diff > limn / limd1)
	{
		intended = orig + limn / limd1;
		if (intended < 0.5) hit++;
	}
	if (diff < -limn / limd2)
	{
		intended = orig - limn / limd2;
		if (intended > 0.5) hit++;
	}
	return intended;
}

// 1 / x on the solver's inner loop: frcp (~1 ulp) where x is an ordinary number, the IEEE
// division where it is not, so that 1 / 0 = inf and non-finite gradients keep steering the bisection as they do in
// the reference (:4131, :4144).
__device__ __forceinline__ double ms_rcp(double x)
{
	const unsigned ex = ((unsigned)__double2hiint(x) >> 20) & 0x7ffu;
	return (ex - 64u < 1920u) ? frcp(x) : 1.0 / x;  // |x| in [2^-959, 2^961): x and 1 / x are both ordinary numbers
}

// ms_log: see cnf2_tables.cuh
// fmin(fmax(v, lo), hi) of the reference for lo <= hi (a NaN v gives lo, as fmax does)
__device__ __forceinline__ double ms_clamp(double v, double lo, double hi)
{
	v = !(v >= lo) ? lo : v;
	return v > hi ? hi : v;
}

// The data term of the gradients at :4275 and :4688.  The reference writes it as a 15-term polynomial in c (current
// value), h (the statistic), s (its total), x, log x and log(1 - x) over (c h + c s x - c s - h x)^2; the two logs only
// enter through their difference, and collecting the terms gives
//   numerator = A et + B + C x,  et = log(1 - x) - log(x) = log(1 / x - 1)   (the entropy term's own logarithm, :4288)
//   A = c h (c - 1)(h - s),  B = c (h - s)(h - c s),  C = -(c s - h)^2,  denominator = (c (h - s) + (c s - h) x)^2
// with A, B, C and the two denominator coefficients fixed per solve.  One logarithm and a dozen operations per
// evaluation instead of two logarithms and ~60: fp64 log and the polynomial are what these kernels spent their time on.
// The factored coefficients carry fewer cancellations than the expanded sum, so the value differs from the reference's
// by the reference's own rounding of that sum (the parity tests hold the new parameters to 1e-9).
__device__ __forceinline__ void ms_data_term(double c, double h, double s, double& A, double& B, double& C, double& d0, double& d1)
{
	const double hs = h - s, csh = c * s - h;
	A = c * h * (c - 1) * hs;
	B = c * hs * (h - c * s);
	C = -(csh * csh);
	d0 = c * hs;
	d1 = csh;
}

// ---- the warp-cooperative form of cappedgd ----
// Each lane owns one solve (or none).  A bisection step needs the gradient at the midpoint (every lane evaluates its
// own) and, for about a third of the steps, a 15-point integral.  Lanes take their integral steps at different
// iterations, so a lane-private solver makes the warp pay 16 evaluations on nearly every step; here the integrals
// pending after a step are served two at a time, one per half-warp, its 15 points on 15 lanes with the owner's
// parameters read from shared memory, and the weighted values are summed over the half-warp by a shuffle tree (the
// reference adds the eight weighted pairs one after the other; the tree differs from that by rounding only).
struct MsParams
{
	double A, B, C, d0, d1;  // ms_data_term of (curprob, hzygcorred, sum) or (haploweight, haplobase, haplocount)
	double p0, p1, p2, p3;   // MODE 0: etf, etf * priord; MODE 1: (1 - similarity) * 1, ef, relskewterm, descendants
	double eps;
};
__device__ __forceinline__ MsParams ms_params(double c, double h, double s, double p0, double p1, double p2, double p3, double eps)
{
	MsParams P;
	ms_data_term(c, h, s, P.A, P.B, P.C, P.d0, P.d1);
	P.p0 = p0, P.p1 = p1, P.p2 = p2, P.p3 = p3, P.eps = eps;
	return P;
}
template <int MODE>
__device__ __forceinline__ double ms_actual(const MsParams& P, double val)
{
	val = ms_clamp(val, P.eps, 1 - P.eps);  // eps > 0: val, 1 - val and val - val^2 are ordinary positive numbers
	const double den = fma(P.d1, val, P.d0), D = den * den;
	double et, c;  // gradient = -N / D + c with N = A et + B + C x; what the solver integrates is 1 / gradient = D / (c D - N):
	               // one reciprocal, and the same 0 / inf / NaN as the reference's two divisions when D or the gradient is 0
	if (MODE == 0)
	{
		et = ms_log((1 - val) * frcp(val));
		c = fma(P.p0, et, P.p1);  // entropy term and prior, :4288-4289
	}
	else
	{
		const double t = frcp(val - val * val);  // 1 / (x (1 - x)); 1 / x - 1 = (1 - x)^2 t
		et = ms_log((1 - val) * (1 - val) * t);
		c = fma(P.p0, P.p1 * et, (P.p2 - val) * t * P.p3);  // :4689-4692
	}
	const double N = fma(P.A, et, fma(P.C, val, P.B));
	return D * ms_rcp(fma(c, D, -N));
}
// A solve's parameters are staged in shared memory by their owner (once per solve) and read back by the half-warp
// that serves one of its integrals: [parameter][thread of the block], so the owner's stores are conflict-free and a
// half-warp's loads are broadcasts.
constexpr int MS_BLOCK = 128, MS_NPAR = 10;
__device__ __forceinline__ void ms_stage(double* sp, const MsParams& P)
{
	sp[0 * MS_BLOCK] = P.A, sp[1 * MS_BLOCK] = P.B, sp[2 * MS_BLOCK] = P.C, sp[3 * MS_BLOCK] = P.d0, sp[4 * MS_BLOCK] = P.d1;
	sp[5 * MS_BLOCK] = P.p0, sp[6 * MS_BLOCK] = P.p1, sp[7 * MS_BLOCK] = P.p2, sp[8 * MS_BLOCK] = P.p3, sp[9 * MS_BLOCK] = P.eps;
}
template <int MODE>
__device__ __forceinline__ MsParams ms_fetch(const double* sp)
{
	MsParams Q;
	Q.A = sp[0 * MS_BLOCK], Q.B = sp[1 * MS_BLOCK], Q.C = sp[2 * MS_BLOCK], Q.d0 = sp[3 * MS_BLOCK], Q.d1 = sp[4 * MS_BLOCK];
	Q.p0 = sp[5 * MS_BLOCK], Q.p1 = sp[6 * MS_BLOCK], Q.eps = sp[9 * MS_BLOCK];
	Q.p2 = MODE ? sp[7 * MS_BLOCK] : 0.0, Q.p3 = MODE ? sp[8 * MS_BLOCK] : 0.0;
	return Q;
}
// All 32 lanes call this together; `active` lanes get their solve's result (the others get `orig` back).
// count_only: the caller throws the result away and only needs the final caplogitchange's hitnnn increment
// (processinfprobs of an `empty` individual or one without priors, :4301-4311).  That increment needs the end point
// beyond the cap (:4022-4032), i.e. every bisection decision towards the cap; the first decision away from it (or a
// stop at the first midpoint, which lies halfway to the cap) rules it out, so such a solve ends there.  Exact: the
// count is what the full solve would give.
#ifdef CNF2_MS_HIST
__device__ unsigned long long ms_hist[2][64];
#endif
template <int MODE>
__device__ __forceinline__ double ms_cappedgd_warp(const MsParams& P, double orig, double scalefactor, bool active, unsigned long long& hit,
                                                   double* stage, MsCount& cnt, bool count_only = false)
{
	const unsigned FULL = 0xffffffffu;
	const int lane = threadIdx.x & 31, half = lane >> 4, q = lane & 15;
	double* const wstage = stage + (threadIdx.x & ~31);  // this warp's 32 columns of the block's staging array
	__syncwarp();
	ms_stage(wstage + lane, P);
	__syncwarp();
	const double GX[8] = { 0.0,
		                   0.2011940939974345223006283033946,
		                   0.3941513470775633698972073709810,
		                   0.5709721726085388475372267372539,
		                   0.7244177313601700474161860546139,
		                   0.8482065834104272162006483207742,
		                   0.9372733924007059043077589477102,
		                   0.9879925180204854284895657185866 };
	const double GW[8] = { 0.2025782419255612728806201999675,
		                   0.1984314853271115764561183264438,
		                   0.1861610000155622110268005618664,
		                   0.1662692058169939335532008604812,
		                   0.1395706779261543144478047945110,
		                   0.1071592204671719350118695466859,
		                   0.0703660474881081247092674164507,
		                   0.0307532419961172683546283935772 };
	// this lane's quadrature point: q = 0 the centre, 1..7 centre + scale x_q, 8..14 centre - scale x_(q-7), 15 idle
	double myx = 0, myw = 0;
#pragma unroll
	for (int k = 0; k < 8; k++)
	{
		if (q == k) myx = GX[k], myw = GW[k];
		if (k && q == k + 7) myx = -GX[k];
	}
	unsigned long long dump = 0;
	const double epsilon = P.eps;
	double lolim = 0, hilim = 0, lo = 0, hi = 0;
	bool lowside = false, fin = !active;
	if (active)
	{
		lolim = ms_caplogit(epsilon, orig, epsilon, dump);
		lo = lolim - epsilon * 0.125;
		hilim = ms_caplogit(1 - epsilon, orig, epsilon, dump);
		hi = hilim + epsilon * 0.125;
		orig = ms_caplogit(orig, orig, epsilon, dump);
		const double gradval = ms_actual<MODE>(P, orig);
		cnt.evals++;
		if (!isfinite(gradval) || !scalefactor)
		{
			lo = orig;
			hi = orig;
		}
		lowside = gradval < 0;
		if (lowside) hi = orig;
		else lo = orig;
		if (!scalefactor) fin = true;
		// count_only: all that is wanted is whether the final caplogitchange counts a hit (:4022-4032) -- the end point beyond
		// the cap on the side the solve moves to, AND that cap on the near side of 0.5.  A cap at or beyond 0.5 (every solve
		// that starts from 0.5, e.g. the untyped F1 rows of an F2 design) cannot count: the solve is over before it starts.
		if (count_only && (lowside ? lolim <= 0.5 : hilim >= 0.5))
		{
			lo = hi = orig;
			fin = true;
		}
	}
	// Integral-free steps.  A step whose midpoint lies on the near side of the gradient's sign change costs the reference a
	// 15-point integral only to learn that the flow has not yet run for `scalefactor` (prel < scalefactor -> move on).  Where
	// the flow converges on a stationary point of the objective, or runs into the cap, EVERY such step of the solve says
	// so, because the integral diverges towards the stationary point -- and in the later EM iterations that is four solves
	// in five, each with ~25 integrals.  Once a step has fallen well short (prel < scalefactor / 4) or overshot the sign
	// change, the solve therefore goes on WITHOUT the integrals, taking "short" for granted: the same midpoints, the same
	// interval updates, bit for bit the reference's.  At its end one integral from the start to the near end of the final
	// interval -- at least as long as any the skipped steps would have taken -- settles it: short by a margin (< 0.9
	// scalefactor, every sample of the integrand of one sign) confirms the result; otherwise the solve returns to the state
	// in which it left the reference's schedule and repeats those steps with their integrals.  (Measured and dropped: a
	// secant-slope estimate of how close the flow gets to the stationary point, to keep the ordinary solves out of this
	// mode -- the solves sent back cost what they cost either way; and one integral over the whole way to the cap as a
	// short cut for the solves that end there -- it misses a stationary point hidden within ~1e-6 of the cap.)
	bool skip = false, banned = count_only, verify = false, skipped = false;
	int myi = 0, smyi = 0;
	double slo = 0, shi = 0;
#ifdef CNF2_MS_NOSKIP
	banned = true;
#endif
#ifdef CNF2_MS_HIST
	int myit = 0;
#endif
#pragma unroll 1
	for (int round = 0; round < 120; round++)
	{
		// lo == hi: the interval has collapsed (a non-finite 1 / gradient at the start, :4131, or scalefactor 0); every
		// further step of the reference sets lo or hi to their common midpoint again, so the result is already fixed
		if (!fin && !verify && (myi >= 51 || lo > hilim || hi < lolim || lo == hi))
		{
			if (skipped) verify = true;
			else fin = true;
		}
		if (__all_sync(FULL, fin)) break;
		double mid = 0, prel = 0, start = 0, end = 0;
		bool pending = false, typea = false, stepping = !fin && !verify;
		if (stepping)
		{
#ifdef CNF2_MS_HIST
			myit++;
#endif
			myi++;
			mid = (lo + hi) / 2;
			const double gmid = ms_actual<MODE>(P, mid);
			cnt.evals++;
			if (((gmid < 0) != lowside) || !isfinite(gmid)) prel = (scalefactor + 0.1) * 1.1, typea = true;
			else
			{
				start = orig, end = mid;
				if (start > end)
				{
					const double t = start;
					start = end;
					end = t;
				}
				if (end - start < 1e-10)
				{
					stepping = false;
					if (skipped) verify = true;
					else fin = true;
				}
				else if (skip) skipped = true;  // prel = 0: taken to fall short of scalefactor, settled at the end
				else pending = true, cnt.evals += 15;
			}
		}
		else if (verify)
		{
			mid = lowside ? hi : lo;  // the near end of the final interval
			start = orig, end = mid;
			if (start > end)
			{
				const double t = start;
				start = end;
				end = t;
			}
			pending = true, cnt.evals += 15;
		}
		// integrals of the lanes that need one: two per round (one per half-warp), two rounds at a time while there are
		// enough of them -- a round is one long dependent fp64 chain per lane, a second independent one fills its gaps
		unsigned mask = __ballot_sync(FULL, pending);
		bool samesign = false;
		auto rounds = [&](auto nu, const int* own) {
			constexpr int NU = decltype(nu)::value;
			double f[NU], tq[NU];
			unsigned pos[NU], neg[NU];
#pragma unroll
			for (int u = 0; u < NU; u++)
			{
				const int owner = half ? own[2 * u + 1] : own[2 * u];  // the solve this half-warp serves (-1: none)
				const int src = owner < 0 ? lane : owner;
				const MsParams Q = ms_fetch<MODE>(wstage + src);
				const double a0 = __shfl_sync(FULL, start, src), b0 = __shfl_sync(FULL, end, src);
				const double avg = (a0 + b0) * 0.5, scale = (b0 - a0) * 0.5;
				const double v = ms_actual<MODE>(Q, avg + scale * myx);
				f[u] = (owner >= 0 && q < 15) ? v : 0.0;
				pos[u] = __ballot_sync(FULL, f[u] > 0);
				neg[u] = __ballot_sync(FULL, f[u] < 0);
			}
			// (f(+x_q) + f(-x_q)) w_q on lanes q = 0..7 of each half, summed over those 8 lanes, handed to the owner
#pragma unroll
			for (int u = 0; u < NU; u++)
			{
				tq[u] = (f[u] + __shfl_down_sync(FULL, f[u], 7, 16)) * myw;
				if (q == 0) tq[u] = f[u] * myw;
			}
#pragma unroll
			for (int d = 1; d < 8; d <<= 1)
#pragma unroll
				for (int u = 0; u < NU; u++) tq[u] += __shfl_xor_sync(FULL, tq[u], d);
#pragma unroll
			for (int u = 0; u < NU; u++)
			{
				const bool mine = lane == own[2 * u] || lane == own[2 * u + 1];
				const bool second = lane == own[2 * u + 1];
				const double result = __shfl_sync(FULL, tq[u], second ? 16 : 0);
				if (mine)
				{
					prel = result * ((end - start) * 0.5);
					if (end != mid) prel = -prel;
					if (!isfinite(prel)) prel = (scalefactor + 0.1) * 1.1;
					const unsigned sh = second ? 16 : 0;
					samesign = ((pos[u] >> sh) & 0x7fffu) == 0x7fffu || ((neg[u] >> sh) & 0x7fffu) == 0x7fffu;
				}
			}
		};
		while (mask)
		{
			int own[4];
#pragma unroll
			for (int u = 0; u < 4; u++)
			{
				own[u] = mask ? __ffs(mask) - 1 : -1;
				mask &= mask - 1;
			}
			if (own[2] >= 0) rounds(std::integral_constant<int, 2>(), own);
			else rounds(std::integral_constant<int, 1>(), own);
		}
		const bool shortby = samesign && prel >= 0 && prel < 0.9 * scalefactor;
		if (verify)
		{
			verify = false;
			if (shortby) fin = true, cnt.confirmed++;
			else
			{
				cnt.rejected++;
				lo = slo, hi = shi, myi = smyi;  // back to where the solve left the reference's schedule
				skip = skipped = false;
				banned = true;
			}
		}
		else if (stepping)
		{
			if (fabs(prel - scalefactor) < scalefactor * 1e-3) fin = true;
			else if ((prel < scalefactor) != lowside)
			{
				lo = mid;
				if (count_only && lowside) fin = true;  // away from the lower cap
			}
			else
			{
				hi = mid;
				if (count_only && !lowside) fin = true;  // away from the upper cap
			}
			if (!fin && !skip && !banned && (typea || prel < 0.25 * scalefactor))
			{
				skip = true;
				slo = lo, shi = hi, smyi = myi;
			}
		}
	}
#ifdef CNF2_MS_HIST
	if (active) atomicAdd(&ms_hist[MODE][myit], 1ull);
#endif
	if (!active) return orig;
	if (!scalefactor)
	{
		lo = orig;
		hi = orig;
	}
	return ms_caplogit((lo + hi) / 2, orig, epsilon, hit);
}

__device__ __forceinline__ StatsCell ms_cell(const MstepArgs& a, int row, int m)
{
	return row < a.n_shared ? stats_cell(a.stats_shared, a.n_shared, a.Mc, row, m)
	                        : stats_cell(a.stats_local, a.n_rows - a.n_shared, a.Mc, row - a.n_shared, m);
}

// processinfprobs (:4179-4323) for both sides of one (row, marker).  A statistic that is exactly 0 is an absent key
// of the reference's small_map (the E-step never adds a zero).  infprobs/homozyg are not cleared here: the next
// E-step of the chromosome zeroes its block.
__global__ void __launch_bounds__(MS_BLOCK, 5) mstep_infprobs_kernel(MstepArgs a)
{
	__shared__ double stage[MS_NPAR][MS_BLOCK];
	const size_t idx0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x, cells = (size_t)a.n_rows * a.Mc;
	const bool valid = idx0 < cells;  // every lane stays for the warp-cooperative solver
	const size_t idx = valid ? idx0 : cells - 1;
	const int row = (int)(idx / a.Mc), m = (int)(idx % a.Mc);
	const MstepRow r = a.rows[row];
	const StatsCell cell = ms_cell(a, row, m);
	uchar2 al = a.alleles[idx];
	double2 su = a.sure[idx];
	const uchar2 pal = r.hasprior ? a.prioralleles[idx] : make_uchar2(0, 0);
	const double2 psu = r.hasprior ? a.priorsure[idx] : make_double2(0, 0);
	const double scalefactor = a.ctl->sf_pass;
	const double epsilon = ms_epsilon(r.children), etf = 1 * a.ctl->entropyfactor;
	unsigned long long hit = 0;
	MsCount cnt;
#pragma unroll 1
	for (int side = 0; side < 2; side++)
	{
		double ip[2] = { cell.infprobs[side * 2], cell.infprobs[side * 2 + 1] };
		double sum = 0;
		for (int k = 0; k < 2; k++)
			if (ip[k] != 0) sum += ip[k];
		const int priorval = side ? pal.y : pal.x;
		const int curmarker = side ? al.y : al.x;
		const double cursure = side ? su.y : su.x;
#pragma unroll 1
		for (int k = 0; k < 2; k++)
		{
			const bool active = valid && ip[k] != 0;
			const int allele = k + 1;
			double curprob = 0.5;
			if (curmarker != 0) curprob = fabs((curmarker == allele ? 1 : 0) - cursure);
			double priord = 0;
			if (active && priorval != 0)
			{
				double priorprob = 1.0 - (side ? psu.y : psu.x);
				if (allele != priorval) priorprob = 1.0 - priorprob;
				if (priorprob == 0) priord -= 10000;
				else if (priorprob == 1) priord += 10000;
				else
				{
					priorprob = fmin(fmax(priorprob, 1e-14), 1 - 1e-14);
					priord += log(priorprob) - log(1 - priorprob);
				}
			}
			const MsParams P = ms_params(curprob, ip[k], sum, etf, etf * priord, 0.0, 0.0, epsilon);
			const double res = ms_cappedgd_warp<0>(P, curprob, scalefactor, active, hit, &stage[0][0], cnt, r.empty || !r.hasprior);
			if (active) ip[k] = res;
		}
		double bestprob = 0;
		int bestmarker = 0;
		for (int k = 0; k < 2; k++)
		{
			if (ip[k] == 0) continue;
			if (ip[k] > bestprob - (side ? 1e-30 : 0))
			{
				bestmarker = k + 1;
				bestprob = ip[k];
			}
		}
		if (!r.empty && (bestmarker != 0 || bestprob > 0) && r.hasprior)
		{
			if (side) al.y = (unsigned char)bestmarker, su.y = 1.0 - bestprob;
			else al.x = (unsigned char)bestmarker, su.x = 1.0 - bestprob;
		}
	}
	if (valid)
	{
		a.alleles[idx] = al;
		a.sure[idx] = su;
		if (hit) atomicAdd(a.ctl->hit + (row < a.n_shared ? 1 : 0), hit);
	}
	ms_count_flush(a.ctl, 0, cnt);
}

// relskewhmm (:4325-4467, realhmm): one thread per row walks the chromosome forward (stores the two forward values per
// marker) and backward (ratio = P(state 1)); anyinfo = some haplocount of the chromosome is non-zero (:4560-4565).
__global__ void __launch_bounds__(128) mstep_relskew_kernel(MstepPass pass)
{
	const MstepArgs a = ms_view(pass, blockIdx.y);
	const int row = blockIdx.x * blockDim.x + threadIdx.x;
	if (row >= a.n_rows) return;
	const int Mc = a.Mc;
	bool any = false;
	for (int m0 = 0; m0 < Mc && !any; m0 += 16)  // 16 independent loads per round: a row without information is scanned to its end
	{
		double hc[16];
#pragma unroll
		for (int j = 0; j < 16; j++)
		{
			const int m = min(m0 + j, Mc - 1);
			hc[j] = a.fresh ? *ms_cell(a, row, m).haplocount : a.carry_count[(size_t)row * Mc + m];
		}
#pragma unroll
		for (int j = 0; j < 16; j++)
			if (hc[j]) any = true;
	}
	a.anyinfo[row] = any;
	if (!any) return;
	const double* hw = a.hw + (size_t)row * Mc;
	const double* rh = a.relhaplo + (size_t)row * Mc;
	double2* fw = reinterpret_cast<double2*>(a.fw) + (size_t)row * Mc;
	double* ratio = a.ratio + (size_t)row * Mc;
	// A thread walks its own row, so its loads are a dependent chain of cache misses unless they are issued ahead:
	// RB markers are loaded at once, the next RB while the current ones are worked through.
	constexpr int RB = 8;
	double s0 = 0.5, s1 = 0.5;
	double wc[RB], nc[RB];
#pragma unroll
	for (int j = 0; j < RB; j++)
	{
		wc[j] = j < Mc ? hw[j] : 0.0;
		nc[j] = j < Mc ? rh[j] : 0.0;
	}
	for (int m0 = 0; m0 < Mc; m0 += RB)
	{
		double wn[RB], nn[RB];
#pragma unroll
		for (int j = 0; j < RB; j++)
		{
			const int m = m0 + RB + j;
			wn[j] = m < Mc ? hw[m] : 0.0;
			nn[j] = m < Mc ? rh[m] : 0.0;
		}
#pragma unroll
		for (int j = 0; j < RB; j++)
		{
			const int m = m0 + j;
			if (m < Mc)
			{
				const double w = wc[j];
				s0 *= fabs(1 - w);
				s1 *= fabs(0 - w);
				fw[m] = make_double2(s0, s1);
				if (s0 + s1 < 1e-10) s0 *= 1e20, s1 *= 1e20;
				const double n = nc[j], nb = 1 - n;
				const double t0 = s0 * n + s1 * nb, t1 = s1 * n + s0 * nb;
				s0 = t0, s1 = t1;
			}
		}
#pragma unroll
		for (int j = 0; j < RB; j++) wc[j] = wn[j], nc[j] = nn[j];
	}
	s0 = s1 = 0.5;
	{
		const double2 f = fw[Mc - 1];
		ratio[Mc - 1] = f.y / (f.x + f.y);
	}
	// backwards: step t handles marker m = Mc - 2 - t with hw[m + 1], relhaplo[m] and the stored forward pair of m
	const int steps = Mc - 1;
	double2 fc[RB];
#pragma unroll
	for (int j = 0; j < RB; j++)
	{
		const int m = Mc - 2 - j;
		wc[j] = j < steps ? hw[m + 1] : 0.0;
		nc[j] = j < steps ? rh[m] : 0.0;
		fc[j] = j < steps ? fw[m] : make_double2(0.0, 0.0);
	}
	for (int t0 = 0; t0 < steps; t0 += RB)
	{
		double wn[RB], nn[RB];
		double2 fn[RB];
#pragma unroll
		for (int j = 0; j < RB; j++)
		{
			const int t = t0 + RB + j, m = Mc - 2 - t;
			wn[j] = t < steps ? hw[m + 1] : 0.0;
			nn[j] = t < steps ? rh[m] : 0.0;
			fn[j] = t < steps ? fw[m] : make_double2(0.0, 0.0);
		}
#pragma unroll
		for (int j = 0; j < RB; j++)
		{
			const int t = t0 + j, m = Mc - 2 - t;
			if (t < steps)
			{
				const double w = wc[j];
				s0 *= fabs(1 - w);
				s1 *= fabs(0 - w);
				const double n = nc[j], nb = 1 - n;
				const double t0v = s0 * n + s1 * nb, t1v = s1 * n + s0 * nb;
				s0 = t0v, s1 = t1v;
				if (s0 + s1 < 1e-10) s0 *= 1e20, s1 *= 1e20;
				const double2 f = fc[j];
				const double r0 = s0 * f.x, r1 = s1 * f.y;
				ratio[m] = r1 / (r0 + r1);
			}
		}
#pragma unroll
		for (int j = 0; j < RB; j++) wc[j] = wn[j], nc[j] = nn[j], fc[j] = fn[j];
	}
}

// updatehaploweights (:4533-4734) for one (row, marker); the adjusted haplobase/haplocount are kept in the carry
// arrays because the reference re-runs this on the chromosome after every later chromosome of the iteration.
__global__ void __launch_bounds__(MS_BLOCK, 6) mstep_haploweight_kernel(MstepPass pass)
{
	__shared__ double stage[MS_NPAR][MS_BLOCK];
	const MstepArgs a = ms_view(pass, blockIdx.y);
	if ((size_t)blockIdx.x * blockDim.x >= (size_t)a.n_rows * a.Mc) return;  // the grid is sized for the longest chromosome of the pass
	const size_t idx0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x, cells = (size_t)a.n_rows * a.Mc;
	const bool valid = idx0 < cells;  // every lane stays for the warp-cooperative solver
	const size_t idx = valid ? idx0 : cells - 1;
	const int row = (int)(idx / a.Mc), m = (int)(idx % a.Mc);
	double hb, hc;
	if (a.fresh)
	{
		const StatsCell cell = ms_cell(a, row, m);
		hb = *cell.haplobase, hc = *cell.haplocount;
	}
	else hb = a.carry_base[idx], hc = a.carry_count[idx];
	const double w = a.hw[idx];
	const bool active = valid && a.anyinfo[row] && w && w != 1;
	const MstepRow r = a.rows[row];
	double similarity = 0;
	if (active)
	{
		const uchar2 al = a.alleles[idx];
		const double2 su = a.sure[idx];
		const double scorea = 1.0 - su.x;
		double scoreb = 1.0 - su.y;
		if (al.x != al.y) scoreb = 1 - scoreb;
		similarity = scorea * scoreb + (1 - scorea) * (1 - scoreb);
		if (!hc || similarity == 1.0)
		{
			hc = fmax(1.0, hc);
			hb = w * hc;
		}
		else
		{
			if (similarity >= MS_ONE_MINUS_MAXDIFF) similarity = MS_ONE_MINUS_MAXDIFF;
			double count = hc;
			hb -= count * w;
			count = count - similarity * count;
			hb += count * w;
			hb *= hc / count;
			if (hb < 0) hb = 0;
			if (hb >= hc) hb = hc;
		}
	}
	const MsParams P = ms_params(w, hb, hc, (1 - similarity) * 1, a.ctl->entropyfactor, active ? a.ratio[idx] : 0.0, (double)r.descendants,
	                             ms_epsilon(r.children));
	unsigned long long hit = 0;
	MsCount cnt;
	const double res = ms_cappedgd_warp<1>(P, w, a.ctl->sf_pass, active, hit, &stage[0][0], cnt);
	ms_count_flush(a.ctl, 1, cnt);
	if (active)
	{
		a.hw[idx] = res;
		if (hit) atomicAdd(a.ctl->hit + (row < a.n_shared ? 1 : 0), hit);
	}
	if (valid)
	{
		a.carry_base[idx] = hb;
		a.carry_count[idx] = hc;
	}
}

// ---- EM control on the device ----
// Start of a pass: the effective scalefactor (0 while phase inversions are applied, :6339-6341) and zeroed hit counters.
// explicit_sf >= 0... the synchronous cnf2_mstep hands scalefactor / entropyfactor in (set = 1).
__global__ void ms_begin_kernel(EmCtl* c, int any, int set, double sf, double ent)
{
	if (set) c->scalefactor = sf, c->entropyfactor = ent;
	c->any = any;
	c->sf_pass = any ? 0.0 : c->scalefactor;
	c->hit[0] = c->hit[1] = 0;
	c->hit_local_d = 0;
}
__global__ void ms_pack_kernel(EmCtl* c) { c->hit_local_d = (double)c->hit[0]; }
// End of a pass: the step-size control of cnF2freq.cpp:6373-6390.  reduced: hit_local_d holds the sum over ranks of the
// local counts; control = 0 leaves scalefactor and the history alone (cnf2_mstep: the caller runs the control).
__global__ void ms_end_kernel(EmCtl* c, int reduced, int control)
{
	const long long hitnnn = (reduced ? (long long)c->hit_local_d : (long long)c->hit[0]) + (long long)c->hit[1];
	c->hitnnn_last = hitnnn;
	if (!control) return;
	const double oldscalefactor = c->scalefactor;
	double sf = c->any ? 0.0 : oldscalefactor;  // :6341
	const bool badhit = hitnnn > max(c->oldhitnnn, c->oldhitnnn2);
	if (badhit) sf /= 1.1;
	const bool goodhit = (double)hitnnn < max(min(c->oldhitnnn, c->oldhitnnn2), c->n_dous_total / NSLOTS) * 0.99;
	if (goodhit) sf *= 1.21;
	sf *= 0.997;
	if (c->any) sf = oldscalefactor;
	else
	{
		c->oldhitnnn2 = c->oldhitnnn;
		c->oldhitnnn = (int32_t)hitnnn;
	}
	c->scalefactor = sf;
}

// ---- phase inversions (row f2): candidates and their application ----
// Per row: the most negative phase-inversion score of the chromosome below -1e-10 and the first marker holding it
// (cnF2freq.cpp:6272-6284); minstart = Mc when there is none.  One warp per row.  "Holding it": within the rounding noise
// of the scores (differences of logarithms of O(1) sums: 1e-15 absolute, 1e-12 relative) of the minimum.  The reference
// adds an individual's scores in one fixed order at every marker, so markers carrying the same information tie (or differ
// in the last bit the same way every run) and the first one wins; here the contributions arrive through atomics in any
// order.  Documented tie: minima closer than that are taken at the first marker.
__global__ void __launch_bounds__(128) negshift_minima_kernel(const double* stats_shared, const double* stats_local, int n_rows, int n_shared,
                                                              int Mc, double* minval, int32_t* minstart)
{
	const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
	if (row >= n_rows) return;
	const bool sh = row < n_shared;
	const size_t nr = sh ? n_shared : n_rows - n_shared;
	const double* ns = (sh ? stats_shared : stats_local) + 8 * nr * Mc + (size_t)(sh ? row : row - n_shared) * Mc;
	double best = -1e-10;
	for (int m = lane; m < Mc; m += 32) best = fmin(best, ns[m]);
	for (int o = 16; o > 0; o >>= 1) best = fmin(best, __shfl_xor_sync(0xffffffffu, best, o));
	int at = Mc;
	double val = best;
	if (best < -1e-10)
	{
		const double thr = best + fmax(1e-15, -best * 1e-12);  // best < 0: slightly above the minimum
		for (int m0 = 0; m0 < Mc && at == Mc; m0 += 32)
		{
			const int m = m0 + lane;
			const double v = m < Mc ? ns[m] : 0.0;
			const unsigned hit = __ballot_sync(0xffffffffu, m < Mc && v <= thr);
			if (hit)
			{
				const int src = __ffs(hit) - 1;
				at = m0 + src;
				val = __shfl_sync(0xffffffffu, v, src);
			}
		}
	}
	if (lane == 0) minval[row] = val, minstart[row] = at;
}
// negshifter (:3437-3459): from the marker after minstart to the chromosome's end the phase is inverted: haploweight
// -> 1 - haploweight, haplobase -> haplocount - haplobase.  One block per candidate.
__global__ void __launch_bounds__(128) apply_inversions_kernel(MstepArgs a, const int32_t* rows, const int32_t* minstart, int n)
{
	const int k = blockIdx.x;
	if (k >= n) return;
	const int row = rows[k];
	for (int m = minstart[k] + 1 + threadIdx.x; m < a.Mc; m += blockDim.x)
	{
		const size_t idx = (size_t)row * a.Mc + m;
		a.hw[idx] = 1.0 - a.hw[idx];
		if (a.fresh)
		{
			const StatsCell cell = ms_cell(a, row, m);
			*cell.haplobase = *cell.haplocount - *cell.haplobase;
		}
		else a.carry_base[idx] = a.carry_count[idx] - a.carry_base[idx];
	}
}
}  // namespace cnf2
