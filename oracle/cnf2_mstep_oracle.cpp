// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// CPU restatement of the reference's M-step for one call of the block at cnF2freq.cpp:6356-6367 (row a6 / f1 of
// SURVEY.md section 8): per individual, processinfprobs (:4179-4323) for every marker and side, then
// updatehaploweights (:4533-4734) with its relskewhmm (:4325-4467); the step solver cappedgd (:4040-4177, the live
// `else` branch) with caplogitchange (:4006-4038).  Scalar C++, one individual after the other, markers in order.
//
// Third-party arithmetic on this path: boost::math::quadrature::gauss<double, 15>::integrate (:4150).  Boost is not
// vendored by the reference and absent here; the published 15-point Gauss-Legendre rule is restated below
// (abscissas/weights of the positive half, symmetric evaluation around the interval's centre), the same rule the
// reference build in oracle/_ref uses through oracle/refshim.
//
// Pinned by tests/test_mstep_oracle.py against the reference binary's own before/after dumps of whole EM iterations
// (tests/golden/mstep_*.npz) for iterations in which the reference made no phase inversion (negshift > -1e-10
// everywhere, so `any` at :6339 stays false and lastinved stays -1, :6290).  Deviation, documented: the dense
// [2 alleles] statistics cannot tell an absent infprobs key from a key holding exactly 0; a zero is treated as absent.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

namespace
{
const float MAXDIFF = 0.000005f;  // const float maxdiff, cnF2freq.cpp:228: `1 - maxdiff` and `maxdiff / int` are FLOAT arithmetic in the reference

struct Ctl
{
	double scalefactor;  // :3573
	long long hitnnn;
};

// caplogitchange, :4006-4038 (nnn = 3)
double caplogit(double intended, double orig, double epsilon, long long& hitnnn, bool breakathalf)
{
	const double nnn = 3;
	const double limn = (nnn - 1.0) * orig * (-1 + orig);
	const double limd1 = -1 - (nnn - 1.0) * orig;
	const double limd2 = (nnn - 1.0) * orig - nnn;
	intended = std::min(intended, 1.0 - epsilon);
	intended = std::max(intended, epsilon);
	const double diff = intended - orig;
	if (diff > limn / limd1)
	{
		intended = orig + limn / limd1;
		if (intended < 0.5) hitnnn++;
	}
	if (diff < -limn / limd2)
	{
		intended = orig - limn / limd2;
		if (intended > 0.5) hitnnn++;
	}
	if (breakathalf && (intended - 0.5) * (orig - 0.5) < 0) intended = 0.5 * (0.5 + orig);
	return intended;
}

// 15-point Gauss-Legendre on [a, b] (boost::math::quadrature::gauss<double, 15>, :4150)
template <class F>
double gauss15(F f, double a, double b)
{
	static const double x[8] = { 0.0,
		                         0.2011940939974345223006283033946,
		                         0.3941513470775633698972073709810,
		                         0.5709721726085388475372267372539,
		                         0.7244177313601700474161860546139,
		                         0.8482065834104272162006483207742,
		                         0.9372733924007059043077589477102,
		                         0.9879925180204854284895657185866 };
	static const double w[8] = { 0.2025782419255612728806201999675,
		                         0.1984314853271115764561183264438,
		                         0.1861610000155622110268005618664,
		                         0.1662692058169939335532008604812,
		                         0.1395706779261543144478047945110,
		                         0.1071592204671719350118695466859,
		                         0.0703660474881081247092674164507,
		                         0.0307532419961172683546283935772 };
	const double avg = (a + b) * 0.5, scale = (b - a) * 0.5;
	double result = f(avg) * w[0];
	for (int i = 1; i < 8; i++) result += (f(avg + scale * x[i]) + f(avg - scale * x[i])) * w[i];
	return result * scale;
}

// cappedgd, :4040-4177 (the branch compiled in): bisection on the end point so that the integral of 1 / gradient
// from orig to it equals scalefactor, inside the caplogitchange limits.
template <class G>
double cappedgd(G gradient, double orig, double epsilon, Ctl& ctl, bool breakathalf)
{
	const double scalefactor = ctl.scalefactor;
	long long dump = 0;
	auto actual = [&](double val) {
		val = std::min(std::max(val, epsilon), 1 - epsilon);
		return 1. / gradient(val);
	};
	const double lolim = caplogit(epsilon, orig, epsilon, dump, breakathalf);
	double lo = lolim - epsilon * 0.125;
	const double hilim = caplogit(1 - epsilon, orig, epsilon, dump, breakathalf);
	double hi = hilim + epsilon * 0.125;
	orig = caplogit(orig, orig, epsilon, dump, breakathalf);
	const double gradval = actual(orig);
	if (!std::isfinite(gradval) || !scalefactor)
	{
		lo = orig;
		hi = orig;
	}
	const bool lowside = gradval < 0;
	(lowside ? hi : lo) = orig;
	for (int i = 0; i < 51 && scalefactor; i++)
	{
		if (lo > hilim || hi < lolim) break;
		const double mid = (lo + hi) / 2;
		double prel = 0;
		const double g = actual(mid);
		if (((g < 0) ^ lowside) || !std::isfinite(g)) prel = (scalefactor + 0.1) * 1.1;
		else
		{
			double start = orig, end = mid;
			if (start > end) std::swap(start, end);
			if (end - start < 1e-10) break;
			prel = gauss15(actual, start, end);
			if (end != mid) prel = -prel;
			if (!std::isfinite(prel)) prel = (scalefactor + 0.1) * 1.1;
		}
		if (std::fabs(prel - scalefactor) < scalefactor * 1e-3) break;
		if ((prel < scalefactor) ^ lowside) lo = mid;
		else hi = mid;
	}
	if (!scalefactor)
	{
		lo = orig;
		hi = orig;
	}
	return caplogit((lo + hi) / 2, orig, epsilon, ctl.hitnnn, breakathalf);
}

inline double sq(double v) { return v * v; }

// The common part of both gradients (:4275 with (curprob, hzygcorred, sum), :4688 with (haploweight, haplobase,
// haplocount)), terms in the reference's order.
inline double gcore(double c, double h, double s, double x)
{
	const double lx = std::log(x), l1 = std::log(1 - x);
	return -(-sq(c * h) * lx + sq(c * h) * l1 + sq(c) * h * s * lx - sq(c) * h * s * l1 - sq(c) * h * s - sq(c * s) * x + sq(c * s) +
	         c * sq(h) * lx - c * sq(h) * l1 + c * sq(h) + 2 * c * h * s * x - c * h * s * lx + c * h * s * l1 - c * h * s - sq(h) * x) /
	       sq(c * h + c * s * x - c * s - h * x);
}
}  // namespace

namespace
{
// relskewhmm (:4325-4467, realhmm = true): two-state forward/backward over one chromosome of one individual;
// emission |!k - haploweight[m]|, transition relhaplo[m]; ratio[m] = P(state 1 | all markers).
struct RelSkew
{
	std::vector<double> ratio;
	RelSkew(int first, int end, const double* hw, const double* relhaplo)
	{
		const int n = end - first;
		ratio.assign(n, 0.0);
		std::vector<double> f0(n), f1(n);
		double s[2] = { 0.5, 0.5 };
		auto emit = [&](int m) {
			const double w = hw[m];
			for (int k = 0; k < 2; k++) s[k] *= std::fabs(!k - w);
		};
		auto trans = [&](int m) {
			const double nn = relhaplo[m], nb = 1 - nn;
			const double a = s[0] * nn + s[1] * nb, b = s[1] * nn + s[0] * nb;
			s[0] = a;
			s[1] = b;
		};
		auto renorm = [&] {
			if (s[0] + s[1] < 1e-10)
			{
				s[0] *= 1e20;
				s[1] *= 1e20;
			}
		};
		for (int m = first; m < end; m++)
		{
			emit(m);
			f0[m - first] = s[0];
			f1[m - first] = s[1];
			renorm();
			trans(m);
		}
		s[0] = s[1] = 0.5;
		ratio[n - 1] = f1[n - 1] / (f0[n - 1] + f1[n - 1]);
		for (int m = end - 2; m >= first; m--)
		{
			emit(m + 1);
			trans(m);
			renorm();
			const double r0 = s[0] * f0[m - first], r1 = s[1] * f1[m - first];
			ratio[m - first] = r1 / (r0 + r1);
		}
	}
};

struct Ind
{
	uint8_t* alleles;  // [M][2]
	double* sure;      // [M][2]
	double* hw;        // [M]
	const double* relhaplo;
	const uint8_t* prioralleles;  // [M][2] or null
	const double* priorsure;
	double *haplobase, *haplocount, *infprobs /* [M][2][2] */, *homozyg /* [M][2] */;
	bool empty;
	int children;
	double descendants;
};

// processinfprobs, :4179-4323
void processinfprobs(Ind& d, int j, int side, double entropyfactor, Ctl& ctl)
{
	double* ip = d.infprobs + ((size_t)j * 2 + side) * 2;
	double sum = 0;
	for (int a = 0; a < 2; a++)
		if (ip[a] != 0) sum += ip[a];
	const int priorval = d.prioralleles ? d.prioralleles[j * 2 + side] : 0;
	const double ef = entropyfactor;
	for (int a = 0; a < 2; a++)
	{
		if (ip[a] == 0) continue;
		const int allele = a + 1;
		double curprob = 0.5;
		const int curmarker = d.alleles[j * 2 + side];
		if (curmarker != 0) curprob = std::fabs((curmarker == allele ? 1 : 0) - d.sure[j * 2 + side]);
		const double hzygcorred = ip[a];
		const double etf = 1 * ef;
		const double epsilon = MAXDIFF / (d.children + 1);  // float division, :4242
		double priord = 0, priorprob = 0.5;
		if (priorval != 0)
		{
			priorprob = 1.0 - d.priorsure[j * 2 + side];
			if (allele != priorval) priorprob = 1.0 - priorprob;
			if (priorprob == 0) priord -= 10000;
			else if (priorprob == 1) priord += 10000;
			else
			{
				priorprob = std::min(std::max(priorprob, 1e-14), 1 - 1e-14);
				priord += std::log(priorprob) - std::log(1 - priorprob);
			}
		}
		auto gradient = [&](double x) {
			double g = gcore(curprob, hzygcorred, sum, x);
			const double et = std::log(1 / x - 1);
			g += etf * et;      // entropy term
			g += etf * priord;  // prior
			return g;
		};
		ip[a] = cappedgd(gradient, curprob, epsilon, ctl, false);
	}
	double bestprob = 0;
	int bestmarker = 0;
	for (int a = 0; a < 2; a++)
	{
		if (ip[a] == 0) continue;
		if (ip[a] > bestprob - (side ? 1e-30 : 0))
		{
			bestmarker = a + 1;
			bestprob = ip[a];
		}
	}
	if (!d.empty && (bestmarker != 0 || bestprob > 0) && d.prioralleles)
	{
		d.alleles[j * 2 + side] = (uint8_t)bestmarker;
		d.sure[j * 2 + side] = 1.0 - bestprob;
	}
	ip[0] = ip[1] = 0;
	if (side == 1) d.homozyg[j * 2] = d.homozyg[j * 2 + 1] = 0;
}

// updatehaploweights, :4533-4734 (RELSKEWS = true; lastinved = -1, see the header)
void updatehaploweights(Ind& d, int C, const uint32_t* chromstarts, double entropyfactor, Ctl& ctl)
{
	for (int c = 0; c < C; c++)
	{
		const int first = (int)chromstarts[c], end = (int)chromstarts[c + 1];
		bool anyinfo = false;
		for (int k = first; k < end && !anyinfo; k++)
			if (d.haplocount[k]) anyinfo = true;
		if (!anyinfo) continue;
		RelSkew rs(first, end, d.hw, d.relhaplo);
		for (int j = first; j < end; j++)
		{
			if (!(d.hw[j] && d.hw[j] != 1)) continue;
			const double relskewterm = rs.ratio[j - first];
			const double scorea = 1.0 - d.sure[j * 2];
			double scoreb = 1.0 - d.sure[j * 2 + 1];
			if (d.alleles[j * 2] != d.alleles[j * 2 + 1]) scoreb = 1 - scoreb;
			double similarity = scorea * scoreb + (1 - scorea) * (1 - scoreb);
			if (!d.haplocount[j] || similarity == 1.0)
			{
				d.haplocount[j] = std::max(1.0, d.haplocount[j]);
				d.haplobase[j] = d.hw[j] * d.haplocount[j];
			}
			else
			{
				if (similarity >= 1 - MAXDIFF) similarity = 1 - MAXDIFF;
				double count = d.haplocount[j];
				d.haplobase[j] -= count * d.hw[j];
				count = count - similarity * count;
				d.haplobase[j] += count * d.hw[j];
				d.haplobase[j] *= d.haplocount[j] / count;
				if (d.haplobase[j] < 0) d.haplobase[j] = 0;
				if (d.haplobase[j] >= d.haplocount[j]) d.haplobase[j] = d.haplocount[j];
			}
			const double ef = entropyfactor, w = d.hw[j], hb = d.haplobase[j], hc = d.haplocount[j];
			auto gradient = [&](double in) {
				double out = gcore(w, hb, hc, in);
				out += ((1 - similarity) * 1 * (ef * std::log(1 / in - 1)) + (relskewterm - in) / (in - in * in) * d.descendants);
				return out;
			};
			d.hw[j] = cappedgd(gradient, w, MAXDIFF / (d.children + 1), ctl, false);
		}
	}
}
}  // namespace

extern "C"
{
// One pass of :6356-6367 over the individuals with active[i] != 0 (those with statistics vectors), in id order.
// In/out: alleles, sure, haploweight (the new parameters), haplobase/haplocount (modified in place as the reference
// does), infprobs/homozyg (cleared).  prior*: per individual when hasprior[i].  Returns hitnnn.
long long cnf2_oracle_mstep(int K, int M, int C, const uint32_t* chromstarts, const uint8_t* active, const uint8_t* empty,
                            const int32_t* children, const int32_t* descendants, uint8_t* alleles, double* sure,
                            double* haploweight, const double* relhaplo, const uint8_t* hasprior, const uint8_t* prioralleles,
                            const double* priorsure, double* haplobase, double* haplocount, double* infprobs, double* homozyg,
                            double scalefactor, double entropyfactor)
{
	Ctl ctl{ scalefactor, 0 };
	for (int i = 0; i < K; i++)
	{
		if (!active[i]) continue;
		const size_t o = (size_t)i * M;
		Ind d{ alleles + o * 2, sure + o * 2, haploweight + o, relhaplo + o, hasprior[i] ? prioralleles + o * 2 : nullptr,
			   priorsure + o * 2, haplobase + o, haplocount + o, infprobs + o * 4, homozyg + o * 2, empty[i] != 0, children[i],
			   (double)descendants[i] };
		for (int j = 0; j < M; j++)
			for (int side = 0; side < 2; side++) processinfprobs(d, j, side, entropyfactor, ctl);
		updatehaploweights(d, C, chromstarts, entropyfactor, ctl);
	}
	return ctl.hitnnn;
}

// The step-size control after the pass, :6373-6390 (no inversion candidates): returns the new scalefactor and
// shifts the hitnnn history.
double cnf2_oracle_scalefactor(double scalefactor, long long hitnnn, int n_dous, int32_t* oldhitnnn, int32_t* oldhitnnn2)
{
	const int TURNBITS = 7;
	const bool badhit = hitnnn > std::max(*oldhitnnn, *oldhitnnn2);
	if (badhit) scalefactor /= 1.1;
	const bool goodhit = hitnnn < std::max(std::min(*oldhitnnn, *oldhitnnn2), n_dous / TURNBITS) * 0.99;
	if (goodhit) scalefactor *= 1.21;
	scalefactor *= 0.997;
	*oldhitnnn2 = *oldhitnnn;
	*oldhitnnn = (int32_t)hitnnn;
	return scalefactor;
}
}
