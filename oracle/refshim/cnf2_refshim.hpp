// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// Minimal std-only stand-ins for the Boost facilities that the reference translation unit
// (/root/reference/cnF2freq.cpp, lines 1-6794) touches, so that the reference's own hot-path code
// can be compiled here WITHOUT Boost (no Boost headers exist in this image).
// Every stub header under oracle/refshim/boost/ includes this file.
//
// Containers are mapped onto std containers with identical ordered-iteration semantics
// (boost::container::flat_map iterates in key order, exactly like std::map).  The only
// *arithmetic* supplied here is
//   * boost::math::quadrature::gauss<double,15>::integrate  (used by cappedgd, cnF2freq.cpp:4150)
//   * boost::math::pdf(binomial_distribution)               (used by readalphadata, cnF2freq.cpp:6599-6611)
// Both follow Boost's published algorithms (15-point Gauss-Legendre, symmetric evaluation order
// f(0)*w0 + sum_i (f(x_i)+f(-x_i))*w_i, scaled by (b-a)/2; binomial pmf).
#pragma once
#include <algorithm>
#include <atomic>
#include <bit>
#include <cmath>
#include <cstddef>
#include <iostream>
#include <map>
#include <random>
#include <set>
#include <sstream>
#include <string>
#include <tuple>
#include <utility>
#include <vector>

namespace boost
{
namespace container
{
template <class T, std::size_t N>
class small_vector : public std::vector<T>
{
public:
	using std::vector<T>::vector;
};

template <class T, std::size_t N>
class static_vector : public std::vector<T>
{
public:
	using std::vector<T>::vector;
};

template <class K, class T, class C = std::less<K>, class Ignored = void>
class flat_map : public std::map<K, T, C>
{
public:
	using std::map<K, T, C>::map;
};

template <class K, class C = std::less<K>, class Ignored = void>
class flat_set : public std::set<K, C>
{
public:
	using std::set<K, C>::set;
};
} // namespace container

template <class... Ts>
struct tuple : std::tuple<Ts...>
{
	using std::tuple<Ts...>::tuple;
	tuple() = default;
	tuple(const std::tuple<Ts...>& t) : std::tuple<Ts...>(t) {}

	template <int N>
	auto& get()
	{
		return std::get<N>(static_cast<std::tuple<Ts...>&>(*this));
	}

	template <int N>
	const auto& get() const
	{
		return std::get<N>(static_cast<const std::tuple<Ts...>&>(*this));
	}
};

template <class... Ts>
tuple<std::decay_t<Ts>...> make_tuple(Ts&&... ts)
{
	return tuple<std::decay_t<Ts>...>(std::forward<Ts>(ts)...);
}

template <class Target, class Source>
Target lexical_cast(const Source& s)
{
	std::ostringstream os;
	os << s;
	return os.str();
}

namespace random
{
using mt19937 = std::mt19937;
}

template <class I = int>
struct uniform_int
{
	std::uniform_int_distribution<I> d;
	uniform_int(I lo, I hi) : d(lo, hi) {}
	template <class G>
	I operator()(G& g)
	{
		return d(g);
	}
};

namespace math
{
template <class R = double>
struct binomial_distribution
{
	R n, p;
	explicit binomial_distribution(R n, R p = R(0.5)) : n(n), p(p) {}
};

template <class R>
R pdf(const binomial_distribution<R>& d, R k)
{
	if (k < 0 || k > d.n) return 0;
	R lc = std::lgamma(d.n + 1) - std::lgamma(k + 1) - std::lgamma(d.n - k + 1);
	return std::exp(lc + k * std::log(d.p) + (d.n - k) * std::log1p(-d.p));
}

template <class R, class K>
R pdf(const binomial_distribution<R>& d, K k)
{
	return pdf(d, (R)k);
}

namespace quadrature
{
template <class Real, unsigned N>
struct gauss;

// 15-point Gauss-Legendre; abscissas/weights as tabulated (QUADPACK qk31's embedded Gauss rule).
template <>
struct gauss<double, 15>
{
	template <class F>
	static double integrate(F f, double a, double b)
	{
		static const double x[8] = {
			0.0,
			0.2011940939974345223006283033946,
			0.3941513470775633698972073709810,
			0.5709721726085388475372267372539,
			0.7244177313601700474161860546139,
			0.8482065834104272162006483207742,
			0.9372733924007059043077589477102,
			0.9879925180204854284895657185866 };
		static const double w[8] = {
			0.2025782419255612728806201999675,
			0.1984314853271115764561183264438,
			0.1861610000155622110268005618664,
			0.1662692058169939335532008604812,
			0.1395706779261543144478047945110,
			0.1071592204671719350118695466859,
			0.0703660474881081247092674164507,
			0.0307532419961172683546283935772 };
		const double avg = (a + b) * 0.5;
		const double scale = (b - a) * 0.5;
		double result = f(avg) * w[0];
		for (int i = 1; i < 8; i++)
		{
			double fp = f(avg + scale * x[i]);
			double fm = f(avg - scale * x[i]);
			result += (fp + fm) * w[i];
		}
		return result * scale;
	}
};
} // namespace quadrature
} // namespace math

namespace iostreams
{
}
namespace spirit
{
}
namespace program_options
{
}
} // namespace boost
