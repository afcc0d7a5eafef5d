// Stand-in for boost::math::tools root bracketing (used only by reactor types
// that are NOT on the batched path: src/zeroD/Reactor.cpp:113, MoleReactor.cpp:97).
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>
#include <stdexcept>
#include <utility>
namespace boost {
typedef std::uintmax_t uintmax_t;
namespace math { namespace tools {
template <class T>
struct eps_tolerance {
    T eps;
    explicit eps_tolerance(unsigned bits) {
        eps = std::max(T(std::ldexp(1.0, 1 - (int) bits)), 4 * std::numeric_limits<T>::epsilon());
    }
    bool operator()(const T& a, const T& b) const {
        return std::fabs(a - b) <= eps * std::min(std::fabs(a), std::fabs(b));
    }
};
template <class F, class T, class Tol>
std::pair<T, T> bisect(F f, T lo, T hi, Tol tol, boost::uintmax_t& max_iter) {
    T flo = f(lo), fhi = f(hi);
    if (flo == 0) return {lo, lo};
    if (fhi == 0) return {hi, hi};
    if ((flo < 0) == (fhi < 0)) throw std::domain_error("bisect: root not bracketed");
    boost::uintmax_t n = 0;
    while (n < max_iter && !tol(lo, hi)) {
        T mid = lo + (hi - lo) / 2, fm = f(mid);
        ++n;
        if (fm == 0) { lo = hi = mid; break; }
        if ((fm < 0) == (flo < 0)) { lo = mid; flo = fm; } else { hi = mid; fhi = fm; }
    }
    max_iter = n;
    return {lo, hi};
}
template <class F, class T, class Tol>
std::pair<T, T> bracket_and_solve_root(F f, const T& guess, T factor, bool rising, Tol tol,
                                       boost::uintmax_t& max_iter) {
    T a = guess, b = guess, fa = f(a), fb = fa;
    boost::uintmax_t left = max_iter;
    if (fa == 0) { max_iter = 0; return {a, a}; }
    bool grow = ((fa < 0) == rising);
    while ((fb < 0) == (fa < 0)) {
        if (left == 0) throw std::domain_error("bracket_and_solve_root: no bracket");
        a = b; fa = fb;
        b = grow ? b * factor : b / factor;
        fb = f(b);
        --left;
    }
    auto r = bisect(f, std::min(a, b), std::max(a, b), tol, left);
    max_iter = left;
    return r;
}
}}}
