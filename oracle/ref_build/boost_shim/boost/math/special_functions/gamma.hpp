#pragma once
#include <cmath>
namespace boost { namespace math { inline double tgamma(double x) { return std::tgamma(x); } } }
