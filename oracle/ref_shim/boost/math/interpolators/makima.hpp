/* TEST INFRASTRUCTURE ONLY: empty stand-in so the reference headers parse (Boost is not installed). */
#pragma once
#include <vector>
namespace boost { namespace math { namespace interpolators {
template <class V> class makima { public: makima(V&&, V&&) {} makima() {} double operator()(double) const { return 0; } };
}}}
