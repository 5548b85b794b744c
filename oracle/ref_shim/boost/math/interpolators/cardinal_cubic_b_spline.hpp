/* TEST INFRASTRUCTURE ONLY: empty stand-in so the reference headers parse (Boost is not installed). */
#pragma once
namespace boost { namespace math { namespace interpolators {
template <class R> class cardinal_cubic_b_spline { public: cardinal_cubic_b_spline() {} template<class... A> cardinal_cubic_b_spline(A...) {} R operator()(R) const { return 0; } R prime(R) const { return 0; } };
}}}
