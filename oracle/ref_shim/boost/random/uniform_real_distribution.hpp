/* TEST INFRASTRUCTURE ONLY: maps the Boost distribution name onto the C++ standard one (Boost is not installed). */
#pragma once
#include <random>
namespace boost { namespace random { template <class R = double> using uniform_real_distribution = std::uniform_real_distribution<R>; } }
