/* TEST INFRASTRUCTURE ONLY: maps the Boost generator name onto the C++ standard one (Boost is not installed).
 * The random table is not used on the short-range gravity path. */
#pragma once
#include <random>
namespace boost { namespace random { using ranlux48 = std::ranlux48; } }
