// TEST INFRASTRUCTURE ONLY (oracle build). Stand-in for <Tools/Math.h>.
// Only FluidSystem::Advect3 (FluidSystem.cpp:965-966, experimental, out of
// scope) calls ScaleSqrt; its real definition is not vendored, this is a guess
// that is never executed by the ORIGINAL pipeline.
#pragma once
#include <cmath>
namespace Tools { namespace Math {
inline float ScaleSqrt(float v) { return v < 0 ? -std::sqrt(-v) : std::sqrt(v); }
} }
