// TEST INFRASTRUCTURE ONLY (oracle build). Not product code.
//
// Stand-in for the un-vendored "AedoroLib" header <Tools/Include.h> that the
// reference includes from FluidSim/src/Inc.h:3 (include path at
// FluidSim/FluidSim.vcxproj:107).  Nothing in here is solver arithmetic: the
// hot path only uses Tools::Fill (FluidSystem.cpp:104-108,994), which is
// std::fill.  Everything else exists so that the UNMODIFIED FluidSystem.cpp
// parses and links with g++ on Linux.
#pragma once
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <iomanip>
#include <iostream>
#include <set>
#include <vector>

typedef unsigned char byte;

#ifndef __forceinline
#define __forceinline inline __attribute__((always_inline))
#endif

namespace Tools {

template <class T>
inline void Fill(std::vector<T>& v, T value) { std::fill(v.begin(), v.end(), value); }

template <class R, class A>
class UpdatableR {
public:
	virtual ~UpdatableR() {}
	virtual R Update(A) = 0;
};

template <class T>
inline void RemoveOneReverse(std::vector<T>& v, const T& item)
{
	for (size_t k = v.size(); k-- > 0;) {
		if (v[k] == item) { v.erase(v.begin() + k); return; }
	}
}

template <class T>
struct Surface2D {
	T width, height;
	T Area() const { return width * height; }
};

template <class T>
struct Point2D {
	T x, y;
};

template <class T> inline T Min(T a, T b) { return a < b ? a : b; }
template <class T> inline T Max(T a, T b) { return a > b ? a : b; }

} // namespace Tools

// Trace.h:117 names trace_line before declaring it (accepted by MSVC's
// one-phase lookup only); declare it up front so g++ parses the header.
namespace trace {
template <class T> struct V2;
template <class PointType>
std::set<V2<int>> trace_line(const PointType&, const PointType&, size_t);
}
