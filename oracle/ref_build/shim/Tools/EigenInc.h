// TEST INFRASTRUCTURE ONLY (oracle build). Minimal stand-in for Eigen 3, which
// the reference pulls in through <Tools/EigenInc.h> (Inc.h:4) but the ORIGINAL
// solver path never touches (only the unused RayTrace, FluidSystem.cpp:1357-1360,
// and the collider headers do).
#pragma once
#include <cstddef>
namespace Eigen {
const int Dynamic = -1;
template <class T, int R, int C>
struct Matrix {
	T v[(R > 0 ? R : 1) * (C > 0 ? C : 1)];
	Matrix() : v{} {}
	Matrix(T a, T b) : v{} { v[0] = a; if (R * C > 1) v[1] = b; }
	Matrix(T a, T b, T c) : v{} { v[0] = a; if (R * C > 1) v[1] = b; if (R * C > 2) v[2] = c; }
	T& operator()(int i) { return v[i]; }
	T operator()(int i) const { return v[i]; }
	T dot(const Matrix& o) const { T s = 0; for (int i = 0; i < R * C; i++) s += v[i] * o.v[i]; return s; }
	Matrix operator*(T s) const { Matrix r; for (int i = 0; i < R * C; i++) r.v[i] = v[i] * s; return r; }
	Matrix operator/(T s) const { Matrix r; for (int i = 0; i < R * C; i++) r.v[i] = v[i] / s; return r; }
	Matrix& operator+=(const Matrix& o) { for (int i = 0; i < R * C; i++) v[i] += o.v[i]; return *this; }
};
template <class T, int R, int C> struct Array {};
typedef Matrix<float, 3, 3> Matrix3f;
}
