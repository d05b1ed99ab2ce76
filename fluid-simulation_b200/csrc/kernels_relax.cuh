// Relaxation kernels: LinearSolve1 in the reference's lexicographic order
// (tile wavefront), the red-black fast path, and Blur's Jacobi passes.
#pragma once
#include "grid.cuh"

namespace fluid {

// ---------------------------------------------------------------------------
// Wavefront Gauss-Seidel, bitwise equal to LinearSolve1 (FluidSystem.cpp:534-548).
//
// In-place lexicographic GS updates cell (i,j) of sweep k from the NEW west and
// south neighbours (sweep k) and the OLD east and north ones (sweep k-1).  Cut
// the interior into TxT tiles (I,J): tile (I,J) of sweep k then depends only on
// tiles (I-1,J,k), (I,J-1,k), (I+1,J,k-1), (I,J+1,k-1), all of which lie on the
// hyperplane  I + J + 2k = s-1.  Launch number s therefore processes every
// (I,J,k) with I+J+2k = s, all `iters` sweeps in flight at once, 2*nt+2*iters-3
// launches per solve.  No two tiles of one launch are adjacent, so each CTA may
// read its one-cell halo from global memory and write its interior back freely.
//
// Inside a tile the same argument is applied per cell: thread r owns tile row r
// and on step d updates column d-r (the cell anti-diagonal), from shared
// memory.  SetBorder after every sweep is fused: a border cell is only ever
// read by the interior cell it mirrors, so the thread that updates an edge
// cell rewrites its border cells (mirror_border) and the next sweep sees them.
// ---------------------------------------------------------------------------
template <int T>
__global__ void __launch_bounds__(T) k_gs_wavefront(GridDesc g, int b, float* __restrict__ xb,
						     const float* __restrict__ x0b, float a, float c, int s, int ntx,
						     int nty)
{
	constexpr int PS = T + 2; // (PS-1) odd multiple pattern: thread r touches r*(PS-1)+const -> conflict free
	__shared__ float sx[(T + 2) * PS];
	__shared__ float s0[T * PS];

	const int I = blockIdx.x;
	const int k = blockIdx.y;
	const int J = s - 2 * k - I;
	if (J < 0 || J >= nty) return;

	const long long off = (long long) blockIdx.z * g.plane;
	float* x = xb + off;
	const float* x0 = x0b + off;
	const int i0 = I * T + 1, j0 = J * T + 1;
	const int w = min(T, g.N - i0 + 1), h = min(T, g.NY - j0 + 1);
	const int r = threadIdx.x;

	for (int rr = 0; rr < h + 2; rr++) {
		const long long row = cell(g, i0 - 1, j0 - 1 + rr);
		for (int cc = r; cc < w + 2; cc += T) sx[rr * PS + cc] = x[row + cc];
	}
	for (int rr = 0; rr < h; rr++) {
		const long long row = cell(g, i0, j0 + rr);
		for (int cc = r; cc < w; cc += T) s0[rr * PS + cc] = __ldg(x0 + row + cc);
	}
	__syncthreads();

	const int nd = w + h - 1;
	float* me = &sx[(r + 1) * PS + 1]; // tile row r, column 0
	for (int d = 0; d < nd; d++) {
		const int col = d - r;
		if (r < h && col >= 0 && col < w) {
			const float v = relax(s0[r * PS + col], a, me[col - 1], me[col + 1], me[col - PS], me[col + PS], c);
			me[col] = v;
		}
		__syncthreads();
	}

	for (int rr = 0; rr < h; rr++) {
		const int j = j0 + rr;
		const long long row = cell(g, i0, j);
		for (int cc = r; cc < w; cc += T) {
			const float v = sx[(rr + 1) * PS + cc + 1];
			x[row + cc] = v;
			mirror_border(g, x, b, i0 + cc, j, v);
		}
	}
}

// ---------------------------------------------------------------------------
// Red-black Gauss-Seidel, one colour per launch (baseline fast path; the
// temporally blocked kernel in kernels_rb_tma.cuh supersedes it where it
// applies).  colour 0 = (i+j) even.  One thread = one float4 group of a row:
// two cells of the active colour.  Border cells are rewritten by the thread
// that updates the cell they mirror, which is equivalent to SetBorder after
// the full sweep because nothing else reads them in between.
// Algorithmic bytes per full sweep (both colours): 12 B per cell.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_rb_half(GridDesc g, int b, float* __restrict__ xb, const float* __restrict__ x0b,
						  float a, float c, int colour)
{
	const long long off = (long long) blockIdx.z * g.plane;
	float* x = xb + off;
	const float* x0 = x0b + off;
	const int q = blockIdx.x * blockDim.x + threadIdx.x;
	const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
	const int i0 = q * 4 - XO;
	if (q * 4 >= g.pitch || j > g.NY || i0 + 3 < 1 || i0 > g.N) return;
	const long long base = (long long) j * g.pitch + q * 4;
	float4 xc = *reinterpret_cast<const float4*>(x + base);
	const float4 xn = *reinterpret_cast<const float4*>(x + base + g.pitch);
	const float4 xs = *reinterpret_cast<const float4*>(x + base - g.pitch);
	const float4 r0 = __ldg(reinterpret_cast<const float4*>(x0 + base));
	const float xw = x[base - 1], xe = x[base + 4];
	// lane l is updated when (i0 + l + j) has the active parity
	const int first = (i0 + j + g.j0 + colour) & 1; // 0: lanes 0,2 ; 1: lanes 1,3 (colour = parity of i + global j)
	float vals[4] = { xc.x, xc.y, xc.z, xc.w };
	if (first == 0) {
		vals[0] = relax(r0.x, a, xw, xc.y, xs.x, xn.x, c);
		vals[2] = relax(r0.z, a, xc.y, xc.w, xs.z, xn.z, c);
	} else {
		vals[1] = relax(r0.y, a, xc.x, xc.z, xs.y, xn.y, c);
		vals[3] = relax(r0.w, a, xc.z, xe, xs.w, xn.w, c);
	}
	if (i0 >= 1 && i0 + 3 <= g.N) {
		*reinterpret_cast<float4*>(x + base) = make_float4(vals[0], vals[1], vals[2], vals[3]);
#pragma unroll
		for (int l = 0; l < 4; l++)
			if (((l ^ first) & 1) == 0) mirror_border(g, x, b, i0 + l, j, vals[l]);
	} else {
#pragma unroll
		for (int l = 0; l < 4; l++) {
			const int i = i0 + l;
			if (((l ^ first) & 1) == 0 && i >= 1 && i <= g.N) {
				x[base + l] = vals[l];
				mirror_border(g, x, b, i, j, vals[l]);
			}
		}
	}
}

// ---------------------------------------------------------------------------
// One Jacobi pass of Blur (FluidSystem.cpp:729-740): dst = relax(src) on the
// interior, reading src's border as stored (the caller has applied SetBorder
// to src, or a previous pass has via out_border).  out_border: 0 = leave dst's
// border alone, 1 = write dst's border as SetBorder(b) would (it becomes the
// next pass's source), 2 = ZeroBorder (last pass, FluidSystem.cpp:742).
// Algorithmic bytes: 8 B per cell per pass.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_blur_pass(GridDesc g, int b, float* __restrict__ dstb,
						    const float* __restrict__ srcb, float a, DivC dv, int out_border)
{
	const float c = dv.c;
	const long long off = (long long) blockIdx.z * g.plane;
	float* dst = dstb + off;
	const float* src = srcb + off;
	const int q = blockIdx.x * blockDim.x + threadIdx.x;
	const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
	const int i0 = q * 4 - XO;
	if (q * 4 >= g.pitch || j > g.NY || i0 + 3 < 1 || i0 > g.N) return;
	const long long base = (long long) j * g.pitch + q * 4;
	const float4 xc = __ldg(reinterpret_cast<const float4*>(src + base));
	const float4 xn = __ldg(reinterpret_cast<const float4*>(src + base + g.pitch));
	const float4 xs = __ldg(reinterpret_cast<const float4*>(src + base - g.pitch));
	const float xw = __ldg(src + base - 1), xe = __ldg(src + base + 4);
	float vals[4];
	bool bad = (dv.range == 0);
	if (!bad) { // verified short division (grid.cuh); any dividend outside the proven window redoes the group exactly
		vals[0] = relax_fast(xc.x, a, xw, xc.y, xs.x, xn.x, dv, bad);
		vals[1] = relax_fast(xc.y, a, xc.x, xc.z, xs.y, xn.y, dv, bad);
		vals[2] = relax_fast(xc.z, a, xc.y, xc.w, xs.z, xn.z, dv, bad);
		vals[3] = relax_fast(xc.w, a, xc.z, xe, xs.w, xn.w, dv, bad);
	}
	if (bad) {
		vals[0] = relax(xc.x, a, xw, xc.y, xs.x, xn.x, c);
		vals[1] = relax(xc.y, a, xc.x, xc.z, xs.y, xn.y, c);
		vals[2] = relax(xc.z, a, xc.y, xc.w, xs.z, xn.z, c);
		vals[3] = relax(xc.w, a, xc.z, xe, xs.w, xn.w, c);
	}
	const bool full = (i0 >= 1 && i0 + 3 <= g.N);
	if (full) *reinterpret_cast<float4*>(dst + base) = make_float4(vals[0], vals[1], vals[2], vals[3]);
	// only the groups along the walls have border cells to write (one test per thread, not per cell)
	if (full && !(i0 == 1 || i0 + 3 == g.N || j == 1 || j == g.NY)) return;
#pragma unroll
	for (int l = 0; l < 4; l++) {
		const int i = i0 + l;
		if (i < 1 || i > g.N) continue;
		if (!full) dst[base + l] = vals[l];
		if (out_border == 1) mirror_border(g, dst, b, i, j, vals[l]);
	}
}

} // namespace fluid
