// Small grids (N <= kSmallMaxN): a whole field pair fits in one SM's shared memory, so a complete
// lin_solve -- all sweeps -- or a complete Blur -- all passes -- is ONE launch of ONE CTA per
// ensemble member instead of tens of launches that each touch a few KB (N=64 was launch-latency
// bound: 2.3 ms per step on the wavefront path, slower than the CPU).
//
// Values are identical to the large-grid kernels and to the oracle: same expressions, same order.
#pragma once
#include "grid.cuh"

namespace fluid {

constexpr int kSmallMaxN = 128;
constexpr int kSmallThreads = 1024;

__host__ __device__ inline int small_pitch(int N) { return ((N + 2) & 1) ? N + 3 : N + 2; } // even: diagonal walks hit distinct banks
inline size_t small_smem_bytes(int N) { return 2 * (size_t) small_pitch(N) * (N + 2) * sizeof(float); }

// smem image of one field: rows 0..N+1, pitch P; cell (i,j) at j*P + i
__device__ __forceinline__ void small_load(const GridDesc& g, const float* __restrict__ src, float* dst, int P)
{
	const int W = g.N + 2;
	for (int idx = threadIdx.x; idx < W * W; idx += blockDim.x) {
		const int j = idx / W, i = idx - j * W;
		dst[j * P + i] = src[cell(g, i, j)];
	}
}

__device__ __forceinline__ void small_store(const GridDesc& g, float* __restrict__ dst, const float* src, int P)
{
	const int W = g.N + 2;
	for (int idx = threadIdx.x; idx < W * W; idx += blockDim.x) {
		const int j = idx / W, i = idx - j * W;
		dst[cell(g, i, j)] = src[j * P + i];
	}
}

// SetBorder for the cell just written, inside the smem image (sides only; corners are done once at the end)
__device__ __forceinline__ void small_mirror(float* x, int P, int N, int b, int i, int j, float v)
{
	if (b < 0 || b > 2) return;
	const float sx = (b == 1) ? -v : v, sy = (b == 2) ? -v : v;
	if (i == 1) x[j * P] = sx;
	if (i == N) x[j * P + N + 1] = sx;
	if (j == 1) x[i] = sy;
	if (j == N) x[(N + 1) * P + i] = sy;
}

__device__ __forceinline__ void small_corners(float* x, int P, int N)
{
	if (threadIdx.x == 0) {
		x[0] = 0.5f * (x[1] + x[P]);
		x[(N + 1) * P] = 0.5f * (x[(N + 1) * P + 1] + x[N * P]);
		x[N + 1] = 0.5f * (x[N] + x[P + N + 1]);
		x[(N + 1) * P + N + 1] = 0.5f * (x[(N + 1) * P + N] + x[N * P + N + 1]);
	}
}

// LinearSolve1 (FluidSystem.cpp:534-548), all `limit` sweeps in one launch.
//
// REDBLACK = false: the reference's lexicographic order.  Cell (i,j) of sweep k depends on cells that
// all lie on the hyperplane i + j + 2k = t - 1, so step t updates every (i,j,k) with i+j+2k = t at
// once -- up to `limit` anti-diagonals, two apart, in flight -- and the whole solve takes
// 2N + 2*limit - 3 barrier-separated steps instead of limit * (2N-1).  Cells written in a step lie on
// diagonals of one parity and read the other parity only, so the update is in place.
// REDBLACK = true: colour passes, 2*limit steps.
template <bool REDBLACK>
__global__ void __launch_bounds__(kSmallThreads, 1) k_small_lin_solve(GridDesc g, int b, float* __restrict__ xb,
								       const float* __restrict__ x0b, float a, float c, int limit)
{
	extern __shared__ float sm[];
	const int N = g.N, P = small_pitch(N);
	float* x = sm;
	float* x0 = sm + P * (N + 2);
	const long long off = (long long) blockIdx.x * g.plane;
	small_load(g, xb + off, x, P);
	small_load(g, x0b + off, x0, P);
	__syncthreads();

	if (REDBLACK) {
		const int half = (N + 1) / 2; // cells of one colour per row (at most)
		for (int hs = 0; hs < 2 * limit; hs++) {
			const int colour = hs & 1;
			for (int idx = threadIdx.x; idx < N * half; idx += blockDim.x) {
				const int j = idx / half + 1;
				const int i = 1 + ((j + 1 + colour) & 1) + 2 * (idx - (j - 1) * half);
				if (i > N) continue;
				const int o = j * P + i;
				const float v = relax(x0[o], a, x[o - 1], x[o + 1], x[o - P], x[o + P], c);
				x[o] = v;
				small_mirror(x, P, N, b, i, j, v);
			}
			__syncthreads();
		}
	} else {
		const int last = 2 * N + 2 * (limit - 1);
		for (int t = 2; t <= last; t++) {
			// sweeps with a cell on this hyperplane: 2 <= t - 2k <= 2N, at most N of them whatever `limit` is
			// (a solve run to convergence has thousands of sweeps but never more than N in flight)
			const int kmin = max(0, (t - 2 * N + 1) / 2), kmax = min(limit - 1, (t - 2) / 2);
			const int nk = kmax - kmin + 1;
			for (int idx = threadIdx.x; idx < nk * N; idx += blockDim.x) {
				const int kk = idx / N, j = idx - kk * N + 1;
				const int i = t - 2 * (kmin + kk) - j;
				if (i < 1 || i > N) continue;
				const int o = j * P + i;
				const float v = relax(x0[o], a, x[o - 1], x[o + 1], x[o - P], x[o + P], c);
				x[o] = v;
				small_mirror(x, P, N, b, i, j, v);
			}
			__syncthreads();
		}
	}
	small_corners(x, P, N);
	__syncthreads();
	small_store(g, xb + off, x, P);
}

// Blur (FluidSystem.cpp:722-743), all passes in one launch: result to xb, penultimate iterate (with
// its border set) to x0b, exactly as the pass-per-launch path leaves them.
__global__ void __launch_bounds__(kSmallThreads, 1) k_small_blur(GridDesc g, int b, float* __restrict__ xb,
								  float* __restrict__ x0b, float a, float c, int limit)
{
	extern __shared__ float sm[];
	const int N = g.N, P = small_pitch(N);
	float* dst = sm;
	float* src = sm + P * (N + 2);
	const long long off = (long long) blockIdx.x * g.plane;
	small_load(g, xb + off, dst, P);  // only its border survives into the result's intermediate states
	small_load(g, x0b + off, src, P);
	__syncthreads();
	for (int k = 0; k < limit; k++) {
		if (k > 0) {
			float* t = dst;
			dst = src;
			src = t;
		}
		// SetBorder(N, b, src)
		if (b >= 0 && b <= 2) {
			for (int q = threadIdx.x + 1; q <= N; q += blockDim.x) {
				const float l = src[q * P + 1], r = src[q * P + N], lo = src[P + q], hi = src[N * P + q];
				src[q * P] = (b == 1) ? -l : l;
				src[q * P + N + 1] = (b == 1) ? -r : r;
				src[q] = (b == 2) ? -lo : lo;
				src[(N + 1) * P + q] = (b == 2) ? -hi : hi;
			}
		}
		__syncthreads();
		small_corners(src, P, N);
		__syncthreads();
		for (int idx = threadIdx.x; idx < N * N; idx += blockDim.x) {
			const int j = idx / N + 1, i = idx - (j - 1) * N + 1;
			const int o = j * P + i;
			dst[o] = relax(src[o], a, src[o - 1], src[o + 1], src[o - P], src[o + P], c);
		}
		__syncthreads();
	}
	// ZeroBorder(N, x)
	for (int q = threadIdx.x; q <= N + 1; q += blockDim.x) {
		dst[q] = 0.0f;
		dst[(N + 1) * P + q] = 0.0f;
		dst[q * P] = 0.0f;
		dst[q * P + N + 1] = 0.0f;
	}
	__syncthreads();
	// `limit` is odd, so dst is the image that was loaded from xb and src the one loaded from x0b
	small_store(g, xb + off, dst, P);
	small_store(g, x0b + off, src, P);
}

} // namespace fluid
