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
// One tile (I, J) of one sweep: load the tile of x with its one-cell halo and the tile of x0, walk the cell
// anti-diagonals in shared memory, store the tile (and the border cells that mirror its wall cells).
// x is read past L1 (ld.global.cg): in the persistent kernel below the halo was written moments ago by CTAs
// on other SMs.
template <int T>
__device__ __forceinline__ void gs_wavefront_tile(const GridDesc& g, int b, float* __restrict__ x, const float* __restrict__ x0,
						  float a, float c, int I, int J, float* sx, float* s0)
{
	constexpr int PS = T + 2; // (PS-1) odd multiple pattern: thread r touches r*(PS-1)+const -> conflict free
	const int i0 = I * T + 1, j0 = J * T + 1;
	const int w = min(T, g.N - i0 + 1), h = min(T, g.NY - j0 + 1);
	const int r = threadIdx.x;

	if (w == T) {
		// full-width tile: the interior of a row is T/4 aligned float4 groups (i0 sits on a 128-byte line);
		// T/4 threads per row, 4 rows per step, plus the two halo columns.  All loads are issued before
		// the first value is used: a tile lives for a few microseconds and must not pay one memory
		// latency per row.
		constexpr int GPR = T / 4, RPS = 4;          // groups per row, rows per step
		constexpr int NX = (T + 2 + RPS - 1) / RPS;  // x rows per thread (incl. the two halo rows)
		constexpr int N0 = T / RPS;                  // x0 rows per thread
		const int q4 = r % GPR, rsub = r / GPR;
		float4 vx[NX], v0[N0];
		float hl[(T + 2 + T - 1) / T], hr[(T + 2 + T - 1) / T];
#pragma unroll
		for (int it = 0; it < NX; it++) {
			const int rr = rsub + RPS * it;
			if (rr < h + 2) vx[it] = __ldcg(reinterpret_cast<const float4*>(x + cell(g, i0 + 4 * q4, j0 - 1 + rr)));
		}
#pragma unroll
		for (int it = 0; it < N0; it++) {
			const int rr = rsub + RPS * it;
			if (rr < h) v0[it] = __ldg(reinterpret_cast<const float4*>(x0 + cell(g, i0 + 4 * q4, j0 + rr)));
		}
#pragma unroll
		for (int it = 0; it < (T + 2 + T - 1) / T; it++) {
			const int rr = r + T * it;
			if (rr < h + 2) {
				hl[it] = __ldcg(x + cell(g, i0 - 1, j0 - 1 + rr));
				hr[it] = __ldcg(x + cell(g, i0 + T, j0 - 1 + rr));
			}
		}
#pragma unroll
		for (int it = 0; it < NX; it++) {
			const int rr = rsub + RPS * it;
			if (rr < h + 2) {
				float* d = &sx[rr * PS + 1 + 4 * q4];
				d[0] = vx[it].x; d[1] = vx[it].y; d[2] = vx[it].z; d[3] = vx[it].w;
			}
		}
#pragma unroll
		for (int it = 0; it < N0; it++) {
			const int rr = rsub + RPS * it;
			if (rr < h) {
				float* d = &s0[rr * PS + 4 * q4];
				d[0] = v0[it].x; d[1] = v0[it].y; d[2] = v0[it].z; d[3] = v0[it].w;
			}
		}
#pragma unroll
		for (int it = 0; it < (T + 2 + T - 1) / T; it++) {
			const int rr = r + T * it;
			if (rr < h + 2) {
				sx[rr * PS] = hl[it];
				sx[rr * PS + T + 1] = hr[it];
			}
		}
	} else {
		for (int rr = 0; rr < h + 2; rr++) {
			const long long row = cell(g, i0 - 1, j0 - 1 + rr);
			for (int cc = r; cc < w + 2; cc += T) sx[rr * PS + cc] = __ldcg(x + row + cc);
		}
		for (int rr = 0; rr < h; rr++) {
			const long long row = cell(g, i0, j0 + rr);
			for (int cc = r; cc < w; cc += T) s0[rr * PS + cc] = __ldg(x0 + row + cc);
		}
	}
	__syncthreads();

	const int nd = w + h - 1;
	float* me = &sx[(r + 1) * PS + 1]; // tile row r, column 0
	for (int d = 0; d < nd; d++) {
		const int col = d - r;
		if (r < h && col >= 0 && col < w) {
			// plain IEEE division: this loop is bound by the latency of the dependent chain, and the part of
			// div.rn that depends on the dividend is as short as the verified sequence (measured: no gain)
			me[col] = relax(s0[r * PS + col], a, me[col - 1], me[col + 1], me[col - PS], me[col + PS], c);
		}
		__syncthreads();
	}

	if (w == T && !(I == 0 || J == 0 || i0 + T - 1 == g.N || j0 + h - 1 == g.NY)) {
		constexpr int GPR = T / 4, RPS = 4;
		const int q4 = r % GPR, rsub = r / GPR;
		for (int rr = rsub; rr < h; rr += RPS) {
			const float* sp = &sx[(rr + 1) * PS + 1 + 4 * q4];
			*reinterpret_cast<float4*>(x + cell(g, i0 + 4 * q4, j0 + rr)) = make_float4(sp[0], sp[1], sp[2], sp[3]);
		}
	} else {
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
}


template <int T>
__global__ void __launch_bounds__(T) k_gs_wavefront(GridDesc g, int b, float* __restrict__ xb,
						     const float* __restrict__ x0b, float a, float c, int s, int ntx,
						     int nty)
{
	constexpr int PS = T + 2;
	__shared__ float sx[(T + 2) * PS];
	__shared__ float s0[T * PS];
	const int I = blockIdx.x;
	const int k = blockIdx.y;
	const int J = s - 2 * k - I;
	if (J < 0 || J >= nty) return;
	const long long off = (long long) blockIdx.z * g.plane;
	gs_wavefront_tile<T>(g, b, xb + off, x0b + off, a, c, I, J, sx, s0);
}

// The same schedule WITHOUT a launch per hyperplane: one launch per solve.  A CTA takes a ticket = one tile
// column I of one sweep k of one member and walks it from J = 0 to nty-1; before tile (I, J, k) it waits until
//   (I-1, J, k) is done      -- its new west halo,
//   (I+1, J, k-1) is done    -- its old east halo,
//   (I, J+1, k-1) is done    -- its old north halo (the last tile of a column waits for the whole column below:
//                               it reads the border row the previous sweep mirrored),
// which it reads off per-column progress counters (tiles done, published with release semantics after the
// tile's stores).  Tickets are dealt in the order sweep-major, column-minor, and every wait is on a ticket with a
// smaller number: whoever a CTA waits for has taken its ticket earlier, so it is running or finished -- CTAs that
// have not started yet are never waited for, and the kernel cannot deadlock however few CTAs are resident.
// Round 1 needed 549 launches of a few microseconds each per solve at N = 8192 (2 ntx + 2 sweeps - 3).
// sync[0] = ticket counter, sync[1 + (sim*limit + k)*ntx + I] = progress of that column; zeroed by the host.
template <int T>
__global__ void __launch_bounds__(T, (T <= 32) ? 24 : 12) k_gs_wavefront_persistent(GridDesc g, int b, float* __restrict__ xb, const float* __restrict__ x0b,
								float a, float c, int limit, int ntx, int nty, int batch, int* __restrict__ sync)
{
	constexpr int PS = T + 2;
	__shared__ float sx[(T + 2) * PS];
	__shared__ float s0[T * PS];
	__shared__ int s_ticket;
	const int total = batch * limit * ntx;
	volatile int* progress = sync + 1;
	for (;;) {
		__syncthreads();
		if (threadIdx.x == 0) s_ticket = atomicAdd(&sync[0], 1);
		__syncthreads();
		const int ticket = s_ticket;
		if (ticket >= total) return;
		const int I = ticket % ntx, k = (ticket / ntx) % limit, sim = ticket / (ntx * limit);
		const long long off = (long long) sim * g.plane;
		volatile int* mine = progress + ticket;
		volatile int* west = (I > 0) ? progress + ticket - 1 : nullptr;
		volatile int* below = (k > 0) ? progress + ticket - ntx : nullptr;                    // (I, *, k-1)
		volatile int* east = (k > 0 && I + 1 < ntx) ? progress + ticket - ntx + 1 : nullptr;  // (I+1, *, k-1)
		for (int J = 0; J < nty; J++) {
			if (threadIdx.x == 0) {
				if (west) while (*west < J + 1) __nanosleep(64);
				if (east) while (*east < J + 1) __nanosleep(64);
				if (below) while (*below < min(J + 2, nty)) __nanosleep(64);
				__threadfence();
			}
			__syncthreads();
			gs_wavefront_tile<T>(g, b, xb + off, x0b + off, a, c, I, J, sx, s0);
			__threadfence(); // this thread's stores before the counter
			__syncthreads();
			if (threadIdx.x == 0) *mine = J + 1;
		}
	}
}

// ---------------------------------------------------------------------------
// Red-black Gauss-Seidel, one colour per launch (baseline fast path; the
// temporally blocked kernel in kernels_rb_tile.cuh supersedes it where it
// applies).  colour 0 = (i+j) even.  One thread = one float4 group of a row:
// two cells of the active colour.  Border cells are rewritten by the thread
// that updates the cell they mirror, which is equivalent to SetBorder after
// the full sweep because nothing else reads them in between.
// Algorithmic bytes per full sweep (both colours): 12 B per cell.
// ---------------------------------------------------------------------------
static __global__ void __launch_bounds__(256) k_rb_half(GridDesc g, int b, float* __restrict__ xb, const float* __restrict__ x0b,
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
static __global__ void __launch_bounds__(256) k_blur_pass(GridDesc g, int b, float* __restrict__ dstb,
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
	if (bad && dv.range != 0) {
		// a verified divisor is positive and finite; the group is on the subnormal fringe of the field: exact without a division
		vals[0] = div_bracket(xc.x + a * (((xw + xc.y) + xs.x) + xn.x), dv);
		vals[1] = div_bracket(xc.y + a * (((xc.x + xc.z) + xs.y) + xn.y), dv);
		vals[2] = div_bracket(xc.z + a * (((xc.y + xc.w) + xs.z) + xn.z), dv);
		vals[3] = div_bracket(xc.w + a * (((xc.z + xe) + xs.w) + xn.w), dv);
	} else if (bad) {
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
