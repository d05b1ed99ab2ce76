// FSType::ORIGINAL_BORDERED (FluidSystem.cpp:79-82): the stencils of VelStep1B / DensStep1BS, masked per cell
// by cellInfo (FluidCode.h:61-73).  A cell takes part only if it is FLUID; a neighbour that cannot be reached
// (direction bit clear: a wall, an object, an edge switched off by a FluidBorder) is replaced by the cell's
// own value, negated for the velocity component normal to that wall.  None of these kernels applies
// SetBorder: the reference's bordered pipeline does not (except Blur's, which keeps it).
//
// The cell map is a byte per cell in the same pitched layout as the fields (cell(g,i,j) indexes both),
// one map per context, shared by the members of an ensemble.
#pragma once
#include "grid.cuh"

namespace fluid {

enum : unsigned char { CI_LEFT = 1, CI_RIGHT = 2, CI_TOP = 4, CI_BOTTOM = 8, CI_FLUID = 16, CI_OBJECT = 32, CI_NORMAL = 31, CI_DIRS = 15 };

// ResetCellInfo (FluidSystem.cpp:1215-1232): every cell a normal fluid cell, the outer ring not FLUID.
__global__ void __launch_bounds__(256) k_cell_info_reset(GridDesc g, unsigned char* __restrict__ info)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	const int j = blockIdx.y;
	if (i > g.N + 1 || j > g.NY + 1) return;
	const bool ring = (i == 0 || i == g.N + 1 || j == 0 || j == g.NY + 1);
	info[cell(g, i, j)] = ring ? (unsigned char) (CI_NORMAL & ~CI_FLUID) : (unsigned char) CI_NORMAL;
}

// CalcBorderFromContent (FluidSystem.cpp:1185-1213): a direction bit of an interior cell is set when the
// neighbour there is as FLUID as the cell itself.  Reads FLUID bits, writes direction bits: order free.
__global__ void __launch_bounds__(256) k_cell_info_calc(GridDesc g, unsigned char* __restrict__ info)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
	const int j = blockIdx.y + 1;
	if (i > g.N || j > g.NY) return;
	const long long id = cell(g, i, j);
	unsigned char m = info[id];
	const bool mine = (m & CI_FLUID) != 0;
	m &= (unsigned char) ~CI_DIRS;
	if (((info[id - 1] & CI_FLUID) != 0) == mine) m |= CI_LEFT;
	if (((info[id + 1] & CI_FLUID) != 0) == mine) m |= CI_RIGHT;
	if (((info[id + g.pitch] & CI_FLUID) != 0) == mine) m |= CI_TOP;
	if (((info[id - g.pitch] & CI_FLUID) != 0) == mine) m |= CI_BOTTOM;
	info[id] = m;
}

// DisableEdge (FluidSystem.cpp:1388-1408) for a list of edges (x0, y0, x1, y1): the edge between two adjacent
// cells becomes a wall.  Several edges may touch one cell, and four cells share a word: atomic AND on the word.
__global__ void __launch_bounds__(256) k_cell_info_disable_edges(GridDesc g, unsigned char* __restrict__ info, const int* __restrict__ e, int n)
{
	const int k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n) return;
	const int x0 = e[4 * k], y0 = e[4 * k + 1], x1 = e[4 * k + 2], y1 = e[4 * k + 3];
	unsigned char ba = 0, bb = 0;
	if (x1 > x0) { ba = CI_RIGHT; bb = CI_LEFT; }
	else if (x0 > x1) { ba = CI_LEFT; bb = CI_RIGHT; }
	else if (y1 > y0) { ba = CI_TOP; bb = CI_BOTTOM; }
	else if (y0 > y1) { ba = CI_BOTTOM; bb = CI_TOP; }
	auto clear = [&](int x, int y, unsigned char bit) {
		const long long id = cell(g, x, y);
		unsigned int* w = reinterpret_cast<unsigned int*>(info + (id & ~3LL));
		atomicAnd(w, ~((unsigned int) bit << (8 * (int) (id & 3))));
	};
	clear(x0, y0, ba);
	clear(x1, y1, bb);
}

// One masked relaxation (FluidSystem.cpp:565-573): neighbours in the order left, right, bottom, top; the
// multiplications by +-1 are kept as multiplications.
__device__ __forceinline__ float relax_b(unsigned char m, float x0, float a, float mine, float w, float e, float s, float n, float c,
					 float bScX, float bScY)
{
	float add = (m & CI_LEFT) ? w : (bScX * mine);
	add += (m & CI_RIGHT) ? e : (bScX * mine);
	add += (m & CI_BOTTOM) ? s : (bScY * mine);
	add += (m & CI_TOP) ? n : (bScY * mine);
	return div_exact(x0 + a * add, c);
}

// The three masked relaxations of the bordered solver types:
//   MODE_B    LinearSolve1B   (FluidSystem.cpp:550-578)  scalar a, c
//   MODE_BSO  LinearSolve1BSO (FluidSystem.cpp:673-720)  a, c per cell (viscosity mixed from the densities, k_mf_coeffs)
//   MODE_BO   LinearSolve1BO  (FluidSystem.cpp:580-615)  neighbours in the order right, left, top, bottom; an
//             unreachable neighbour is replaced by bScX * own value (bScX for all four sides, as the reference has
//             it) plus the neighbour cell's own value when that cell is not FLUID (an object's velocity)
enum { MODE_B = 0, MODE_BSO = 1, MODE_BO = 2 };

__device__ __forceinline__ float relax_bo(unsigned char m, unsigned char me, unsigned char mw, unsigned char mn, unsigned char ms, float x0, float a,
					  float mine, float w, float e, float s, float n, float c, float bScX)
{
	float add = (m & CI_RIGHT) ? e : (bScX * mine + ((me & CI_FLUID) ? 0.0f : e));
	add += (m & CI_LEFT) ? w : (bScX * mine + ((mw & CI_FLUID) ? 0.0f : w));
	add += (m & CI_TOP) ? n : (bScX * mine + ((mn & CI_FLUID) ? 0.0f : n));
	add += (m & CI_BOTTOM) ? s : (bScX * mine + ((ms & CI_FLUID) ? 0.0f : s));
	return div_exact(x0 + a * add, c);
}

// Per-cell coefficients of LinearSolve1BSO (FluidSystem.cpp:700-716): a = aSc * (sum_n d_n*coef_n) / sum_n d_n where the
// summed density exceeds 1e-9, else aSc * a_fallback; c = 1 + 4a.  The densities do not change during VelStep, so the
// pair is computed once per velocity step instead of once per cell and sweep.
struct MfFluids {
	const float* dens[8];
	float coef[8];
	int n;
};

__global__ void __launch_bounds__(256) k_mf_coeffs(GridDesc g, MfFluids fl, float aSc, float a_fallback, float* __restrict__ ab, float* __restrict__ cb)
{
	const long long off = (long long) blockIdx.z * g.plane;
	const int i = blockIdx.x * blockDim.x + threadIdx.x + 1, j = blockIdx.y + 1;
	if (i > g.N || j > g.NY) return;
	const long long id = off + cell(g, i, j);
	float c = 1.0f, a = 0.0f, total = 0.0f;
	for (int n = 0; n < fl.n; n++) {
		const float d = __ldg(fl.dens[n] + id);
		a += d * fl.coef[n];
		total += d;
	}
	if (total > 0.000000001f) a = div_exact(aSc * a, total);
	else a = aSc * a_fallback;
	c += 4 * a;
	ab[id] = a;
	cb[id] = c;
}

// The masked relaxations in the reference's lexicographic order: the tile hyperplane schedule of k_gs_wavefront
// (kernels_relax.cuh) -- launch s runs every tile (I, J) of sweep k with I + J + 2k = s; inside a tile thread r walks
// row r along the cell anti-diagonal.  A masked stencil reads a subset of the plain one's neighbours (or the cell
// itself, or a non-FLUID cell that no sweep writes), so the same schedule respects every dependence.
template <int T, int MODE>
__global__ void __launch_bounds__(T) k_gs_wavefront_b(GridDesc g, int b, float* __restrict__ xb, const float* __restrict__ x0b,
						       const unsigned char* __restrict__ info, float a, float c, const float* __restrict__ ab,
						       const float* __restrict__ cb, int s, int ntx, int nty)
{
	constexpr int PS = T + 2;
	__shared__ float sx[(T + 2) * PS];
	__shared__ float s0[T * PS];
	__shared__ float sa[(MODE == MODE_BSO) ? T * PS : 1], sc[(MODE == MODE_BSO) ? T * PS : 1];
	__shared__ unsigned char sm[(T + 2) * PS]; // with halo (MODE_BO looks at the neighbours' FLUID bits)
	const int I = blockIdx.x, k = blockIdx.y, J = s - 2 * k - I;
	if (J < 0 || J >= nty) return;
	const long long off = (long long) blockIdx.z * g.plane;
	float* x = xb + off;
	const float* x0 = x0b + off;
	const int i0 = I * T + 1, j0 = J * T + 1;
	const int w = min(T, g.N - i0 + 1), h = min(T, g.NY - j0 + 1);
	const int r = threadIdx.x;
	for (int rr = 0; rr < h + 2; rr++) {
		const long long row = cell(g, i0 - 1, j0 - 1 + rr);
		for (int cc = r; cc < w + 2; cc += T) {
			sx[rr * PS + cc] = x[row + cc];
			sm[rr * PS + cc] = info[row + cc];
		}
	}
	for (int rr = 0; rr < h; rr++) {
		const long long row = cell(g, i0, j0 + rr);
		for (int cc = r; cc < w; cc += T) {
			s0[rr * PS + cc] = __ldg(x0 + row + cc);
			if (MODE == MODE_BSO) {
				sa[rr * PS + cc] = __ldg(ab + off + row + cc);
				sc[rr * PS + cc] = __ldg(cb + off + row + cc);
			}
		}
	}
	__syncthreads();
	const float bScX = (b == 1) ? -1.0f : 1.0f, bScY = (b == 2) ? -1.0f : 1.0f;
	float* me = &sx[(r + 1) * PS + 1];
	const unsigned char* mm = &sm[(r + 1) * PS + 1];
	const int nd = w + h - 1;
	for (int d = 0; d < nd; d++) {
		const int col = d - r;
		if (r < h && col >= 0 && col < w) {
			const unsigned char m = mm[col];
			if (m & CI_FLUID) {
				const float ca = (MODE == MODE_BSO) ? sa[r * PS + col] : a, ccc = (MODE == MODE_BSO) ? sc[r * PS + col] : c;
				if (MODE == MODE_BO)
					me[col] = relax_bo(m, mm[col + 1], mm[col - 1], mm[col + PS], mm[col - PS], s0[r * PS + col], ca, me[col], me[col - 1],
							   me[col + 1], me[col - PS], me[col + PS], ccc, bScX);
				else
					me[col] = relax_b(m, s0[r * PS + col], ca, me[col], me[col - 1], me[col + 1], me[col - PS], me[col + PS], ccc, bScX, bScY);
			}
		}
		__syncthreads();
	}
	for (int rr = 0; rr < h; rr++) {
		const long long row = cell(g, i0, j0 + rr);
		for (int cc = r; cc < w; cc += T) x[row + cc] = sx[(rr + 1) * PS + cc + 1];
	}
}

// Red-black ordering of the same relaxations, one colour per launch, one thread per cell of that colour
// (colour 0 = i + j even).  No counterpart in the reference; bit-identical to the oracle's red-black statements.
template <int MODE>
__global__ void __launch_bounds__(256) k_rb_half_b(GridDesc g, int b, float* __restrict__ xb, const float* __restrict__ x0b,
						    const unsigned char* __restrict__ info, float a, float c, const float* __restrict__ ab,
						    const float* __restrict__ cb, int colour)
{
	const long long off = (long long) blockIdx.z * g.plane;
	float* x = xb + off;
	const float* x0 = x0b + off;
	const int j = blockIdx.y + 1;
	const int i = 1 + ((j + 1 + colour) & 1) + 2 * (blockIdx.x * blockDim.x + threadIdx.x);
	if (i > g.N || j > g.NY) return;
	const long long id = cell(g, i, j);
	const unsigned char m = info[id];
	if (!(m & CI_FLUID)) return;
	const float bScX = (b == 1) ? -1.0f : 1.0f, bScY = (b == 2) ? -1.0f : 1.0f;
	if (MODE == MODE_BSO) {
		a = __ldg(ab + off + id);
		c = __ldg(cb + off + id);
	}
	if (MODE == MODE_BO)
		x[id] = relax_bo(m, info[id + 1], info[id - 1], info[id + g.pitch], info[id - g.pitch], __ldg(x0 + id), a, x[id], x[id - 1], x[id + 1],
				 x[id - g.pitch], x[id + g.pitch], c, bScX);
	else
		x[id] = relax_b(m, __ldg(x0 + id), a, x[id], x[id - 1], x[id + 1], x[id - g.pitch], x[id + g.pitch], c, bScX, bScY);
}

// One pass of BlurB (FluidSystem.cpp:758-772): dst = masked stencil of src on FLUID cells; the others keep
// what dst held.  The caller applies SetBorder to src before the pass, as the reference does.
__global__ void __launch_bounds__(256) k_blur_pass_b(GridDesc g, float* __restrict__ dstb, const float* __restrict__ srcb,
						      const unsigned char* __restrict__ info, float a, float c)
{
	const long long off = (long long) blockIdx.z * g.plane;
	const float* src = srcb + off;
	const int i = blockIdx.x * blockDim.x + threadIdx.x + 1, j = blockIdx.y + 1;
	if (i > g.N || j > g.NY) return;
	const long long id = cell(g, i, j);
	const unsigned char m = info[id];
	if (!(m & CI_FLUID)) return;
	const float mine = __ldg(src + id);
	(dstb + off)[id] = relax_b(m, mine, a, mine, __ldg(src + id - 1), __ldg(src + id + 1), __ldg(src + id - g.pitch), __ldg(src + id + g.pitch), c,
				   1.0f, 1.0f);
}

// ProjectB part 1 (FluidSystem.cpp:1014-1032) fused with the whole-array p = 0 (:1034).
__global__ void __launch_bounds__(256) k_divergence_b(GridDesc g, const float* __restrict__ ub, const float* __restrict__ vb,
						       float* __restrict__ divb, float* __restrict__ pb, const unsigned char* __restrict__ info)
{
	const long long off = (long long) blockIdx.z * g.plane;
	const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y; // every stored cell: p is zeroed everywhere
	if (i > g.N + 1 || j > g.NY + 1) return;
	const long long id = cell(g, i, j);
	(pb + off)[id] = 0.0f;
	if (i < 1 || i > g.N || j < 1 || j > g.NY) return;
	const float* u = ub + off;
	const float* v = vb + off;
	const unsigned char m = info[id];
	float r = 0.0f;
	if (m & CI_FLUID) {
		const float uc = __ldg(u + id), vc = __ldg(v + id);
		float val = 0.0f;
		val += (m & CI_RIGHT) ? __ldg(u + id + 1) : -uc;
		val -= (m & CI_LEFT) ? __ldg(u + id - 1) : -uc;
		val += (m & CI_TOP) ? __ldg(v + id + g.pitch) : -vc;
		val -= (m & CI_BOTTOM) ? __ldg(v + id - g.pitch) : -vc;
		r = div_exact(-0.5f * val, (float) g.N);
	}
	(divb + off)[id] = r;
}

// ProjectB part 3 (FluidSystem.cpp:1038-1058): no SetBorder afterwards.
__global__ void __launch_bounds__(256) k_subtract_gradient_b(GridDesc g, float* __restrict__ ub, float* __restrict__ vb,
							      const float* __restrict__ pb, const unsigned char* __restrict__ info)
{
	const long long off = (long long) blockIdx.z * g.plane;
	const int i = blockIdx.x * blockDim.x + threadIdx.x + 1, j = blockIdx.y + 1;
	if (i > g.N || j > g.NY) return;
	const long long id = cell(g, i, j);
	float* u = ub + off;
	float* v = vb + off;
	const float* p = pb + off;
	const unsigned char m = info[id];
	if (m & CI_FLUID) {
		const float mid = __ldg(p + id);
		const float l = (m & CI_LEFT) ? __ldg(p + id - 1) : mid;
		const float r = (m & CI_RIGHT) ? __ldg(p + id + 1) : mid;
		const float t = (m & CI_TOP) ? __ldg(p + id + g.pitch) : mid;
		const float bo = (m & CI_BOTTOM) ? __ldg(p + id - g.pitch) : mid;
		const float h = 0.5f * (float) g.N;
		u[id] -= h * (r - l);
		v[id] -= h * (t - bo);
	} else {
		u[id] = 0.0f;
		v[id] = 0.0f;
	}
}

// ProjectBO part 1 (FluidSystem.cpp:1066-1092) fused with the whole-array p = 0 (:1094): the divergence of EVERY interior
// cell (the reference's FLUID test is commented out), an unreachable neighbour replaced by minus the own value plus the
// neighbour cell's velocity when that cell is not FLUID.
__global__ void __launch_bounds__(256) k_divergence_bo(GridDesc g, const float* __restrict__ ub, const float* __restrict__ vb,
							float* __restrict__ divb, float* __restrict__ pb, const unsigned char* __restrict__ info)
{
	const long long off = (long long) blockIdx.z * g.plane;
	const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
	if (i > g.N + 1 || j > g.NY + 1) return;
	const long long id = cell(g, i, j);
	(pb + off)[id] = 0.0f;
	if (i < 1 || i > g.N || j < 1 || j > g.NY) return;
	const float* u = ub + off;
	const float* v = vb + off;
	const unsigned char m = info[id];
	const float uc = __ldg(u + id), vc = __ldg(v + id);
	const float ue = __ldg(u + id + 1), uw = __ldg(u + id - 1), vn = __ldg(v + id + g.pitch), vs = __ldg(v + id - g.pitch);
	float val = 0.0f;
	val += (m & CI_RIGHT) ? ue : (-uc + ((info[id + 1] & CI_FLUID) ? 0.0f : ue));
	val -= (m & CI_LEFT) ? uw : (-uc + ((info[id - 1] & CI_FLUID) ? 0.0f : uw));
	val += (m & CI_TOP) ? vn : (-vc + ((info[id + g.pitch] & CI_FLUID) ? 0.0f : vn));
	val -= (m & CI_BOTTOM) ? vs : (-vc + ((info[id - g.pitch] & CI_FLUID) ? 0.0f : vs));
	(divb + off)[id] = div_exact(-0.5f * val, (float) g.N);
}

} // namespace fluid
