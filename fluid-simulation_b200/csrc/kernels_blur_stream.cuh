// Blur (FluidSystem.cpp:722-743), L Jacobi passes per launch: the register-resident sliding window of
// kernels_rb_stream.cuh applied to the density diffusion.
//
// One pass per launch (k_blur_pass) moves 8 B per cell per pass and sits at 70 % of HBM bandwidth;
// Blur runs 21 of them per step.  Here a warp owns a strip of 128 columns (4 per lane) and streams up
// through the rows.  A Jacobi pass needs the previous iterate of a row and of both its vertical
// neighbours, so for every level h = 0..L-1 the lane keeps a ring of the three newest rows of iterate
// h; when row t arrives (iterate 0), iterate h of row t-h is computed for h = 1..L in turn, and
// iterate L goes to memory: 8 B per cell per L passes, 12 L registers.  Horizontal neighbours are the
// lane's own columns plus two shuffles per row and level.  Rows at the ends of a chunk and columns at
// the ends of a strip lack neighbours; what is computed from the gap is wrong one cell further per
// level, hence L extra rows below and above and L (rounded up to a lane) extra columns per side.
//
// SetBorder(src) before every pass means: at every level the border cells mirror that level's wall
// cells.  The wall strips do that with a shuffle, the bottom and top chunks by writing the mirrored
// row into the level's ring.  The launch leaves its output with the border set (and the corners),
// exactly like k_blur_pass with out_border = 1, so launches and single passes can be mixed.
//
// Values are bit-identical to oracle_blur.  The division is the verified short sequence; segments of
// rows in which a dividend left the proven window are listed for k_blur_redo (exact division).
#pragma once
#include <type_traits>

#include "grid.cuh"
#include "kernels_rb_stream.cuh"

namespace fluid {

template <int L>
struct BlurGeom {
	static constexpr int HALO = ((L + 3) / 4) * 4; // untrusted columns on each side of a strip
	static constexpr int OUT = 128 - 2 * HALO;     // columns a strip produces
};

// b is 0, 1 or 2 here (the host keeps other values on the pass-per-launch path: without mirroring the
// two buffers' stored borders would alternate as the passes' inputs).
// INL: as in kernels_rb_stream.cuh -- the scaled sequence for every update; only NaN rows are listed.
template <int L, bool WALL, bool INL>
__device__ __forceinline__ void blur_stream_task(const GridDesc& g, const float* __restrict__ xin, float* __restrict__ xout, float a,
						 const DivC& dv, int lane0_col, int out_lo, int out_hi, int col_lo, int col_hi,
						 int b, float* ring, int* __restrict__ redo, int redo_base, int seg_rows)
{
	const int lane = threadIdx.x & 31;
	const int col = min(lane0_col + 4 * lane, g.pitch - 4); // lanes past the row are folded back (see kernels_rb_stream.cuh)
	const int row_lo = out_lo - L, row_hi = out_hi + L;     // rows that pass through the window; some may be virtual
	const int ylo = max(row_lo + 1, 1), yhi = min(row_hi - 1, g.NY); // rows that are computed
	// border rows inside the window (the chunk that owns row 1 / NY, and a neighbour whose halo reaches them)
	const bool has_bot = (row_lo <= 0), has_top = (row_hi >= g.NY + 1);
	const bool wallB = has_bot && (g.walls & 1), wallT = has_top && (g.walls & 2);
	const bool store_lane = (col >= col_lo) && (col + 4 <= col_hi) && (lane >= BlurGeom<L>::HALO / 4) && (lane < 32 - BlurGeom<L>::HALO / 4);
	const bool ghostL = WALL && (col + 3 == XO);
	const bool ghostR = WALL && (lane0_col + 4 * lane == XO + g.N + 1);
	const float* pin = xin + (long long) row_lo * g.pitch + col;
	float* pout = xout + (long long) out_lo * g.pitch + col;

	float4 lev[L][3]; // lev[h][k]: iterate h of the row whose index is k modulo 3 (relative to row_lo)
#pragma unroll
	for (int h = 0; h < L; h++)
#pragma unroll
		for (int k = 0; k < 3; k++) lev[h][k] = make_float4(1.0f, 1.0f, 1.0f, 1.0f); // harmless stand-ins for rows that do not exist

	uint32_t low = 0xffffffffu;
	bool bad = false, bad_before = false;
	int seg = 0, seg_left = seg_rows;
	auto close_segment = [&]() { // a dividend met while segment k was stored can spoil segments k and k+1 only
		const bool now = __any_sync(0xffffffffu, bad || (low < 2u * dv.lo - 1u));
		if ((now || bad_before) && lane == 0) redo[1 + atomicAdd(&redo[0], 1)] = redo_base + seg;
		bad_before = now;
		bad = false;
		low = 0xffffffffu;
		seg++;
		seg_left = seg_rows;
	};

	constexpr int kDepth = kStreamDepth;
	const uint32_t ring_base = (uint32_t) __cvta_generic_to_shared(ring) + 16u * lane;
	auto fetch = [&](int slot, bool real) {
		if (real) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring_base + 512u * slot), "l"(pin) : "memory");
		pin += g.pitch;
	};
	int issued = row_lo;
#pragma unroll
	for (int d = 0; d < kDepth; d++) {
		if (issued <= row_hi) {
			fetch(d, issued >= 0 && issued <= g.NY + 1);
			issued++;
		}
		asm volatile("cp.async.commit_group;" ::: "memory");
	}
	int slot = 0;

	auto mirror_y = [&](const float4& v) { return (b == 2) ? make_float4(-v.x, -v.y, -v.z, -v.w) : v; };
	// border cells of a freshly computed row: they mirror the wall cells next to them
	auto side_walls = [&](float4& o) {
		const float v1 = __shfl_down_sync(0xffffffffu, o.x, 1); // cell i = 1, seen from the lane holding i = 0
		const float vN = __shfl_up_sync(0xffffffffu, o.w, 1);   // cell i = N, seen from the lane holding i = N+1
		if (ghostL) o.w = (b == 1) ? -v1 : v1;
		if (ghostR) o.x = (b == 1) ? -vN : vN;
	};
	auto store_row = [&](const float4& v, float* p) {
		if (store_lane) *reinterpret_cast<float4*>(p) = v;
		if (WALL && ghostL) p[3] = v.w;
		if (WALL && ghostR) p[0] = v.x;
	};
	// border row next to row 1 / NY of the output, with the corners (average of the two adjacent border cells)
	auto store_border_row = [&](const float4& vi, int y) {
		float4 v = mirror_y(vi);
		if (WALL) {
			const float h1 = __shfl_down_sync(0xffffffffu, v.x, 1);
			const float hN = __shfl_up_sync(0xffffffffu, v.w, 1);
			if (ghostL) v.w = 0.5f * (h1 + vi.w);
			if (ghostR) v.x = 0.5f * (hN + vi.x);
		}
		store_row(v, xout + (long long) y * g.pitch + col);
	};

	auto block = [&](int t0, auto steady_tag) {
		constexpr bool STEADY = decltype(steady_tag)::value;
#pragma unroll
		for (int u = 0; u < 3; u++) {
			const int t = t0 + u; // row entering the window: iterate 0, ring position u
			if (!STEADY && t > row_hi) break;
			asm volatile("cp.async.wait_group %0;" ::"n"(kDepth - 1) : "memory");
			if (STEADY || t <= g.NY + 1) { // (a virtual row must not replace the top border row, see below)
				float4& d = lev[0][u];
				asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(d.x), "=f"(d.y), "=f"(d.z), "=f"(d.w) : "r"(ring_base + 512u * slot) : "memory");
			}
			if (STEADY || issued <= row_hi) {
				fetch(slot, STEADY || (issued >= 0 && issued <= g.NY + 1));
				issued++;
			}
			asm volatile("cp.async.commit_group;" ::: "memory");
			slot = (slot + 1 == kDepth) ? 0 : slot + 1;
#pragma unroll
			for (int h = 1; h <= L; h++) {
				const int y = t - h;
				if (!STEADY && (y < ylo || y > yhi)) continue;
				const int sy = (u - h + 30) % 3, sn = (u - h + 31) % 3, ss = (u - h + 29) % 3;
				const float4 c4 = lev[h - 1][sy], s4 = lev[h - 1][ss];
				float4 n4 = lev[h - 1][sn];
				// Top border row of iterate h-1 >= 1: the mirror of row NY, or the stored row where the slab goes
				// on.  It cannot be kept in the ring like the bottom one: its position is still occupied by
				// row NY-2, which the next level reads in this very row step.
				if (!STEADY && has_top && y == g.NY && h > 1) n4 = wallT ? mirror_y(c4) : lev[0][sn];
				const float wl = __shfl_up_sync(0xffffffffu, c4.w, 1);
				const float er = __shfl_down_sync(0xffffffffu, c4.x, 1);
				float4 o;
				if (INL) {
					o.x = stream_relax_full(c4.x, a, wl, c4.y, s4.x, n4.x, dv);
					o.y = stream_relax_full(c4.y, a, c4.x, c4.z, s4.y, n4.y, dv);
					o.z = stream_relax_full(c4.z, a, c4.y, c4.w, s4.z, n4.z, dv);
					o.w = stream_relax_full(c4.w, a, c4.z, er, s4.w, n4.w, dv);
				} else {
					o.x = stream_relax(c4.x, a, wl, c4.y, s4.x, n4.x, dv, low);
					o.y = stream_relax(c4.y, a, c4.x, c4.z, s4.y, n4.y, dv, low);
					o.z = stream_relax(c4.z, a, c4.y, c4.w, s4.z, n4.z, dv, low);
					o.w = stream_relax(c4.w, a, c4.z, er, s4.w, n4.w, dv, low);
				}
				if (WALL) side_walls(o);
				if (h < L) {
					lev[h < L ? h : 0][sy] = o;
					if (!STEADY) { // the bottom border row of this iterate: mirrored at a wall, unchanged where the slab goes on
						if (has_bot && y == 1) lev[h < L ? h : 0][ss] = wallB ? mirror_y(o) : s4;
					}
				} else if (STEADY ? (y >= out_lo) : (y >= out_lo && y <= out_hi)) {
					bad = bad || !(o.x == o.x) || !(o.y == o.y) || !(o.z == o.z) || !(o.w == o.w);
					store_row(o, pout);
					pout += g.pitch;
					if ((--seg_left == 0 || (!STEADY && y == out_hi))) close_segment();
					if (!STEADY && wallB && y == 1) store_border_row(o, 0);
					if (!STEADY && wallT && y == g.NY) store_border_row(o, g.NY + 1);
				}
			}
		}
	};
	for (int t0 = row_lo; t0 <= row_hi; t0 += 3) {
		// steady: every row up to the one fetched now is real, and no border row is involved
		if ((!has_bot || t0 > row_lo + 2 * L) && t0 + 2 + kDepth <= min(row_hi, g.NY + 1)) block(t0, std::true_type{});
		else block(t0, std::false_type{});
	}
}

template <int L, bool INL = false>
__global__ void __launch_bounds__(128, 4) k_blur_stream(GridDesc g, StreamRect rc, const float* __restrict__ xinb, float* __restrict__ xoutb,
							float a, DivC dv, int ntasks, int* __restrict__ redo, int b)
{
	const int task = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (task >= ntasks) return;
	__shared__ __align__(16) float ring_all[4][kStreamDepth][128];
	float* ring = &ring_all[threadIdx.x >> 5][0][0];
	const StreamTask k = stream_task(rc, task);
	const long long off = (long long) k.sim * g.plane;
	const int lane0_col = rc.col0 + k.strip * BlurGeom<L>::OUT - BlurGeom<L>::HALO;
	const bool wall = k.wall || lane0_col <= XO || lane0_col + 128 > XO + g.N + 1; // lanes reach a border column
	if (wall) blur_stream_task<L, true, INL>(g, xinb + off, xoutb + off, a, dv, lane0_col, k.out_lo, k.out_hi, rc.col0, rc.col1, b, ring, redo, task * rc.segs, rc.seg_rows);
	else blur_stream_task<L, false, INL>(g, xinb + off, xoutb + off, a, dv, lane0_col, k.out_lo, k.out_hi, rc.col0, rc.col1, b, ring, redo, task * rc.segs, rc.seg_rows);
}

// The listed (task, segment) rectangles again, with exact division: one CTA per entry does the L passes on
// the rectangle plus L cells all round in shared memory.  SEG is the segment height (rc.seg_rows).
template <int L, int SEG>
__global__ void __launch_bounds__(1024) k_blur_redo(GridDesc g, StreamRect rc, const float* __restrict__ xinb, float* __restrict__ xoutb,
						    float a, float c, const int* __restrict__ redo, int b)
{
	constexpr int OUT = BlurGeom<L>::OUT, WW = OUT + 2 * L, WH = SEG + 2 * L, P = WW + 1;
	__shared__ float bufA[WH * P], bufB[WH * P];
	const int nredo = redo[0];
	for (int e = blockIdx.x; e < nredo; e += gridDim.x) {
		const int entry = redo[1 + e];
		const StreamTask st = stream_task(rc, entry / rc.segs);
		const int ya = st.out_lo + (entry % rc.segs) * SEG, yb = min(ya + SEG - 1, st.out_hi);
		const int ia = rc.col0 + st.strip * OUT - XO, ib = min(ia + OUT - 1, g.N); // interior columns ia..ib
		const float* xin = xinb + (long long) st.sim * g.plane;
		float* xout = xoutb + (long long) st.sim * g.plane;
		const int wi0 = ia - L, wy0 = ya - L; // grid coordinates of window cell (0,0)
		float *A = bufA, *B = bufB;
		__syncthreads(); // the previous entry's stores have read the buffers
		for (int k = threadIdx.x; k < WH * WW; k += blockDim.x) {
			const int wy = k / WW, wx = k - wy * WW;
			const int i = wi0 + wx, y = wy0 + wy;
			A[wy * P + wx] = (i >= 0 && i <= g.N + 1 && y >= 0 && y <= g.NY + 1) ? xin[cell(g, i, y)] : 0.0f;
		}
		__syncthreads();
		for (int h = 1; h <= L; h++) {
			for (int k = threadIdx.x; k < WH * WW; k += blockDim.x) {
				const int wy = k / WW, wx = k - wy * WW;
				const int i = wi0 + wx, y = wy0 + wy, o = wy * P + wx;
				float v = A[o]; // cells that are not computed keep their value (rows beyond a slab's end, the window's rim)
				if (wx >= 1 && wx < WW - 1 && wy >= 1 && wy < WH - 1 && i >= 1 && i <= g.N && y >= 1 && y <= g.NY)
					v = div_wide(A[o] + a * (((A[o - 1] + A[o + 1]) + A[o - P]) + A[o + P]), c);
				B[o] = v;
			}
			__syncthreads();
			// SetBorder of this iterate (sides; corners are not read by any pass)
			for (int k = threadIdx.x; k < WH * WW; k += blockDim.x) {
				const int wy = k / WW, wx = k - wy * WW;
				const int i = wi0 + wx, y = wy0 + wy, o = wy * P + wx;
				if (y >= 1 && y <= g.NY) {
					if (i == 0 && wx + 1 < WW) B[o] = (b == 1) ? -B[o + 1] : B[o + 1];
					if (i == g.N + 1 && wx >= 1) B[o] = (b == 1) ? -B[o - 1] : B[o - 1];
				}
				if (i >= 1 && i <= g.N) {
					if (y == 0 && (g.walls & 1) && wy + 1 < WH) B[o] = (b == 2) ? -B[o + P] : B[o + P];
					if (y == g.NY + 1 && (g.walls & 2) && wy >= 1) B[o] = (b == 2) ? -B[o - P] : B[o - P];
				}
			}
			__syncthreads();
			float* t = A;
			A = B;
			B = t;
		}
		// the rectangle, the border cells next to it, and a corner if it has one
		const int i_lo = (ia == 1) ? 0 : ia, i_hi = (ib == g.N) ? g.N + 1 : ib;
		const int y_lo = (ya == 1 && (g.walls & 1)) ? 0 : ya, y_hi = (yb == g.NY && (g.walls & 2)) ? g.NY + 1 : yb;
		const int nw = i_hi - i_lo + 1, nh = y_hi - y_lo + 1;
		for (int k = threadIdx.x; k < nw * nh; k += blockDim.x) {
			const int ry = k / nw, i = i_lo + (k - ry * nw), y = y_lo + ry;
			const int o = (y - wy0) * P + (i - wi0);
			float v = A[o];
			const bool cx = (i == 0 || i == g.N + 1), cy = (y == 0 || y == g.NY + 1);
			if (cx && cy) v = 0.5f * (A[o + (i == 0 ? 1 : -1)] + A[o + (y == 0 ? P : -P)]);
			xout[cell(g, i, y)] = v;
		}
	}
}

} // namespace fluid
