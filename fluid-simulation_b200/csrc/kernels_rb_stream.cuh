// Red-black Gauss-Seidel for the interior of large grids: a register-resident sliding window.
//
// k_rb_tile (shared-memory tiles) ended up bound by instruction issue: 4 LDS + 1 STS of ~19
// instructions per cell update, plus per-tile overheads (ncu: 32.7 warp-level instructions per
// cell and sweep in total).  Here a WARP owns a strip of 128 columns (4 adjacent columns per lane)
// and streams up through the rows.  A row lives in registers from the moment it is loaded until
// it has received all 2S colour passes (S sweeps) and is stored -- no shared memory at all:
//   * vertical neighbours are other rows in the same thread's registers,
//   * horizontal neighbours are the thread's own columns, except one value per two updates that
//     comes from the adjacent lane with a shuffle.
// When row t arrives, colour pass h is applied to row t-h for h = 1..2S, in ascending h.  That
// is in place and needs no second copy of any row: pass h changes only cells of colour h, every
// neighbour of such a cell has the other colour, and the other colour of rows t-h+1 / t-h-1 was
// last changed by pass h-1 (a moment ago / one row step ago) -- exactly the values the red-black
// order prescribes.  Rows at the ends of a chunk and columns at the ends of a strip lack
// neighbours and are not trusted: wrong values spread one cell per pass, so a chunk loads 2S
// extra rows below and above, and a strip 2S (rounded up to a lane) extra columns left and right.
// Redundant arithmetic: 128/112 * (CY+16)/CY for S = 4; HBM traffic: 13 B per cell per FOUR sweeps.
//
// The kernel handles the rows that are at least one tile away from the bottom and top walls
// (k_rb_tile keeps those two bands and the corners) and, when rows are whole float4 groups, the full
// width including the left and right walls' set_bnd.  Values are bit-identical to
// oracle_lin_solve_rb: same expression per cell, same neighbours.
#pragma once
#include <type_traits>

#include "grid.cuh"

#ifndef STREAM_B4
#define STREAM_B4 4
#endif
namespace fluid {

constexpr int kStreamDepth = 4; // rows in flight per warp (cp.async ring)

struct StreamRect {
	int col0, col1; // storage columns [col0, col1) to produce, multiples of 4
	int row0, row1; // rows [row0, row1) to produce
	int chunk_rows; // rows per warp task
	int nstrips, nchunks;
	// Strips that contain a wall run a longer instruction sequence (and fetch it alone), so they are cut into
	// shorter tasks: wall_chunks tasks of wall_rows rows for each of the two; 0 = no strip touches a wall.
	int wall_rows, wall_chunks;
	int per_sim;    // tasks per ensemble member
};

struct StreamTask { int sim, strip, out_lo, out_hi; bool wall; };

// Task id -> rectangle; the tile kernel uses the same mapping for the redo list.  Wall tasks come first.
__device__ __forceinline__ StreamTask stream_task(const StreamRect& rc, int task)
{
	StreamTask k;
	k.sim = task / rc.per_sim;
	int q = task - k.sim * rc.per_sim, rows;
	k.wall = q < 2 * rc.wall_chunks;
	if (k.wall) {
		const int side = q / rc.wall_chunks, sub = q - side * rc.wall_chunks;
		k.strip = side ? rc.nstrips - 1 : 0;
		rows = rc.wall_rows;
		k.out_lo = rc.row0 + sub * rows;
	} else {
		q -= 2 * rc.wall_chunks;
		const int first = rc.wall_chunks ? 1 : 0, inner = rc.nstrips - 2 * first;
		const int chunk = q / inner;
		k.strip = first + q - chunk * inner;
		rows = rc.chunk_rows;
		k.out_lo = rc.row0 + chunk * rows;
	}
	k.out_hi = min(k.out_lo + rows, rc.row1) - 1;
	return k;
}

template <int S>
struct StreamGeom {
	static constexpr int L = 2 * S;                      // colour passes per residency
	static constexpr int W = L + 2;                      // rows resident per lane
	static constexpr int HALO = ((L + 3) / 4) * 4;       // untrusted columns on each side of a strip
	static constexpr int OUT = 128 - 2 * HALO;           // columns a strip produces
};

// The update with the verified short division and a 3-instruction guard instead of per-update
// predicates: `low` keeps the unsigned minimum of 2*bits(t)-1 over every dividend (the doubling drops
// the sign, the -1 sends zero -- always exact -- to the top), compared once per task against the
// lower end of the proven window; the sign of the quotient is copied from the dividend, which is a
// no-op except that it turns the sequence's +0 for t = -0 into the correct -0.  The upper end of the
// window is the largest finite float (the host checks), so only an infinite or NaN dividend can
// exceed it, and those come out of the sequence as NaN and are caught when the rows are stored.
__device__ __forceinline__ float stream_relax(float x0, float a, float w, float e, float s, float n, const DivC& dv, uint32_t& low)
{
	const float t = x0 + a * (((w + e) + s) + n);
	const float q = div_seq(t, dv);
	const uint32_t tb = __float_as_uint(t);
	low = min(low, 2u * tb - 1u);
	return __uint_as_float((__float_as_uint(q) & 0x7fffffffu) | (tb & 0x80000000u));
}

// One warp task: rows [out_lo, out_hi] x storage columns [lane0_col + HALO, ... + OUT) of one member,
// with the verified short division.  Returns (warp-uniform) whether some dividend fell outside the
// proven window; the caller then hands the task's rectangle to the tile kernel.
// WALL: the strip contains the left and/or the right wall of the grid.  The border cell next to a wall
// cell mirrors it (SetBorder: negated for b == 1) and sits in the neighbouring lane, so after every
// colour pass that updated the wall cell its new value is shuffled across and written into the border
// cell; in the other passes the border cell -- which the plain update has just treated like an interior
// cell -- gets its previous value back.  The border column is stored with the rows.
template <int S, bool WALL>
__device__ __forceinline__ bool rb_stream_task(const GridDesc& g, const float* __restrict__ xin, const float* __restrict__ x0g,
					       float* __restrict__ xout, float a, const DivC& dv, int lane0_col, int out_lo,
					       int out_hi, int col_lo, int col_hi, int b, float* ring)
{
	constexpr int L = StreamGeom<S>::L, W = StreamGeom<S>::W;
	const int lane = threadIdx.x & 31;
	// storage column of this lane's first cell; lanes of the last strip that would reach past the row
	// are folded back (they are farther than HALO from every produced column, so what they compute
	// never matters, but their loads must stay inside the array)
	const int col = min(lane0_col + 4 * lane, g.pitch - 4);
	const int row_lo = out_lo - L, row_hi = out_hi + L;    // rows that are loaded
	const bool store_lane = (col >= col_lo) && (col + 4 <= col_hi) && (lane >= StreamGeom<S>::HALO / 4) &&
				(lane < 32 - StreamGeom<S>::HALO / 4);
	// colour of cell (i, y) is (i + y + j0) & 1 with i = storage column - XO; col is a multiple of 4
	const int par0 = (col - XO + g.j0) & 1;
	// lanes holding the border cells: left border i = 0 is component .w of the lane before the one with i = 1;
	// right border i = N+1 (N % 4 == 0) is component .x of the lane after the one whose .w is i = N
	const bool mirror = (b >= 0 && b <= 2);
	const bool ghostL = WALL && (col + 3 == XO);
	const bool ghostR = WALL && (lane0_col + 4 * lane == XO + g.N + 1); // not a folded-back lane
	const float* pin = xin + (long long) row_lo * g.pitch + col;
	const float* p0 = x0g + (long long) row_lo * g.pitch + col;
	float* pout = xout + (long long) out_lo * g.pitch + col;
	float4 x[W], r[W];
	uint32_t low = 0xffffffffu;
	bool bad = false;
	// Rows arrive through a small per-lane ring in shared memory filled by cp.async kDepth rows ahead of
	// their use: DRAM latency is covered without holding the rows in flight in registers (one row ahead
	// in registers left the warps waiting on the loads for a third of their time).  A lane reads back only
	// the 2 x 16 bytes it copied itself, so no synchronisation between lanes is involved.
	constexpr int kDepth = kStreamDepth;
	constexpr int B = (S >= 4) ? STREAM_B4 : W; // row steps per unrolled block
	const uint32_t ring_base = (uint32_t) __cvta_generic_to_shared(ring) + 16u * lane;
	auto fetch = [&](int slot) { // next row of x and x0 into ring slot `slot`; always one group
		asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring_base + 1024u * slot), "l"(pin) : "memory");
		asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring_base + 1024u * slot + 512u), "l"(p0) : "memory");
		pin += g.pitch;
		p0 += g.pitch;
	};
	int issued = row_lo; // next row to fetch
#pragma unroll
	for (int d = 0; d < kDepth; d++) {
		if (issued <= row_hi) {
			fetch(d);
			issued++;
		}
		asm volatile("cp.async.commit_group;" ::: "memory");
	}
	int slot = 0;

	// B row steps with compile-time register slots.  Unrolling all W steps needs no register moves, but for
	// S = 4 that loop is 35 KB of code, more than the 32 KB instruction cache level that 12 warps at 12
	// different places in the loop can share -- ncu showed more warps waiting for instructions than
	// issuing.  Four steps and one rotation of the window (80 moves, +9 % instructions) fit: 1.81 -> 1.60 ms
	// per 20-sweep solve at N = 8192 (six steps: 1.64, two: 1.62).  STEADY: the window is full and stays inside the
	// chunk, so every pass applies and every finished row is stored -- no per-row conditions at all.
	// In steady blocks the colour phase of every row step is a compile-time constant as well (W is even, so
	// all blocks of a task start on the same parity, P0): half the code, and -- the host makes chunks an
	// even number of rows -- every warp of the launch runs the SAME half, which is what keeps the unrolled
	// loop inside the instruction cache.
	auto block = [&](int t0, auto steady_tag, auto phase_tag) {
		constexpr bool STEADY = decltype(steady_tag)::value;
		constexpr int P0 = decltype(phase_tag)::value;
#pragma unroll
		for (int u = 0; u < B; u++) {
			const int t = t0 + u; // row entering the window; it goes to slot u
			if (!STEADY && t > row_hi) break;
			asm volatile("cp.async.wait_group %0;" ::"n"(kDepth - 1) : "memory");
			{
				const uint32_t sa = ring_base + 1024u * slot;
				asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x[u].x), "=f"(x[u].y), "=f"(x[u].z), "=f"(x[u].w) : "r"(sa) : "memory");
				asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r[u].x), "=f"(r[u].y), "=f"(r[u].z), "=f"(r[u].w) : "r"(sa + 512u) : "memory");
			}
			if (STEADY || issued <= row_hi) { // refill the slot just read
				fetch(slot);
				issued++;
			}
			asm volatile("cp.async.commit_group;" ::: "memory");
			slot = (slot + 1 == kDepth) ? 0 : slot + 1;
			const int ph = STEADY ? ((P0 + u) & 1) : ((par0 + t) & 1); // parity of (column + row t)
#pragma unroll
			for (int h = 1; h <= L; h++) {
				if (!STEADY && t - h <= row_lo) continue; // the chunk's first row has no lower neighbour
				constexpr int WW = 4 * W;
				const int sy = (u - h + WW) % W, sn = (u - h + 1 + WW) % W, ss = (u - h - 1 + WW) % W;
				// row y = t - h, colour (h - 1) & 1: (par0 + y + colour) & 1 = (ph + h + h - 1) & 1 = (ph + 1) & 1,
				// the same for every pass of this row step
				if (ph) { // cells .x and .z of the row have the active colour
					const float wl = __shfl_up_sync(0xffffffffu, x[sy].w, 1);
					const float keep = x[sy].x;
					x[sy].x = stream_relax(r[sy].x, a, wl, x[sy].y, x[ss].x, x[sn].x, dv, low);
					x[sy].z = stream_relax(r[sy].z, a, x[sy].y, x[sy].w, x[ss].z, x[sn].z, dv, low);
					if (WALL) {
						if (ghostR) x[sy].x = keep; // the right border cell is not an interior cell
						const float v1 = __shfl_down_sync(0xffffffffu, x[sy].x, 1); // cell i = 1 was updated
						if (ghostL && mirror) x[sy].w = (b == 1) ? -v1 : v1;
					}
				} else {  // cells .y and .w
					const float er = __shfl_down_sync(0xffffffffu, x[sy].x, 1);
					const float keep = x[sy].w;
					x[sy].y = stream_relax(r[sy].y, a, x[sy].x, x[sy].z, x[ss].y, x[sn].y, dv, low);
					x[sy].w = stream_relax(r[sy].w, a, x[sy].z, er, x[ss].w, x[sn].w, dv, low);
					if (WALL) {
						if (ghostL) x[sy].w = keep; // the left border cell is not an interior cell
						const float vN = __shfl_up_sync(0xffffffffu, x[sy].w, 1);   // cell i = N was updated
						if (ghostR && mirror) x[sy].x = (b == 1) ? -vN : vN;
					}
				}
			}
			if (STEADY || (t - L >= out_lo && t - L <= out_hi)) { // this row has now received all L passes
				const float4 v = x[(u - L + 4 * W) % W];
				bad = bad || !(v.x == v.x) || !(v.y == v.y) || !(v.z == v.z) || !(v.w == v.w);
				if (store_lane) *reinterpret_cast<float4*>(pout) = v;
				if (WALL && ghostL) pout[3] = v.w;
				if (WALL && ghostR) pout[0] = v.x;
				pout += g.pitch;
			}
		}
		if (B != W) { // rotate the window so that the next block starts at slot 0 again
			float4 tx[W], tr[W];
#pragma unroll
			for (int i = 0; i < W; i++) {
				tx[i] = x[(i + B) % W];
				tr[i] = r[(i + B) % W];
			}
#pragma unroll
			for (int i = 0; i < W; i++) {
				x[i] = tx[i];
				r[i] = tr[i];
			}
		}
	};
	static_assert(B % 2 == 0 && B <= W, "blocks must start on one parity");
	const int phase0 = (par0 + row_lo) & 1;
	for (int t0 = row_lo; t0 <= row_hi; t0 += B) {
		if (t0 >= row_lo + 2 * L && t0 + B - 1 + kDepth <= row_hi) {
			if (phase0) block(t0, std::true_type{}, std::integral_constant<int, 1>{});
			else block(t0, std::true_type{}, std::integral_constant<int, 0>{});
		} else block(t0, std::false_type{}, std::integral_constant<int, 0>{});
	}
	bad = bad || (low < 2u * dv.lo - 1u);
	return __any_sync(0xffffffffu, bad);
}

// redo[0] counts, redo[1..] lists the tasks whose rectangle the tile kernel must recompute exactly.
template <int S>
__global__ void __launch_bounds__(128, (S >= 4) ? 3 : 4) k_rb_stream(GridDesc g, StreamRect rc, const float* __restrict__ xinb,
							const float* __restrict__ x0b, float* __restrict__ xoutb, float a, DivC dv,
							int ntasks, int* __restrict__ redo, int b)
{
	const int task = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (task >= ntasks) return;
	__shared__ __align__(16) float ring_all[4][kStreamDepth][2][128]; // per warp: kStreamDepth rows of x and x0
	float* ring = &ring_all[threadIdx.x >> 5][0][0][0];
	const StreamTask k = stream_task(rc, task);
	const long long off = (long long) k.sim * g.plane;
	const int lane0_col = rc.col0 + k.strip * StreamGeom<S>::OUT - StreamGeom<S>::HALO;
	const bool bad = k.wall ? rb_stream_task<S, true>(g, xinb + off, x0b + off, xoutb + off, a, dv, lane0_col, k.out_lo, k.out_hi, rc.col0, rc.col1, b, ring)
				: rb_stream_task<S, false>(g, xinb + off, x0b + off, xoutb + off, a, dv, lane0_col, k.out_lo, k.out_hi, rc.col0, rc.col1, b, ring);
	// A dividend outside the proven window (the subnormal fringe of a field, an overflow): what this warp
	// stored may be wrong.  The tile kernel, which spreads such slow cells over 16 warps per tile and
	// balances tiles dynamically, recomputes the rectangle with IEEE division right after this kernel.
	if (bad && (threadIdx.x & 31) == 0) redo[1 + atomicAdd(&redo[0], 1)] = task;
}

} // namespace fluid
