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
// The kernel handles all rows including the bottom and top walls' set_bnd and, when rows are whole
// float4 groups, the full width including the left and right walls and the corners (otherwise
// k_rb_tile keeps one tile column along each side wall).  Values are bit-identical to
// oracle_lin_solve_rb: same expression per cell, same neighbours.
#pragma once
#include <type_traits>

#include "grid.cuh"

#ifndef STREAM_B4
#define STREAM_B4 4
#endif
#ifndef STREAM_INL_B
#define STREAM_INL_B 2
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
	int row_shift;  // the first and last chunk are this much shorter: their wall handling costs extra row steps
	int seg_rows;   // a task reports trouble per segment of this many rows (the tile kernel's tile height)
	int segs;       // segments per task at most; a redo entry is task * segs + segment
	// Ensembles (nsims > 1): all members' wall tasks first, then all members' inner tasks.  The two kinds run different
	// instruction sequences of 15-30 KB each, and warps of both kinds on one SM at the same time miss the instruction cache
	// on most fetches (ncu, 512 members of N=512, members' tasks interleaved: 7 warps waiting for instructions per issue,
	// 22 % issue utilisation; two of a member's five strips are wall strips).
	int nsims;
};

struct StreamTask { int sim, strip, out_lo, out_hi; bool wall; };

// Task id -> rectangle; the tile kernel uses the same mapping for the redo list.  Wall tasks come first.
__device__ __forceinline__ StreamTask stream_task(const StreamRect& rc, int task)
{
	StreamTask k;
	int q, rows;
	if (rc.nsims > 1) {
		const int nwall = 2 * rc.wall_chunks, wall_total = nwall * rc.nsims;
		if (task < wall_total) {
			k.sim = task / nwall;
			q = task - k.sim * nwall;
		} else {
			const int t2 = task - wall_total, ninner = rc.per_sim - nwall;
			k.sim = t2 / ninner;
			q = nwall + t2 - k.sim * ninner;
		}
	} else {
		k.sim = task / rc.per_sim;
		q = task - k.sim * rc.per_sim;
	}
	k.wall = q < 2 * rc.wall_chunks;
	if (k.wall) {
		const int side = q / rc.wall_chunks, sub = q - side * rc.wall_chunks;
		k.strip = side ? rc.nstrips - 1 : 0;
		rows = rc.wall_rows;
		k.out_lo = rc.row0 - rc.row_shift + sub * rows;
	} else {
		q -= 2 * rc.wall_chunks;
		const int first = rc.wall_chunks ? 1 : 0, inner = rc.nstrips - 2 * first;
		const int chunk = q / inner;
		k.strip = first + q - chunk * inner;
		rows = rc.chunk_rows;
		k.out_lo = rc.row0 - rc.row_shift + chunk * rows;
	}
	k.out_hi = min(k.out_lo + rows, rc.row1) - 1;
	k.out_lo = max(k.out_lo, rc.row0);
	return k;
}

template <int S>
struct StreamGeom {
	static constexpr int L = 2 * S;                      // colour passes per residency
	static constexpr int W = L + 2;                      // rows resident per lane
	static constexpr int HALO = ((L + 3) / 4) * 4;       // untrusted columns on each side of a strip
	static constexpr int OUT = 128 - 2 * HALO;           // columns a strip produces
};

// The update with the verified short division and a 2-instruction guard instead of per-update
// predicates: `low` keeps the unsigned minimum of 2*bits(t)-1 over every dividend (the doubling drops
// the sign, the -1 sends zero -- always exact -- to the top), compared once per task against the
// lower end of the proven window (zeros keep their sign inside div_seq).  The upper end of the
// window is the largest finite float (the host checks), so only an infinite or NaN dividend can
// exceed it, and those come out of the sequence as NaN and are caught when the rows are stored.
// POW2: a == 1 and c is a power of two -- the pressure solve of Project (a = 1, c = 4).  1.0f * s is s and
// t / 2^k is t * 2^-k, the same real number rounded the same way, for every t (subnormal results, infinities
// and NaN included): five instructions instead of twelve, no window, nothing to redo.
template <bool POW2 = false>
__device__ __forceinline__ float stream_relax(float x0, float a, float w, float e, float s, float n, const DivC& dv, uint32_t& low)
{
	if (POW2) return (x0 + (((w + e) + s) + n)) * dv.y;
	const float t = x0 + a * (((w + e) + s) + n);
	low = min(low, 2u * __float_as_uint(t) - 1u);
	return div_seq(t, dv);
}

// The update with the scaled sequence (grid.cuh: div_scaled), exact for every dividend below 2^64: no window, no guard.
__device__ __forceinline__ float stream_relax_full(float x0, float a, float w, float e, float s, float n, const DivC& dv)
{
	const float t = x0 + a * (((w + e) + s) + n);
	return div_scaled(t * 0x1p64f, dv);
}

// ... with x0 already scaled by 2^64 (the right-hand side of a red-black solve is the same in every pass, so a row is
// scaled once when it enters the window): RN(p * 2^64 + x0 * 2^64) = 2^64 * RN(p + x0) -- both sums round at the same
// relative position, and where p + x0 is subnormal it is exact in both -- so one fused multiply-add forms the scaled
// dividend from the rounded product p = a * sum.  (An x0 that overflows when scaled makes the dividend infinite or NaN:
// listed.)
__device__ __forceinline__ float stream_relax_full_pre(float x0s, float a, float w, float e, float s, float n, const DivC& dv)
{
	const float p = a * (((w + e) + s) + n);
	return div_scaled(__fmaf_rn(p, 0x1p64f, x0s), dv);
}

// One warp task: rows [out_lo, out_hi] x storage columns [lane0_col + HALO, ... + OUT) of one member,
// with the verified short division.  Segments of rows in which some dividend fell outside the proven
// window (the subnormal fringe of a field, an overflow) may have been stored wrong: they go on the redo
// list and the tile kernel -- which spreads such slow cells over 16 warps per tile and balances tiles
// dynamically -- recomputes them with IEEE division right after this kernel.  A dividend met while
// segment k was being stored belongs to a row at most 2S ahead of the stored one, so it can spoil
// segments k and k+1, never an earlier one.
// WALL: the strip contains the left and/or the right wall of the grid.  The border cell next to a wall
// cell mirrors it (SetBorder: negated for b == 1) and sits in the neighbouring lane, so after every
// colour pass that updated the wall cell its new value is shuffled across and written into the border
// cell; in the other passes the border cell -- which the plain update has just treated like an interior
// cell -- gets its previous value back.  The border column is stored with the rows.
// XZERO: x is known to be zero everywhere (Project has just "set" p = 0 without writing it): it is not read.
// INL: every update uses the scaled sequence (stream_relax_full_pre), which is exact on the subnormal fringe of a field
// as well: nothing is listed for it.  Nine instructions more per update than the windowed sequence with its guard, so
// it is the choice where the fringe is most of the field (ensembles of small grids: a point source decaying through the subnormals in every member; the host
// switches over when the redo counts say so) and not elsewhere.  Only infinite, NaN or >= 2^64 dividends are still
// listed (they reach the stored rows as NaN).
template <int S, bool WALL, bool POW2, bool XZERO, bool INL, int BS>
__device__ __forceinline__ void rb_stream_task(const GridDesc& g, const float* __restrict__ xin, const float* __restrict__ x0g,
					       float* __restrict__ xout, float a, const DivC& dv, int lane0_col, int out_lo,
					       int out_hi, int col_lo, int col_hi, int b, float* ring, int* __restrict__ redo, int redo_base, int seg_rows)
{
	constexpr int L = StreamGeom<S>::L, W = StreamGeom<S>::W;
	const int lane = threadIdx.x & 31;
	// storage column of this lane's first cell; lanes of the last strip that would reach past the row
	// are folded back (they are farther than HALO from every produced column, so what they compute
	// never matters, but their loads must stay inside the array)
	const int col = min(lane0_col + 4 * lane, g.pitch - 4);
	// Rows that pass through the window.  At the bottom and top of the grid the range runs past the stored
	// rows 0..NY+1: those row steps are virtual -- nothing is fetched, no pass applies -- and keep every
	// task on the same schedule (and the same colour phase).
	const int row_lo = out_lo - L, row_hi = out_hi + L;
	const int ylo = max(row_lo + 1, 1), yhi = min(row_hi - 1, g.NY); // rows that receive colour passes
	const bool botw = (out_lo == 1), topw = (out_hi == g.NY);         // the task owns the bottom / top row (and stores the border row)
	// The border row itself may also lie in the halo of the NEXT chunk (edge chunks are short): every task whose
	// window contains it must keep it mirrored, or its halo goes wrong from the wall inwards.
	const bool has_bot = (row_lo <= 0), has_top = (row_hi >= g.NY + 1);
	const bool store_lane = (col >= col_lo) && (col + 4 <= col_hi) && (lane >= StreamGeom<S>::HALO / 4) &&
				(lane < 32 - StreamGeom<S>::HALO / 4);
	// colour of cell (i, y) is (i + y + j0) & 1 with i = storage column - XO; col is a multiple of 4
	const int par0 = (col - XO + g.j0) & 1;
	// lanes holding the border cells: left border i = 0 is component .w of the lane before the one with i = 1;
	// right border i = N+1 (N % 4 == 0) is component .x of the lane after the one whose .w is i = N
	const bool mirror = (b >= 0 && b <= 2);
	// bottom / top border rows mirror rows 1 / NY the same way (negated for b == 2) where the slab ends in a wall
	const bool wallB = has_bot && (g.walls & 1), wallT = has_top && (g.walls & 2);
	const bool ghostL = WALL && (col + 3 == XO);
	const bool ghostR = WALL && (lane0_col + 4 * lane == XO + g.N + 1); // not a folded-back lane
	const float* pin = xin + (long long) row_lo * g.pitch + col;
	const float* p0 = x0g + (long long) row_lo * g.pitch + col;
	float* pout = xout + (long long) out_lo * g.pitch + col;
	float4 x[W], r[W];
	// Rows below the first loaded one do not exist, but the fast blocks apply every pass from the first row
	// step on (no per-row conditions): give the empty slots harmless values.  What is computed from them
	// stays within the 2S untrusted rows, like everything else that depends on the missing neighbours.
#pragma unroll
	for (int i = 0; i < W; i++) {
		x[i] = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
		r[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
	}
	uint32_t low = 0xffffffffu;
	bool bad = false, bad_before = false;
	int seg = 0, seg_left = seg_rows;
	auto close_segment = [&]() {
		const bool now = __any_sync(0xffffffffu, bad || (low < 2u * dv.lo - 1u));
		if ((now || bad_before) && lane == 0) redo[1 + atomicAdd(&redo[0], 1)] = redo_base + seg;
		bad_before = now;
		bad = false;
		low = 0xffffffffu;
		seg++;
		seg_left = seg_rows;
	};
	// Rows arrive through a small per-lane ring in shared memory filled by cp.async kDepth rows ahead of
	// their use: DRAM latency is covered without holding the rows in flight in registers (one row ahead
	// in registers left the warps waiting on the loads for a third of their time).  A lane reads back only
	// the 2 x 16 bytes it copied itself, so no synchronisation between lanes is involved.
	constexpr int kDepth = kStreamDepth;
	// (INL: two steps -- its passes are longer, and it runs where wall strips and inner strips share the SMs; BS: the
	// host's choice, two steps for ensembles for the same reason)
	constexpr int B = (S >= 4) ? (BS ? BS : (INL ? STREAM_INL_B : STREAM_B4)) : W; // row steps per unrolled block (POW2 too: all W steps unrolled measured 200 instead of 148 us per pass)
	const uint32_t ring_base = (uint32_t) __cvta_generic_to_shared(ring) + 16u * lane;
	auto fetch = [&](int slot, bool real) { // next row of x and x0 into ring slot `slot`
		if (real) {
			if (!XZERO) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring_base + 1024u * slot), "l"(pin) : "memory");
			asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring_base + 1024u * slot + 512u), "l"(p0) : "memory");
		}
		pin += g.pitch;
		p0 += g.pitch;
	};
	int issued = row_lo; // next row to fetch
#pragma unroll
	for (int d = 0; d < kDepth; d++) {
		if (issued <= row_hi) {
			fetch(d, issued >= 0 && issued <= g.NY + 1);
			issued++;
		}
		asm volatile("cp.async.commit_group;" ::: "memory");
	}
	// border row next to row 1 / NY, written when that row is final; the corners average their two
	// neighbouring border cells (SetBorder, FluidSystem.cpp:496-499), whatever b is
	auto store_border_row = [&](const float4& vg, const float4& vi, int y, bool is_wall) {
		float4 v = vg;
		if (WALL && is_wall) {
			const float h1 = __shfl_down_sync(0xffffffffu, vg.x, 1); // border-row cell above/below i = 1
			const float hN = __shfl_up_sync(0xffffffffu, vg.w, 1);   // ... i = N
			if (ghostL) v.w = 0.5f * (h1 + vi.w);
			if (ghostR) v.x = 0.5f * (hN + vi.x);
		}
		float* pb = xout + (long long) y * g.pitch + col;
		if (store_lane) *reinterpret_cast<float4*>(pb) = v;
		if (WALL && ghostL) pb[3] = v.w;
		if (WALL && ghostR) pb[0] = v.x;
	};
	int slot = 0;

	// B row steps with compile-time register slots.  Unrolling all W steps needs no register moves, but for
	// S = 4 that loop is 35 KB of code, more than the 32 KB instruction cache level that 12 warps at 12
	// different places in the loop can share -- ncu showed more warps waiting for instructions than
	// issuing.  Four steps and one rotation of the window (80 moves, +9 % instructions) fit: 1.81 -> 1.60 ms
	// per 20-sweep solve at N = 8192 (six steps: 1.64, two: 1.62).  STEADY: every pass applies and
	// every row is fetched unconditionally; the only condition left is whether the finished row is stored.
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
				if (XZERO) x[u] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
				else asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x[u].x), "=f"(x[u].y), "=f"(x[u].z), "=f"(x[u].w) : "r"(sa) : "memory");
				asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r[u].x), "=f"(r[u].y), "=f"(r[u].z), "=f"(r[u].w) : "r"(sa + 512u) : "memory");
				if (INL && !POW2) { // stream_relax_full_pre wants the right-hand side scaled
					r[u].x *= 0x1p64f;
					r[u].y *= 0x1p64f;
					r[u].z *= 0x1p64f;
					r[u].w *= 0x1p64f;
				}
			}
			if (STEADY || issued <= row_hi) { // refill the slot just read
				fetch(slot, STEADY || (issued >= 0 && issued <= g.NY + 1));
				issued++;
			}
			asm volatile("cp.async.commit_group;" ::: "memory");
			slot = (slot + 1 == kDepth) ? 0 : slot + 1;
			const int ph = STEADY ? ((P0 + u) & 1) : ((par0 + t) & 1); // parity of (column + row t)
#pragma unroll
			for (int h = 1; h <= L; h++) {
				if (!STEADY && (t - h < ylo || t - h > yhi)) continue; // no such row, or it lacks a neighbour
				constexpr int WW = 4 * W;
				const int sy = (u - h + WW) % W, sn = (u - h + 1 + WW) % W, ss = (u - h - 1 + WW) % W;
				// row y = t - h, colour (h - 1) & 1: (par0 + y + colour) & 1 = (ph + h + h - 1) & 1 = (ph + 1) & 1,
				// the same for every pass of this row step
				if (ph) { // cells .x and .z of the row have the active colour
					const float wl = __shfl_up_sync(0xffffffffu, x[sy].w, 1);
					const float keep = x[sy].x;
					if (INL && !POW2) {
						x[sy].x = stream_relax_full_pre(r[sy].x, a, wl, x[sy].y, x[ss].x, x[sn].x, dv);
						x[sy].z = stream_relax_full_pre(r[sy].z, a, x[sy].y, x[sy].w, x[ss].z, x[sn].z, dv);
					} else {
						x[sy].x = stream_relax<POW2>(r[sy].x, a, wl, x[sy].y, x[ss].x, x[sn].x, dv, low);
						x[sy].z = stream_relax<POW2>(r[sy].z, a, x[sy].y, x[sy].w, x[ss].z, x[sn].z, dv, low);
					}
					if (WALL) {
						if (ghostR) x[sy].x = keep; // the right border cell is not an interior cell
						const float v1 = __shfl_down_sync(0xffffffffu, x[sy].x, 1); // cell i = 1 was updated
						if (ghostL && mirror) x[sy].w = (b == 1) ? -v1 : v1;
					}
					if (!STEADY && mirror) {
						if (wallB && t - h == 1) {
							x[ss].x = (b == 2) ? -x[sy].x : x[sy].x;
							x[ss].z = (b == 2) ? -x[sy].z : x[sy].z;
						}
						if (wallT && t - h == g.NY) {
							x[sn].x = (b == 2) ? -x[sy].x : x[sy].x;
							x[sn].z = (b == 2) ? -x[sy].z : x[sy].z;
						}
					}
				} else {  // cells .y and .w
					const float er = __shfl_down_sync(0xffffffffu, x[sy].x, 1);
					const float keep = x[sy].w;
					if (INL && !POW2) {
						x[sy].y = stream_relax_full_pre(r[sy].y, a, x[sy].x, x[sy].z, x[ss].y, x[sn].y, dv);
						x[sy].w = stream_relax_full_pre(r[sy].w, a, x[sy].z, er, x[ss].w, x[sn].w, dv);
					} else {
						x[sy].y = stream_relax<POW2>(r[sy].y, a, x[sy].x, x[sy].z, x[ss].y, x[sn].y, dv, low);
						x[sy].w = stream_relax<POW2>(r[sy].w, a, x[sy].z, er, x[ss].w, x[sn].w, dv, low);
					}
					if (WALL) {
						if (ghostL) x[sy].w = keep; // the left border cell is not an interior cell
						const float vN = __shfl_up_sync(0xffffffffu, x[sy].w, 1);   // cell i = N was updated
						if (ghostR && mirror) x[sy].x = (b == 1) ? -vN : vN;
					}
					if (!STEADY && mirror) {
						if (wallB && t - h == 1) {
							x[ss].y = (b == 2) ? -x[sy].y : x[sy].y;
							x[ss].w = (b == 2) ? -x[sy].w : x[sy].w;
						}
						if (wallT && t - h == g.NY) {
							x[sn].y = (b == 2) ? -x[sy].y : x[sy].y;
							x[sn].w = (b == 2) ? -x[sy].w : x[sy].w;
						}
					}
				}
			}
			if (STEADY ? (t - L >= out_lo) : (t - L >= out_lo && t - L <= out_hi)) { // this row has now received all L passes
				const float4 v = x[(u - L + 4 * W) % W];
				if (!POW2) bad = bad || !(v.x == v.x) || !(v.y == v.y) || !(v.z == v.z) || !(v.w == v.w);
				if (store_lane) *reinterpret_cast<float4*>(pout) = v;
				if (WALL && ghostL) pout[3] = v.w;
				if (WALL && ghostR) pout[0] = v.x;
				pout += g.pitch;
				if (!POW2 && (--seg_left == 0 || (!STEADY && t - L == out_hi))) close_segment();
				if (!STEADY && botw && t - L == 1) store_border_row(x[(u - L - 1 + 4 * W) % W], v, 0, wallB);
				if (!STEADY && topw && t - L == g.NY) store_border_row(x[(u - L + 1 + 4 * W) % W], v, g.NY + 1, wallT);
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
		// steady: every row up to the one fetched now is real, no border row to mirror or store -- the whole
		// task except its last few rows and, in the bottom and top chunks, the rows next to the wall
		if ((!has_bot || t0 > row_lo + 2 * L) && t0 + B - 1 + kDepth <= min(row_hi, g.NY + 1)) {
			if (phase0) block(t0, std::true_type{}, std::integral_constant<int, 1>{});
			else block(t0, std::true_type{}, std::integral_constant<int, 0>{});
		} else block(t0, std::false_type{}, std::integral_constant<int, 0>{});
	}
}

// redo[0] counts, redo[1..] lists the (task, segment) rectangles the tile kernel must recompute exactly.
template <int S, bool POW2, bool XZERO = false, bool INL = false, int BS = 0>
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
	// the wall variant for every strip whose lanes reach a border column: the first and the last strip, and the
	// last but one when the last is narrower than the halo
	const bool wall = k.wall || lane0_col <= XO || lane0_col + 128 > XO + g.N + 1;
	if (wall) rb_stream_task<S, true, POW2, XZERO, INL, BS>(g, xinb + off, x0b + off, xoutb + off, a, dv, lane0_col, k.out_lo, k.out_hi, rc.col0, rc.col1, b, ring, redo, task * rc.segs, rc.seg_rows);
	else rb_stream_task<S, false, POW2, XZERO, INL, BS>(g, xinb + off, x0b + off, xoutb + off, a, dv, lane0_col, k.out_lo, k.out_hi, rc.col0, rc.col1, b, ring, redo, task * rc.segs, rc.seg_rows);
}

} // namespace fluid
