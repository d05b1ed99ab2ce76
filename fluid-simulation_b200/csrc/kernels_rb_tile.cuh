// Red-black Gauss-Seidel, fast path: TMA-staged shared-memory tiles with both
// colours -- and S whole sweeps -- per tile residency (temporal blocking).
//
// Why: the one-colour-per-launch kernel (k_rb_half) moves 11.4 B/cell per
// half-sweep through HBM (ncu, profiles/) against 6 B algorithmic, because each
// colour pass streams x, x0 and x again.  Here a CTA loads an (RY x RX) window
// of x once with one TMA box copy (halo H = 2S cells: one full red-black sweep
// needs the other colour at distance 1 and this colour at distance 2), keeps
// x0 of "its" cells in registers, runs 2S colour passes in shared memory and
// stores the central (RY-2H) x (RX-2H) cells.  Reading a neighbour tile's cells
// while that tile is being rewritten would race, so a pass reads x_in and
// writes x_out (ping-pong with one scratch field; the host arranges an even
// number of passes so the result lands in the caller's x).
//
// The values are identical to oracle_lin_solve_rb (CPU red-black statement):
// within a colour every update is independent, each cell sees exactly the same
// neighbours, and the arithmetic is the same expression with -fmad=false.
// Cells in the outer rings of the window are computed from incomplete data;
// the error front moves one cell per colour pass, i.e. 2S = H cells in total,
// and never reaches the stored centre.
#pragma once
#include <cuda.h> // CUtensorMap (types only; the encode function is fetched at run time)

#include "grid.cuh"
#include "kernels_rb_stream.cuh"

namespace fluid {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
	uint32_t done;
	do {
		asm volatile(
			"{\n"
			".reg .pred p;\n"
			"mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
			"selp.u32 %0, 1, 0, p;\n"
			"}\n"
			: "=r"(done)
			: "r"(smem_u32(bar)), "r"(parity)
			: "memory");
	} while (!done);
}

// 3-D tiled TMA load (x, y, ensemble member) global -> shared, completion on an mbarrier.
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, int cx, int cy, int cz, uint64_t* bar)
{
	asm volatile(
		"cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
		::"r"(smem_u32(dst)), "l"(map), "r"(cx), "r"(cy), "r"(cz), "r"(smem_u32(bar))
		: "memory");
}

// One colour pass over the KB cells of that colour this thread owns.
// q points at the thread's first cell of the active colour.  MASKED: only cells
// whose bit is set in m are real interior cells away from the window ring.
// WALL (implies MASKED): also keep the border cells inside the window current,
// i.e. SetBorder after every sweep -- a border cell mirrors one interior cell
// and is read by no other.
template <int RX, int CB, int KB, bool MASKED, bool WALL, bool SECOND>
__device__ __forceinline__ void rb_colour_pass(float* q, const float2 (&x0)[KB], uint32_t m, float a, float c, int b,
					       const GridDesc& g, int i_first, int j_first, const DivC& dv)
{
	// Cells this thread really updates (bit k): everything for an unmasked tile.
	const uint32_t live = MASKED ? m : 0xffffffffu;
	uint32_t redo = live; // cells still to be computed with the exact division
	if (dv.range != 0) {
		// verified short division first; a cell whose dividend lies outside the verified window (the
		// subnormal fringe of a field, overflow) stays in `redo` and is recomputed exactly below.  Its
		// inputs are the other colour's cells, x0 and -- along a wall -- border cells, none of which
		// this loop modifies (border cells are written at the very end).
		uint32_t bad = 0;
#pragma unroll
		for (int k = 0; k < KB; k++) {
			const int o = 2 * (k / CB) * RX + 32 * (k % CB);
			bool kbad = false;
			const float v = relax_fast(SECOND ? x0[k].y : x0[k].x, a, q[o - 1], q[o + 1], q[o - RX], q[o + RX], dv, kbad);
			if (!MASKED || ((m >> k) & 1)) q[o] = v;
			bad |= (uint32_t) kbad << k;
		}
		redo &= bad;
	}
	if (redo) {
#pragma unroll
		for (int k = 0; k < KB; k++) {
			if (!((redo >> k) & 1)) continue;
			const int o = 2 * (k / CB) * RX + 32 * (k % CB);
			// a verified divisor is positive and finite: the cells left over are the subnormal fringe, cheaper in double
			const float t = (SECOND ? x0[k].y : x0[k].x) + a * (((q[o - 1] + q[o + 1]) + q[o - RX]) + q[o + RX]);
			q[o] = (dv.range != 0) ? div_bracket(t, dv) : div_exact(t, c);
		}
	}
	if (WALL) {
		// SetBorder inside the window: a border cell mirrors one interior cell and is read by no other
#pragma unroll
		for (int k = 0; k < KB; k++) {
			if (!((live >> k) & 1)) continue;
			const int i = i_first + 32 * (k % CB), j = j_first + 2 * (k / CB);
			const bool L = (i == 1), R = (i == g.N), B = (j == 1 && (g.walls & 1)), T = (j == g.NY && (g.walls & 2));
			if (!(L || R || B || T)) continue;
			const int o = 2 * (k / CB) * RX + 32 * (k % CB);
			const float v = q[o];
			const float sx = (b == 1) ? -v : v, sy = (b == 2) ? -v : v;
			if (L) q[o - 1] = sx;
			if (R) q[o + 1] = sx;
			if (B) q[o - RX] = sy;
			if (T) q[o + RX] = sy;
		}
	}
}

template <int S, int RX, int CB, int KB, bool MASKED, bool WALL>
__device__ __forceinline__ void rb_tile_sweeps(float* base, const float2 (&x0)[KB], uint32_t mask0, uint32_t mask1, int par,
					       float a, float c, int b, const GridDesc& g, int i_first, int j_first,
					       const DivC& dv)
{
	// after the conditional swap in the caller x0[k].x belongs to the red cell of pair k, .y to the black one
	float* q0 = base + par;       // red cell (i+j even) of each pair
	float* q1 = base + (par ^ 1); // black cell
#pragma unroll 1
	for (int s = 0; s < S; s++) {
		rb_colour_pass<RX, CB, KB, MASKED, WALL, false>(q0, x0, mask0, a, c, b, g, i_first + par, j_first, dv);
		__syncthreads();
		rb_colour_pass<RX, CB, KB, MASKED, WALL, true>(q1, x0, mask1, a, c, b, g, i_first + (par ^ 1), j_first, dv);
		__syncthreads();
	}
}

// Persistent kernel: one CTA per SM walks over tiles t = blockIdx.x, +gridDim.x, ...
// While the CTA relaxes tile n in one shared-memory buffer, the TMA engine fills
// the other x buffer and the x0 staging buffer with tile n+1, so HBM latency is
// hidden behind the arithmetic instead of being paid at the start of every CTA
// (ncu on the non-persistent version: long-scoreboard stalls 6 of 11 cycles per
// issue, DRAM at 35 %).  x0 of the thread's own cells is copied from the staging
// buffer into registers at the start of a tile, which frees the staging buffer
// for the next TMA at once and keeps the register budget for the relaxation loop
// (an earlier version prefetched x0 with ld.global into a second register set:
// 64 pinned registers, spills inside the hot loop, two copies of the loop body).
//
// S sweeps per residency; window RY x RX floats (vertical halo 2S, horizontal
// halo padded to 4: the global start address of a TMA box must be 16-byte
// aligned).  RX % 32 == 0 so that rows r and r+1 of a 32-column block sit in the
// same banks and the two half-warps, which work on opposite colours' columns,
// never collide.  NT threads, each owning KB = RY*RX/(2*NT) adjacent cell pairs.
// Shared memory: 2 x (RY+2) x RX (x, one guard row above and below) + RY x RX (x0).
template <int S, int RY, int RX, int NT>
__global__ void __launch_bounds__(NT, 1) k_rb_tile(const __grid_constant__ CUtensorMap tm_in,
						    const __grid_constant__ CUtensorMap tm_x0, GridDesc g, int b,
						    float* __restrict__ xoutb, float a, DivC dv, int ntx, int nty, int ntiles,
						    int ftx0, int ftx1, int fty0, int fty1, const int* __restrict__ redo, StreamRect rc,
						    int stream_out, int stream_halo)
{
	constexpr int HY = 2 * S, HX = ((2 * S + 3) / 4) * 4, TY = RY - 2 * HY, TX = RX - 2 * HX;
	constexpr int CB = RX / 32;               // 32-column blocks per window row
	constexpr int NW = NT / 32;               // warps
	constexpr int NB = (RY / 2) * CB;         // warp-blocks (2 rows x 32 cols)
	constexpr int KB = NB / NW;               // blocks per warp = pairs per thread
	static_assert(RX % 32 == 0 && RY % 2 == 0 && NB % NW == 0, "bad tile shape");
	static_assert(KB % CB == 0 && KB <= 32, "a warp must own whole row pairs; masks are 32 bits");
	static_assert(TX % 4 == 0 && HX % 4 == 0, "float4 store path");
	constexpr int RPW = KB / CB;              // row pairs per warp
	constexpr int BUF = (RY + 2) * RX;        // one guard row above and below each window
	const float c = dv.c;

	extern __shared__ __align__(128) float smem[]; // [2][BUF] x windows, then [RY*RX] x0 staging
	__shared__ __align__(8) uint64_t bar[3];       // x buffer 0, x buffer 1, x0 staging
	float* x0s = smem + 2 * BUF;

	const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, sub = lane >> 4, p = lane & 15;
	const int r_first = 2 * RPW * w + sub; // window row of this thread's first pair; later pairs add 2 per row pair
	const int cc_first = 2 * p;            // window column of the pair; later pairs add 32 per column block

	// guard rows are read by the ring cells of unmasked tiles: keep them finite
	for (int k = tid; k < RX; k += NT) {
		smem[k] = 0.f;
		smem[(RY + 1) * RX + k] = 0.f;
		smem[BUF + k] = 0.f;
		smem[BUF + (RY + 1) * RX + k] = 0.f;
	}
	if (tid == 0) {
		mbar_init(&bar[0], 1);
		mbar_init(&bar[1], 1);
		mbar_init(&bar[2], 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();

	// Tile coordinates (column, row, ensemble member), advanced by gridDim.x tiles per iteration with
	// two divisions by small constants-per-launch (a subtract loop took 49 rounds per step on an
	// ensemble of 3-tile-wide grids; decoding the linear index afresh costs three divisions).
	struct TilePos { int tx, ty, sim, tx0, ty0; }; // tile grid position and the cell offsets of its output
	auto advance = [&](TilePos q, int by) {
		q.tx += by;
		if (q.tx >= ntx) {
			const int carry = q.tx / ntx;
			q.tx -= carry * ntx;
			q.ty += carry;
			if (q.ty >= nty) {
				const int c2 = q.ty / nty;
				q.ty -= c2 * nty;
				q.sim += c2;
			}
		}
		q.tx0 = q.tx * TX;
		q.ty0 = q.ty * TY;
		return q;
	};
	// Frame mode (ftx1 > 0): the tiles with ftx0 <= tx < ftx1 and fty0 <= ty < fty1 belong to k_rb_stream; this
	// launch enumerates only the frame around them -- bottom tile rows, top tile rows, left columns and
	// right columns, possibly none at all -- from a linear index (few tiles, so the divisions do not matter).
	const int frame_mid = ftx0 + (ntx - ftx1), frame_bot = fty0 * ntx, frame_top = (nty - fty1) * ntx;
	const int frame_per_sim = frame_bot + frame_top + frame_mid * (fty1 - fty0);
	// After the frame tiles come the rectangles k_rb_stream could not finish with the short division
	// (redo[0] of them, task ids in redo[1..]): `redo_rt` tiles stacked over each rectangle.  A redo tile
	// is wider than the rectangle and may overlap finished neighbours; it stores the same values there.
	const int redo_rt = (ftx1 > 0) ? 1 : 0; // one tile per listed (task, segment): segments are TY rows
	if (ftx1 > 0) ntiles += redo_rt * redo[0]; // written by the preceding kernel in stream order
	auto frame_pos = [&](int t) {
		TilePos q;
		const int nframe = frame_per_sim * (int) (g.plane ? 1 : 1);
		(void) nframe;
		q.tx = q.ty = 0;
		const int frame_total = ntiles - redo_rt * redo[0];
		if (t >= frame_total) { // a redo tile
			const int e = (t - frame_total) / redo_rt, k = (t - frame_total) - e * redo_rt;
			const int entry = redo[1 + e], k_seg = entry % rc.segs;
			const StreamTask st = stream_task(rc, entry / rc.segs);
			q.sim = st.sim;
			q.tx0 = rc.col0 + st.strip * stream_out - (XO + 1);
			// tiles past the rectangle's last row repeat the last useful one (same values, and inside the grid)
			q.ty0 = min(st.out_lo - 1 + (k + k_seg) * TY, max(st.out_lo - 1, st.out_hi - TY));
			(void) stream_halo;
			return q;
		}
		q.sim = t / frame_per_sim;
		int k = t - q.sim * frame_per_sim;
		if (k < frame_bot) { q.ty = k / ntx; q.tx = k - q.ty * ntx; }
		else if (k - frame_bot < frame_top) { k -= frame_bot; q.ty = fty1 + k / ntx; q.tx = k - (q.ty - fty1) * ntx; }
		else {
			k -= frame_bot + frame_top;
			q.ty = fty0 + k / frame_mid;
			const int m = k - (q.ty - fty0) * frame_mid;
			q.tx = (m < ftx0) ? m : ftx1 + (m - ftx0);
		}
		q.tx0 = q.tx * TX;
		q.ty0 = q.ty * TY;
		return q;
	};
	auto edge_tile = [&](int tx0, int ty0) { // window reaches a wall, or the tile is partial
		return tx0 == 0 || ty0 == 0 || tx0 + TX + HX >= g.N || ty0 + TY + HY >= g.NY;
	};
	// both boxes of a tile: same window, one from x (into x buffer `buf`), one from x0 (staging)
	auto issue_tma = [&](TilePos tp, int buf) {
		const int cx = XO + 1 + tp.tx0 - HX, cy = 1 + tp.ty0 - HY;
		mbar_expect_tx(&bar[buf], RY * RX * sizeof(float));
		tma_load_3d(smem + buf * BUF + RX, &tm_in, cx, cy, tp.sim, &bar[buf]);
		mbar_expect_tx(&bar[2], RY * RX * sizeof(float));
		tma_load_3d(x0s, &tm_x0, cx, cy, tp.sim, &bar[2]);
	};

	int t = blockIdx.x;
	if (t >= ntiles) return;
	TilePos tp = (ftx1 > 0) ? frame_pos(t) : advance(TilePos{ 0, 0, 0, 0, 0 }, t);
	if (tid == 0) issue_tma(tp, 0);

#pragma unroll 1
	for (int n = 0; t < ntiles; n++, t += gridDim.x) {
		const int buf = n & 1;
		const int tx0 = tp.tx0, ty0 = tp.ty0, sim = tp.sim;
		const int ci0 = 1 + tx0 - HX, cj0 = 1 + ty0 - HY;   // interior index of window cell (0,0)
		const bool edge = edge_tile(tx0, ty0);
		const bool wall = edge && (b >= 0 && b <= 2);
		const int par = (tx0 + ty0 + r_first + g.j0) & 1;   // 0: the pair's first cell is red (i + global j even)
		float* tile = smem + buf * BUF + RX;
		float* base = tile + r_first * RX + cc_first;

		// x0 of this thread's cells: staging buffer -> registers; .x = red cell, .y = black cell
		mbar_wait(&bar[2], n & 1);
		float2 x0[KB];
#pragma unroll
		for (int k = 0; k < KB; k++) {
			const float2 v = *reinterpret_cast<const float2*>(x0s + (r_first + 2 * (k / CB)) * RX + cc_first + 32 * (k % CB));
			x0[k] = par ? make_float2(v.y, v.x) : v;
		}
		// staging buffer and the other x buffer (stored from in iteration n-1) are free: refill them
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
		__syncthreads();
		const TilePos tpn = (ftx1 > 0) ? frame_pos(min(t + (int) gridDim.x, ntiles - 1)) : advance(tp, gridDim.x);
		if (tid == 0 && t + (int) gridDim.x < ntiles) issue_tma(tpn, buf ^ 1);

		mbar_wait(&bar[buf], (n >> 1) & 1);

		if (!edge) {
			rb_tile_sweeps<S, RX, CB, KB, false, false>(base, x0, 0, 0, par, a, c, b, g, 0, 0, dv);
		} else {
			uint32_t mask0 = 0, mask1 = 0; // real interior cells inside the ring, per colour
#pragma unroll
			for (int k = 0; k < KB; k++) {
				const int r = r_first + 2 * (k / CB), cc = cc_first + 32 * (k % CB);
				const int j = cj0 + r, i = ci0 + cc;
				const bool rowok = (r >= 1 && r <= RY - 2 && j >= 1 && j <= g.NY);
				const bool ok_first = rowok && cc >= 1 && i >= 1 && i <= g.N;
				const bool ok_second = rowok && cc + 1 <= RX - 2 && i + 1 >= 1 && i + 1 <= g.N;
				mask0 |= (uint32_t) (par ? ok_second : ok_first) << k;
				mask1 |= (uint32_t) (par ? ok_first : ok_second) << k;
			}
			if (wall) rb_tile_sweeps<S, RX, CB, KB, true, true>(base, x0, mask0, mask1, par, a, c, b, g, ci0 + cc_first, cj0 + r_first, dv);
			else rb_tile_sweeps<S, RX, CB, KB, true, false>(base, x0, mask0, mask1, par, a, c, b, g, 0, 0, dv);
		}

		// --- store the centre ---
		float* xout = xoutb + (long long) sim * g.plane;
		if (!edge) {
			// a warp per row, a lane per float4: no index division, one 64-bit address per row
			constexpr int PER_ROW = TX / 4;
			for (int r = w; r < TY; r += NW) {
				const float* sp = tile + (HY + r) * RX + HX;
				float* d = xout + cell(g, tx0 + 1, ty0 + 1 + r);
#pragma unroll
				for (int cv = lane; cv < PER_ROW; cv += 32)
					*reinterpret_cast<float4*>(d + 4 * cv) = *reinterpret_cast<const float4*>(sp + 4 * cv);
			}
		} else {
			// edge tile: the rows/columns that lie inside the grid, vectorised where whole float4 groups
			// are inside, then the border cells that mirror the wall-adjacent cells of this tile
			const int rows_in = min(TY, g.NY - ty0), cols_in = min(TX, g.N - tx0);
			const int groups = cols_in / 4;
			for (int r = w; r < rows_in; r += NW) { // a warp per row, a lane per float4 / per leftover column
				const float* sp = tile + (HY + r) * RX + HX;
				float* d = xout + cell(g, tx0 + 1, ty0 + 1 + r);
				for (int cv = lane; cv < groups; cv += 32)
					*reinterpret_cast<float4*>(d + 4 * cv) = *reinterpret_cast<const float4*>(sp + 4 * cv);
				const int cc = 4 * groups + lane; // 0..3 leftover columns at the right wall
				if (cc < cols_in) d[cc] = sp[cc];
			}
			// walls: left/right columns and bottom/top rows of the grid that fall into this tile
			const bool wl = (tx0 == 0), wr = (tx0 + cols_in == g.N), wb = (ty0 == 0), wt = (ty0 + rows_in == g.NY);
			if (wl || wr) {
				for (int r = tid; r < rows_in; r += NT) {
					const int j = ty0 + 1 + r;
					if (wl) mirror_border(g, xout, b, 1, j, tile[(HY + r) * RX + HX]);
					if (wr && g.N > 1) mirror_border(g, xout, b, g.N, j, tile[(HY + r) * RX + HX + cols_in - 1]);
				}
			}
			if (wb || wt) {
				for (int cc = tid; cc < cols_in; cc += NT) {
					const int i = tx0 + 1 + cc;
					if (i == 1 || i == g.N) continue; // corners' cells were handled by the column loop above
					if (wb) mirror_border(g, xout, b, i, 1, tile[HY * RX + HX + cc]);
					if (wt && g.NY > 1) mirror_border(g, xout, b, i, g.NY, tile[(HY + rows_in - 1) * RX + HX + cc]);
				}
			}
		}
		// order this thread's generic-proxy accesses to the buffer before the TMA write that refills it
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
		__syncthreads();
		tp = tpn;
	}
}

} // namespace fluid
