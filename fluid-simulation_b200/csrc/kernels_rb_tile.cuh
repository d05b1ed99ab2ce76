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

// The 2S colour passes of one tile residency.  WALL (block uniform) additionally
// keeps the border cells inside the window current, i.e. SetBorder after every
// sweep: a border cell mirrors one interior cell and is read by no other.
template <int S, int RX, int CB, int KB, bool WALL>
__device__ __forceinline__ void rb_tile_sweeps(float* base, const float (&x0a)[KB], const float (&x0c)[KB], uint32_t mask0,
					       uint32_t mask1, int par, float a, float c, int b, const GridDesc& g, int i_first,
					       int j_first)
{
#pragma unroll 1
	for (int hs = 0; hs < 2 * S; hs++) {
		const int colour = hs & 1;
		float* q = base + (colour ^ par); // the cell of each pair that has the active colour
		const uint32_t m = colour ? mask1 : mask0;
#pragma unroll
		for (int k = 0; k < KB; k++) {
			const int o = 2 * (k / CB) * RX + 32 * (k % CB);
			const float v = relax(colour ? x0c[k] : x0a[k], a, q[o - 1], q[o + 1], q[o - RX], q[o + RX], c);
			if ((m >> k) & 1) {
				q[o] = v;
				if (WALL) {
					const int i = i_first + 32 * (k % CB) + (colour ^ par), j = j_first + 2 * (k / CB);
					const float sx = (b == 1) ? -v : v, sy = (b == 2) ? -v : v;
					if (i == 1) q[o - 1] = sx;
					if (i == g.N) q[o + 1] = sx;
					if (j == 1) q[o - RX] = sy;
					if (j == g.NY) q[o + RX] = sy;
				}
			}
		}
		__syncthreads();
	}
}

// S  sweeps per residency (halo H = 2S);  window RY x RX floats, RX % 32 == 0 so
// that rows r and r+1 of a 32-column block sit in the same banks and the two
// half-warps (which work on opposite colours' columns) never collide;
// NT threads, each owning KB = RY*RX/(2*NT) horizontally adjacent cell pairs.
template <int S, int RY, int RX, int NT>
__global__ void __launch_bounds__(NT, 2) k_rb_tile(const __grid_constant__ CUtensorMap tm_in, GridDesc g, int b,
						    float* __restrict__ xoutb, const float* __restrict__ x0b, float a,
						    float c)
{
	// vertical halo 2S; the horizontal one is padded to a multiple of 4 floats because the
	// global start address of a TMA box must be 16-byte aligned
	constexpr int HY = 2 * S, HX = ((2 * S + 3) / 4) * 4, TY = RY - 2 * HY, TX = RX - 2 * HX;
	constexpr int CB = RX / 32;               // 32-column blocks per window row
	constexpr int NW = NT / 32;               // warps
	constexpr int NB = (RY / 2) * CB;         // warp-blocks (2 rows x 32 cols)
	constexpr int KB = NB / NW;               // blocks per warp = pairs per thread
	static_assert(RX % 32 == 0 && RY % 2 == 0 && NB % NW == 0, "bad tile shape");
	static_assert(KB % CB == 0, "a warp must own whole row pairs");
	constexpr int RPW = KB / CB;              // row pairs per warp

	extern __shared__ __align__(128) float smem[]; // (RY + 2) rows of RX: one guard row above and below
	__shared__ __align__(8) uint64_t bar;
	float* tile = smem + RX;

	const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, sub = lane >> 4, p = lane & 15;
	const int tx0 = blockIdx.x * TX, ty0 = blockIdx.y * TY; // output cells: i in tx0+1..tx0+TX, j in ty0+1..ty0+TY
	const int sim = blockIdx.z;
	const int ci0 = 1 + tx0 - HX, cj0 = 1 + ty0 - HY;       // interior index of window cell (0,0)

	if (tid == 0) {
		mbar_init(&bar, 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	if (tid == 0) {
		mbar_expect_tx(&bar, RY * RX * sizeof(float));
		tma_load_3d(tile, &tm_in, XO + ci0, cj0, sim, &bar);
	}

	// --- x0 of this thread's cells, straight from global into registers ---
	const long long off = (long long) sim * g.plane;
	const float* x0 = x0b + off;
	const int r_first = 2 * RPW * w + sub; // window row of this thread's first pair; later pairs add 2 per row pair
	const int cc_first = 2 * p;            // window column of the pair; later pairs add 32 per column block
	const int par = (tx0 + ty0 + r_first) & 1; // 0: the pair's first cell is red (i+j even)
	float x0a[KB], x0c[KB];                // x0 of the cell updated in colour passes 0 / 1
	uint32_t mask0 = 0, mask1 = 0;         // which of those cells are real interior cells inside the ring
#pragma unroll
	for (int k = 0; k < KB; k++) {
		const int r = r_first + 2 * (k / CB), cc = cc_first + 32 * (k % CB);
		const int j = cj0 + r, i = ci0 + cc;
		float2 v = make_float2(0.f, 0.f);
		if (j >= 1 && j <= g.NY && i + 1 >= 1 && i <= g.N) v = __ldg(reinterpret_cast<const float2*>(x0 + cell(g, i, j)));
		const bool rowok = (r >= 1 && r <= RY - 2 && j >= 1 && j <= g.NY);
		const bool ok_first = rowok && cc >= 1 && i >= 1 && i <= g.N;
		const bool ok_second = rowok && cc + 1 <= RX - 2 && i + 1 >= 1 && i + 1 <= g.N;
		x0a[k] = par ? v.y : v.x;
		x0c[k] = par ? v.x : v.y;
		mask0 |= (uint32_t) (par ? ok_second : ok_first) << k;
		mask1 |= (uint32_t) (par ? ok_first : ok_second) << k;
	}

	// tiles that contain a wall-adjacent cell take the (block-uniform) slow path that maintains border cells
	const bool wall = (b >= 0 && b <= 2) &&
			  (tx0 == 0 || ty0 == 0 || tx0 + TX + HX >= g.N || ty0 + TY + HY >= g.NY);

	mbar_wait(&bar, 0);

	float* base = tile + r_first * RX + cc_first;
	if (wall) rb_tile_sweeps<S, RX, CB, KB, true>(base, x0a, x0c, mask0, mask1, par, a, c, b, g, ci0 + cc_first, cj0 + r_first);
	else rb_tile_sweeps<S, RX, CB, KB, false>(base, x0a, x0c, mask0, mask1, par, a, c, b, g, 0, 0);

	// --- store the centre ---
	float* xout = xoutb + off;
	const bool full = (tx0 + TX <= g.N) && (ty0 + TY <= g.NY);
	if (!wall && full) {
		static_assert(TX % 4 == 0 && HX % 4 == 0, "float4 store path");
		constexpr int PER_ROW = TX / 4;
		for (int idx = tid; idx < TY * PER_ROW; idx += NT) {
			const int r = idx / PER_ROW, cv = idx - r * PER_ROW;
			const float* s = tile + (HY + r) * RX + HX + 4 * cv;
			float* d = xout + cell(g, tx0 + 1 + 4 * cv, ty0 + 1 + r);
			*reinterpret_cast<float4*>(d) = *reinterpret_cast<const float4*>(s);
		}
	} else {
		for (int idx = tid; idx < TY * TX; idx += NT) {
			const int r = idx / TX, cc = idx - r * TX;
			const int i = tx0 + 1 + cc, j = ty0 + 1 + r;
			if (i > g.N || j > g.NY) continue;
			const float v = tile[(HY + r) * RX + HX + cc];
			xout[cell(g, i, j)] = v;
			mirror_border(g, xout, b, i, j, v);
		}
	}
}

} // namespace fluid
