// Advect1 (FluidSystem.cpp:850-862) for large grids: a register-resident row window instead of gathers.
//
// The one-thread-per-cell kernels (kernels_basic.cuh: k_advect, k_advect_velocity) spend 170-190
// instructions and 12-16 scalar gathers per cell; ncu showed them bound by instruction issue with the
// load/store unit close behind (one LDG costs the SM 1.8 cycles whatever it fetches), at 36-40 % of HBM
// bandwidth.  Half of those instructions are addressing, bounds tests and conversions, not the
// reference's arithmetic.
//
// Here a WARP owns a strip of 128 storage columns (4 adjacent cells per lane, one aligned float4) and
// walks up through the rows of its chunk.  For every field it samples it keeps rows j-1, j, j+1 (and the
// prefetched row j+2) in registers: the lane's float4 plus the one value left and right of it, fetched
// from the neighbouring lanes by shuffle.  A backtrace that moves less than one cell -- |dt*N*v| < 1,
// the normal case of a stable-fluids step away from a jet -- lands with its 2x2 interpolation stencil
// inside that window, at an integer offset (dx, dy) in {-1,0} x {-1,0} from the cell.  When ALL cells
// of the warp's row agree on (dx, dy) (one vote), the four stencil values of every cell are registers
// picked at compile time: no addresses, no loads, only the reference's arithmetic.  The mid-point
// sample (MidpointBackward, FluidSystem.h:185-206) and the final sample vote independently.  A row
// whose cells disagree, or that reaches further, takes the gather path for that sample -- same
// expressions, so results are bit-identical whichever path runs -- and reach stays unbounded.
//
// Per row of 128 cells: one LDG.128 + two shuffles per sampled field, two STG.128.
#pragma once
#include "grid.cuh"
#include "kernels_basic.cuh"

namespace fluid {

constexpr int kAdvectWarps = 4; // warps (tasks) per CTA

// One row of one field as the window keeps it: w[0] = storage column 4q-1, w[1..4] = the lane's own
// float4 (columns 4q .. 4q+3), w[5] = column 4q+4.
struct RowWin {
	float w[6];
};

// Rows are FETCHED one row step ahead with cp.async into a small per-warp ring in shared memory (two slots per
// field) and read back into the window when their turn comes.  History: (1) plain loads followed at once by
// the neighbour shuffles made every warp wait out the memory latency at the top of each row step (ncu: 42 % of
// the stall samples on the first shuffle); (2) loads kept in registers for one row step fixed the velocity
// kernel but not the scalar one -- with three fields the loads in flight share scoreboard slots with the ones
// being consumed, and the first use of an OLD row waited for a load issued a moment ago (31 % of the samples
// on one select).  cp.async has its own completion groups (wait_group), so consuming row j+1 never waits for
// row j+2; and the neighbours' values are simply the adjacent floats of the ring slot: no shuffles, no
// predicated scalar loads into registers, no selects.
// Slot layout (floats): [3] = storage column 4q0-1, [4 .. 131] = the strip's 128 columns, [132] = column 4q0+128.
constexpr int kAdvectSlot = 136;

__device__ __forceinline__ void advect_fetch_row(uint32_t slot, const float* __restrict__ pf, int y, int pitch, int lane, int off_l, int off_r)
{
	const float* p = pf + y * pitch;
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(slot + 16u + 16u * lane), "l"(p) : "memory");
	if (lane == 0) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(slot + 12u), "l"(p + off_l) : "memory");
	if (lane == 31) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(slot + 4u * 132u), "l"(p + off_r) : "memory");
}

__device__ __forceinline__ RowWin advect_finish_row(uint32_t slot, int lane)
{
	RowWin r;
	const uint32_t a = slot + 16u + 16u * lane;
	asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.w[1]), "=f"(r.w[2]), "=f"(r.w[3]), "=f"(r.w[4]) : "r"(a) : "memory");
	asm volatile("ld.shared.f32 %0, [%1];" : "=f"(r.w[0]) : "r"(a - 4u) : "memory");
	asm volatile("ld.shared.f32 %0, [%1];" : "=f"(r.w[5]) : "r"(a + 16u) : "memory");
	return r;
}

// BilinearCoeffs::calc (FluidCode.h:91-112) for one sample: clamped position -> cell and weights.  The cell is
// kept as floats (fi0 = trunc(px), exactly what (float)(int)px is for 0.5 <= px < 2^24; fjl = the LOCAL row):
// the window path never needs them as integers, only their distance from the thread's own cell.
struct BilinW {
	float fi0, fjl;
	float s0, s1, t0, t1;
};

template <bool SLAB>
__device__ __forceinline__ BilinW advect_calc(const GridDesc& g, float px, float py)
{
	BilinW r;
	px = fminf(fmaxf(px, 0.5f), g.N + 0.5f);
	py = fminf(fmaxf(py, 0.5f), g.NYG + 0.5f);
	r.fi0 = truncf(px);
	const float fj0 = truncf(py);
	r.s1 = px - r.fi0;
	r.s0 = 1 - r.s1;
	r.t1 = py - fj0;
	r.t0 = 1 - r.t1;
	r.fjl = SLAB ? fj0 - (float) g.j0 : fj0; // small integers: exact.  (Rows off the slab are the gather path's business:
	return r;                                // the window path only ever takes rows j-1 .. j+1.)
}

// A sample whose rows are not stored on this slab (backtrace longer than the halo) sets *oob and is clamped into the
// slab so that the access stays legal; the host reports it after the step.
template <bool SLAB>
__device__ __forceinline__ void advect_slab_rows(const GridDesc& g, BilinW& c, int* oob)
{
	if (SLAB && (c.fjl < 0.0f || c.fjl > (float) g.NY)) {
		if (oob) *oob = 1;
		c.fjl = fminf(fmaxf(c.fjl, 0.0f), (float) g.NY);
	}
}

// BilinearCoeffs::Eval (FluidCode.h:88) from memory: the general path.
__device__ __forceinline__ float advect_gather(const BilinW& c, const float* __restrict__ f, int pitch)
{
	const float* r0 = f + ((int) c.fjl * pitch + XO + (int) c.fi0);
	const float* r1 = r0 + pitch;
	const float d00 = __ldg(r0), d10 = __ldg(r0 + 1);
	const float d01 = __ldg(r1), d11 = __ldg(r1 + 1);
	return c.s0 * (c.t0 * d00 + c.t1 * d01) + c.s1 * (c.t0 * d10 + c.t1 * d11);
}

// ... and from the window: cell k of the lane, stencil offset (DX, DY) in {-1,0}^2, all compile time.
template <int DX, int DY>
__device__ __forceinline__ float advect_window(const BilinW& c, const RowWin (&win)[3], int k)
{
	const float d00 = win[DY + 1].w[k + DX + 1], d10 = win[DY + 1].w[k + DX + 2];
	const float d01 = win[DY + 2].w[k + DX + 1], d11 = win[DY + 2].w[k + DX + 2];
	return c.s0 * (c.t0 * d00 + c.t1 * d01) + c.s1 * (c.t0 * d10 + c.t1 * d11);
}

// The warp's vote: code (dx+1) | (dy+1) << 1 when every valid cell of every lane has the same stencil
// offset inside the window, else -1.  The offsets are differences of small integer-valued floats, so they
// are exactly 0.0f or -1.0f when in range, and "all equal" is an OR of XORs of their bit patterns.
template <bool PARTIAL>
__device__ __forceinline__ int advect_vote(const BilinW (&c)[4], float fi, float fj, uint32_t valid)
{
	uint32_t dx[4], dy[4];
#pragma unroll
	for (int k = 0; k < 4; k++) {
		dx[k] = __float_as_uint(c[k].fi0 - (fi + k));
		dy[k] = __float_as_uint(c[k].fjl - fj);
	}
	uint32_t diff = 0;
#pragma unroll
	for (int k = 1; k < 4; k++)
		if (!PARTIAL || ((valid >> k) & 1)) diff |= (dx[k] ^ dx[0]) | (dy[k] ^ dy[0]);
	const uint32_t m1 = 0xbf800000u; // -1.0f
	const bool inx = (dx[0] == 0u) || (dx[0] == m1), iny = (dy[0] == 0u) || (dy[0] == m1);
	const int code = ((dx[0] == 0u) ? 1 : 0) | ((dy[0] == 0u) ? 2 : 0);
	const bool mine = (diff == 0u) && inx && iny;
	const int c0 = __shfl_sync(0xffffffffu, code, 0); // lane 0 always holds the strip's first cell, which exists
	const bool agree = PARTIAL ? ((valid == 0u) || (mine && code == c0)) : (mine && code == c0);
	return __all_sync(0xffffffffu, agree) ? c0 : -1;
}

template <int NF, bool PARTIAL, bool SLAB>
__device__ __forceinline__ void advect_sample(int code, const GridDesc& g, BilinW (&c)[4], const RowWin (&w0)[3], const RowWin (&w1)[3],
					       const float* __restrict__ f0, const float* __restrict__ f1, int pitch, uint32_t valid,
					       int* oob, float (&o0)[4], float (&o1)[4])
{
	switch (code) {
#define ADVECT_CASE(CODE, DX, DY)                                                                                 \
	case CODE:                                                                                                \
		_Pragma("unroll") for (int k = 0; k < 4; k++) {                                                   \
			o0[k] = advect_window<DX, DY>(c[k], w0, k);                                               \
			if (NF > 1) o1[k] = advect_window<DX, DY>(c[k], w1, k);                                   \
		}                                                                                                 \
		break;
		ADVECT_CASE(0, -1, -1)
		ADVECT_CASE(1, 0, -1)
		ADVECT_CASE(2, -1, 0)
		ADVECT_CASE(3, 0, 0)
#undef ADVECT_CASE
	default:
#pragma unroll
		for (int k = 0; k < 4; k++) {
			if (PARTIAL && !((valid >> k) & 1)) { // cells past the row's end: nothing to fetch, nothing is stored
				o0[k] = 0.0f;
				if (NF > 1) o1[k] = 0.0f;
				continue;
			}
			advect_slab_rows<SLAB>(g, c[k], oob);
			o0[k] = advect_gather(c[k], f0, pitch);
			if (NF > 1) o1[k] = advect_gather(c[k], f1, pitch);
		}
	}
}

struct AdvectTasks {
	int nstrips, nchunks, chunk_rows;
};

// VEL: VelStep1's self-advection (FluidSystem.cpp:372-373), Advect1(1,u,u0,u0,v0) and Advect1(2,v,v0,u0,v0) in one
// pass (outa = u, outb = v; b = -2 leaves the border alone, as Advect1B does, anything else means SetBorder 1 / 2).
// Otherwise Advect1(b, outa, d0, u0, v0) for one scalar field.
// MODE 0 = MidpointBackward, 1 = EulerBack (FluidSystem.h:185-219).
// PARTIAL: the row is not a whole number of 128-cell strips (N % 128 != 0): cells past its end are masked.
template <bool SLAB, int MODE, bool VEL, bool PARTIAL>
__global__ void __launch_bounds__(32 * kAdvectWarps) k_advect_rows(GridDesc g, AdvectTasks tk, int b, float* __restrict__ outab,
									   float* __restrict__ outbb, const float* __restrict__ d0b,
									   const float* __restrict__ u0b, const float* __restrict__ v0b,
									   float dtN, int* oob_flag, int own_lo, int own_hi)
{
	const int task = blockIdx.x * kAdvectWarps + (threadIdx.x >> 5);
	if (task >= tk.nstrips * tk.nchunks) return;
	const int lane = threadIdx.x & 31;
	const int chunk = task / tk.nstrips, strip = task - chunk * tk.nstrips;
	const long long off = (long long) blockIdx.y * g.plane;
	const float* u0 = u0b + off;
	const float* v0 = v0b + off;
	const float* d0 = VEL ? nullptr : d0b + off;
	float* outa = outab + off;
	float* outb = VEL ? outbb + off : nullptr;

	const int pitch = g.pitch;
	const int col = 32 + 128 * strip + 4 * lane;        // storage column of the lane's first cell
	const int colc = PARTIAL ? min(col, pitch - 4) : col; // lanes past the row's end read its last group (and store nothing)
	const int i_first = col - XO;
	uint32_t valid = 0xfu;
	if (PARTIAL) {
		valid = 0;
#pragma unroll
		for (int k = 0; k < 4; k++)
			if (i_first + k <= g.N) valid |= 1u << k;
	}
	const int off_l = -1, off_r = min(4, pitch - 1 - colc);
	const float* pu = u0 + colc;
	const float* pv = v0 + colc;
	const float* pd = VEL ? nullptr : d0 + colc;
	const float fi = (float) i_first;

	const int j_lo = 1 + chunk * tk.chunk_rows, j_hi = min(g.NY, j_lo + tk.chunk_rows - 1);
	constexpr int NF = VEL ? 2 : 3; // fields sampled: u, v (and d0)
	__shared__ __align__(16) float ring_all[kAdvectWarps][2][NF][kAdvectSlot];
	const uint32_t ring = (uint32_t) __cvta_generic_to_shared(&ring_all[threadIdx.x >> 5][0][0][0]);
	auto slot_of = [&](int parity, int f) { return ring + 4u * (uint32_t) ((parity * NF + f) * kAdvectSlot); };
	auto fetch_rows = [&](int y, int parity) { // one completion group per row
		advect_fetch_row(slot_of(parity, 0), pu, y, pitch, lane, off_l, off_r);
		advect_fetch_row(slot_of(parity, 1), pv, y, pitch, lane, off_l, off_r);
		if (!VEL) advect_fetch_row(slot_of(parity, 2), pd, y, pitch, lane, off_l, off_r);
		asm volatile("cp.async.commit_group;" ::: "memory");
	};
	auto arrived = [&]() { // every copy this warp has issued so far is in shared memory and visible to all its lanes
		asm volatile("cp.async.wait_group 0;" ::: "memory");
		__syncwarp();
	};
	RowWin wu[3], wv[3], wd[3];
	fetch_rows(j_lo - 1, 0);
	fetch_rows(j_lo, 1);
	arrived();
	wu[1] = advect_finish_row(slot_of(0, 0), lane);
	wv[1] = advect_finish_row(slot_of(0, 1), lane);
	wu[2] = advect_finish_row(slot_of(1, 0), lane);
	wv[2] = advect_finish_row(slot_of(1, 1), lane);
	if (!VEL) {
		wd[1] = advect_finish_row(slot_of(0, 2), lane);
		wd[2] = advect_finish_row(slot_of(1, 2), lane);
	}
	__syncwarp(); // all lanes have read slot 0 before it is refilled
	fetch_rows(j_lo + 1, 0);
	const bool edge_lane = (i_first == 1) || (i_first + 3 >= g.N && i_first <= g.N);

#pragma unroll 1
	for (int j = j_lo; j <= j_hi; j++) {
		// slide: rows j-1, j, j+1 <- j, j+1 and the row fetched one step ago (ring slot (j - j_lo) & 1); then fetch
		// row j+2 into the slot that held row j
		const int par = (j - j_lo) & 1;
		arrived();
		wu[0] = wu[1]; wu[1] = wu[2]; wu[2] = advect_finish_row(slot_of(par, 0), lane);
		wv[0] = wv[1]; wv[1] = wv[2]; wv[2] = advect_finish_row(slot_of(par, 1), lane);
		if (!VEL) { wd[0] = wd[1]; wd[1] = wd[2]; wd[2] = advect_finish_row(slot_of(par, 2), lane); }
		__syncwarp();
		fetch_rows(min(j + 2, g.NY + 1), par ^ 1);

		int* oob = (SLAB && j >= own_lo && j <= own_hi) ? oob_flag : nullptr; // halo rows may legitimately trace off the slab
		const float fjl = (float) j;                            // local row, for the vote
		const float fj = (float) (SLAB ? j + g.j0 : j);         // global row, for the positions
		float px[4], py[4];
		if (MODE == 1) {
#pragma unroll
			for (int k = 0; k < 4; k++) {
				px[k] = (fi + k) - dtN * wu[1].w[1 + k];
				py[k] = fj - dtN * wv[1].w[1 + k];
			}
		} else {
			BilinW m[4];
#pragma unroll
			for (int k = 0; k < 4; k++) {
				const float dirx = -dtN * wu[1].w[1 + k];
				const float diry = -dtN * wv[1].w[1 + k];
				const float mx = (fi + k) + 0.5f * dirx;
				const float my = fj + 0.5f * diry;
				m[k] = advect_calc<SLAB>(g, mx, my);
			}
			float vx[4], vy[4];
			advect_sample<2, PARTIAL, SLAB>(advect_vote<PARTIAL>(m, fi, fjl, valid), g, m, wu, wv, u0, v0, pitch, valid, oob, vx, vy);
#pragma unroll
			for (int k = 0; k < 4; k++) {
				px[k] = (fi + k) - dtN * vx[k];
				py[k] = fj - dtN * vy[k];
			}
		}
		BilinW c[4];
#pragma unroll
		for (int k = 0; k < 4; k++) c[k] = advect_calc<SLAB>(g, px[k], py[k]);
		const int code = advect_vote<PARTIAL>(c, fi, fjl, valid);
		float ra[4], rb[4];
		if (VEL) advect_sample<2, PARTIAL, SLAB>(code, g, c, wu, wv, u0, v0, pitch, valid, oob, ra, rb);
		else advect_sample<1, PARTIAL, SLAB>(code, g, c, wd, wd, d0, d0, pitch, valid, oob, ra, rb);

		float* pa = outa + (j * pitch + col);
		if (!PARTIAL || valid == 0xfu) {
			*reinterpret_cast<float4*>(pa) = make_float4(ra[0], ra[1], ra[2], ra[3]);
			if (VEL) *reinterpret_cast<float4*>(outb + (j * pitch + col)) = make_float4(rb[0], rb[1], rb[2], rb[3]);
		} else {
#pragma unroll
			for (int k = 0; k < 4; k++) {
				if (!((valid >> k) & 1)) continue;
				pa[k] = ra[k];
				if (VEL) outb[j * pitch + col + k] = rb[k];
			}
		}
		// SetBorder (FluidSystem.cpp:861): only the groups along a wall have border cells to write
		if (edge_lane || ((j == 1) && (g.walls & 1)) || ((j == g.NY) && (g.walls & 2))) {
#pragma unroll
			for (int k = 0; k < 4; k++) {
				if (PARTIAL && !((valid >> k) & 1)) continue;
				if (VEL) { // b = -2: VelStep1B's pair (no SetBorder); anything else: VelStep1's (SetBorder 1 and 2)
					mirror_border(g, outa, b == -2 ? -2 : 1, i_first + k, j, ra[k]);
					mirror_border(g, outb, b == -2 ? -2 : 2, i_first + k, j, rb[k]);
				} else {
					mirror_border(g, outa, b, i_first + k, j, ra[k]);
				}
			}
		}
	}
}

} // namespace fluid
