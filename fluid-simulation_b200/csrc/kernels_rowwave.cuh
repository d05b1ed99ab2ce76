// LinearSolve1 (FluidSystem.cpp:534-548) in the REFERENCE's order -- lexicographic Gauss-Seidel, bit-identical to the
// CPU -- as a pipeline of warps instead of a launch per hyperplane of tiles.
//
// Cell (i, j) of sweep k needs (i-1, j) and (i, j-1) of sweep k and (i+1, j), (i, j+1) of sweep k-1, and the new west
// neighbour is the innermost operand of the reference's sum: a row is a strictly sequential recurrence along i, one
// dependent chain of ~8 operations per cell.  The tile kernels (kernels_relax.cuh) walk anti-diagonals of 16x16 or
// 32x32 tiles with a barrier per diagonal: 2T-1 barrier-separated steps per tile, half-empty on average, one launch
// (or one flag handshake) per tile -- 47 ms per step at N = 8192.
//
// Here a WARP owns a band of 30 rows for one sweep and streams along i.  Lane l holds row j0 - 1 + l: lanes 1..30
// compute, one column behind the lane below (the value a lane needs from the row below was computed one step ago and
// arrives by shuffle), and every step of the warp is one link of 30 chains.  (Two columns apart the shuffle would be off
// the critical path, but a band would reach its top row twice as late, and how soon the next band can start is what a
// solve's duration is made of.)
// Lane 0 and lane 31 are FEEDERS: they stream the row below the band (new values of this sweep) and the row above it
// (old values) exactly as the computing lanes stream theirs, so that no lane needs a special path -- lane 0 hands its
// row's values up the shuffle as if it had computed them, lane 31's ring is simply the north row of lane 30.
// Old values (x of sweep k-1 and the right-hand side) stream through per-lane rings in shared memory, filled by
// cp.async 16 bytes at a time a few dozen columns ahead; the old north neighbour of a lane is the next lane's ring.
// New values leave as aligned float4 stores.  Bands of one sweep and sweeps of one band are different warps, possibly
// on different SMs, which meet through global memory: each task publishes how many steps it has completed and waits for
//     the same band of sweep k-1 to be P columns ahead            -- its rows' old values,
//     the band above of sweep k-1 to have passed its bottom row   -- the old north row (lane 31),
//     the band below of sweep k to have finished its top row      -- the new south row (lane 0),
// which also covers every write-after-read hazard (see the notes at the waits).  Tasks are dealt as tickets in the order
// (band + sweep, sweep): every wait is on a ticket with a smaller number, so whoever a warp waits for is running or
// finished and the kernel cannot deadlock however few warps are resident; and the warps that are running at any time
// form a diagonal front across bands and sweeps whose working set (a few dozen MB) stays in L2: a solve moves x and x0
// through HBM about once, not once per sweep.
//
// The computing warp never touches the counters in global memory: a fence that makes its stores visible to the GPU
// costs ~3000 cycles with the row stores in flight (ncu: 1.6 us per MEMBAR.SC.GPU) and a poll of a counter another L2
// round trip.  Each CTA has a second warp whose lane 0 is the HELPER: it polls the three producers' counters into shared
// memory and publishes the worker's progress, fence included; the worker only reads and writes shared memory.
//
// SetBorder after every sweep (FluidSystem.cpp:546): a task writes the border cells next to the interior cells it has
// just produced (mirror_border).  During sweep k a border cell is read once, by the interior cell next to it, before
// that cell is updated: the value it holds then is the mirror from sweep k-1, as in the reference.
//
// x is read with cp.async.cg (L2 only): it was written moments ago by warps on other SMs.
#pragma once
#include "grid.cuh"

namespace fluid {

constexpr int kRwRows = 30;     // rows a band computes (lanes 1..30)
constexpr int kRwRing = 64;     // columns per ring (power of two): a ring is 256 bytes
#ifndef RW_AHEAD
#define RW_AHEAD 16
#endif
#ifndef RW_PENDING
#define RW_PENDING 2
#endif
constexpr int kRwAhead = RW_AHEAD;     // a lane fetches the group that starts kRwAhead .. +3 columns ahead of its block (first read kRwAhead - 4 steps later)
constexpr int kRwPending = RW_PENDING; // cp.async groups (one per block of 4 steps) that may still be in flight: 4 * (kRwPending + 1) <= kRwAhead - 4
constexpr int kRwSkew = 1;      // columns between neighbouring lanes
constexpr int kRwPublish = 8;   // steps between progress updates (a multiple of 4)
constexpr int kRwLead = kRwAhead + 8;     // steps before lane 0 reaches the first interior column (the first groups are fetched on the way; a multiple of 4)

// steps a task runs (a multiple of 4): lane 31 reaches the last interior column at step kRwLead + N - 1 + 31 * kRwSkew, and a
// lane stores its last float4 group up to one block after it has computed the last column
__host__ __device__ inline int rowwave_steps(int N) { return (kRwLead + N + 31 * kRwSkew + 3 + 4) & ~3; }
__host__ __device__ inline int rowwave_bands(int NY) { return (NY + kRwRows - 1) / kRwRows; }

__device__ __forceinline__ float rw_lds(uint32_t addr)
{
	float v;
	asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
	return v;
}

__device__ __forceinline__ void rw_fetch(uint32_t saddr, const float* gptr)
{
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(saddr), "l"(gptr) : "memory");
}

// MODE 0: IEEE division (div_exact); 1: the scaled short sequence, proven for this divisor (DivC::full), with the exact
// division for dividends that overflow it; 2: c is a power of two, the division is a multiplication by 1/c.
template <int MODE>
__device__ __forceinline__ float rw_quotient(float rhs, float p, const DivC& dv)
{
	if (MODE == 2) return (rhs + p) * dv.y;
	if (MODE == 1) {
		const float ts = __fmaf_rn(p, 0x1p64f, rhs * 0x1p64f); // = 2^64 * RN(rhs + p), see stream_relax_full_pre
		float q = div_scaled(ts, dv);
		if (!(fabsf(ts) < __int_as_float(0x7f800000))) q = div_wide(rhs + p, dv.c); // >= 2^64, infinite or NaN
		return q;
	}
	return div_exact(rhs + p, dv.c);
}

// What a lane carries along its row.  Columns are STORAGE columns c = XO + i (aligned float4 groups are c % 4 == 0; the
// first interior cell, i = 1, starts one).  A lane's column at step n is c0(n) - lane with c0 a multiple of 4 at
// every fourth step.
struct RwLane {
	uint32_t ring_x, ring_n, ring_0; // shared-memory byte addresses: own row, the row above (next lane's ring), right-hand side
	float* xrow;                     // the lane's row of x, indexed by storage column
	const float* x0row;
	float h1, h2, h3;                // "results" at columns c-1, c-2, c-3
	int c;                           // storage column at the first step of the block
	int cls;                         // lane & 3: the lane's column is = -cls (mod 4) at the first step of a block
	bool real, computes;
	float* mirror_row;               // the border row that mirrors this lane's row (row 0 under row 1, NY+1 over NY), or null
	uint32_t mirror_flip;            // ... with this sign bit flipped (b == 2)
};

// Four steps of one lane: columns c .. c+3.  The twelve operands come out of the rings first, the four chains follow back
// to back, the lane's one fetch and one float4 store per block are in one place each.
// A step whose column is not an interior column of a computed row -- before the first column, past the last, and every
// step of the two feeder lanes -- "produces" the value the column already holds.  That one rule covers the ends of a row
// and the feeders: the first interior column finds the left border cell as its west neighbour, lane 0 hands the row
// below the band up the shuffle as if it had computed it.  Stores are another matter: only interior columns are written.
template <int MODE>
__device__ __forceinline__ void rw_block(const GridDesc& g, int b, float* __restrict__ x, RwLane& L, float a, const DivC& dv, int j,
					 int c_first, int c_last, int last_group)
{
	constexpr uint32_t M = 4u * kRwRing - 1u;
	const int c = L.c;
	const uint32_t off = 4u * (uint32_t) (c & (kRwRing - 1));
	float e[4], nn[4], rhs[4];
	const float here = rw_lds(L.ring_x + off); // what column c holds
#pragma unroll
	for (int u = 0; u < 4; u++) {
		e[u] = rw_lds(L.ring_x + ((off + 4u * (u + 1)) & M));
		nn[u] = rw_lds(L.ring_n + ((off + 4u * u) & M));
		rhs[u] = rw_lds(L.ring_0 + ((off + 4u * u) & M));
	}
	{ // the group this lane is due to fetch during these four steps: the one that starts kRwAhead + cls columns ahead
		const int fc = c + kRwAhead + L.cls;
		if (L.real && fc >= (XO & ~3) && fc <= last_group) {
			const uint32_t slot = (off + 4u * (uint32_t) (kRwAhead + L.cls)) & M;
			rw_fetch(L.ring_x + slot, L.xrow + fc);
			rw_fetch(L.ring_0 + slot, L.x0row + fc); // (the feeders fetch their rows of x0 as well: harmless, and no branch)
		}
	}
	const int lo = c_first - c, hi = c_last - c; // steps lo .. hi of this block are interior columns
	const bool comp = L.computes;
	// the lane below is one column ahead: its h1 and its first three results of this block are this lane's four south values
	// (each arrives one shuffle after it was produced: the chains of neighbouring lanes run a step apart)
	const float s0 = __shfl_up_sync(0xffffffffu, L.h1, 1);
	float q0 = rw_quotient<MODE>(rhs[0], a * (((L.h1 + e[0]) + s0) + nn[0]), dv);
	if (!(comp && lo <= 0 && hi >= 0)) q0 = here;
	const float s1 = __shfl_up_sync(0xffffffffu, q0, 1);
	float q1 = rw_quotient<MODE>(rhs[1], a * (((q0 + e[1]) + s1) + nn[1]), dv);
	if (!(comp && lo <= 1 && hi >= 1)) q1 = e[0];
	const float s2 = __shfl_up_sync(0xffffffffu, q1, 1);
	float q2 = rw_quotient<MODE>(rhs[2], a * (((q1 + e[2]) + s2) + nn[2]), dv);
	if (!(comp && lo <= 2 && hi >= 2)) q2 = e[1];
	const float s3 = __shfl_up_sync(0xffffffffu, q2, 1);
	float q3 = rw_quotient<MODE>(rhs[3], a * (((q2 + e[3]) + s3) + nn[3]), dv);
	if (!(comp && lo <= 3 && hi >= 3)) q3 = e[2];
	if (comp) {
		// The float4 group this block completes: columns g0 .. g0+3, g0 the multiple of 4 at or below c.  The first interior column starts a
		// group, so a group is never cut by the left end of the row; one cut by the right end goes out column by column.
		const int g0 = c - ((4 - L.cls) & 3); // c = -cls (mod 4)
		float v0 = q0, v1 = q1, v2 = q2, v3 = q3;
		if (L.cls == 3) { v0 = L.h1; v1 = q0; v2 = q1; v3 = q2; }
		if (L.cls == 2) { v0 = L.h2; v1 = L.h1; v2 = q0; v3 = q1; }
		if (L.cls == 1) { v0 = L.h3; v1 = L.h2; v2 = L.h1; v3 = q0; }
		if (g0 >= c_first && g0 + 3 <= c_last) {
			*reinterpret_cast<float4*>(L.xrow + g0) = make_float4(v0, v1, v2, v3);
		} else if (g0 >= c_first && g0 <= c_last) {
			L.xrow[g0] = v0;
			if (g0 + 1 <= c_last) L.xrow[g0 + 1] = v1;
			if (g0 + 2 <= c_last) L.xrow[g0 + 2] = v2;
		}
		// SetBorder: the first and last column (sides, corners), every column of the bottom / top row
		if (L.mirror_row || g.NY == 1 || (lo >= 0 && lo <= 3) || (hi >= 0 && hi <= 3)) {
			const float q[4] = { q0, q1, q2, q3 };
#pragma unroll
			for (int u = 0; u < 4; u++) {
				if (u < lo || u > hi) continue;
				if (u == lo || u == hi || g.NY == 1) mirror_border(g, x, b, c + u - XO, j, q[u]);
				else if (L.mirror_row) L.mirror_row[c + u] = __uint_as_float(__float_as_uint(q[u]) ^ L.mirror_flip);
			}
		}
	}
	L.h3 = q1;
	L.h2 = q2;
	L.h1 = q3;
	L.c = c + 4;
}

// The same four steps where no lane is near an end of its row (the caller knows from the step count): every computing lane's
// four columns and the float4 group it completes are interior, so nothing but the feeder lane needs a select and the
// instruction stream is the chains -- about a third fewer instructions than rw_block, whose predicates and branches a lone
// warp cannot hide (a task is latency-bound: what it issues per step is what a step lasts).
template <int MODE>
__device__ __forceinline__ void rw_block_inner(RwLane& L, float a, const DivC& dv, int last_group, bool feeder)
{
	constexpr uint32_t M = 4u * kRwRing - 1u;
	const int c = L.c;
	const uint32_t off = 4u * (uint32_t) (c & (kRwRing - 1));
	float e[4], nn[4], rhs[4];
	const float here = rw_lds(L.ring_x + off);
#pragma unroll
	for (int u = 0; u < 4; u++) {
		e[u] = rw_lds(L.ring_x + ((off + 4u * (u + 1)) & M));
		nn[u] = rw_lds(L.ring_n + ((off + 4u * u) & M));
		rhs[u] = rw_lds(L.ring_0 + ((off + 4u * u) & M));
	}
	{
		const int fc = c + kRwAhead + L.cls;
		if (L.real && fc <= last_group) {
			const uint32_t slot = (off + 4u * (uint32_t) (kRwAhead + L.cls)) & M;
			rw_fetch(L.ring_x + slot, L.xrow + fc);
			rw_fetch(L.ring_0 + slot, L.x0row + fc);
		}
	}
	const float s0 = __shfl_up_sync(0xffffffffu, L.h1, 1);
	float q0 = rw_quotient<MODE>(rhs[0], a * (((L.h1 + e[0]) + s0) + nn[0]), dv);
	if (feeder) q0 = here;
	const float s1 = __shfl_up_sync(0xffffffffu, q0, 1);
	float q1 = rw_quotient<MODE>(rhs[1], a * (((q0 + e[1]) + s1) + nn[1]), dv);
	if (feeder) q1 = e[0];
	const float s2 = __shfl_up_sync(0xffffffffu, q1, 1);
	float q2 = rw_quotient<MODE>(rhs[2], a * (((q1 + e[2]) + s2) + nn[2]), dv);
	if (feeder) q2 = e[1];
	const float s3 = __shfl_up_sync(0xffffffffu, q2, 1);
	float q3 = rw_quotient<MODE>(rhs[3], a * (((q2 + e[3]) + s3) + nn[3]), dv);
	if (feeder) q3 = e[2];
	if (L.computes) {
		float* px = L.xrow + c;
		if (L.cls == 0) *reinterpret_cast<float4*>(px) = make_float4(q0, q1, q2, q3);
		else if (L.cls == 1) *reinterpret_cast<float4*>(px - 3) = make_float4(L.h3, L.h2, L.h1, q0);
		else if (L.cls == 2) *reinterpret_cast<float4*>(px - 2) = make_float4(L.h2, L.h1, q0, q1);
		else *reinterpret_cast<float4*>(px - 1) = make_float4(L.h1, q0, q1, q2);
		if (L.mirror_row) { // SetBorder for the bottom / top row of the grid
			float* pm = L.mirror_row + c;
			pm[0] = __uint_as_float(__float_as_uint(q0) ^ L.mirror_flip);
			pm[1] = __uint_as_float(__float_as_uint(q1) ^ L.mirror_flip);
			pm[2] = __uint_as_float(__float_as_uint(q2) ^ L.mirror_flip);
			pm[3] = __uint_as_float(__float_as_uint(q3) ^ L.mirror_flip);
		}
	}
	L.h3 = q1;
	L.h2 = q2;
	L.h1 = q3;
	L.c = c + 4;
}

// ... and the blocks before those, while the lanes arrive at the first column one after the other -- what a hand-off to
// the band above waits for, so they get a loop of their own (in one loop with the interior block the extra live values
// slowed THAT block by a tenth).  Nothing is near the right end; a lane may still be left of its row or reach the first
// column inside the block.  Left of the row a lane computes on whatever its ring holds (nobody uses it: the lane above is
// left of its row at those columns too), except that the column of the left border cell "produces" that cell, so that the
// first interior column finds it as its west neighbour; a group is stored once it is interior (the first interior column
// starts a group: all or nothing), and the first column writes its border cell.
template <int MODE>
__device__ __forceinline__ void rw_block_ramp(const GridDesc& g, int b, float* __restrict__ x, RwLane& L, float a, const DivC& dv, int j,
					      int c_first, int last_group, bool feeder)
{
	constexpr uint32_t M = 4u * kRwRing - 1u;
	const int c = L.c;
	const uint32_t off = 4u * (uint32_t) (c & (kRwRing - 1));
	float e[4], nn[4], rhs[4];
	const float here = rw_lds(L.ring_x + off);
#pragma unroll
	for (int u = 0; u < 4; u++) {
		e[u] = rw_lds(L.ring_x + ((off + 4u * (u + 1)) & M));
		nn[u] = rw_lds(L.ring_n + ((off + 4u * u) & M));
		rhs[u] = rw_lds(L.ring_0 + ((off + 4u * u) & M));
	}
	{
		const int fc = c + kRwAhead + L.cls;
		if (L.real && fc >= (XO & ~3) && fc <= last_group) {
			const uint32_t slot = (off + 4u * (uint32_t) (kRwAhead + L.cls)) & M;
			rw_fetch(L.ring_x + slot, L.xrow + fc);
			rw_fetch(L.ring_0 + slot, L.x0row + fc);
		}
	}
	const int ub = c_first - 1 - c; // the step of this block whose column is the left border cell, if 0 .. 3
	const float s0 = __shfl_up_sync(0xffffffffu, L.h1, 1);
	float q0 = rw_quotient<MODE>(rhs[0], a * (((L.h1 + e[0]) + s0) + nn[0]), dv);
	if (feeder || ub == 0) q0 = here;
	const float s1 = __shfl_up_sync(0xffffffffu, q0, 1);
	float q1 = rw_quotient<MODE>(rhs[1], a * (((q0 + e[1]) + s1) + nn[1]), dv);
	if (feeder || ub == 1) q1 = e[0];
	const float s2 = __shfl_up_sync(0xffffffffu, q1, 1);
	float q2 = rw_quotient<MODE>(rhs[2], a * (((q1 + e[2]) + s2) + nn[2]), dv);
	if (feeder || ub == 2) q2 = e[1];
	const float s3 = __shfl_up_sync(0xffffffffu, q2, 1);
	float q3 = rw_quotient<MODE>(rhs[3], a * (((q2 + e[3]) + s3) + nn[3]), dv);
	if (feeder || ub == 3) q3 = e[2];
	if (L.computes && ub <= 2) { // some column of this block is interior
		float* px = L.xrow + c;
		if (c - ((4 - L.cls) & 3) >= c_first) {
			if (L.cls == 0) *reinterpret_cast<float4*>(px) = make_float4(q0, q1, q2, q3);
			else if (L.cls == 1) *reinterpret_cast<float4*>(px - 3) = make_float4(L.h3, L.h2, L.h1, q0);
			else if (L.cls == 2) *reinterpret_cast<float4*>(px - 2) = make_float4(L.h2, L.h1, q0, q1);
			else *reinterpret_cast<float4*>(px - 1) = make_float4(L.h1, q0, q1, q2);
		}
		const float q[4] = { q0, q1, q2, q3 };
#pragma unroll
		for (int u = 0; u < 4; u++) {
			if (u <= ub) continue; // left of the row (or the border cell itself)
			if (u == ub + 1) mirror_border(g, x, b, 1, j, q[u]); // the first column: its border cell, the corners
			else if (L.mirror_row) L.mirror_row[c + u] = __uint_as_float(__float_as_uint(q[u]) ^ L.mirror_flip);
		}
	}
	L.h3 = q1;
	L.h2 = q2;
	L.h1 = q3;
	L.c = c + 4;
}

// What the worker warp and its helper share.  The worker never touches the global counters: a fence that makes its
// stores visible to the GPU costs ~3000 cycles (ncu: 1.6 us per MEMBAR.SC.GPU with the row stores in flight) and a poll of
// a counter another L2 round trip, which at one of each per 16 steps was most of a step's time.
struct RwMail {
	volatile int gen;      // task number + 1, bumped by the worker when idx[] describes a new task
	volatile int idx[4];   // counters of this task: its own, (band, k-1), (band+1, k-1), (band-1, k); -1 = none
	volatile int done;     // worker -> helper: steps completed (stores ordered before it at CTA scope)
	volatile int pub;      // helper -> worker: the value of `done` that has been published to the GPU
	volatile int seen[3];  // helper -> worker: latest values of the three producers' counters
	volatile int ack;      // helper -> worker: the task (gen) that seen[] and pub refer to
	volatile int quit;
};

// Lane 0 of warp 1: polls the producers' counters into shared memory and publishes the worker's progress, fence included.
// Memory order: the worker's row stores are ordered before `done` by asm volatile("fence.acq_rel.cta;" ::: "memory"); the helper reads `done`, fences at
// GPU scope and only then stores the counter, so whoever reads the counter (and fences) sees the rows -- fence-to-fence
// synchronisation composes across the two scopes because both fences include both threads.
__device__ __forceinline__ void rw_helper(RwMail& mail, volatile int* progress, volatile int* abort_flag)
{
	int my_gen = 0, published = 0, i_mine = -1, i_p[3] = { -1, -1, -1 }, last[3] = { -1, -1, -1 };
	unsigned int idle = 50u;
	for (;;) {
		if (mail.quit || *abort_flag) return;
		const int gnow = mail.gen;
		if (gnow != my_gen) {
			asm volatile("fence.acq_rel.cta;" ::: "memory");
			my_gen = gnow;
			published = 0;
			i_mine = mail.idx[0];
			i_p[0] = mail.idx[1];
			i_p[1] = mail.idx[2];
			i_p[2] = mail.idx[3];
			last[0] = last[1] = last[2] = -1;
		}
		if (my_gen == 0) {
			__nanosleep(200);
			continue;
		}
		int v[3];
		bool news = (mail.ack != my_gen);
#pragma unroll
		for (int k = 0; k < 3; k++) {
			v[k] = (i_p[k] >= 0) ? progress[i_p[k]] : 0x7fffffff;
			news = news || (v[k] != last[k]);
		}
		if (news) {
			// (a fence at GPU scope stalls the whole SM's memory pipeline for its duration -- with one per poll the worker
			// ran at a tenth of its speed -- so only when there is something new to pass on)
			asm volatile("fence.acq_rel.gpu;" ::: "memory"); // the counters were read before anything the worker reads on their strength (acq_rel, not the sequentially consistent __threadfence)
#pragma unroll
			for (int k = 0; k < 3; k++) mail.seen[k] = last[k] = v[k];
		}
		if (mail.ack != my_gen) { // first values for this task are in place: the worker may trust seen[] from now on
			asm volatile("fence.acq_rel.cta;" ::: "memory");
			mail.ack = my_gen;
		}
		const int d = mail.done;
		if (d > published && mail.gen == my_gen) {
			asm volatile("fence.acq_rel.gpu;" ::: "memory");
			progress[i_mine] = d;
			published = d;
			mail.pub = d;
		} else if (!news) {
			idle = min(idle + idle / 4u + 1u, 600u);
			__nanosleep(idle);
		}
		if (news) idle = 50u;
	}
}

// sync[0] = ticket counter, sync[1] = abort flag, sync[2 + (sim * limit + k) * nb + band] = steps completed.
// CTA = two warps: warp 0 computes, lane 0 of warp 1 is its helper.
template <int MODE>
__global__ void __launch_bounds__(64) k_gs_rowwave(GridDesc g, int b, float* __restrict__ xb, const float* __restrict__ x0b, float a,
						    DivC dv, int limit, int nb, int batch, int* __restrict__ sync, int* __restrict__ fault,
						    unsigned long long* __restrict__ stamps)
{
	__shared__ __align__(16) float rx[33][kRwRing]; // 0..31: the lanes' rows; 32: never filled (lane 31's "north", unused)
	__shared__ __align__(16) float r0[32][kRwRing]; // right-hand side
	__shared__ RwMail mail;
	const int lane = threadIdx.x & 31;
	const int total = batch * nb * limit;
	const int nsteps = rowwave_steps(g.N);
	volatile int* progress = sync + 2;
	volatile int* abort_flag = sync + 1;
	if (threadIdx.x == 0) {
		mail.gen = 0;
		mail.done = 0;
		mail.pub = 0;
		mail.quit = 0;
		mail.seen[0] = mail.seen[1] = mail.seen[2] = 0;
		mail.ack = 0;
	}
	__syncthreads();
	if (threadIdx.x >= 32) {
		if (threadIdx.x == 32) rw_helper(mail, progress, abort_flag);
		return;
	}
	uint32_t rx_base = (uint32_t) __cvta_generic_to_shared(&rx[0][0]);
	uint32_t r0_base = (uint32_t) __cvta_generic_to_shared(&r0[0][0]);
	asm volatile("" : "+r"(rx_base), "+r"(r0_base)); // opaque: or the compiler re-derives the shared window's base at every use
	const int c_first = XO + 1, c_last = XO + g.N;  // interior columns
	const int last_group = (XO + g.N + 1) & ~3;     // the group that holds the right border cell; interior columns from here on are stored one by one
	// Blocks (first step n, a multiple of 4) in which every computing lane is well inside its row: lane 30, the last to arrive,
	// has its whole float4 group past the first column (its column n - kRwLead + c_first - 30 >= c_first + 3), and lane 1, the
	// first to leave, ends the block before the last column and before the tail that is stored column by column.
	const int inner_lo = (kRwLead + 30 * kRwSkew + 3 + 3) & ~3;
	const int inner_hi = ((kRwLead + min(c_last, last_group) - c_first - 1 + kRwSkew) & ~3) - 4;
	bool dead = false;
	// lane 0: spin on shared memory until cond() holds; gives up after ~3 s (a producer that will never come) or on abort
	auto spin = [&](auto cond) {
		if (lane == 0) {
			const long long t0 = clock64();
			unsigned int polls = 0, nap = 20;
			while (!cond()) {
				__nanosleep(nap); // (a waiting warp that polls at full rate takes issue slots from the working warps of its SM)
				nap = min(nap + nap / 2u, 400u);
				if ((++polls & 1023u) == 0u && (*abort_flag || clock64() - t0 > 6000000000LL)) {
					*abort_flag = 1;
					atomicOr(fault, 4);
					dead = true;
					break;
				}
			}
			asm volatile("fence.acq_rel.cta;" ::: "memory");
		}
		dead = __shfl_sync(0xffffffffu, dead ? 1 : 0, 0) != 0;
	};

	for (;;) {
		int ticket = 0;
		if (lane == 0) ticket = atomicAdd(&sync[0], 1);
		ticket = __shfl_sync(0xffffffffu, ticket, 0);
		if (ticket >= total || *abort_flag) break;
		// ticket -> (member, diagonal d = band + sweep, sweep): members innermost, then sweeps ascending within a diagonal
		const int sim = ticket % batch;
		int t = ticket / batch, k = 0, band = 0;
		for (int d = 0;; d++) {
			const int klo = max(0, d - nb + 1), khi = min(limit - 1, d);
			const int cnt = khi - klo + 1;
			if (t < cnt) {
				k = klo + t;
				band = d - k;
				break;
			}
			t -= cnt;
		}
		float* x = xb + (long long) sim * g.plane;
		const float* x0 = x0b + (long long) sim * g.plane;
		const int j0 = 1 + kRwRows * band, j = j0 - 1 + lane; // lane 0: the row below the band, lane 31: the row above it
		RwLane L;
		L.real = j <= g.NY + 1; // stored row: interior, or a border row
		L.computes = lane >= 1 && lane <= kRwRows && j <= g.NY; // (lanes 0 and 31 are the feeders)
		L.cls = lane & 3;
		L.xrow = x + (long long) j * g.pitch;
		L.x0row = x0 + (long long) j * g.pitch;
		L.ring_x = rx_base + 4u * kRwRing * (uint32_t) lane;
		L.ring_n = rx_base + 4u * kRwRing * (uint32_t) (lane + 1);
		L.ring_0 = r0_base + 4u * kRwRing * (uint32_t) lane;
		L.h1 = L.h2 = L.h3 = 0.0f;
		L.mirror_row = nullptr;
		L.mirror_flip = (b == 2) ? 0x80000000u : 0u;
		if (L.computes && b >= 0 && b <= 2 && g.NY > 1) { // (NY == 1: both border rows mirror the one row; mirror_border does that)
			if (j == 1 && (g.walls & 1)) L.mirror_row = L.xrow - g.pitch;
			if (j == g.NY && (g.walls & 2)) L.mirror_row = L.xrow + g.pitch;
		}
		L.c = c_first - kRwLead - kRwSkew * lane; // lane 0's column is a multiple of 4 whenever the step count is

		const int i_mine = (sim * limit + k) * nb + band;
		const bool has_own = (k > 0), has_above = (k > 0 && band + 1 < nb), has_below = (band > 0);
		if (lane == 0) { // hand the task to the helper
			mail.idx[0] = i_mine;
			mail.idx[1] = has_own ? i_mine - nb : -1;       // (band, k-1)
			mail.idx[2] = has_above ? i_mine - nb + 1 : -1; // (band+1, k-1)
			mail.idx[3] = has_below ? i_mine - 1 : -1;      // (band-1, k)
			mail.done = 0;
			mail.pub = 0;
			asm volatile("fence.acq_rel.cta;" ::: "memory");
			mail.gen = ticket + 1; // (seen[] is the helper's: it is valid once ack == gen)
		}

		// (diagnostics, FLUID_RW_STAMPS: when the task was taken, when its first wait ended, when its steps were done, when
		// its last count was published -- nanoseconds of the global timer)
		auto stamp = [&](int which) {
			if (stamps && lane == 0) {
				unsigned long long tns;
				asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tns));
				stamps[4 * (long long) i_mine + which] = tns;
			}
		};
		stamp(0);
		// before a block: every kRwPublish steps, wait for the producers
		auto before = [&](int n) {
			if (n == kRwPublish) stamp(1);
			if ((n % kRwPublish) == 0) {
				// What the next kRwPublish steps will touch.  Step counts are comparable across tasks (same schedule), so:
				//  own band, sweep k-1: at the first step of a block lane l fetches the group that ends at most 22 columns ahead
				//    of its column, which that task's lane l has stored when ITS block around that column is complete;
				//    also covers the border rows 0 and NY+1 (mirrored by that task) in the bottom / top band.
				//  band above, sweep k-1: lane 31 (31 steps behind lane 0) fetches row j0+30, written by that task's lane 1.
				//  band below, sweep k: lane 0 fetches row j0-1, written by that task's lane 30 (30 steps behind ITS lane 0).
				// Write-after-read: row j0 is read (as north row) by the band below in sweep k, which fetched column c long
				// before the step it must have completed for us to be at c; row j0+29 is read (as south row) by the band above
				// in sweep k-1, which fetched column c at its step c - 16, before the step c + 16 - 31 we wait for.
				// (tests: bitwise against the CPU.)
				const int upto = n + kRwPublish;
				const int need_own = has_own ? min(nsteps, upto + kRwAhead + 8) : 0;
				const int need_above = has_above ? min(nsteps, max(0, upto + kRwAhead + 8 - 31 * kRwSkew + 1)) : 0;
				const int need_below = has_below ? min(nsteps, upto + kRwAhead + 8 + 30 * kRwSkew) : 0;
				spin([&]() { return mail.ack == ticket + 1 && mail.seen[0] >= need_own && mail.seen[1] >= need_above && mail.seen[2] >= need_below; });
			}
		};
		// after a block
		auto after = [&](int n) {
			// what was fetched kRwPending + 1 blocks ago has arrived and is visible to the neighbouring lanes
			asm volatile("cp.async.commit_group;" ::: "memory");
			asm volatile("cp.async.wait_group %0;" ::"n"(kRwPending) : "memory");
			__syncwarp();
			if (((n + 4) % kRwPublish) == 0 || n + 4 >= nsteps) {
				if (lane == 0) {
					asm volatile("fence.acq_rel.cta;" ::: "memory"); // (after the __syncwarp: every lane's stores are ordered before the count; measured free)
					mail.done = min(n + 4, nsteps);
				}
			}
		};
		// The blocks in which the lanes arrive at their rows (what the band above waits for), then the rest.
		const int ramp_end = (g.NY > 1 && inner_hi >= inner_lo) ? inner_lo : 0;
		int n = 0;
		for (; n < ramp_end; n += 4) {
			before(n);
			if (dead) break;
			rw_block_ramp<MODE>(g, b, x, L, a, dv, j, c_first, last_group, lane == 0);
			after(n);
		}
		for (; n < nsteps && !dead; n += 4) {
			before(n);
			if (dead) break;
			if (n >= inner_lo && n <= inner_hi && g.NY > 1) rw_block_inner<MODE>(L, a, dv, last_group, lane == 0); // (NY == 1: see mirror_row)
			else rw_block<MODE>(g, b, x, L, a, dv, j, c_first, c_last, last_group);
			after(n);
		}
		if (dead) break;
		stamp(2);
		// the helper must have published the last count before the mailbox is reused (one fence's latency per task)
		spin([&]() { return mail.ack == ticket + 1 && mail.pub >= nsteps; });
		if (dead) break;
		stamp(3);
	}
	if (lane == 0) mail.quit = 1;
}

} // namespace fluid
