// Streaming / gather kernels of the Stable-Fluids step: add_source, set_bnd,
// divergence, gradient subtraction, advect, forcing, vorticity, damp.
// blockIdx.z always selects the ensemble member.
#pragma once
#include "grid.cuh"

namespace fluid {

// ---------------------------------------------------------------------------
// AddArrDt (FluidSystem.cpp:449-453): x += dt*s on every stored float (border
// included; pad columns are don't-care).  12 B per cell, pure streaming.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_add_source(float4* __restrict__ x, const float4* __restrict__ s, float dt,
						     long long n4)
{
	long long k = (long long) blockIdx.x * blockDim.x + threadIdx.x;
	const long long stride = (long long) gridDim.x * blockDim.x;
	for (; k < n4; k += stride) {
		float4 a = x[k];
		const float4 b = __ldg(&s[k]);
		a.x += dt * b.x;
		a.y += dt * b.y;
		a.z += dt * b.z;
		a.w += dt * b.w;
		x[k] = a;
	}
}

// Two fields in one launch (VelStep1's two AddArrDt calls, FluidSystem.cpp:356-357).
__global__ void __launch_bounds__(256) k_add_source2(float4* __restrict__ x, const float4* __restrict__ s,
						      float4* __restrict__ y, const float4* __restrict__ t, float dt,
						      long long n4)
{
	long long k = (long long) blockIdx.x * blockDim.x + threadIdx.x;
	const long long stride = (long long) gridDim.x * blockDim.x;
	for (; k < n4; k += stride) {
		float4 a = x[k];
		const float4 b = __ldg(&s[k]);
		float4 c = y[k];
		const float4 d = __ldg(&t[k]);
		a.x += dt * b.x; a.y += dt * b.y; a.z += dt * b.z; a.w += dt * b.w;
		c.x += dt * d.x; c.y += dt * d.y; c.z += dt * d.z; c.w += dt * d.w;
		x[k] = a;
		y[k] = c;
	}
}

// vel *= factor on every stored float (Damp, FluidSystem.cpp:1177-1183; the
// powf factor is loop invariant and comes from the host's libm).
__global__ void __launch_bounds__(256) k_scale(float4* __restrict__ x, float f, long long n4)
{
	long long k = (long long) blockIdx.x * blockDim.x + threadIdx.x;
	const long long stride = (long long) gridDim.x * blockDim.x;
	for (; k < n4; k += stride) {
		float4 a = x[k];
		a.x *= f; a.y *= f; a.z *= f; a.w *= f;
		x[k] = a;
	}
}

// ---------------------------------------------------------------------------
// Stand-alone SetBorder / ZeroBorder (FluidSystem.cpp:465-500, 455-463).  The
// fast paths fuse this into the producing kernel (mirror_border); it remains
// for the ABI (fluid_set_bnd) and for the first Blur pass.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_set_bnd(GridDesc g, int b, float* __restrict__ xb)
{
	float* x = xb + (long long) blockIdx.z * g.plane;
	const int k = blockIdx.x * blockDim.x + threadIdx.x + 1;
	const bool sides = (b >= 0 && b <= 2);
	if (sides) {
		if (k <= g.NY) {
			const float l = x[cell(g, 1, k)], r = x[cell(g, g.N, k)];
			x[cell(g, 0, k)] = (b == 1) ? -l : l;
			x[cell(g, g.N + 1, k)] = (b == 1) ? -r : r;
		}
		if (k <= g.N) {
			const float lo = x[cell(g, k, 1)], hi = x[cell(g, k, g.NY)];
			if (g.walls & 1) x[cell(g, k, 0)] = (b == 2) ? -lo : lo;
			if (g.walls & 2) x[cell(g, k, g.NY + 1)] = (b == 2) ? -hi : hi;
		}
	}
	if (k == 1) {
		// corners from the values the side pass writes (recomputed here from
		// their interior sources so no inter-thread ordering is needed)
		const int N = g.N, M = g.NY;
		float b10, b01, b1M1, b0M, bN0, bN11, bNM1, bN1M;
		if (sides) {
			const float c11 = x[cell(g, 1, 1)], c1M = x[cell(g, 1, M)], cN1 = x[cell(g, N, 1)], cNM = x[cell(g, N, M)];
			b10 = (b == 2) ? -c11 : c11;   b01 = (b == 1) ? -c11 : c11;
			b1M1 = (b == 2) ? -c1M : c1M;  b0M = (b == 1) ? -c1M : c1M;
			bN0 = (b == 2) ? -cN1 : cN1;   bN11 = (b == 1) ? -cN1 : cN1;
			bNM1 = (b == 2) ? -cNM : cNM;  bN1M = (b == 1) ? -cNM : cNM;
		} else {
			b10 = x[cell(g, 1, 0)];      b01 = x[cell(g, 0, 1)];
			b1M1 = x[cell(g, 1, M + 1)]; b0M = x[cell(g, 0, M)];
			bN0 = x[cell(g, N, 0)];      bN11 = x[cell(g, N + 1, 1)];
			bNM1 = x[cell(g, N, M + 1)]; bN1M = x[cell(g, N + 1, M)];
		}
		if (g.walls & 1) {
			x[cell(g, 0, 0)] = 0.5f * (b10 + b01);
			x[cell(g, N + 1, 0)] = 0.5f * (bN0 + bN11);
		}
		if (g.walls & 2) {
			x[cell(g, 0, M + 1)] = 0.5f * (b1M1 + b0M);
			x[cell(g, N + 1, M + 1)] = 0.5f * (bNM1 + bN1M);
		}
	}
}

__global__ void __launch_bounds__(256) k_zero_bnd(GridDesc g, float* __restrict__ xb)
{
	float* x = xb + (long long) blockIdx.z * g.plane;
	const int k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k <= g.N + 1) {
		if (g.walls & 1) x[cell(g, k, 0)] = 0.0f;
		if (g.walls & 2) x[cell(g, k, g.NY + 1)] = 0.0f;
	}
	if (k <= g.NY + 1) {
		x[cell(g, 0, k)] = 0.0f;
		x[cell(g, g.N + 1, k)] = 0.0f;
	}
}

// ---------------------------------------------------------------------------
// Project part 1 (FluidSystem.cpp:986-994): div = -0.5f*(uE-uW+vN-vS)/N on the
// interior (div's border is left untouched, as in the reference) fused with
// the whole-array p = 0.  One thread = one float4 group of one row.
// Algorithmic bytes: u,v read, div,p written = 16 B per cell.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_divergence(GridDesc g, const float* __restrict__ ub,
						     const float* __restrict__ vb, float* __restrict__ divb,
						     float* __restrict__ pb)
{
	const long long off = (long long) blockIdx.z * g.plane;
	const float* u = ub + off;
	const float* v = vb + off;
	float* dv = divb + off;
	const int q = blockIdx.x * blockDim.x + threadIdx.x; // float4 group in the row
	const int j = blockIdx.y * blockDim.y + threadIdx.y; // row 0..NY+1
	if (q * 4 >= g.pitch || j > g.NY + 1) return;
	const long long base = (long long) j * g.pitch + q * 4;
	// (pb == nullptr: the solve that follows has been told that p is zero and will not read it)
	if (pb) *reinterpret_cast<float4*>(pb + off + base) = make_float4(0.f, 0.f, 0.f, 0.f);
	const int i0 = q * 4 - XO; // interior index of lane 0
	if (j < 1 || j > g.NY || i0 + 3 < 1 || i0 > g.N) return;
	const float4 vn = __ldg(reinterpret_cast<const float4*>(v + base + g.pitch));
	const float4 vs = __ldg(reinterpret_cast<const float4*>(v + base - g.pitch));
	const float4 uc = __ldg(reinterpret_cast<const float4*>(u + base));
	const float uw = __ldg(u + base - 1), ue = __ldg(u + base + 4);
	const float fN = (float) g.N;
	float4 r;
	r.x = div_exact(-0.5f * (uc.y - uw + vn.x - vs.x), fN);
	r.y = div_exact(-0.5f * (uc.z - uc.x + vn.y - vs.y), fN);
	r.z = div_exact(-0.5f * (uc.w - uc.y + vn.z - vs.z), fN);
	r.w = div_exact(-0.5f * (ue - uc.z + vn.w - vs.w), fN);
	if (i0 >= 1 && i0 + 3 <= g.N) {
		*reinterpret_cast<float4*>(dv + base) = r;
	} else {
		if (i0 >= 1 && i0 <= g.N) dv[base] = r.x;
		if (i0 + 1 >= 1 && i0 + 1 <= g.N) dv[base + 1] = r.y;
		if (i0 + 2 >= 1 && i0 + 2 <= g.N) dv[base + 2] = r.z;
		if (i0 + 3 >= 1 && i0 + 3 <= g.N) dv[base + 3] = r.w;
	}
}

// ---------------------------------------------------------------------------
// Project part 3 (FluidSystem.cpp:998-1006): u -= 0.5f*N*(pE-pW),
// v -= 0.5f*N*(pN-pS), with SetBorder(1,u) and SetBorder(2,v) fused.
// Algorithmic bytes: p read, u,v read+write = 20 B per cell.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_subtract_gradient(GridDesc g, float* __restrict__ ub, float* __restrict__ vb,
							    const float* __restrict__ pb)
{
	const long long off = (long long) blockIdx.z * g.plane;
	float* u = ub + off;
	float* v = vb + off;
	const float* p = pb + off;
	const int q = blockIdx.x * blockDim.x + threadIdx.x;
	const int j = blockIdx.y * blockDim.y + threadIdx.y + 1; // row 1..NY
	const int i0 = q * 4 - XO;
	if (q * 4 >= g.pitch || j > g.NY || i0 + 3 < 1 || i0 > g.N) return;
	const long long base = (long long) j * g.pitch + q * 4;
	const float4 pc = __ldg(reinterpret_cast<const float4*>(p + base));
	const float4 pn = __ldg(reinterpret_cast<const float4*>(p + base + g.pitch));
	const float4 ps = __ldg(reinterpret_cast<const float4*>(p + base - g.pitch));
	const float pw = __ldg(p + base - 1), pe = __ldg(p + base + 4);
	float4 uc = *reinterpret_cast<const float4*>(u + base);
	float4 vc = *reinterpret_cast<const float4*>(v + base);
	const float h = 0.5f * (float) g.N;
	uc.x -= h * (pc.y - pw);   vc.x -= h * (pn.x - ps.x);
	uc.y -= h * (pc.z - pc.x); vc.y -= h * (pn.y - ps.y);
	uc.z -= h * (pc.w - pc.y); vc.z -= h * (pn.z - ps.z);
	uc.w -= h * (pe - pc.z);   vc.w -= h * (pn.w - ps.w);
	const float un[4] = { uc.x, uc.y, uc.z, uc.w }, vn[4] = { vc.x, vc.y, vc.z, vc.w };
	const bool full = (i0 >= 1 && i0 + 3 <= g.N);
	if (full) {
		*reinterpret_cast<float4*>(u + base) = uc;
		*reinterpret_cast<float4*>(v + base) = vc;
	}
#pragma unroll
	for (int l = 0; l < 4; l++) {
		const int i = i0 + l;
		if (i < 1 || i > g.N) continue;
		if (!full) {
			u[base + l] = un[l];
			v[base + l] = vn[l];
		}
		mirror_border(g, u, 1, i, j, un[l]);
		mirror_border(g, v, 2, i, j, vn[l]);
	}
}

// ---------------------------------------------------------------------------
// BilinearCoeffs::calc + Eval (FluidCode.h:91-112), generalised to NY rows.
// ---------------------------------------------------------------------------
// Offsets inside one member's plane are 32-bit (the context refuses planes of 2^31 floats or more): the
// gathers were spending more instructions on 64-bit address arithmetic than on the interpolation.
struct Bilin {
	int o00; // offset of (i0,j0); (i1,j0)=+1, (i0,j1)=+pitch
	float s0, s1, t0, t1;
};

__device__ __forceinline__ int celli(const GridDesc& g, int i, int j) { return j * g.pitch + (XO + i); }

// (px, py) are GLOBAL cell coordinates; the sampled rows are converted to local ones at the end.
// The reference clamps with `if (i < 0.5f) i = 0.5f; if (i > N + 0.5f) i = N + 0.5f` (FluidCode.h:95-103);
// fmaxf/fminf give the same value for every non-NaN input in two instructions, and map a NaN
// coordinate (for which the reference indexes out of bounds) to the lower clamp.
// SLAB: a sample whose rows are not stored on this slab (backtrace longer than the halo) sets *oob
// and is clamped into the slab so that the access stays legal; the host reports it after the step.
template <bool SLAB>
__device__ __forceinline__ Bilin bilin_calc(const GridDesc& g, float px, float py, int* oob)
{
	Bilin r;
	px = fminf(fmaxf(px, 0.5f), g.N + 0.5f);
	py = fminf(fmaxf(py, 0.5f), g.NYG + 0.5f);
	const int i0 = (int) px;
	const int j0 = (int) py;
	r.s1 = px - i0;
	r.s0 = 1 - r.s1;
	r.t1 = py - j0;
	r.t0 = 1 - r.t1;
	int jl = j0;
	if (SLAB) {
		jl = j0 - g.j0;
		if (jl < 0 || jl > g.NY) {
			if (oob) *oob = 1;
			jl = min(max(jl, 0), g.NY);
		}
	}
	r.o00 = celli(g, i0, jl);
	return r;
}

// `fc` points at the THREAD'S OWN cell (offset `id`) of the field, the sample is addressed relative to it:
// a per-thread 64-bit pointer plus a 32-bit difference is one IMAD.WIDE, whereas kernel-parameter base +
// member offset + cell offset made the compiler rebuild a 64-bit sum (shifts, carries: ~5 instructions) for
// every one of the 12-16 gathers of a cell -- more integer work than the interpolation has arithmetic.
// (opaque_ptr: left to itself the compiler folds the pointer back into base + index and does the 64-bit sum again)
__device__ __forceinline__ void opaque_ptr(const float*& p) { asm("" : "+l"(p)); }

__device__ __forceinline__ float bilin_eval(const Bilin& c, const float* __restrict__ fc, int id, int pitch)
{
	const float* r0 = fc + (c.o00 - id);
	const float* r1 = r0 + pitch;
	const float d00 = __ldg(r0), d10 = __ldg(r0 + 1);
	const float d01 = __ldg(r1), d11 = __ldg(r1 + 1);
	return c.s0 * (c.t0 * d00 + c.t1 * d01) + c.s1 * (c.t0 * d10 + c.t1 * d11);
}

// Backtraced sample position of cell (i,j): MidpointBackward (FluidSystem.h:185-206)
// or EulerBack (FluidSystem.h:208-219).
// (i, j) local; the returned position is in global coordinates.
template <bool SLAB, int MODE>
__device__ __forceinline__ void backtrace(const GridDesc& g, const float* __restrict__ uc_p, const float* __restrict__ vc_p,
					  int i, int j, int id, float dtN, float& px, float& py, int* oob)
{
	const float uc = __ldg(uc_p), vc = __ldg(vc_p); // pointers at the thread's own cell
	const int jg = SLAB ? j + g.j0 : j;
	if (MODE == 1) {
		px = i - dtN * uc;
		py = jg - dtN * vc;
	} else {
		const float dirx = -dtN * uc;
		const float diry = -dtN * vc;
		const float mx = i + 0.5f * dirx;
		const float my = jg + 0.5f * diry;
		const Bilin m = bilin_calc<SLAB>(g, mx, my, oob);
		const float vx = bilin_eval(m, uc_p, id, g.pitch);
		const float vy = bilin_eval(m, vc_p, id, g.pitch);
		px = i - dtN * vx;
		py = jg - dtN * vy;
	}
}

// ---------------------------------------------------------------------------
// Advect1 (FluidSystem.cpp:850-862) for one scalar field, SetBorder fused.
// Gather kernel: u,v,d0 read, d written = 16 B per cell algorithmic.
// ---------------------------------------------------------------------------
template <bool SLAB, int MODE>
__global__ void __launch_bounds__(256) k_advect(GridDesc g, int b, float* __restrict__ db, const float* __restrict__ d0b,
						 const float* __restrict__ ub, const float* __restrict__ vb, float dtN,
						 int* oob_flag, int own_lo, int own_hi)
{
	const long long off = (long long) blockIdx.z * g.plane;
	const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
	const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
	if (i > g.N || j > g.NY) return;
	int* oob = (SLAB && j >= own_lo && j <= own_hi) ? oob_flag : nullptr; // halo rows may legitimately trace off the slab
	float px, py;
	const int id = celli(g, i, j);
	const long long mine = off + id;
	const float* d0c = d0b + mine;
	const float* uc = ub + mine;
	const float* vc = vb + mine;
	opaque_ptr(d0c);
	opaque_ptr(uc);
	opaque_ptr(vc);
	// The sample of d0 is taken three dependent memory round trips from now, most often within a cell or two
	// of this one: touch the two lines now, together with the loads of u and v.  Nothing uses the values, and
	// a prefetch hint or an unused load ends up scheduled right before the sample (or removed), so they are
	// tied into the address of the mid-point gathers through a term that is always zero (g.walls < 4) but that the
	// compiler cannot evaluate.
	const int warm = (__float_as_int(__ldg(d0c)) | __float_as_int(__ldg(d0c + g.pitch))) & (g.walls >> 2);
	backtrace<SLAB, MODE>(g, uc, vc, i, j, id + warm, dtN, px, py, oob); // (id only offsets the mid-point gathers)
	const Bilin c = bilin_calc<SLAB>(g, px, py, oob);
	const float val = bilin_eval(c, d0c, id, g.pitch);
	db[mine] = val;
	mirror_border(g, db + off, b, i, j, val);
}

// VelStep1's self-advection (FluidSystem.cpp:372-373): Advect1(1,u,u0,u0,v0) and
// Advect1(2,v,v0,u0,v0) share the backtrace, so both components are produced
// by one pass over (u0,v0): 16 B per cell instead of 2 x 12 B.
template <bool SLAB, int MODE>
__global__ void __launch_bounds__(256) k_advect_velocity(GridDesc g, float* __restrict__ ub, float* __restrict__ vb,
							  const float* __restrict__ u0b, const float* __restrict__ v0b,
							  float dtN, int* oob_flag, int own_lo, int own_hi, int bu, int bv)
{
	const long long off = (long long) blockIdx.z * g.plane;
	const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
	const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
	if (i > g.N || j > g.NY) return;
	int* oob = (SLAB && j >= own_lo && j <= own_hi) ? oob_flag : nullptr;
	float px, py;
	const int id = celli(g, i, j);
	const long long mine = off + id;
	const float* u0c = u0b + mine;
	const float* v0c = v0b + mine;
	opaque_ptr(u0c);
	opaque_ptr(v0c);
	backtrace<SLAB, MODE>(g, u0c, v0c, i, j, id, dtN, px, py, oob);
	const Bilin c = bilin_calc<SLAB>(g, px, py, oob);
	const float un = bilin_eval(c, u0c, id, g.pitch);
	const float vn = bilin_eval(c, v0c, id, g.pitch);
	float* u = ub + off;
	float* v = vb + off;
	ub[mine] = un;
	vb[mine] = vn;
	mirror_border(g, u, bu, i, j, un); // (1, 2) for VelStep1; no SetBorder (-2, -2) for VelStep1B
	mirror_border(g, v, bv, i, j, vn);
}

// ---------------------------------------------------------------------------
// Forcing: what the reference's updaters do between Update's zero fill and
// VelStep (FluidSystem.cpp:104-118, FluidUpdaters.cpp:45-67).  Entries of one
// ensemble member are applied in list order by one thread (as the CPU loop
// does), members in parallel.
// ---------------------------------------------------------------------------
__global__ void k_apply_sources(GridDesc g, FieldPtrs f, const SourceDev* __restrict__ src,
				const int* __restrict__ first, int batch)
{
	const int sim = blockIdx.x * blockDim.x + threadIdx.x;
	if (sim >= batch) return;
	const long long off = (long long) sim * g.plane;
	for (int k = first[sim]; k < first[sim + 1]; k++) {
		const SourceDev e = src[k];
		const int jl = e.j - g.j0; // entries carry global rows; a slab applies those it stores (halo rows included)
		if (jl < 0 || jl > g.NY + 1) continue;
		const long long id = off + cell(g, e.i, jl);
		if (e.mode == 0) f.prev[e.field][id] = e.value;
		else if (e.mode == 1) f.cur[e.field][id] += e.value;
		else f.cur[e.field][id] = e.value;
	}
}

// ---------------------------------------------------------------------------
// AddArrDt (FluidSystem.cpp:449-453) when the source array is zero except for a short list of cells -- the
// arrays Update has just zero-filled plus what the updaters wrote (FluidSystem.cpp:104-118).  x += dt*0 leaves
// every value as it is except -0, which becomes +0 (for dt >= 0; for dt < 0 the addend is -0 and nothing
// changes at all): k_canon_zero reads the field once and writes only the float4 groups that hold a -0 -- 4
// instead of 12 bytes per cell, and the zero-filled array is not read.  The listed cells are different: they
// get x + dt*source, the reference's expression on the reference's operands, from k_add_source_cells BEFORE
// the sweep (which might turn a sum of -0 into +0) and are put back after it by k_restore_source_cells.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_canon_zero(float4* __restrict__ x, float4* __restrict__ y, long long n4)
{
	long long k = (long long) blockIdx.x * blockDim.x + threadIdx.x;
	const long long stride = (long long) gridDim.x * blockDim.x;
	auto fix = [](float4* p) {
		float4 a = *p;
		const uint32_t nz = 0x80000000u;
		if (__float_as_uint(a.x) == nz || __float_as_uint(a.y) == nz || __float_as_uint(a.z) == nz || __float_as_uint(a.w) == nz) {
			a.x += 0.0f; a.y += 0.0f; a.z += 0.0f; a.w += 0.0f; // every other value is unchanged by + 0
			*p = a;
		}
	};
	for (; k < n4; k += stride) {
		fix(x + k);
		if (y) fix(y + k);
	}
}

// slots [slot_lo, slot_hi) of the field table: 0 = u, 1 = v, 2 + k = density of fluid k
__global__ void k_add_source_cells(GridDesc g, FieldPtrs f, const SourceDev* __restrict__ src, const int* __restrict__ first,
				   int batch, float* __restrict__ saved, float dt, int slot_lo, int slot_hi, int restore)
{
	const int sim = blockIdx.x * blockDim.x + threadIdx.x;
	if (sim >= batch) return;
	const long long off = (long long) sim * g.plane;
	for (int k = first[sim]; k < first[sim + 1]; k++) {
		const SourceDev e = src[k];
		const int jl = e.j - g.j0;
		if (e.mode != 0 || !e.last || e.field < slot_lo || e.field >= slot_hi || jl < 0 || jl > g.NY + 1) continue;
		float* p = f.cur[e.field] + (off + cell(g, e.i, jl));
		if (restore) {
			*p = saved[k];
		} else {
			const float y = *p + dt * e.value;
			*p = y;
			saved[k] = y;
		}
	}
}

// ---------------------------------------------------------------------------
// VortConfinement (FluidSystem.cpp:1124-1175) in three order-independent
// phases; SetBorder(0) of the intermediate fields fused via mirror_border.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_vort_curl(GridDesc g, const float* __restrict__ ub, const float* __restrict__ vb,
						    float* __restrict__ wb, float* __restrict__ awb)
{
	const long long off = (long long) blockIdx.z * g.plane;
	const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
	const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
	if (i > g.N || j > g.NY) return;
	const float* u = ub + off;
	const float* v = vb + off;
	const long long id = cell(g, i, j);
	const float w = 0.5f * (__ldg(v + id + 1) - __ldg(v + id - 1) - __ldg(u + id + g.pitch) + __ldg(u + id - g.pitch));
	const float aw = (w >= 0.0f) ? w : -w;
	(wb + off)[id] = w;
	(awb + off)[id] = aw;
	mirror_border(g, wb + off, 0, i, j, w);
	mirror_border(g, awb + off, 0, i, j, aw);
}

__global__ void __launch_bounds__(256) k_vort_dir(GridDesc g, const float* __restrict__ awb, float* __restrict__ fxb,
						   float* __restrict__ fyb)
{
	const long long off = (long long) blockIdx.z * g.plane;
	const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
	const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
	if (i > g.N || j > g.NY) return;
	const float* aw = awb + off;
	const long long id = cell(g, i, j);
	const float gx = 0.5f * (__ldg(aw + id + 1) - __ldg(aw + id - 1));
	const float gy = 0.5f * (__ldg(aw + id + g.pitch) - __ldg(aw + id - g.pitch));
	const float len = sqrtf(gx * gx + gy * gy);
	float fx = 0.0f, fy = 0.0f;
	if (!(len < 0.01f)) {
		fx = div_exact(gx, len);
		fy = div_exact(gy, len);
	}
	(fxb + off)[id] = fx;
	(fyb + off)[id] = fy;
	mirror_border(g, fxb + off, 0, i, j, fx);
	mirror_border(g, fyb + off, 0, i, j, fy);
}

__global__ void __launch_bounds__(256) k_vort_apply(GridDesc g, float* __restrict__ ub, float* __restrict__ vb,
						     const float* __restrict__ wb, const float* __restrict__ fxb,
						     const float* __restrict__ fyb, float vorticity)
{
	const long long off = (long long) blockIdx.z * g.plane;
	const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
	const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
	if (i > g.N || j > g.NY) return;
	const long long id = cell(g, i, j);
	const float w = __ldg(wb + off + id);
	float* u = ub + off;
	float* v = vb + off;
	const float un = u[id] + vorticity * (__ldg(fyb + off + id) * w);
	const float vn = v[id] + vorticity * (-__ldg(fxb + off + id) * w);
	u[id] = un;
	v[id] = vn;
	mirror_border(g, u, 0, i, j, un);
	mirror_border(g, v, 0, i, j, vn);
}

// ---------------------------------------------------------------------------
// Headless stand-in for the display path: the per-cell colourise loop and the
// velocity line list of FluidRenderer::DoBuffer (FluidRenderer.cpp:56-84, 97-132), written to dense device
// buffers instead of a GL texture / vertex buffer (graphics APIs are unavailable on this box).
// FLUID cells: every fluid's colour weighted by its density, normalised where the summed density exceeds 1
// (:106-119); OBJECT cells magenta (:121-123); anything else green.  info == nullptr: the map ResetCellInfo
// leaves (FluidSystem.cpp:1215-1232: outer ring not FLUID, no OBJECT cells), which the ORIGINAL path never changes.
// pixels: (N+2)^2 x 3 floats, lines: (N+2)^2 x 4 floats (x0, y0, x1, y1), both indexed by ID(i,j).
// ---------------------------------------------------------------------------
struct RenderFluids {
	const float* dens[8];
	float col[8][3];
	int n;
};

__global__ void __launch_bounds__(256) k_colourise(GridDesc g, RenderFluids fl, const unsigned char* __restrict__ info,
						    const float* __restrict__ ub, const float* __restrict__ vb,
						    float* __restrict__ pixels, float* __restrict__ lines)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	const int j = blockIdx.y * blockDim.y + threadIdx.y;
	if (i > g.N + 1 || j > g.NY + 1) return;
	const long long id = cell(g, i, j);
	const long long out = (long long) i + (long long) (g.N + 2) * j;
	float r = 0.f, gg = 0.f, b = 0.f;
	const unsigned char m = info ? info[id] : (unsigned char) ((i >= 1 && i <= g.N && j >= 1 && j <= g.NY) ? 31 : 15);
	if (m & 16) {
		float d_total = 0.f;
		for (int n = 0; n < fl.n; n++) {
			const float d = __ldg(fl.dens[n] + id);
			d_total += d;
			r += fl.col[n][0] * d;
			gg += fl.col[n][1] * d;
			b += fl.col[n][2] * d;
		}
		if (d_total > 1) {
			r = div_exact(r, d_total);
			gg = div_exact(gg, d_total);
			b = div_exact(b, d_total);
		}
	} else if (m & 32) {
		b = r = 1.f;
	} else {
		gg = 1.f;
	}
	pixels[3 * out + 0] = r;
	pixels[3 * out + 1] = gg;
	pixels[3 * out + 2] = b;
	if (lines) {
		const float vx0 = __ldg(ub + id), vy0 = __ldg(vb + id);
		const float mag = sqrtf(vx0 * vx0 + vy0 * vy0);
		const int N2 = g.N + 2;
		const float s = N2 * 1.05f;
		float vx = (vx0 / mag) / s, vy = (vy0 / mag) / s;
		if ((double) mag < 0.000000000001) vx = vy = 0.f;
		const float lx = (i + 0.5f) / (float) N2, ly = (j + 0.5f) / (float) N2;
		lines[4 * out + 0] = lx;
		lines[4 * out + 1] = ly;
		lines[4 * out + 2] = lx + vx;
		lines[4 * out + 3] = ly + vy;
	}
}

} // namespace fluid
