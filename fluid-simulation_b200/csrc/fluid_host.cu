// Stam-signature host-pointer entry points (include/fluid_cuda.h): one cached context per host
// thread; arrays are copied in, the primitive runs on the GPU, every array the reference writes is
// copied back.  The signatures are the reference's with list -> float* (FluidCode.h:4).
#include "ctx.h"

extern "C" int fluid_ctx_destroy(fluid_ctx* c);

struct HostCtxCache {
	fluid_ctx* c = nullptr;
	~HostCtxCache() { /* process teardown: the CUDA runtime may already be gone, do not touch it */ }
};
static thread_local HostCtxCache t_cache;

static thread_local int t_gs_order = FLUID_GS_WAVEFRONT;
static thread_local int t_gs_iters = 20;
static thread_local int t_advect_mode = FLUID_ADVECT_MIDPOINT;
static thread_local int t_scratch_download = 1;

extern "C" int fluid_set_gs_order(int order)
{
	if (order != FLUID_GS_WAVEFRONT && order != FLUID_GS_REDBLACK) return fail(FLUID_ERR_INVALID, "unknown gs order %d", order);
	t_gs_order = order;
	return FLUID_OK;
}

extern "C" int fluid_set_gs_iters(int iters)
{
	if (iters < 1) return fail(FLUID_ERR_INVALID, "gs_iters must be >= 1 (got %d)", iters);
	t_gs_iters = iters;
	return FLUID_OK;
}

extern "C" int fluid_set_advect_mode(int mode)
{
	if (mode != FLUID_ADVECT_MIDPOINT && mode != FLUID_ADVECT_EULER) return fail(FLUID_ERR_INVALID, "unknown advect mode %d", mode);
	t_advect_mode = mode;
	return FLUID_OK;
}

extern "C" int fluid_set_scratch_download(int on)
{
	t_scratch_download = on ? 1 : 0;
	return FLUID_OK;
}

extern "C" int fluid_set_device(int device)
{
	int ndev = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || ndev == 0)
		return fail(FLUID_ERR_CUDA, "no CUDA device available (%s); this library has no CPU fallback",
			    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
	if (device < 0 || device >= ndev) return fail(FLUID_ERR_INVALID, "device %d does not exist (%d devices)", device, ndev);
	CU_TRY(cudaSetDevice(device));
	if (t_cache.c && t_cache.c->device != device) {
		fluid_ctx_destroy(t_cache.c);
		t_cache.c = nullptr;
		CU_TRY(cudaSetDevice(device));
	}
	return FLUID_OK;
}
// ---------------------------------------------------------------------------
// Stam-signature host-pointer entry points.  One cached context per host
// thread (re-created when N changes); arrays are copied in, the primitive runs
// on the GPU, and every array the reference writes is copied back.
// ---------------------------------------------------------------------------
static int host_ctx(int N, fluid_ctx** out)
{
	if (N < 1) return fail(FLUID_ERR_INVALID, "N=%d must be >= 1", N);
	if (t_cache.c && t_cache.c->g.N != N) {
		fluid_ctx_destroy(t_cache.c);
		t_cache.c = nullptr;
	}
	if (!t_cache.c) {
		fluid_params p;
		fluid_params_default(&p);
		p.N = N;
		FL_TRY(fluid_ctx_create(&p, &t_cache.c));
	}
	fluid_ctx* c = t_cache.c;
	FL_TRY(ctx_bind(c));
	c->gs_order = t_gs_order;
	c->gs_iters = t_gs_iters;
	c->advect_mode = t_advect_mode;
	*out = c;
	return FLUID_OK;
}

static int up(fluid_ctx* c, int field, const float* h) { return fluid_ctx_upload(c, field, 0, h); }

static int down_async(fluid_ctx* c, int field, float* h)
{
	const size_t w = (size_t) (c->g.N + 2) * sizeof(float);
	CU_TRY(cudaMemcpy2DAsync(h, w, c->f[field] + XO, (size_t) c->g.pitch * sizeof(float), w, c->g.NY + 2,
				 cudaMemcpyDeviceToHost, c->stream));
	return FLUID_OK;
}

static int finish(fluid_ctx* c)
{
	CU_TRY(cudaStreamSynchronize(c->stream));
	CU_TRY(cudaGetLastError());
	return FLUID_OK;
}

#define NEED(p)                                                                                                   \
	do {                                                                                                      \
		if (!(p)) return fail(FLUID_ERR_INVALID, "null array argument '%s'", #p);                         \
	} while (0)

extern "C" int fluid_add_source(int N, float* x, const float* s, float dt)
{
	NEED(x); NEED(s);
	fluid_ctx* c;
	FL_TRY(host_ctx(N, &c));
	FL_TRY(up(c, 0, x));
	FL_TRY(up(c, 1, s));
	FL_TRY(dev_add_source(c, c->f[0], c->f[1], dt));
	FL_TRY(down_async(c, 0, x));
	return finish(c);
}

extern "C" int fluid_set_bnd(int N, int b, float* x)
{
	NEED(x);
	fluid_ctx* c;
	FL_TRY(host_ctx(N, &c));
	FL_TRY(up(c, 0, x));
	FL_TRY(dev_set_bnd(c, b, c->f[0]));
	FL_TRY(down_async(c, 0, x));
	return finish(c);
}

extern "C" int fluid_lin_solve(int N, int b, float* x, const float* x0, float a, float cc)
{
	NEED(x); NEED(x0);
	if (x == x0) return fail(FLUID_ERR_INVALID, "lin_solve: x and x0 must not alias");
	fluid_ctx* c;
	FL_TRY(host_ctx(N, &c));
	FL_TRY(up(c, 0, x));
	FL_TRY(up(c, 1, x0));
	float* px = c->f[0];
	FL_TRY(dev_lin_solve(c, b, px, c->f[1], a, cc));
	FL_TRY(down_async(c, 0, x));
	return finish(c);
}

extern "C" int fluid_diffuse(int N, int b, float* x, const float* x0, float diff, float dt)
{
	NEED(x); NEED(x0);
	if (x == x0) return fail(FLUID_ERR_INVALID, "diffuse: x and x0 must not alias");
	fluid_ctx* c;
	FL_TRY(host_ctx(N, &c));
	FL_TRY(up(c, 0, x));
	FL_TRY(up(c, 1, x0));
	const float a = diffusion_a(c, diff, dt);
	float* px = c->f[0];
	FL_TRY(dev_lin_solve(c, b, px, c->f[1], a, 1 + 4 * a));
	FL_TRY(down_async(c, 0, x));
	return finish(c);
}

extern "C" int fluid_diffuse_blur(int N, int b, float* x, float* x0, float diff, float dt)
{
	NEED(x); NEED(x0);
	if (x == x0) return fail(FLUID_ERR_INVALID, "diffuse_blur: x and x0 must not alias");
	fluid_ctx* c;
	FL_TRY(host_ctx(N, &c));
	FL_TRY(up(c, 0, x));
	FL_TRY(up(c, 1, x0));
	const float a = diffusion_a(c, diff, dt);
	FL_TRY(dev_blur(c, b, c->f[0], c->f[1], a, 1 + 4 * a));
	FL_TRY(down_async(c, 0, x));
	if (t_scratch_download) FL_TRY(down_async(c, 1, x0));
	return finish(c);
}

extern "C" int fluid_advect(int N, int b, float* d, const float* d0, const float* u, const float* v, float dt)
{
	NEED(d); NEED(d0); NEED(u); NEED(v);
	if (d == d0 || d == u || d == v) return fail(FLUID_ERR_INVALID, "advect: the output d must not alias an input");
	fluid_ctx* c;
	FL_TRY(host_ctx(N, &c));
	// d0 may alias u or v (VelStep1 does exactly that, FluidSystem.cpp:372-373)
	FL_TRY(up(c, 0, d));
	FL_TRY(up(c, 2, u));
	FL_TRY(up(c, 3, v));
	int f0 = 1;
	if (d0 == u) f0 = 2;
	else if (d0 == v) f0 = 3;
	else FL_TRY(up(c, 1, d0));
	FL_TRY(dev_advect(c, b, c->f[0], c->f[f0], c->f[2], c->f[3], dt));
	FL_TRY(down_async(c, 0, d));
	return finish(c);
}

extern "C" int fluid_project(int N, float* u, float* v, float* p, float* div)
{
	NEED(u); NEED(v); NEED(p); NEED(div);
	fluid_ctx* c;
	FL_TRY(host_ctx(N, &c));
	FL_TRY(up(c, 0, u));
	FL_TRY(up(c, 1, v));
	FL_TRY(up(c, 3, div)); // div's border cells survive the call (FluidSystem.cpp:986-990 write the interior only)
	float* pp = c->f[2];
	FL_TRY(dev_project(c, c->f[0], c->f[1], pp, c->f[3]));
	FL_TRY(down_async(c, 0, u));
	FL_TRY(down_async(c, 1, v));
	if (t_scratch_download) {
		FL_TRY(down_async(c, 2, p));
		FL_TRY(down_async(c, 3, div));
	}
	return finish(c);
}

extern "C" int fluid_vel_step(int N, float* u, float* v, float* u0, float* v0, float visc, float dt)
{
	NEED(u); NEED(v); NEED(u0); NEED(v0);
	fluid_ctx* c;
	FL_TRY(host_ctx(N, &c));
	FL_TRY(up(c, FLUID_U, u));
	FL_TRY(up(c, FLUID_V, v));
	FL_TRY(up(c, FLUID_U_PREV, u0));
	FL_TRY(up(c, FLUID_V_PREV, v0));
	FL_TRY(dev_vel_step(c, c->f[FLUID_U], c->f[FLUID_V], c->f[FLUID_U_PREV], c->f[FLUID_V_PREV], visc, dt));
	FL_TRY(down_async(c, FLUID_U, u));
	FL_TRY(down_async(c, FLUID_V, v));
	if (t_scratch_download) {
		FL_TRY(down_async(c, FLUID_U_PREV, u0));
		FL_TRY(down_async(c, FLUID_V_PREV, v0));
	}
	return finish(c);
}

extern "C" int fluid_dens_step(int N, float* x, float* x0, const float* u, const float* v, float diff, float dt)
{
	NEED(x); NEED(x0); NEED(u); NEED(v);
	fluid_ctx* c;
	FL_TRY(host_ctx(N, &c));
	FL_TRY(up(c, FLUID_DENS, x));
	FL_TRY(up(c, FLUID_DENS_PREV, x0));
	FL_TRY(up(c, FLUID_U, u));
	FL_TRY(up(c, FLUID_V, v));
	FL_TRY(dev_dens_step(c, c->f[FLUID_DENS], c->f[FLUID_DENS_PREV], c->f[FLUID_U], c->f[FLUID_V], diff, dt));
	FL_TRY(down_async(c, FLUID_DENS, x));
	if (t_scratch_download) FL_TRY(down_async(c, FLUID_DENS_PREV, x0));
	return finish(c);
}

// FluidSystem::Update (FluidSystem.cpp:101-128) on host arrays: the reference's public per-frame call.
// Only the three fields cross the bus; the source arrays are zero-filled and forced on the device.
extern "C" int fluid_update(int N, float* u, float* v, float* dens, int nsrc, const fluid_source* src, float visc, float diff,
			    float vorticity, float damp, float dt)
{
	NEED(u); NEED(v); NEED(dens);
	fluid_ctx* c;
	FL_TRY(host_ctx(N, &c));
	FL_TRY(fluid_ctx_set_sources(c, -1, nsrc, src));
	c->visc = visc;
	c->diff = diff;
	c->vorticity = vorticity;
	c->damp = damp;
	// The density is not needed before DensStep and the velocity is final after VelStep (no damping): with
	// pinned host arrays the density goes up while VelStep runs and the velocity comes down while DensStep
	// runs, on a second stream.  Forcing that writes the current density needs it on the device first.
	bool overlap = (damp == 0.0f);
	for (int k = 0; k < nsrc; k++)
		if (src[k].mode != FLUID_SRC_PREV_SET && src[k].field >= FLUID_DENS) overlap = false;
	if (!overlap) {
		FL_TRY(up(c, FLUID_U, u));
		FL_TRY(up(c, FLUID_V, v));
		FL_TRY(up(c, FLUID_DENS, dens));
		FL_TRY(dev_update(c, dt));
		FL_TRY(down_async(c, FLUID_U, u));
		FL_TRY(down_async(c, FLUID_V, v));
		FL_TRY(down_async(c, FLUID_DENS, dens));
		return finish(c);
	}
	if (!c->copy_stream) {
		CU_TRY(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
		CU_TRY(cudaEventCreateWithFlags(&c->ev_dens_up, cudaEventDisableTiming));
		CU_TRY(cudaEventCreateWithFlags(&c->ev_vel_done, cudaEventDisableTiming));
	}
	const size_t w = (size_t) (c->g.N + 2) * sizeof(float), dp = (size_t) c->g.pitch * sizeof(float);
	auto enqueue = [&]() -> int {
		FL_TRY(up(c, FLUID_U, u));
		FL_TRY(up(c, FLUID_V, v));
		CU_TRY(cudaMemcpy2DAsync(c->f[FLUID_DENS] + XO, dp, dens, w, w, c->g.NY + 2, cudaMemcpyHostToDevice, c->copy_stream));
		CU_TRY(cudaEventRecord(c->ev_dens_up, c->copy_stream));
		auto velocity_home = [&]() -> int { // VelStep is enqueued: its results may leave while DensStep runs
			CU_TRY(cudaEventRecord(c->ev_vel_done, c->stream));
			CU_TRY(cudaStreamWaitEvent(c->copy_stream, c->ev_vel_done, 0));
			CU_TRY(cudaMemcpy2DAsync(u, w, c->f[FLUID_U] + XO, dp, w, c->g.NY + 2, cudaMemcpyDeviceToHost, c->copy_stream));
			CU_TRY(cudaMemcpy2DAsync(v, w, c->f[FLUID_V] + XO, dp, w, c->g.NY + 2, cudaMemcpyDeviceToHost, c->copy_stream));
			return FLUID_OK;
		};
		FL_TRY(dev_update_hooks(c, dt, c->ev_dens_up, velocity_home));
		return down_async(c, FLUID_DENS, dens);
	};
	const int rc = enqueue();
	if (rc != FLUID_OK) { // nothing may still be writing the caller's arrays when the error is returned
		cudaStreamSynchronize(c->copy_stream);
		cudaStreamSynchronize(c->stream);
		return rc;
	}
	CU_TRY(cudaStreamSynchronize(c->copy_stream));
	return finish(c);
}

extern "C" int fluid_vort_confinement(int N, float* u, float* v, float vorticity)
{
	NEED(u); NEED(v);
	fluid_ctx* c;
	FL_TRY(host_ctx(N, &c));
	FL_TRY(up(c, 0, u));
	FL_TRY(up(c, 1, v));
	FL_TRY(dev_vort_confinement(c, c->f[0], c->f[1], vorticity));
	FL_TRY(down_async(c, 0, u));
	FL_TRY(down_async(c, 1, v));
	return finish(c);
}

extern "C" int fluid_damp(int N, float* vel, float amount, float dt)
{
	NEED(vel);
	fluid_ctx* c;
	FL_TRY(host_ctx(N, &c));
	FL_TRY(up(c, 0, vel));
	FL_TRY(dev_damp(c, c->f[0], amount, dt));
	FL_TRY(down_async(c, 0, vel));
	return finish(c);
}
