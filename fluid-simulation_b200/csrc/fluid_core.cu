// libfluid_cuda.so -- the C ABI declared in include/fluid_cuda.h: library-level calls and the
// resident context (life cycle, transfers, forcing lists, modes, profiling, single primitives on
// resident fields).  Without a CUDA device every entry point fails; there is no CPU fallback.
#include "ctx.h"

#include <stdlib.h>

thread_local char t_err[512] = "";
thread_local unsigned long long t_launches = 0;

int fail(int code, const char* fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(t_err, sizeof t_err, fmt, ap);
	va_end(ap);
	return code;
}

// ---------------------------------------------------------------------------
// library-level entry points
// ---------------------------------------------------------------------------
extern "C" const char* fluid_last_error(void) { return t_err; }

extern "C" const char* fluid_version(void)
{
	return "fluid_cuda 0.1 (sm_100a, -fmad=false -prec-div=true -prec-sqrt=true -ftz=false)";
}

extern "C" unsigned long long fluid_launch_count(void) { return t_launches; }
extern "C" void fluid_launch_count_reset(void) { t_launches = 0; }

extern "C" int fluid_ctx_destroy(fluid_ctx* c);
extern "C" int fluid_host_alloc(void** p, size_t bytes)
{
	if (!p) return fail(FLUID_ERR_INVALID, "null pointer");
	CU_TRY(cudaHostAlloc(p, bytes, cudaHostAllocDefault));
	return FLUID_OK;
}

extern "C" int fluid_host_free(void* p)
{
	CU_TRY(cudaFreeHost(p));
	return FLUID_OK;
}

extern "C" void fluid_params_default(fluid_params* p)
{
	if (!p) return;
	memset(p, 0, sizeof *p);
	p->N = 64;
	p->batch = 1;
	p->gs_iters = 20;
	p->gs_order = FLUID_GS_WAVEFRONT;
	p->advect_mode = FLUID_ADVECT_MIDPOINT;
	p->device = -1;
}

extern "C" int fluid_ctx_destroy(fluid_ctx* c)
{
	if (!c) return FLUID_OK;
	cudaSetDevice(c->device);
	if (c->stream) cudaStreamSynchronize(c->stream);
	for (auto& p : c->f) if (p) cudaFree(p);
	for (auto& p : c->vscr) if (p) cudaFree(p);
	if (c->rb_scratch) cudaFree(c->rb_scratch);
	if (c->rb_scratch2) cudaFree(c->rb_scratch2);
	if (c->lane_stream) cudaStreamDestroy(c->lane_stream);
	if (c->ev_fork2) cudaEventDestroy(c->ev_fork2);
	if (c->ev_join2) cudaEventDestroy(c->ev_join2);
	if (c->d_oob) cudaFree(c->d_oob);
	if (c->peer_block) cudaFree(c->peer_block);
	for (auto& e : c->pace_ev) if (e) cudaEventDestroy(e);
	if (c->peer_counter) cudaFree(c->peer_counter);
	if (c->div_pending.stream) {
		cudaStreamSynchronize(c->div_pending.stream);
		cudaStreamDestroy(c->div_pending.stream);
		cudaEventDestroy(c->div_pending.done);
		cudaFreeHost(c->div_pending.h_bins);
	}
	if (c->d_divbins) cudaFree(c->d_divbins);
	for (auto& p : c->d_redo) if (p) cudaFree(p);
	if (c->side_stream) cudaStreamDestroy(c->side_stream);
	if (c->ev_fork) cudaEventDestroy(c->ev_fork);
	if (c->ev_join) cudaEventDestroy(c->ev_join);
	if (c->d_pixels) cudaFree(c->d_pixels);
	if (c->d_lines) cudaFree(c->d_lines);
	if (c->d_src) cudaFree(c->d_src);
	if (c->d_src_saved) cudaFree(c->d_src_saved);
	if (c->d_info) cudaFree(c->d_info);
	for (float* p : { c->mf_a, c->mf_c, c->mf_p }) if (p) cudaFree(p);
	for (auto& p : c->d_wf_sync) if (p) cudaFree(p);
	if (c->d_first) cudaFree(c->d_first);
	for (auto& sg : c->prof_segs) { cudaEventDestroy(sg.a); cudaEventDestroy(sg.b); }
	for (auto& e : c->prof_pool) cudaEventDestroy(e);
	if (c->ev0) cudaEventDestroy(c->ev0);
	if (c->ev1) cudaEventDestroy(c->ev1);
	for (auto& fb : c->feedback) {
		if (fb.h_count) cudaFreeHost(fb.h_count);
		if (fb.arrived) cudaEventDestroy(fb.arrived);
	}
	if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
	if (c->ev_dens_up) cudaEventDestroy(c->ev_dens_up);
	if (c->ev_vel_done) cudaEventDestroy(c->ev_vel_done);
	if (c->own_stream) cudaStreamDestroy(c->own_stream);
	delete c;
	return FLUID_OK;
}

// Row slabs: every kernel computes on all stored rows and stale values creep in from the outer halo edge by
// one row per stencil application.  Between two halo refreshes the longest chain is a projection: divergence
// (1 row), gs_iters red-black sweeps (2 rows each), gradient (1 row) -- and Blur's gs_iters + 1 passes.
static int check_slab_iters(int gs_iters, int halo)
{
	if (2 * gs_iters + 2 > halo)
		return fail(FLUID_ERR_UNSUPPORTED, "row slabs: gs_iters=%d needs a halo of at least %d rows (2 per red-black sweep + 2), slab_halo is %d",
			    gs_iters, 2 * gs_iters + 2, halo);
	return FLUID_OK;
}

extern "C" int fluid_ctx_create(const fluid_params* p, fluid_ctx** out)
{
	if (!p || !out) return fail(FLUID_ERR_INVALID, "null argument");
	*out = nullptr;
	if (p->N < 1 || p->N > 32768) return fail(FLUID_ERR_INVALID, "N=%d out of range [1,32768]", p->N);
	if (p->batch < 1 || p->batch > 65535) return fail(FLUID_ERR_INVALID, "batch=%d out of range [1,65535]", p->batch);
	if (p->gs_iters < 1) return fail(FLUID_ERR_INVALID, "gs_iters=%d must be >= 1", p->gs_iters);
	if (p->gs_order != FLUID_GS_WAVEFRONT && p->gs_order != FLUID_GS_REDBLACK) return fail(FLUID_ERR_INVALID, "unknown gs_order %d", p->gs_order);
	if (p->advect_mode != FLUID_ADVECT_MIDPOINT && p->advect_mode != FLUID_ADVECT_EULER) return fail(FLUID_ERR_INVALID, "unknown advect_mode %d", p->advect_mode);
	if (p->fs_type != FLUID_TYPE_ORIGINAL && p->fs_type != FLUID_TYPE_ORIGINAL_BORDERED && p->fs_type != FLUID_TYPE_ORIGINAL_BORDERED_MF) return fail(FLUID_ERR_INVALID, "unknown fs_type %d", p->fs_type);
	if (p->fs_type != FLUID_TYPE_ORIGINAL && (p->slab_count > 1 || p->slab_rows > 0))
		return fail(FLUID_ERR_UNSUPPORTED, "the bordered solver types run on whole square grids only (no row slabs)");
	if (p->n_fluids < 0 || p->n_fluids > FLUID_MAX_FLUIDS) return fail(FLUID_ERR_INVALID, "n_fluids=%d out of range [1,%d]", p->n_fluids, FLUID_MAX_FLUIDS);
	if (p->slab_rows < 0 || p->slab_rows > 1048576) return fail(FLUID_ERR_INVALID, "slab_rows=%d out of range", p->slab_rows);
	if (p->slab_count < 0 || (p->slab_count > 1 && (p->slab_index < 0 || p->slab_index >= p->slab_count)))
		return fail(FLUID_ERR_INVALID, "slab_index=%d / slab_count=%d is not a valid slab", p->slab_index, p->slab_count);
	if (p->slab_count > 1) {
		const int halo = p->slab_halo > 0 ? p->slab_halo : 48;
		if (p->slab_rows < halo + 1) return fail(FLUID_ERR_INVALID, "slab_rows=%d must exceed the halo depth %d", p->slab_rows, halo);
		if (p->batch != 1) return fail(FLUID_ERR_UNSUPPORTED, "row slabs and ensembles cannot be combined");
		if (p->gs_order != FLUID_GS_REDBLACK) return fail(FLUID_ERR_UNSUPPORTED, "row slabs need FLUID_GS_REDBLACK (the reference order is a pipeline across slabs)");
		FL_TRY(check_slab_iters(p->gs_iters, halo));
	}

	int ndev = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || ndev == 0)
		return fail(FLUID_ERR_CUDA, "no CUDA device available (%s); this library has no CPU fallback",
			    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
	int dev = p->device;
	if (dev < 0) CU_TRY(cudaGetDevice(&dev));
	if (dev >= ndev) return fail(FLUID_ERR_INVALID, "device %d does not exist (%d devices)", dev, ndev);

	fluid_ctx* c = new (std::nothrow) fluid_ctx();
	if (!c) return fail(FLUID_ERR_NOMEM, "host allocation failed");
	c->device = dev;
	c->batch = p->batch;
	c->gs_iters = p->gs_iters;
	c->gs_order = p->gs_order;
	c->advect_mode = p->advect_mode;
	c->visc = p->visc;
	c->diff = p->diff;
	c->vorticity = p->vorticity;
	c->damp = p->damp;
	c->n_fluids = p->n_fluids > 0 ? p->n_fluids : 1;
	c->n_fields = FLUID_FIELD_COUNT + 2 * (c->n_fluids - 1);
	c->g.N = p->N;
	c->g.NY = p->N;
	c->g.j0 = 0;
	c->g.NYG = p->N;
	c->g.walls = 3;
	if (p->slab_count <= 1 && p->slab_rows > 0) {
		// a whole (non-square) grid of N columns x slab_rows rows on one device; h = 1/N as for slabs
		c->g.NY = p->slab_rows;
		c->g.NYG = p->slab_rows;
	}
	if (p->slab_count > 1) {
		// rank `slab_index` of `slab_count` owns `slab_rows` rows of a global N x (slab_rows*slab_count) grid
		c->slab_index = p->slab_index;
		c->slab_count = p->slab_count;
		c->slab_rows = p->slab_rows;
		c->slab_halo = p->slab_halo > 0 ? p->slab_halo : 48;
		const bool lo_nb = c->slab_index > 0, hi_nb = c->slab_index < c->slab_count - 1;
		c->g.NY = c->slab_rows + (lo_nb ? c->slab_halo : 0) + (hi_nb ? c->slab_halo : 0);
		c->g.NYG = c->slab_rows * c->slab_count;
		c->g.j0 = c->slab_index * c->slab_rows - (lo_nb ? c->slab_halo : 0);
		c->g.walls = (lo_nb ? 0 : 1) | (hi_nb ? 0 : 2);
		c->own_lo = lo_nb ? c->slab_halo + 1 : 1;
	}
	c->g.pitch = ((XO + p->N + 2 + 31) / 32) * 32;
	c->g.plane = (long long) c->g.pitch * (c->g.NY + 2);
	if (c->g.plane >= (1LL << 31)) { // kernels index one member's plane with 32-bit offsets
		const int rows = c->g.NY;
		delete c;
		return fail(FLUID_ERR_UNSUPPORTED, "a field of %d x %d cells has 2^31 floats or more; split it into row slabs", p->N, rows);
	}
	c->rb_rx = choose_rb_rx(p->N);
	c->field_bytes = (size_t) c->g.plane * c->batch * sizeof(float);
	c->src.resize(c->batch);

	int rc = FLUID_OK;
	do {
		if (cudaSetDevice(dev) != cudaSuccess) { rc = fail(FLUID_ERR_CUDA, "cudaSetDevice(%d) failed", dev); break; }
		cudaDeviceGetAttribute(&c->num_sms, cudaDevAttrMultiProcessorCount, dev);
		if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { rc = fail(FLUID_ERR_CUDA, "cudaStreamCreate failed: %s", cudaGetErrorString(cudaGetLastError())); break; }
		c->own_stream = c->stream;
		if (cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess) { rc = fail(FLUID_ERR_CUDA, "cudaEventCreate failed"); break; }
		for (int k = 0; k < c->n_fields && rc == FLUID_OK; k++) {
			e = cudaMalloc(&c->f[k], c->field_bytes);
			if (e != cudaSuccess) {
				rc = fail(FLUID_ERR_NOMEM, "cudaMalloc of %zu bytes failed: %s", c->field_bytes, cudaGetErrorString(e));
				cudaGetLastError();
				break;
			}
			e = cudaMemsetAsync(c->f[k], 0, c->field_bytes, c->stream);
			if (e != cudaSuccess) rc = fail(FLUID_ERR_CUDA, "cudaMemset failed: %s", cudaGetErrorString(e));
		}
		if (rc != FLUID_OK) break;
		if (cudaMalloc(&c->d_oob, sizeof(int)) != cudaSuccess || cudaMemsetAsync(c->d_oob, 0, sizeof(int), c->stream) != cudaSuccess) {
			rc = fail(FLUID_ERR_NOMEM, "cudaMalloc failed: %s", cudaGetErrorString(cudaGetLastError()));
			break;
		}
		c->src_dirty = true;
		rc = sync_sources(c);
		if (rc != FLUID_OK) break;
		c->fs_type = p->fs_type;
		if (c->fs_type != FLUID_TYPE_ORIGINAL) {
			rc = bordered_alloc(c);
			if (rc != FLUID_OK) break;
		}
		if (cudaStreamSynchronize(c->stream) != cudaSuccess) { rc = fail(FLUID_ERR_CUDA, "stream sync failed: %s", cudaGetErrorString(cudaGetLastError())); break; }
	} while (0);
	if (rc != FLUID_OK) {
		char keep[sizeof t_err];
		memcpy(keep, t_err, sizeof keep);
		fluid_ctx_destroy(c);
		memcpy(t_err, keep, sizeof keep);
		return rc;
	}
	*out = c;
	return FLUID_OK;
}

int check_field_n(int f, int n) { return (f >= 0 && f < n) ? FLUID_OK : fail(FLUID_ERR_INVALID, "field id %d out of range [0,%d)", f, n); }

int check_distinct_n(int n_fields, std::initializer_list<int> ids)
{
	std::vector<int> v(ids);
	for (size_t a = 0; a < v.size(); a++) {
		FL_TRY(check_field_n(v[a], n_fields));
		for (size_t b = a + 1; b < v.size(); b++)
			if (v[a] == v[b]) return fail(FLUID_ERR_INVALID, "field %d passed twice where distinct arrays are required", v[a]);
	}
	return FLUID_OK;
}

extern "C" int fluid_ctx_clear(fluid_ctx* c)
{
	CHECK_CTX(c);
	for (int k = 0; k < c->n_fields; k++) CU_TRY(cudaMemsetAsync(c->f[k], 0, c->field_bytes, c->stream));
	return FLUID_OK;
}

extern "C" int fluid_ctx_upload(fluid_ctx* c, int field, int sim, const float* host)
{
	CHECK_CTX(c);
	FL_TRY(check_field(field));
	if (!host) return fail(FLUID_ERR_INVALID, "null host pointer");
	if (sim < 0 || sim >= c->batch) return fail(FLUID_ERR_INVALID, "sim %d out of range [0,%d)", sim, c->batch);
	// slabs: `host` is the GLOBAL dense array; every stored row (halo included) is taken from it
	const size_t w = (size_t) (c->g.N + 2) * sizeof(float);
	CU_TRY(cudaMemcpy2DAsync(c->f[field] + (long long) sim * c->g.plane + XO, (size_t) c->g.pitch * sizeof(float),
				 host + (size_t) c->g.j0 * (c->g.N + 2), w, w, c->g.NY + 2, cudaMemcpyHostToDevice, c->stream));
	return FLUID_OK;
}

extern "C" int fluid_ctx_download(fluid_ctx* c, int field, int sim, float* host)
{
	CHECK_CTX(c);
	FL_TRY(check_field(field));
	if (!host) return fail(FLUID_ERR_INVALID, "null host pointer");
	if (sim < 0 || sim >= c->batch) return fail(FLUID_ERR_INVALID, "sim %d out of range [0,%d)", sim, c->batch);
	// slabs: `host` is the GLOBAL dense array; only the rows this slab owns (plus the wall rows) are written
	const size_t w = (size_t) (c->g.N + 2) * sizeof(float);
	int lo = 0, n = c->g.NY + 2;
	if (c->slab_count > 1) {
		lo = (c->g.walls & 1) ? 0 : c->own_lo;
		const int hi = (c->g.walls & 2) ? c->g.NY + 1 : c->own_lo + c->slab_rows - 1;
		n = hi - lo + 1;
	}
	CU_TRY(cudaMemcpy2DAsync(host + (size_t) (c->g.j0 + lo) * (c->g.N + 2), w,
				 c->f[field] + (long long) sim * c->g.plane + (long long) lo * c->g.pitch + XO,
				 (size_t) c->g.pitch * sizeof(float), w, n, cudaMemcpyDeviceToHost, c->stream));
	CU_TRY(cudaStreamSynchronize(c->stream));
	return FLUID_OK;
}

// Slab-local transfers: `host` is a dense (N+2) x (NY_local+2) array of exactly the rows this
// context stores (halo rows included); for a whole-grid context identical to upload/download.
extern "C" int fluid_ctx_upload_local(fluid_ctx* c, int field, const float* host)
{
	CHECK_CTX(c);
	FL_TRY(check_field(field));
	if (!host) return fail(FLUID_ERR_INVALID, "null host pointer");
	const size_t w = (size_t) (c->g.N + 2) * sizeof(float);
	CU_TRY(cudaMemcpy2DAsync(c->f[field] + XO, (size_t) c->g.pitch * sizeof(float), host, w, w, c->g.NY + 2, cudaMemcpyHostToDevice,
				 c->stream));
	return FLUID_OK;
}

extern "C" int fluid_ctx_download_local(fluid_ctx* c, int field, float* host)
{
	CHECK_CTX(c);
	FL_TRY(check_field(field));
	if (!host) return fail(FLUID_ERR_INVALID, "null host pointer");
	const size_t w = (size_t) (c->g.N + 2) * sizeof(float);
	CU_TRY(cudaMemcpy2DAsync(host, w, c->f[field] + XO, (size_t) c->g.pitch * sizeof(float), w, c->g.NY + 2, cudaMemcpyDeviceToHost,
				 c->stream));
	CU_TRY(cudaStreamSynchronize(c->stream));
	return FLUID_OK;
}

extern "C" int fluid_ctx_local_rows(fluid_ctx* c, int* rows_stored, int* first_global_row)
{
	CHECK_CTX(c);
	if (rows_stored) *rows_stored = c->g.NY + 2;
	if (first_global_row) *first_global_row = c->g.j0;
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_sources(fluid_ctx* c, int sim, int n, const fluid_source* src)
{
	CHECK_CTX(c);
	if (n < 0 || (n > 0 && !src)) return fail(FLUID_ERR_INVALID, "bad source list");
	if (sim < -1 || sim >= c->batch) return fail(FLUID_ERR_INVALID, "sim %d out of range", sim);
	std::vector<SourceDev> v;
	for (int k = 0; k < n; k++) {
		const fluid_source& s = src[k];
		int slot = -1; // 0 = u, 1 = v, 2 + fluid = a density field
		if (s.field == FLUID_U || s.field == FLUID_V) slot = s.field;
		for (int fl = 0; fl < c->n_fluids; fl++)
			if (s.field == FLUID_DENS_OF(fl)) slot = 2 + fl;
		if (slot < 0) return fail(FLUID_ERR_INVALID, "source %d: field %d is not U, V or a density field of this context", k, s.field);
		if (s.i < 0 || s.i > c->g.N + 1 || s.j < 0 || s.j > c->g.NYG + 1) return fail(FLUID_ERR_INVALID, "source %d: cell (%d,%d) outside the grid", k, s.i, s.j);
		if (s.mode < FLUID_SRC_PREV_SET || s.mode > FLUID_SRC_CURR_SET) return fail(FLUID_ERR_INVALID, "source %d: unknown mode %d", k, s.mode);
		v.push_back({ slot, s.i, s.j, s.mode, s.value, 0 });
	}
	auto same = [](const std::vector<SourceDev>& a, const std::vector<SourceDev>& b) {
		return a.size() == b.size() && (a.empty() || memcmp(a.data(), b.data(), a.size() * sizeof(SourceDev)) == 0);
	};
	bool changed = false;
	if (sim < 0) {
		for (auto& e : c->src)
			if (!same(e, v)) { e = v; changed = true; }
	} else if (!same(c->src[sim], v)) {
		c->src[sim] = v;
		changed = true;
	}
	if (!changed) return FLUID_OK; // a per-frame caller re-sending the same forcing costs nothing
	c->src_dirty = true;
	CU_TRY(cudaStreamSynchronize(c->stream)); // the old list may still be in use
	return sync_sources(c);
}

extern "C" int fluid_ctx_set_modes(fluid_ctx* c, int gs_order, int gs_iters, int advect_mode)
{
	CHECK_CTX(c);
	if (gs_order != FLUID_GS_WAVEFRONT && gs_order != FLUID_GS_REDBLACK) return fail(FLUID_ERR_INVALID, "unknown gs_order %d", gs_order);
	if (gs_iters < 1) return fail(FLUID_ERR_INVALID, "gs_iters=%d must be >= 1", gs_iters);
	if (advect_mode != FLUID_ADVECT_MIDPOINT && advect_mode != FLUID_ADVECT_EULER) return fail(FLUID_ERR_INVALID, "unknown advect_mode %d", advect_mode);
	if (c->slab_count > 1) {
		if (gs_order != FLUID_GS_REDBLACK) return fail(FLUID_ERR_UNSUPPORTED, "row slabs need FLUID_GS_REDBLACK (the reference order is a pipeline across slabs)");
		FL_TRY(check_slab_iters(gs_iters, c->slab_halo));
	}
	c->gs_order = gs_order;
	c->gs_iters = gs_iters;
	c->advect_mode = advect_mode;
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_tuning(fluid_ctx* c, int rb_tile_sweeps)
{
	CHECK_CTX(c);
	if (rb_tile_sweeps < 0 || rb_tile_sweeps > 2) return fail(FLUID_ERR_INVALID, "rb_tile_sweeps=%d must be 0, 1 or 2", rb_tile_sweeps);
	c->rb_tile_sweeps = rb_tile_sweeps;
	return FLUID_OK;
}

extern "C" int fluid_ctx_render(fluid_ctx* c, int sim, const float* color_rgb, float* host_pixels, float* host_lines)
{
	CHECK_CTX(c);
	return dev_render(c, sim, color_rgb, host_pixels, host_lines);
}

extern "C" int fluid_ctx_set_stream_kernel(fluid_ctx* c, int on)
{
	CHECK_CTX(c);
	if (on < 0 || on > 2) return fail(FLUID_ERR_INVALID, "stream kernel mode %d must be 0 (off), 1 (on) or 2 (on, always the inline-exact variant)", on);
	c->rb_stream = on != 0;
	c->stream_force_inline = on == 2;
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_advect_kernel(fluid_ctx* c, int which)
{
	CHECK_CTX(c);
	if (which < 0 || which > 2) return fail(FLUID_ERR_INVALID, "advect kernel %d must be 0 (automatic), 1 (gather per cell) or 2 (row window)", which);
	c->advect_kernel = which;
	return FLUID_OK;
}

#define CHECK_BORDERED(c)                                                                                         \
	do {                                                                                                      \
		CHECK_CTX(c);                                                                                     \
		if ((c)->fs_type == FLUID_TYPE_ORIGINAL)                                                          \
			return fail(FLUID_ERR_INVALID, "this context is FLUID_TYPE_ORIGINAL: it has no cell map (create it with fs_type = FLUID_TYPE_ORIGINAL_BORDERED)"); \
	} while (0)

extern "C" int fluid_ctx_reset_cell_info(fluid_ctx* c)
{
	CHECK_BORDERED(c);
	return bordered_reset(c);
}

extern "C" int fluid_ctx_disable_edges(fluid_ctx* c, int n, const int* edges)
{
	CHECK_BORDERED(c);
	if (n < 0 || (n > 0 && !edges)) return fail(FLUID_ERR_INVALID, "bad edge list");
	for (int k = 0; k < n; k++) {
		const int* e = edges + 4 * k;
		const int dx = e[2] - e[0], dy = e[3] - e[1];
		if (e[0] < 0 || e[0] > c->g.N + 1 || e[2] < 0 || e[2] > c->g.N + 1 || e[1] < 0 || e[1] > c->g.NY + 1 || e[3] < 0 || e[3] > c->g.NY + 1 ||
		    abs(dx) + abs(dy) != 1)
			return fail(FLUID_ERR_INVALID, "edge %d: (%d,%d)-(%d,%d) does not join two adjacent cells of the grid", k, e[0], e[1], e[2], e[3]);
	}
	return bordered_disable_edges(c, n, edges);
}

// SquareBorder::WriteBorder (FluidBorder.cpp:4-19): bounds in float, truncated, exactly as the scene computes them.
extern "C" int fluid_ctx_add_square_border(fluid_ctx* c)
{
	CHECK_BORDERED(c);
	const int N = c->g.N;
	const float size = 0.3;
	const int low = 1 + N / 2 - size * N / 2;
	const int high = 1 + N / 2 + size * N / 2;
	std::vector<int> e;
	for (int i = low; i <= high; i++) {
		const int four[4][4] = { { i, low, i, low - 1 }, { i, high, i, high + 1 }, { low, i, low - 1, i }, { high, i, high + 1, i } };
		for (auto& q : four) e.insert(e.end(), q, q + 4);
	}
	return fluid_ctx_disable_edges(c, (int) e.size() / 4, e.data());
}

extern "C" int fluid_ctx_set_cell_info(fluid_ctx* c, const unsigned char* host_info, int recalc)
{
	CHECK_BORDERED(c);
	if (!host_info) return fail(FLUID_ERR_INVALID, "null pointer");
	const size_t w = (size_t) c->g.N + 2;
	CU_TRY(cudaMemcpy2DAsync(c->d_info + XO, (size_t) c->g.pitch, host_info, w, w, c->g.NY + 2, cudaMemcpyHostToDevice, c->stream));
	if (recalc) FL_TRY(bordered_calc(c));
	CU_TRY(cudaStreamSynchronize(c->stream));
	return FLUID_OK;
}

extern "C" int fluid_ctx_get_cell_info(fluid_ctx* c, unsigned char* host_info)
{
	CHECK_BORDERED(c);
	if (!host_info) return fail(FLUID_ERR_INVALID, "null pointer");
	const size_t w = (size_t) c->g.N + 2;
	CU_TRY(cudaMemcpy2DAsync(host_info, w, c->d_info + XO, (size_t) c->g.pitch, w, c->g.NY + 2, cudaMemcpyDeviceToHost, c->stream));
	CU_TRY(cudaStreamSynchronize(c->stream));
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_sparse_sources(fluid_ctx* c, int on)
{
	CHECK_CTX(c);
	c->sparse_sources = on != 0;
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_overlap(fluid_ctx* c, int on)
{
	CHECK_CTX(c);
	c->overlap_density = on != 0;
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_wavefront_tile(fluid_ctx* c, int tile)
{
	CHECK_CTX(c);
	if (tile != 0 && tile != 16 && tile != 32 && tile != 64) return fail(FLUID_ERR_INVALID, "wavefront tile %d must be 0 (automatic), 16, 32 or 64", tile);
	c->wavefront_tile = tile;
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_wavefront_persistent(fluid_ctx* c, int on)
{
	CHECK_CTX(c);
	if (on < 0 || on > 3) return fail(FLUID_ERR_INVALID, "wavefront_persistent=%d must be 0 (never), 1 (automatic), 2 (tile tickets) or 3 (row pipeline)", on);
	c->wavefront_persistent = on;
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_small_grid_kernels(fluid_ctx* c, int on)
{
	CHECK_CTX(c);
	c->small_kernels = on != 0;
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_fast_division(fluid_ctx* c, int mode)
{
	CHECK_CTX(c);
	if (mode < 0 || mode > 2) return fail(FLUID_ERR_INVALID, "fast division mode %d must be 0 (never), 1 (auto) or 2 (always)", mode);
	c->fast_div_mode = mode;
	return FLUID_OK;
}

// Exponent window [2^lo_exp, 2^(hi_exp+1)) of dividends for which the short division by `divisor` was proven exact.
extern "C" int fluid_ctx_division_window(fluid_ctx* c, float divisor, int* lo_exp, int* hi_exp)
{
	CHECK_CTX(c);
	if (!lo_exp || !hi_exp) return fail(FLUID_ERR_INVALID, "null pointer");
	const int keep = c->fast_div_mode;
	c->fast_div_mode = 2;
	DivC dv;
	const int rc = division_for(c, divisor, 0, &dv);
	c->fast_div_mode = keep;
	FL_TRY(rc);
	if (dv.range == 0) {
		*lo_exp = 0;
		*hi_exp = -1;
	} else {
		*lo_exp = (int) (dv.lo >> 23) - 127;
		*hi_exp = (int) ((dv.lo + dv.range) >> 23) - 128;
	}
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_coeffs(fluid_ctx* c, float visc, float diff, float vorticity, float damp)
{
	CHECK_CTX(c);
	c->visc = visc;
	c->diff = diff;
	c->fl_set[0] = false; // fluid 0 follows these again
	c->vorticity = vorticity;
	c->damp = damp;
	return FLUID_OK;
}

extern "C" int fluid_ctx_vel_step(fluid_ctx* c, float dt)
{
	CHECK_CTX(c);
	return dev_vel_step(c, c->f[FLUID_U], c->f[FLUID_V], c->f[FLUID_U_PREV], c->f[FLUID_V_PREV], visc_avg(c), dt);
}

extern "C" int fluid_ctx_dens_step(fluid_ctx* c, float dt)
{
	CHECK_CTX(c);
	for (int fl = 0; fl < c->n_fluids; fl++)
		FL_TRY(dev_dens_step(c, c->f[FLUID_DENS_OF(fl)], c->f[FLUID_DENS_PREV_OF(fl)], c->f[FLUID_U], c->f[FLUID_V], fluid_diff(c, fl), dt));
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_visc_air(fluid_ctx* c, float visc_air)
{
	CHECK_CTX(c);
	c->visc_air = visc_air;
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_fluid(fluid_ctx* c, int k, float diff, float visc)
{
	CHECK_CTX(c);
	if (k < 0 || k >= c->n_fluids) return fail(FLUID_ERR_INVALID, "fluid %d out of range [0,%d)", k, c->n_fluids);
	c->fl_diff[k] = diff;
	c->fl_visc[k] = visc;
	c->fl_set[k] = true;
	return FLUID_OK;
}

extern "C" int fluid_ctx_step(fluid_ctx* c, float dt, int nsteps)
{
	CHECK_CTX(c);
	if (nsteps < 0) return fail(FLUID_ERR_INVALID, "nsteps=%d must be >= 0", nsteps);
	CU_TRY(cudaEventRecord(c->ev0, c->stream));
	// Row slabs on the peer-memory transport: the host enqueues frame k only when frame k-2 has finished.  With the host
	// many frames ahead of the GPU a frame took 10.0 ms instead of 6.07 (2 GPUs, N=8192 rows each; bounded to 1, 2 or 3
	// frames: 6.12 / 6.07 / 6.07): each refresh waits on the neighbour GPU, and the deeper the queue of launches behind
	// such a wait, the later the refresh kernels of BOTH sides got to start.  Two frames of run-ahead still hide every
	// launch latency.  (FLUID_PACE_FRAMES overrides: 0 = unbounded.)
	static const int pace_env = getenv("FLUID_PACE_FRAMES") ? atoi(getenv("FLUID_PACE_FRAMES")) : -1;
	const int pace = (c->slab_count > 1 && c->peer_connected) ? (pace_env >= 0 ? std::min(pace_env, 3) : 2) : 0;
	for (int s = 0; s < nsteps; s++) {
		if (pace > 0) {
			cudaEvent_t& e = c->pace_ev[s % 4];
			if (!e) CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
			if (s >= pace) CU_TRY(cudaEventSynchronize(c->pace_ev[(s - pace) % 4]));
			FL_TRY(dev_update(c, dt));
			CU_TRY(cudaEventRecord(e, c->stream));
		} else {
			FL_TRY(dev_update(c, dt));
		}
	}
	CU_TRY(cudaEventRecord(c->ev1, c->stream));
	c->timed = true;
	CU_TRY(cudaGetLastError());
	return FLUID_OK;
}

extern "C" int fluid_ctx_sync(fluid_ctx* c)
{
	CHECK_CTX(c);
	CU_TRY(cudaStreamSynchronize(c->stream));
	CU_TRY(cudaGetLastError());
	if (c->slab_count > 1 || c->rowwave_used) {
		int oob = 0;
		CU_TRY(cudaMemcpy(&oob, c->d_oob, sizeof(int), cudaMemcpyDeviceToHost));
		c->rowwave_used = false;
		if (oob & 4) {
			CU_TRY(cudaMemset(c->d_oob, 0, sizeof(int)));
			return fail(FLUID_ERR_CUDA, "reference-order solve: a warp of the row pipeline waited ~3 s for a neighbour and gave up; results are invalid");
		}
		if (oob & 2) {
			CU_TRY(cudaMemset(c->d_oob, 0, sizeof(int)));
			return fail(FLUID_ERR_CUDA, "halo refresh: a neighbouring slab did not deliver its rows within ~2 s (its process ended, or the slabs "
						    "did not make the same sequence of calls); results are invalid");
		}
		if (oob) {
			CU_TRY(cudaMemset(c->d_oob, 0, sizeof(int)));
			return fail(FLUID_ERR_UNSUPPORTED, "advect: a backtrace reached rows that are not stored on this slab "
							   "(dt*N*|v| exceeded the %d-row halo); results are invalid -- use a deeper slab_halo", c->slab_halo);
		}
	}
	return FLUID_OK;
}

extern "C" int fluid_ctx_halo_refresh(fluid_ctx* c, int nfields, const int* fields)
{
	CHECK_CTX(c);
	if (!fields) return fail(FLUID_ERR_INVALID, "null pointer");
	for (int k = 0; k < nfields; k++) FL_TRY(check_field(fields[k]));
	return dev_halo_refresh(c, nfields, fields);
}

extern "C" int fluid_ctx_set_exchange(fluid_ctx* c, fluid_exchange_fn fn, void* user)
{
	CHECK_CTX(c);
	c->exchange = fn;
	c->exchange_user = user;
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_stream(fluid_ctx* c, void* stream, int exchange_stream_ordered)
{
	CHECK_CTX(c);
	CU_TRY(cudaStreamSynchronize(c->stream));
	c->stream = stream ? (cudaStream_t) stream : c->own_stream;
	c->exchange_stream_ordered = exchange_stream_ordered != 0;
	return FLUID_OK;
}

// Rows a transport must move for one field of a slab: (slab_halo + 1) whole rows of `pitch` floats each way.
extern "C" int fluid_ctx_halo_regions(fluid_ctx* c, int field, float** send_down, float** recv_down, float** send_up, float** recv_up,
				      size_t* count)
{
	CHECK_CTX(c);
	FL_TRY(check_field(field));
	if (!send_down || !recv_down || !send_up || !recv_up || !count) return fail(FLUID_ERR_INVALID, "null pointer");
	*send_down = *recv_down = *send_up = *recv_up = nullptr;
	*count = 0;
	if (c->slab_count <= 1) return FLUID_OK;
	const long long rows = c->slab_halo + 1, P = c->g.pitch;
	*count = (size_t) (rows * P);
	float* f = c->f[field];
	if (!(c->g.walls & 1)) {
		*recv_down = f;                                  // local rows 0 .. halo
		*send_down = f + (long long) c->own_lo * P;      // my first halo+1 owned rows -> lower neighbour's upper halo
	}
	if (!(c->g.walls & 2)) {
		const long long top = c->own_lo + c->slab_rows;  // first row above the owned ones
		*recv_up = f + top * P;                          // local rows top .. top + halo  (= NY + 1)
		*send_up = f + (top - rows) * P;                 // my last halo+1 owned rows -> upper neighbour's lower halo
	}
	return FLUID_OK;
}

extern "C" int fluid_ctx_last_step_ms(fluid_ctx* c, float* ms)
{
	CHECK_CTX(c);
	if (!ms) return fail(FLUID_ERR_INVALID, "null pointer");
	if (!c->timed) return fail(FLUID_ERR_INVALID, "no fluid_ctx_step call to time yet");
	CU_TRY(cudaEventSynchronize(c->ev1));
	CU_TRY(cudaEventElapsedTime(ms, c->ev0, c->ev1));
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_profiling(fluid_ctx* c, int on)
{
	CHECK_CTX(c);
	c->prof_on = on != 0;
	return FLUID_OK;
}

extern "C" int fluid_ctx_profile(fluid_ctx* c, float* ms, unsigned long long* launches, int reset)
{
	CHECK_CTX(c);
	if (!ms || !launches) return fail(FLUID_ERR_INVALID, "null pointer");
	CU_TRY(cudaStreamSynchronize(c->stream));
	for (int k = 0; k < FLUID_PROF_COUNT; k++) { ms[k] = 0; launches[k] = 0; }
	for (auto& sg : c->prof_segs) {
		float t = 0;
		CU_TRY(cudaEventElapsedTime(&t, sg.a, sg.b));
		ms[sg.cls] += t;
		launches[sg.cls] += sg.launches;
	}
	if (reset) {
		for (auto& sg : c->prof_segs) { c->prof_pool.push_back(sg.a); c->prof_pool.push_back(sg.b); }
		c->prof_segs.clear();
	}
	return FLUID_OK;
}

extern "C" int fluid_ctx_add_source(fluid_ctx* c, int x, int s, float dt)
{
	CHECK_CTX(c);
	FL_TRY(check_distinct_n(c->n_fields, { x, s }));
	return dev_add_source(c, c->f[x], c->f[s], dt);
}

extern "C" int fluid_ctx_set_bnd(fluid_ctx* c, int b, int x)
{
	CHECK_CTX(c);
	FL_TRY(check_field(x));
	return dev_set_bnd(c, b, c->f[x]);
}

extern "C" int fluid_ctx_lin_solve(fluid_ctx* c, int b, int x, int x0, float a, float cc)
{
	CHECK_CTX(c);
	FL_TRY(check_distinct_n(c->n_fields, { x, x0 }));
	float* px = c->f[x];
	return dev_lin_solve(c, b, px, c->f[x0], a, cc);
}

extern "C" int fluid_ctx_blur(fluid_ctx* c, int b, int x, int x0, float a, float cc)
{
	CHECK_CTX(c);
	FL_TRY(check_distinct_n(c->n_fields, { x, x0 }));
	return dev_blur(c, b, c->f[x], c->f[x0], a, cc);
}

extern "C" int fluid_ctx_advect(fluid_ctx* c, int b, int d, int d0, int u, int v, float dt)
{
	CHECK_CTX(c);
	FL_TRY(check_field(d0));
	FL_TRY(check_distinct_n(c->n_fields, { d, u, v }));
	if (d == d0) return fail(FLUID_ERR_INVALID, "advect: d and d0 must be different arrays");
	return dev_advect(c, b, c->f[d], c->f[d0], c->f[u], c->f[v], dt);
}

extern "C" int fluid_ctx_advect_velocity(fluid_ctx* c, int u, int v, int u0, int v0, float dt)
{
	CHECK_CTX(c);
	FL_TRY(check_distinct_n(c->n_fields, { u, v, u0, v0 }));
	return dev_advect_velocity(c, c->f[u], c->f[v], c->f[u0], c->f[v0], dt);
}

extern "C" int fluid_ctx_project(fluid_ctx* c, int u, int v, int p, int div)
{
	CHECK_CTX(c);
	FL_TRY(check_distinct_n(c->n_fields, { u, v, p, div }));
	float* pp = c->f[p];
	return dev_project(c, c->f[u], c->f[v], pp, c->f[div]);
}

extern "C" int fluid_ctx_time_kernel(fluid_ctx* c, int which, int reps, float* ms)
{
	CHECK_CTX(c);
	if (!ms || reps < 1) return fail(FLUID_ERR_INVALID, "bad arguments");
	const float a = diffusion_a(c, c->visc > 0 ? c->visc : 1e-4f, 0.1f);
	CU_TRY(cudaEventRecord(c->ev0, c->stream));
	for (int r = 0; r < reps; r++) {
		// (a red-black solve may move a field to another array, so the table is read afresh every time)
		float *U = c->f[FLUID_U], *V = c->f[FLUID_V], *D = c->f[FLUID_DENS];
		float *U0 = c->f[FLUID_U_PREV], *V0 = c->f[FLUID_V_PREV], *D0 = c->f[FLUID_DENS_PREV];
		switch (which) {
		case FLUID_K_LIN_SOLVE: FL_TRY(dev_lin_solve(c, 1, U0, U, a, 1 + 4 * a)); break;
		case FLUID_K_BLUR: FL_TRY(dev_blur(c, 0, D0, D, a, 1 + 4 * a)); break;
		case FLUID_K_ADVECT: FL_TRY(dev_advect(c, 0, D0, D, U, V, 0.1f)); break;
		case FLUID_K_ADVECT_VEL: FL_TRY(dev_advect_velocity(c, U0, V0, U, V, 0.1f)); break;
		case FLUID_K_PROJECT: FL_TRY(dev_project(c, U, V, U0, V0)); break;
		case FLUID_K_ADD_SOURCE: FL_TRY(dev_add_source(c, D0, D, 0.1f)); break;
		default: return fail(FLUID_ERR_INVALID, "unknown kernel id %d", which);
		}
	}
	CU_TRY(cudaEventRecord(c->ev1, c->stream));
	CU_TRY(cudaEventSynchronize(c->ev1));
	float t = 0;
	CU_TRY(cudaEventElapsedTime(&t, c->ev0, c->ev1));
	*ms = t / reps;
	return FLUID_OK;
}
