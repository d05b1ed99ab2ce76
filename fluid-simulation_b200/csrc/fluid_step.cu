// The streaming and gather kernels of the Stable-Fluids step and their composition:
// VelStep1 (FluidSystem.cpp:354-376), DensStep1 (:289-299), Update (:101-128), the headless render.
#include "ctx.h"
#include "kernels_basic.cuh"
#include "kernels_advect.cuh"

#include <tuple>

int dev_add_source(fluid_ctx* c, float* x, const float* s, float dt)
{
	ProfScope ps(c, FLUID_PROF_ADD_SOURCE);
	LAUNCH(k_add_source, flat_blocks(c), 256, 0, c->stream, (float4*) x, (const float4*) s, dt, n4_of(c));
	return FLUID_OK;
}

int dev_add_source2(fluid_ctx* c, float* x, const float* s, float* y, const float* t, float dt)
{
	ProfScope ps(c, FLUID_PROF_ADD_SOURCE);
	LAUNCH(k_add_source2, flat_blocks(c), 256, 0, c->stream, (float4*) x, (const float4*) s, (float4*) y,
	       (const float4*) t, dt, n4_of(c));
	return FLUID_OK;
}

int dev_set_bnd(fluid_ctx* c, int b, float* x)
{
	const int n = std::max(c->g.N, c->g.NY);
	LAUNCH(k_set_bnd, dim3(cdiv(n, 256), 1, c->batch), 256, 0, c->stream, c->g, b, x);
	return FLUID_OK;
}

int dev_zero_bnd(fluid_ctx* c, float* x)
{
	const int n = std::max(c->g.N, c->g.NY) + 2;
	LAUNCH(k_zero_bnd, dim3(cdiv(n, 256), 1, c->batch), 256, 0, c->stream, c->g, x);
	return FLUID_OK;
}
// Which advect kernel: the row-window kernel (kernels_advect.cuh) once the grid is large enough to give every
// SM several warp tasks, the one-thread-per-cell gather kernel below that.  Bit-identical either way.
static bool advect_rows_geometry(const fluid_ctx* c, AdvectTasks* tk)
{
	if (c->advect_kernel == 1) return false;
	if (c->advect_kernel == 0 && (long long) c->g.N * c->g.NY * c->batch < (1LL << 20)) return false;
	tk->nstrips = cdiv(c->g.N, 128);
	const long long slots = (long long) c->num_sms * 16;
	int rows = 32;
	while (rows > 8 && (long long) tk->nstrips * cdiv(c->g.NY, rows) * c->batch < 2 * slots) rows /= 2;
	tk->chunk_rows = rows;
	tk->nchunks = cdiv(c->g.NY, rows);
	return true;
}

template <bool VEL, bool PARTIAL>
static void launch_advect_rows_p(fluid_ctx* c, const AdvectTasks& tk, int b, float* oa, float* ob, const float* d0, const float* u, const float* v,
				 float dtN)
{
	const dim3 grid(cdiv((long long) tk.nstrips * tk.nchunks, kAdvectWarps), c->batch);
	const int nt = 32 * kAdvectWarps;
	const bool euler = (c->advect_mode == FLUID_ADVECT_EULER);
	const int lo = c->own_lo, hi = c->own_lo + c->slab_rows - 1;
	if (c->slab_count > 1) {
		if (euler) LAUNCH((k_advect_rows<true, 1, VEL, PARTIAL>), grid, nt, 0, c->stream, c->g, tk, b, oa, ob, d0, u, v, dtN, c->d_oob, lo, hi);
		else LAUNCH((k_advect_rows<true, 0, VEL, PARTIAL>), grid, nt, 0, c->stream, c->g, tk, b, oa, ob, d0, u, v, dtN, c->d_oob, lo, hi);
	} else {
		if (euler) LAUNCH((k_advect_rows<false, 1, VEL, PARTIAL>), grid, nt, 0, c->stream, c->g, tk, b, oa, ob, d0, u, v, dtN, c->d_oob, 1, 0);
		else LAUNCH((k_advect_rows<false, 0, VEL, PARTIAL>), grid, nt, 0, c->stream, c->g, tk, b, oa, ob, d0, u, v, dtN, c->d_oob, 1, 0);
	}
}

template <bool VEL>
static void launch_advect_rows(fluid_ctx* c, const AdvectTasks& tk, int b, float* oa, float* ob, const float* d0, const float* u, const float* v,
			       float dtN)
{
	if (c->g.N % 128 == 0) launch_advect_rows_p<VEL, false>(c, tk, b, oa, ob, d0, u, v, dtN); // whole strips: no masking of cells
	else launch_advect_rows_p<VEL, true>(c, tk, b, oa, ob, d0, u, v, dtN);
}

int dev_advect(fluid_ctx* c, int b, float* d, const float* d0, const float* u, const float* v, float dt)
{
	ProfScope ps(c, FLUID_PROF_ADVECT);
	const float dtN = dt * c->g.N;
	const bool euler = (c->advect_mode == FLUID_ADVECT_EULER);
	if (c->fs_type != FLUID_TYPE_ORIGINAL) b = kBndNone; // Advect1B (FluidSystem.cpp:864-874): no SetBorder
	AdvectTasks tk;
	if (advect_rows_geometry(c, &tk)) {
		launch_advect_rows<false>(c, tk, b, d, nullptr, d0, u, v, dtN);
		return FLUID_OK;
	}
	if (c->slab_count > 1) {
		if (euler) LAUNCH((k_advect<true, 1>), cell_grid(c), kBlk2D, 0, c->stream, c->g, b, d, d0, u, v, dtN, c->d_oob, c->own_lo, c->own_lo + c->slab_rows - 1);
		else LAUNCH((k_advect<true, 0>), cell_grid(c), kBlk2D, 0, c->stream, c->g, b, d, d0, u, v, dtN, c->d_oob, c->own_lo, c->own_lo + c->slab_rows - 1);
	} else {
		if (euler) LAUNCH((k_advect<false, 1>), cell_grid(c), kBlk2D, 0, c->stream, c->g, b, d, d0, u, v, dtN, c->d_oob, 1, 0);
		else LAUNCH((k_advect<false, 0>), cell_grid(c), kBlk2D, 0, c->stream, c->g, b, d, d0, u, v, dtN, c->d_oob, 1, 0);
	}
	return FLUID_OK;
}

int dev_advect_velocity(fluid_ctx* c, float* u, float* v, const float* u0, const float* v0, float dt)
{
	ProfScope ps(c, FLUID_PROF_ADVECT);
	const float dtN = dt * c->g.N;
	const bool euler = (c->advect_mode == FLUID_ADVECT_EULER);
	const bool nobnd = c->fs_type != FLUID_TYPE_ORIGINAL;
	const int bu = nobnd ? kBndNone : 1, bv = nobnd ? kBndNone : 2;
	AdvectTasks tk;
	if (advect_rows_geometry(c, &tk)) {
		launch_advect_rows<true>(c, tk, nobnd ? kBndNone : 0, u, v, nullptr, u0, v0, dtN);
		return FLUID_OK;
	}
	if (c->slab_count > 1) {
		if (euler) LAUNCH((k_advect_velocity<true, 1>), cell_grid(c), kBlk2D, 0, c->stream, c->g, u, v, u0, v0, dtN, c->d_oob, c->own_lo, c->own_lo + c->slab_rows - 1, bu, bv);
		else LAUNCH((k_advect_velocity<true, 0>), cell_grid(c), kBlk2D, 0, c->stream, c->g, u, v, u0, v0, dtN, c->d_oob, c->own_lo, c->own_lo + c->slab_rows - 1, bu, bv);
	} else {
		if (euler) LAUNCH((k_advect_velocity<false, 1>), cell_grid(c), kBlk2D, 0, c->stream, c->g, u, v, u0, v0, dtN, c->d_oob, 1, 0, bu, bv);
		else LAUNCH((k_advect_velocity<false, 0>), cell_grid(c), kBlk2D, 0, c->stream, c->g, u, v, u0, v0, dtN, c->d_oob, 1, 0, bu, bv);
	}
	return FLUID_OK;
}

// Project (FluidSystem.cpp:982-1007).
int dev_project(fluid_ctx* c, float* u, float* v, float*& p, float* div)
{
	if (c->fs_type == FLUID_TYPE_ORIGINAL_BORDERED_MF) return dev_project_bo(c, u, v, div); // (p_old is ignored, as in the reference)
	if (c->fs_type != FLUID_TYPE_ORIGINAL) return dev_project_b(c, u, v, p, div);
	// p = 0 is the solve's starting guess.  Where the first pass of the solve can simply be told so, the zeros
	// are neither written here nor read there (2 x 268 MB at N = 8192).
	const int limit = c->gs_iters; // a = 1 > 0
	const bool skip_zero = !use_small_kernels(c) && c->gs_order == FLUID_GS_REDBLACK && c->rb_tile_sweeps > 0 && limit >= 2 &&
			       solve_accepts_zero_x(c, 1.0f, 4.0f, limit);
	{
		ProfScope ps(c, FLUID_PROF_PROJECT_AUX);
		LAUNCH(k_divergence, rowgroup_grid(c, c->g.NY + 2), kBlk2D, 0, c->stream, c->g, u, v, div, skip_zero ? (float*) nullptr : p);
	}
	FL_TRY(dev_lin_solve(c, 0, p, div, 1.0f, 4.0f, skip_zero));
	{
		ProfScope ps(c, FLUID_PROF_PROJECT_AUX);
		LAUNCH(k_subtract_gradient, rowgroup_grid(c, c->g.NY), kBlk2D, 0, c->stream, c->g, u, v, p);
	}
	return FLUID_OK;
}
// Halo exchange between row slabs.  The slabs overlap by slab_halo (+1) rows and every
// kernel computes on all stored rows, so stale values creep in from the outer
// halo edge by one row per stencil application; the fields named here are made
// valid again over the full halo depth by asking the host's transport (NCCL
// send/recv, peer copies ...) to refill the halo rows from the neighbours' owned
// rows.  With a 48-row halo that is needed once per 20-sweep solve instead of
// once per colour pass: 9 exchanges of ~3 MB per step instead of ~200 of 64 KB.
static int halo_exchange(fluid_ctx* c, std::initializer_list<const float*> arrays);

// Independent parts of a frame side by side on extra streams?  Not for ensembles: their solves run on the persistent
// tile kernel (one CTA per SM, 200 KB of shared memory each), two of which cannot share an SM -- measured on 512 members
// of N=512: 41.5 ms per frame with the extra streams, 36.5 ms without.
static inline bool overlap_now(const fluid_ctx* c) { return c->overlap_density && c->batch == 1; }

// One halo refresh of the listed fields, on its own (tests and transport measurements).
int dev_halo_refresh(fluid_ctx* c, int n, const int* fields)
{
	if (n == 1) return halo_exchange(c, { c->f[fields[0]] });
	if (n == 2) return halo_exchange(c, { c->f[fields[0]], c->f[fields[1]] });
	return fail(FLUID_ERR_INVALID, "a refresh carries 1 or 2 fields");
}

static int halo_exchange(fluid_ctx* c, std::initializer_list<const float*> arrays)
{
	if (c->slab_count <= 1) return FLUID_OK;
	if (!c->exchange && !c->peer_connected)
		return fail(FLUID_ERR_INVALID, "this context is a row slab but no halo transport is set (fluid_ctx_set_exchange or fluid_ctx_peer_halo_connect)");
	int ids[fluid_ctx::kMaxFields], n = 0;
	for (const float* a : arrays)
		for (int f = 0; f < c->n_fields; f++)
			if (c->f[f] == a) ids[n++] = f;
	// A refresh is short and its two ends wait for each other: nothing else may be queued on this GPU that could keep
	// its kernels from starting at once.  The densities' diffusion shares the GPU with the velocity diffusion up to
	// here (one-wave streaming launches that leave no room for anything else while they run) and is joined now.
	if (c->side_join_pending) {
		CU_TRY(cudaStreamWaitEvent(c->stream, c->ev_join, 0));
		c->side_join_pending = false;
	}
	ProfScope ps(c, FLUID_PROF_EXCHANGE);
	if (c->peer_connected) return peer_halo_exchange(c, n, ids); // pushes through peer memory, two kernels on this stream
	if (!c->exchange_stream_ordered) CU_TRY(cudaStreamSynchronize(c->stream)); // the transport reads the owned rows
	const int rc = c->exchange(c->exchange_user, n, ids);
	if (rc != 0) return fail(FLUID_ERR_CUDA, "halo exchange callback failed with code %d", rc);
	return FLUID_OK;
}

// VelStep1 (FluidSystem.cpp:354-376).  The four pointers are locals: the
// reference's two Swap() calls per FlipFlopArr become local pointer swaps, so
// the result lands in the arrays passed as u and v.
int dev_vel_step(fluid_ctx* c, float* u, float* v, float* u0, float* v0, float visc, float dt, bool sources_added)
{
	nvtxRangePushA("fluid:VelStep");
	struct Pop { ~Pop() { nvtxRangePop(); } } pop_on_return;
	// slabs: u, v (and the source arrays, which every slab fills itself) are valid over the whole halo on entry
	if (c->fs_type != FLUID_TYPE_ORIGINAL) {
		// VelStep1B (FluidSystem.cpp:378-398) is VelStep1 over the masked primitives, except that its first AddArrDt
		// adds the Y source to vX (:380, AddArrDt(N, vX.Curr(), vY.Prev(), dt)): reproduced as the reference has it
		FL_TRY(dev_add_source2(c, u, v0, v, v0, dt));
	} else if (!sources_added) {
		FL_TRY(dev_add_source2(c, u, u0, v, v0, dt)); // (Update() may already have added its sparse forcing)
	}
	std::swap(u, u0);
	const float a = diffusion_a(c, visc, dt);
	std::swap(v, v0);
	const bool mf = (c->fs_type == FLUID_TYPE_ORIGINAL_BORDERED_MF);
	// VelStep1BSO: the viscosity is mixed per cell from the densities as they are BEFORE this frame's DensStep (Diffuse1BSO,
	// :828-831).  Update() has computed the coefficients already, before it let the densities' own step start on the side stream.
	if (mf && !c->mf_coeffs_ready) FL_TRY(dev_mf_coeffs(c, dt));
	c->mf_coeffs_ready = false;
	auto diffuse = [&](int b, float*& x, const float* x0) -> int {
		return mf ? dev_lin_solve_bso(c, b, x, x0) : dev_lin_solve(c, b, x, x0, a, 1 + 4 * a);
	};
	if (overlap_now(c)) { // the two diffusions are independent: v on a second lane beside u
		if (!c->lane_stream) {
			CU_TRY(cudaStreamCreateWithFlags(&c->lane_stream, cudaStreamNonBlocking));
			CU_TRY(cudaEventCreateWithFlags(&c->ev_fork2, cudaEventDisableTiming));
			CU_TRY(cudaEventCreateWithFlags(&c->ev_join2, cudaEventDisableTiming));
		}
		CU_TRY(cudaEventRecord(c->ev_fork2, c->stream));
		CU_TRY(cudaStreamWaitEvent(c->lane_stream, c->ev_fork2, 0));
		cudaStream_t main_stream = c->stream;
		c->stream = c->lane_stream;
		c->lane = 1;
		const int rc = diffuse(2, v, v0);
		c->lane = 0;
		c->stream = main_stream;
		FL_TRY(rc);
		CU_TRY(cudaEventRecord(c->ev_join2, c->lane_stream));
		FL_TRY(diffuse(1, u, u0));
		CU_TRY(cudaStreamWaitEvent(c->stream, c->ev_join2, 0));
	} else {
		FL_TRY(diffuse(1, u, u0));
		FL_TRY(diffuse(2, v, v0));
	}
	FL_TRY(halo_exchange(c, { u, v }));
	FL_TRY(dev_project(c, u, v, u0, v0));
	FL_TRY(halo_exchange(c, { u, v }));
	std::swap(u, u0);
	std::swap(v, v0);
	FL_TRY(dev_advect_velocity(c, u, v, u0, v0, dt));
	FL_TRY(halo_exchange(c, { u, v }));
	FL_TRY(dev_project(c, u, v, u0, v0));
	FL_TRY(halo_exchange(c, { u, v }));
	return FLUID_OK;
}

// DensStep1 (FluidSystem.cpp:289-299), one fluid.
// First half (sources and diffusion: no velocity involved; the blurred field ends in x0) and second half (advection).
static int dev_dens_diffuse(fluid_ctx* c, float* x, float* x0, float diff, float dt, bool sources_added = false)
{
	if (!sources_added) FL_TRY(dev_add_source(c, x, x0, dt));
	const float a = diffusion_a(c, diff, dt);
	return dev_blur(c, 0, x0, x, a, 1 + 4 * a); // after the reference's Swap: x0 is relaxed, x holds the sum
}

static int dev_dens_advect(fluid_ctx* c, float* x, float* x0, const float* u, const float* v, float dt)
{
	FL_TRY(halo_exchange(c, { x0 })); // 21 blur passes consumed 21 halo rows of the field advect samples
	FL_TRY(dev_advect(c, 0, x, x0, u, v, dt));
	FL_TRY(halo_exchange(c, { x }));
	return FLUID_OK;
}

int dev_dens_step(fluid_ctx* c, float* x, float* x0, const float* u, const float* v, float diff, float dt)
{
	FL_TRY(dev_dens_diffuse(c, x, x0, diff, dt));
	return dev_dens_advect(c, x, x0, u, v, dt);
}

static int ensure_vort_scratch(fluid_ctx* c)
{
	for (int k = 0; k < 4; k++) {
		if (!c->vscr[k]) {
			CU_TRY(cudaMalloc(&c->vscr[k], c->field_bytes));
			CU_TRY(cudaMemsetAsync(c->vscr[k], 0, c->field_bytes, c->stream));
		}
	}
	return FLUID_OK;
}

// VortConfinement (FluidSystem.cpp:1124-1175).
int dev_vort_confinement(fluid_ctx* c, float* u, float* v, float vorticity)
{
	FL_TRY(ensure_vort_scratch(c));
	ProfScope ps(c, FLUID_PROF_OTHER);
	LAUNCH(k_vort_curl, cell_grid(c), kBlk2D, 0, c->stream, c->g, u, v, c->vscr[0], c->vscr[1]);
	LAUNCH(k_vort_dir, cell_grid(c), kBlk2D, 0, c->stream, c->g, c->vscr[1], c->vscr[2], c->vscr[3]);
	LAUNCH(k_vort_apply, cell_grid(c), kBlk2D, 0, c->stream, c->g, u, v, c->vscr[0], c->vscr[2], c->vscr[3], vorticity);
	return FLUID_OK;
}

int dev_damp(fluid_ctx* c, float* vel, float amount, float dt)
{
	const float f = powf(1 - amount, dt); // host libm, as the reference (FluidSystem.cpp:1181)
	ProfScope ps(c, FLUID_PROF_OTHER);
	LAUNCH(k_scale, flat_blocks(c), 256, 0, c->stream, (float4*) vel, f, n4_of(c));
	return FLUID_OK;
}

int sync_sources(fluid_ctx* c)
{
	if (!c->src_dirty) return FLUID_OK;
	std::vector<SourceDev> flat;
	std::vector<int> first(c->batch + 1, 0);
	for (int s = 0; s < c->batch; s++) {
		first[s] = (int) flat.size();
		flat.insert(flat.end(), c->src[s].begin(), c->src[s].end());
	}
	first[c->batch] = (int) flat.size();
	// PREV_SET entries: a later entry on the same cell overwrites an earlier one; the survivor is marked
	for (int s = 0; s < c->batch; s++) {
		std::map<std::tuple<int, int, int>, int> seen;
		for (int k = first[s + 1] - 1; k >= first[s]; k--) {
			SourceDev& e = flat[k];
			e.last = (e.mode == 0 && seen.emplace(std::make_tuple(e.field, e.i, e.j), k).second) ? 1 : 0;
		}
	}
	if (c->d_src) CU_TRY(cudaFree(c->d_src));
	c->d_src = nullptr;
	if (c->d_src_saved) CU_TRY(cudaFree(c->d_src_saved));
	c->d_src_saved = nullptr;
	if (!c->d_first) CU_TRY(cudaMalloc(&c->d_first, sizeof(int) * (c->batch + 1)));
	if (!flat.empty()) {
		CU_TRY(cudaMalloc(&c->d_src, sizeof(SourceDev) * flat.size()));
		CU_TRY(cudaMemcpy(c->d_src, flat.data(), sizeof(SourceDev) * flat.size(), cudaMemcpyHostToDevice));
		CU_TRY(cudaMalloc(&c->d_src_saved, sizeof(float) * flat.size()));
	}
	CU_TRY(cudaMemcpy(c->d_first, first.data(), sizeof(int) * first.size(), cudaMemcpyHostToDevice));
	c->n_src = (int) flat.size();
	c->src_dirty = false;
	return FLUID_OK;
}
// FluidSystem::Update (FluidSystem.cpp:101-128) for FSType::ORIGINAL.
// FluidSystem::Update (FluidSystem.cpp:101-128): forcing, [vorticity], VelStep, DensStep per fluid, [damping].
// The densities' add_source + Blur touch neither velocity array, so on a single-GPU context they are enqueued on
// a side stream (fork after the forcing, join before the advection) and share the GPU with VelStep.
// `dens_ready`: an event (of another stream) after which the current density arrays are valid, or nullptr;
// `after_velocity`: called once VelStep has been enqueued on c->stream, or empty.  fluid_update() uses both to
// slip its bus transfers in; everybody else calls dev_update.
struct NvtxRange {
	explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
	~NvtxRange() { nvtxRangePop(); }
};

int dev_update_hooks(fluid_ctx* c, float dt, cudaEvent_t dens_ready, const std::function<int()>& after_velocity)
{
	NvtxRange frame("fluid:Update");
	c->epoch++; // a frame (the divisor bookkeeping of division_for counts in frames and API calls)
	// The source arrays of a frame are zero except for the listed PREV_SET cells, so AddArrDt over them is known
	// in advance (k_canon_zero, kernels_basic.cuh): the three dense x += dt*s passes and the zero fill of the
	// densities' source arrays (which nothing else reads: Blur overwrites them) are left out.  Small grids keep
	// the plain sequence (they are bound by launches, not bytes); so does a dt whose product with zero is NaN.
	const bool sparse = c->sparse_sources && !use_small_kernels(c) && isfinite(dt) && c->fs_type == FLUID_TYPE_ORIGINAL;
	FieldPtrs fp = {};
	fp.cur[0] = c->f[FLUID_U];
	fp.prev[0] = c->f[FLUID_U_PREV];
	fp.cur[1] = c->f[FLUID_V];
	fp.prev[1] = c->f[FLUID_V_PREV];
	for (int fl = 0; fl < c->n_fluids; fl++) {
		fp.cur[2 + fl] = c->f[FLUID_DENS_OF(fl)];
		fp.prev[2 + fl] = c->f[FLUID_DENS_PREV_OF(fl)];
	}
	{
		ProfScope ps(c, FLUID_PROF_OTHER);
		CU_TRY(cudaMemsetAsync(c->f[FLUID_U_PREV], 0, c->field_bytes, c->stream)); // the diffusion solves start from these
		CU_TRY(cudaMemsetAsync(c->f[FLUID_V_PREV], 0, c->field_bytes, c->stream));
		if (!sparse)
			for (int fl = 0; fl < c->n_fluids; fl++) CU_TRY(cudaMemsetAsync(c->f[FLUID_DENS_PREV_OF(fl)], 0, c->field_bytes, c->stream));
		if (c->n_src > 0) LAUNCH(k_apply_sources, cdiv(c->batch, 64), 64, 0, c->stream, c->g, fp, c->d_src, c->d_first, c->batch);
	}
	// x += dt*s for one group of fields, sparse form, on the current stream
	auto add_sparse = [&](float* x, float* y, int slot_lo, int slot_hi) -> int {
		ProfScope ps(c, FLUID_PROF_ADD_SOURCE);
		const dim3 grid(cdiv(c->batch, 64));
		if (c->n_src > 0)
			LAUNCH(k_add_source_cells, grid, 64, 0, c->stream, c->g, fp, c->d_src, c->d_first, c->batch, c->d_src_saved, dt, slot_lo, slot_hi, 0);
		if (dt >= 0.0f) {
			LAUNCH(k_canon_zero, flat_blocks(c), 256, 0, c->stream, (float4*) x, (float4*) y, n4_of(c));
			if (c->n_src > 0)
				LAUNCH(k_add_source_cells, grid, 64, 0, c->stream, c->g, fp, c->d_src, c->d_first, c->batch, c->d_src_saved, dt, slot_lo, slot_hi, 1);
		}
		return FLUID_OK;
	};
	if (c->fs_type == FLUID_TYPE_ORIGINAL_BORDERED_MF) { // reads the densities: before their side-stream step may touch them
		FL_TRY(dev_mf_coeffs(c, dt));
		c->mf_coeffs_ready = true;
	}
	const bool concurrent = overlap_now(c); // slabs too: the halo exchanges of the densities come after the join
	auto diffuse_all = [&]() -> int {
		for (int fl = 0; fl < c->n_fluids; fl++) // DensStep1's loop over fluid_count (FluidSystem.cpp:291)
		{
			if (sparse) FL_TRY(add_sparse(c->f[FLUID_DENS_OF(fl)], nullptr, 2 + fl, 3 + fl));
			FL_TRY(dev_dens_diffuse(c, c->f[FLUID_DENS_OF(fl)], c->f[FLUID_DENS_PREV_OF(fl)], fluid_diff(c, fl), dt, sparse));
		}
		return FLUID_OK;
	};
	if (concurrent) {
		if (!c->side_stream) {
			CU_TRY(cudaStreamCreateWithFlags(&c->side_stream, cudaStreamNonBlocking));
			CU_TRY(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
			CU_TRY(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
		}
		CU_TRY(cudaEventRecord(c->ev_fork, c->stream));
		CU_TRY(cudaStreamWaitEvent(c->side_stream, c->ev_fork, 0));
		if (dens_ready) CU_TRY(cudaStreamWaitEvent(c->side_stream, dens_ready, 0));
		cudaStream_t main_stream = c->stream;
		c->stream = c->side_stream; // every launch helper and profiling scope follows c->stream
		const int rc = diffuse_all();
		c->stream = main_stream;
		FL_TRY(rc);
		CU_TRY(cudaEventRecord(c->ev_join, c->side_stream));
		c->side_join_pending = true;
	}
	if (c->vorticity != 0.0f) {
		FL_TRY(dev_vort_confinement(c, c->f[FLUID_U], c->f[FLUID_V], c->vorticity));
		FL_TRY(halo_exchange(c, { c->f[FLUID_U], c->f[FLUID_V] })); // three stencil phases consumed 3 halo rows
	}
	if (sparse) FL_TRY(add_sparse(c->f[FLUID_U], c->f[FLUID_V], 0, 2));
	FL_TRY(dev_vel_step(c, c->f[FLUID_U], c->f[FLUID_V], c->f[FLUID_U_PREV], c->f[FLUID_V_PREV], visc_avg(c), dt, sparse));
	if (after_velocity) FL_TRY(after_velocity());
	if (concurrent) {
		if (c->side_join_pending) CU_TRY(cudaStreamWaitEvent(c->stream, c->ev_join, 0));
		c->side_join_pending = false;
	} else {
		if (dens_ready) CU_TRY(cudaStreamWaitEvent(c->stream, dens_ready, 0));
		FL_TRY(diffuse_all());
	}
	for (int fl = 0; fl < c->n_fluids; fl++)
		FL_TRY(dev_dens_advect(c, c->f[FLUID_DENS_OF(fl)], c->f[FLUID_DENS_PREV_OF(fl)], c->f[FLUID_U], c->f[FLUID_V], dt));
	if (c->damp != 0.0f) {
		FL_TRY(dev_damp(c, c->f[FLUID_U], c->damp, dt));
		FL_TRY(dev_damp(c, c->f[FLUID_V], c->damp, dt));
	}
	return FLUID_OK;
}

int dev_update(fluid_ctx* c, float dt) { return dev_update_hooks(c, dt, nullptr, {}); }

// FluidRenderer::DoBuffer without a window (color_rgb: 3 floats per fluid): colourise + velocity lines of ensemble member `sim` into
// device buffers, then (optionally) to dense host arrays.
int dev_render(fluid_ctx* c, int sim, const float* color_rgb, float* host_pixels, float* host_lines)
{
	if (!color_rgb || !host_pixels) return fail(FLUID_ERR_INVALID, "null pointer");
	if (sim < 0 || sim >= c->batch) return fail(FLUID_ERR_INVALID, "sim %d out of range [0,%d)", sim, c->batch);
	if (c->slab_count > 1) return fail(FLUID_ERR_UNSUPPORTED, "render a slab by downloading it; the display buffer is per whole grid");
	const size_t cells = (size_t) (c->g.N + 2) * (c->g.NY + 2);
	if (!c->d_pixels) CU_TRY(cudaMalloc(&c->d_pixels, cells * 3 * sizeof(float)));
	if (host_lines && !c->d_lines) CU_TRY(cudaMalloc(&c->d_lines, cells * 4 * sizeof(float)));
	const long long off = (long long) sim * c->g.plane;
	RenderFluids rf = {};
	rf.n = c->n_fluids;
	for (int fl = 0; fl < c->n_fluids; fl++) {
		rf.dens[fl] = c->f[FLUID_DENS_OF(fl)] + off;
		for (int k = 0; k < 3; k++) rf.col[fl][k] = color_rgb[3 * fl + k];
	}
	LAUNCH(k_colourise, dim3(cdiv(c->g.N + 2, 32), cdiv(c->g.NY + 2, 8), 1), kBlk2D, 0, c->stream, c->g, rf, c->d_info, c->f[FLUID_U] + off,
	       c->f[FLUID_V] + off, c->d_pixels, host_lines ? c->d_lines : nullptr);
	CU_TRY(cudaMemcpyAsync(host_pixels, c->d_pixels, cells * 3 * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
	if (host_lines) CU_TRY(cudaMemcpyAsync(host_lines, c->d_lines, cells * 4 * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
	CU_TRY(cudaStreamSynchronize(c->stream));
	return FLUID_OK;
}
