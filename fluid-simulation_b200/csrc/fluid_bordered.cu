// FSType::ORIGINAL_BORDERED on a resident context (FluidSystem.cpp:79-82: VelStep1B + DensStep1BS): the cell
// map and the masked primitives.  A context created with fluid_params.fs_type = FLUID_TYPE_ORIGINAL_BORDERED
// routes lin_solve, Blur, project and advect through these, exactly as SetType swaps the reference's step
// functions; VelStep/DensStep/Update (fluid_step.cu) are the same sequences over them.
#include "ctx.h"
#include "kernels_bordered.cuh"

static const dim3 kBlkRow(256, 1);
static inline dim3 cells_grid(const fluid_ctx* c, int extra) { return dim3(cdiv(c->g.N + extra, 256), c->g.NY + extra, c->batch); }

int bordered_alloc(fluid_ctx* c)
{
	const size_t bytes = (size_t) c->g.plane;
	CU_TRY(cudaMalloc(&c->d_info, bytes));
	CU_TRY(cudaMemsetAsync(c->d_info, 0, bytes, c->stream));
	return bordered_reset(c);
}

// ResetCellInfo + CalcBorderFromContent (FluidSystem.cpp:1215-1232, 1185-1213): what Update does to the map
// every frame before the borders and colliders have their say (:112-115).
int bordered_reset(fluid_ctx* c)
{
	LAUNCH(k_cell_info_reset, dim3(cdiv(c->g.N + 2, 256), c->g.NY + 2), 256, 0, c->stream, c->g, c->d_info);
	LAUNCH(k_cell_info_calc, dim3(cdiv(c->g.N, 256), c->g.NY), 256, 0, c->stream, c->g, c->d_info);
	return FLUID_OK;
}

int bordered_calc(fluid_ctx* c)
{
	LAUNCH(k_cell_info_calc, dim3(cdiv(c->g.N, 256), c->g.NY), 256, 0, c->stream, c->g, c->d_info);
	return FLUID_OK;
}

int bordered_disable_edges(fluid_ctx* c, int n, const int* edges)
{
	if (n == 0) return FLUID_OK;
	int* d = nullptr;
	CU_TRY(cudaMalloc(&d, sizeof(int) * 4 * (size_t) n));
	cudaError_t e = cudaMemcpyAsync(d, edges, sizeof(int) * 4 * (size_t) n, cudaMemcpyHostToDevice, c->stream);
	if (e == cudaSuccess) {
		LAUNCH(k_cell_info_disable_edges, cdiv(n, 256), 256, 0, c->stream, c->g, c->d_info, d, n);
		e = cudaStreamSynchronize(c->stream); // `edges` is the caller's memory and d is freed below
	}
	cudaFree(d);
	if (e != cudaSuccess) return fail(FLUID_ERR_CUDA, "disable_edges: %s", cudaGetErrorString(e));
	return FLUID_OK;
}

// The masked relaxations in the context's order: MODE_B LinearSolve1B (:550-578), MODE_BSO LinearSolve1BSO (:673-720, per-cell
// coefficient arrays ab / cb, always gs_iters sweeps), MODE_BO LinearSolve1BO (:580-615).
template <int MODE>
static int lin_solve_masked(fluid_ctx* c, int b, float* x, const float* x0, float a, float cc, const float* ab, const float* cb)
{
	ProfScope ps(c, FLUID_PROF_LIN_SOLVE);
	const int limit = (MODE == MODE_BSO) ? c->gs_iters : (a > 0 ? c->gs_iters : 1);
	if (c->gs_order == FLUID_GS_REDBLACK) {
		const dim3 grid(cdiv((c->g.N + 1) / 2, 256), c->g.NY, c->batch);
		for (int k = 0; k < limit; k++) {
			LAUNCH(k_rb_half_b<MODE>, grid, 256, 0, c->stream, c->g, b, x, x0, c->d_info, a, cc, ab, cb, 0);
			LAUNCH(k_rb_half_b<MODE>, grid, 256, 0, c->stream, c->g, b, x, x0, c->d_info, a, cc, ab, cb, 1);
		}
	} else {
		// tiles of 16 or 32 cells (the masked kernels stage up to five arrays per tile in static shared memory)
		const int T = (c->wavefront_tile > 0) ? std::min(c->wavefront_tile, 32) : (c->g.N < 2048 ? 16 : 32);
		const int ntx = cdiv(c->g.N, T), nty = cdiv(c->g.NY, T);
		const int last = (ntx - 1) + (nty - 1) + 2 * (limit - 1);
		for (int s = 0; s <= last; s++) {
			if (T == 16) LAUNCH((k_gs_wavefront_b<16, MODE>), dim3(ntx, limit, c->batch), 16, 0, c->stream, c->g, b, x, x0, c->d_info, a, cc, ab, cb, s, ntx, nty);
			else LAUNCH((k_gs_wavefront_b<32, MODE>), dim3(ntx, limit, c->batch), 32, 0, c->stream, c->g, b, x, x0, c->d_info, a, cc, ab, cb, s, ntx, nty);
		}
	}
	CU_TRY(cudaGetLastError());
	return FLUID_OK;
}

int dev_lin_solve_b(fluid_ctx* c, int b, float* x, const float* x0, float a, float cc)
{
	return lin_solve_masked<MODE_B>(c, b, x, x0, a, cc, nullptr, nullptr);
}

// ---- FSType::ORIGINAL_BORDERED_MF (VelStep1BSO, FluidSystem.cpp:424-447) ----
static int mf_scratch(fluid_ctx* c)
{
	for (float** p : { &c->mf_a, &c->mf_c, &c->mf_p }) {
		if (!*p) {
			CU_TRY(cudaMalloc(p, c->field_bytes));
			CU_TRY(cudaMemsetAsync(*p, 0, c->field_bytes, c->stream));
		}
	}
	return FLUID_OK;
}

// The per-cell (a, c) of LinearSolve1BSO for this velocity step: Diffuse1BSO (:828-831) passes aSc = dt*N*N, the
// fallback visc_air and the fluids' viscosities; the densities are the CURRENT ones.
int dev_mf_coeffs(fluid_ctx* c, float dt)
{
	FL_TRY(mf_scratch(c));
	MfFluids fl = {};
	fl.n = c->n_fluids;
	for (int k = 0; k < c->n_fluids; k++) {
		fl.dens[k] = c->f[FLUID_DENS_OF(k)];
		fl.coef[k] = c->fl_set[k] ? c->fl_visc[k] : c->visc;
	}
	const float aSc = dt * c->g.N * c->g.N;
	ProfScope ps(c, FLUID_PROF_OTHER);
	LAUNCH(k_mf_coeffs, cells_grid(c, 0), kBlkRow, 0, c->stream, c->g, fl, aSc, c->visc_air, c->mf_a, c->mf_c);
	return FLUID_OK;
}

int dev_lin_solve_bso(fluid_ctx* c, int b, float* x, const float* x0)
{
	return lin_solve_masked<MODE_BSO>(c, b, x, x0, 0.0f, 0.0f, c->mf_a, c->mf_c);
}

// ProjectBO (FluidSystem.cpp:1061-1122): the pressure lives in the context's own array (the reference solves into its
// member p and ignores the array passed as p_old, :1096-1098), so the caller's p array is left untouched.
int dev_project_bo(fluid_ctx* c, float* u, float* v, float* div)
{
	FL_TRY(mf_scratch(c));
	{
		ProfScope ps(c, FLUID_PROF_PROJECT_AUX);
		LAUNCH(k_divergence_bo, cells_grid(c, 2), kBlkRow, 0, c->stream, c->g, u, v, div, c->mf_p, c->d_info);
	}
	FL_TRY((lin_solve_masked<MODE_BO>(c, 0, c->mf_p, div, 1.0f, 4.0f, nullptr, nullptr)));
	{
		ProfScope ps(c, FLUID_PROF_PROJECT_AUX);
		LAUNCH(k_subtract_gradient_b, cells_grid(c, 0), kBlkRow, 0, c->stream, c->g, u, v, c->mf_p, c->d_info);
	}
	return FLUID_OK;
}

// BlurB (FluidSystem.cpp:745-777): result in x, penultimate iterate in x0.
int dev_blur_b(fluid_ctx* c, int b, float* x, float* x0, float a, float cc)
{
	ProfScope ps(c, FLUID_PROF_BLUR);
	int limit = a > 0 ? c->gs_iters : 1;
	if (limit % 2 == 0) limit += 1;
	float *dst = x, *src = x0;
	for (int k = 0; k < limit; k++) {
		if (k > 0) std::swap(dst, src);
		FL_TRY(dev_set_bnd(c, b, src));
		LAUNCH(k_blur_pass_b, cells_grid(c, 0), kBlkRow, 0, c->stream, c->g, dst, src, c->d_info, a, cc);
	}
	FL_TRY(dev_zero_bnd(c, dst));
	CU_TRY(cudaGetLastError());
	return FLUID_OK;
}

// ProjectB (FluidSystem.cpp:1009-1059).
int dev_project_b(fluid_ctx* c, float* u, float* v, float* p, float* div)
{
	{
		ProfScope ps(c, FLUID_PROF_PROJECT_AUX);
		LAUNCH(k_divergence_b, cells_grid(c, 2), kBlkRow, 0, c->stream, c->g, u, v, div, p, c->d_info);
	}
	FL_TRY(dev_lin_solve_b(c, 0, p, div, 1.0f, 4.0f));
	{
		ProfScope ps(c, FLUID_PROF_PROJECT_AUX);
		LAUNCH(k_subtract_gradient_b, cells_grid(c, 0), kBlkRow, 0, c->stream, c->g, u, v, p, c->d_info);
	}
	return FLUID_OK;
}
