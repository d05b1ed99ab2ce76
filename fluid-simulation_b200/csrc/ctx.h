// Shared by the translation units of libfluid_cuda.so: error plumbing, the resident context that
// plays the role of the reference's FluidSystem (FSType::ORIGINAL; FluidSystem.cpp:18-128), and the
// device-level primitives each unit provides to the others.  One unit per concern:
//   fluid_core.cu    context life cycle, setters, transfers, profiling, resident-field entry points
//   fluid_relax.cu   lin_solve / Blur scheduling and the relaxation kernels
//   fluid_step.cu    the streaming/gather kernels, vel_step, dens_step, Update, render
//   fluid_host.cu    Stam-signature host-pointer entry points
//   fluid_static.cu  the entry point that never allocates
//   fluid_multi.cu   several GPUs from one C process
// There is no CPU fallback anywhere: every arithmetic step is a kernel launch.
#pragma once
#include "../../include/fluid_cuda.h"

#include <cuda.h> // CUtensorMap (types only; the encode function is fetched at run time)
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <functional>
#include <map>
#include <new>
#include <vector>

#include "grid.cuh"

using namespace fluid;

// ---------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------
extern thread_local char t_err[512];
extern thread_local unsigned long long t_launches;

int fail(int code, const char* fmt, ...);

#define CU_TRY(expr)                                                                                              \
	do {                                                                                                      \
		cudaError_t e__ = (expr);                                                                         \
		if (e__ != cudaSuccess)                                                                           \
			return fail(FLUID_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__),      \
				    __FILE__, __LINE__);                                                          \
	} while (0)

#define FL_TRY(expr)                                                                                              \
	do {                                                                                                      \
		int r__ = (expr);                                                                                 \
		if (r__ != FLUID_OK) return r__;                                                                  \
	} while (0)

// every kernel launch goes through this so that gpu_launches is a real count
#define LAUNCH(kernel, grid, block, smem, stream, ...)                                                            \
	do {                                                                                                      \
		kernel<<<grid, block, smem, stream>>>(__VA_ARGS__);                                               \
		t_launches++;                                                                                     \
	} while (0)

static inline int cdiv(long long a, long long b) { return (int) ((a + b - 1) / b); }

// ---------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------
struct fluid_ctx {
	int device = 0;
	cudaStream_t stream = nullptr;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	bool timed = false;
	GridDesc g{};
	int batch = 1;
	int gs_iters = 20, gs_order = FLUID_GS_WAVEFRONT, advect_mode = FLUID_ADVECT_MIDPOINT;
	float visc = 0, diff = 0, vorticity = 0, damp = 0;
	size_t field_bytes = 0;   // one field, all ensemble members
	static constexpr int kMaxFields = FLUID_FIELD_COUNT + 2 * (FLUID_MAX_FLUIDS - 1);
	float* f[kMaxFields] = {}; // field id -> device array (ids are fixed; steps swap locals; rb_tile_solve may trade an array with rb_scratch)
	int n_fluids = 1, n_fields = FLUID_FIELD_COUNT;
	float fl_diff[FLUID_MAX_FLUIDS] = {}, fl_visc[FLUID_MAX_FLUIDS] = {};
	bool fl_set[FLUID_MAX_FLUIDS] = {}; // fluid k has its own coefficients (else it follows fluid 0)
	float* vscr[4] = {};      // vorticity scratch, allocated on first use
	SourceDev* d_src = nullptr;
	float* d_src_saved = nullptr; // one float per entry (k_apply_sources / k_add_source_cells)
	bool sparse_sources = true;  // Update(): add_source over the listed cells only, when the kernels that follow allow it
	int* d_first = nullptr;
	int n_src = 0;
	std::vector<std::vector<SourceDev>> src; // per ensemble member
	bool src_dirty = false;
	// red-black tile path: sweeps per tile residency (0 = one-colour-per-launch kernels), its
	// ping-pong scratch field and the TMA descriptors of the arrays it has been asked to read
	int num_sms = 148;
	// row slab of a taller global grid (multi-GPU); slab_count == 1 means the whole grid is here
	int slab_index = 0, slab_count = 1, slab_rows = 0, slab_halo = 0;
	int own_lo = 1;            // local row of the first owned row
	fluid_exchange_fn exchange = nullptr;
	void* exchange_user = nullptr;
	bool exchange_stream_ordered = false; // the transport enqueues on c->stream; no host sync around it
	cudaStream_t own_stream = nullptr;    // the stream created with the context (c->stream may be the caller's)
	// fluid_update overlaps its bus transfers with the step: second stream + the two events that order them
	cudaStream_t copy_stream = nullptr;
	cudaEvent_t ev_dens_up = nullptr, ev_vel_done = nullptr;
	// native halo transport (fluid_peer.cu): my inboxes + flags, the neighbours' (peer pointers), refresh counter
	void* peer_block = nullptr;
	unsigned int* peer_counter = nullptr;
	void *peer_lower = nullptr, *peer_upper = nullptr;
	bool peer_connected = false;
	unsigned int peer_seq = 0;
	cudaEvent_t pace_ev[4] = {};
	int* d_oob = nullptr;
	float* d_pixels = nullptr; // headless display buffers, allocated on first fluid_ctx_render
	float* d_lines = nullptr;      // set by advect when a backtrace leaves the rows stored on this slab
	int wavefront_tile = 0;    // 0 = automatic, else 32 or 64
	int wavefront_persistent = 1; // reference-order solve as ONE launch instead of one per hyperplane of tiles: 0 never, 1 automatic (= 3), 2 tile tickets, 3 row pipeline
	bool rowwave_used = false;    // a row-pipeline launch since the last fluid_ctx_sync (which then reads its watchdog flag)
	int* d_wf_sync[2] = { nullptr, nullptr }; // its ticket counter and per-column progress counters, one set per lane (u and v solve side by side)
	long long wf_sync_capacity[2] = { 0, 0 };
	bool small_kernels = true; // N <= kSmallMaxN: whole solve / whole blur in one single-CTA launch per member
	int rb_tile_sweeps = 2;
	bool stream_inline = false; // this solve's streaming passes recompute out-of-window colour passes on the spot (no redo list)
	bool stream_force_inline = false; // tests: every streaming pass in the inline-exact variant
	bool rb_stream = true;     // interior of large grids by the register-resident streaming kernel (k_rb_stream)
	// Feedback for the streaming kernels ([0] relaxation, [1] blur).  They hand every segment that met a
	// dividend outside the verified window to an exact fallback; on fields that are mostly fringe (an ensemble
	// of small grids, each with its own front decaying through the subnormals) most segments come back and the
	// streaming pass was wasted.  The count of the last launch of a solve is copied to pinned memory; when it
	// has arrived (no waiting) and says more than a twelfth of the segments were redone -- an exact tile costs
	// about ten times its share of a streaming pass, measured break-even 8-12 % -- the next solves use the
	// pass-per-launch kernels, and streaming is tried again after kStreamHoldoff of them.  Results are
	// bit-identical either way.
	struct StreamFeedback {
		int* h_count = nullptr; // pinned
		cudaEvent_t arrived = nullptr;
		bool pending = false;
		long long segments = 0; // segments of the launch whose count is in flight
		int holdoff = 0;      // solves left before the windowed sequence is tried again
		int holdoff_len = 0;  // length of the next hold-off: doubles while the trials keep coming back mostly fringe
	} feedback[2];
	// redo lists of the two streaming kernels: [0] = count, [1..] = (task, segment) entries.  One per kernel
	// family (0 relaxation, 1 blur): the density diffusion may run on the side stream while VelStep runs
	// ([2]: the relaxation of the second lane, see `lane`)
	int* d_redo[3] = { nullptr, nullptr, nullptr };
	int redo_capacity[3] = { 0, 0, 0 };
	// VelStep's two diffusion solves (u, v) are independent: the second runs on its own stream ("lane" 1) with its
	// own ping-pong scratch field and redo list, so that each fills the ragged end of the other's launches
	int lane = 0;
	float* rb_scratch2 = nullptr;
	cudaStream_t lane_stream = nullptr;
	cudaEvent_t ev_fork2 = nullptr, ev_join2 = nullptr;
	// Update(): add_source + Blur of the densities do not depend on the velocity, so they run on a second stream
	// beside VelStep and fill the tails of its launches
	cudaStream_t side_stream = nullptr;
	cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
	bool overlap_density = true;
	bool side_join_pending = false; // the densities' side-stream work of this frame has not been joined yet (row slabs join it before the first halo refresh)
	int rb_rx = 256; // tile window width, see choose_rb_rx
	int fs_type = FLUID_TYPE_ORIGINAL; // which of the reference's step function pairs this context runs (SetType, FluidSystem.cpp:69-99)
	float *mf_a = nullptr, *mf_c = nullptr, *mf_p = nullptr; // ORIGINAL_BORDERED_MF: per-cell coefficients of LinearSolve1BSO, the system's own pressure array
	bool mf_coeffs_ready = false;       // Update() has computed mf_a / mf_c for the velocity step that follows
	float visc_air = 0.001f;            // FluidSystem.h:149, the viscosity of cells that hold no fluid
	unsigned char* d_info = nullptr;    // cellInfo, one byte per cell in the fields' pitched layout (bordered types)
	int advect_kernel = 0; // 0 automatic, 1 one thread per cell (gathers), 2 row window (kernels_advect.cuh)
	// divisors whose short division sequence has been verified exhaustively (grid.cuh: DivC)
	// 0 never; 1 automatic: a divisor that comes back in a later frame / call, and has enough work behind it to
	// repay the 1.8 ms proof, is verified in the background (no host wait) and used from the frame after it lands;
	// 2 always, at once, waiting for the proof (tests, one-off solves)
	int fast_div_mode = 1;
	struct DivEntry { DivC dv; unsigned long long last_use; };
	struct DivSeen { double work; unsigned long long first_epoch; };
	std::map<uint32_t, DivEntry> divs;   // proven divisors (bounded: least recently used leaves)
	std::map<uint32_t, DivSeen> div_seen; // not yet proven: cell updates and the epoch of the first sight (bounded)
	unsigned long long epoch = 0;         // advances with every API call on the context and every frame
	unsigned int* d_divbins = nullptr;
	// the one proof that may be in flight: its divisor, pinned result bins, completion event, side stream
	struct DivPending {
		bool active = false;
		uint32_t key = 0;
		DivC dv{};
		unsigned int* h_bins = nullptr; // pinned, 512 counters (window, scaled sequence)
		cudaEvent_t done = nullptr;
		cudaStream_t stream = nullptr;
	} div_pending;
	float* rb_scratch = nullptr;
	std::map<std::pair<const float*, int>, CUtensorMap> tmaps; // (array, window width) -> descriptor
	// optional per-kernel-class device timing (CUDA event pairs on c->stream)
	bool prof_on = false;
	struct ProfSeg { cudaEvent_t a, b; int cls; unsigned long long launches; };
	std::vector<ProfSeg> prof_segs;
	std::vector<cudaEvent_t> prof_pool;
};

// RAII segment: an NVTX range named after the kernel class around the launches of one primitive (header-only NVTX 3:
// a no-op of a few nanoseconds unless a tool such as ncu --nvtx or nsys is attached), and, when profiling is on, two
// events that bracket them on the stream.
static inline const char* prof_class_name(int cls)
{
	static const char* const names[] = { "fluid:lin_solve", "fluid:blur", "fluid:advect", "fluid:project_aux", "fluid:add_source", "fluid:other",
					     "fluid:halo_exchange" };
	return (cls >= 0 && cls < 7) ? names[cls] : "fluid:?";
}

struct ProfScope {
	fluid_ctx* c;
	int idx = -1;
	unsigned long long l0 = 0;
	ProfScope(fluid_ctx* ctx, int cls) : c(ctx)
	{
		nvtxRangePushA(prof_class_name(cls));
		if (!c->prof_on) return;
		cudaEvent_t ev[2];
		for (auto& e : ev) {
			if (!c->prof_pool.empty()) { e = c->prof_pool.back(); c->prof_pool.pop_back(); }
			else if (cudaEventCreate(&e) != cudaSuccess) return;
		}
		cudaEventRecord(ev[0], c->stream);
		l0 = t_launches;
		c->prof_segs.push_back({ ev[0], ev[1], cls, 0 });
		idx = (int) c->prof_segs.size() - 1;
	}
	~ProfScope()
	{
		nvtxRangePop();
		if (idx < 0) return;
		cudaEventRecord(c->prof_segs[idx].b, c->stream);
		c->prof_segs[idx].launches = t_launches - l0;
	}
};

static inline int ctx_bind(fluid_ctx* c)
{
	c->epoch++;
	CU_TRY(cudaSetDevice(c->device));
	return FLUID_OK;
}

static inline long long n4_of(const fluid_ctx* c) { return c->g.plane * c->batch / 4; }

static inline dim3 rowgroup_grid(const fluid_ctx* c, int rows) { return dim3(cdiv(c->g.pitch / 4, 32), cdiv(rows, 8), c->batch); }
static inline dim3 cell_grid(const fluid_ctx* c) { return dim3(cdiv(c->g.N, 32), cdiv(c->g.NY, 8), c->batch); }
static const dim3 kBlk2D(32, 8);

static inline int flat_blocks(const fluid_ctx* c)
{
	long long need = (n4_of(c) + 255) / 256;
	return (int) std::min<long long>(need, 148LL * 16);
}

#define CHECK_CTX(c)                                                                                              \
	do {                                                                                                      \
		if (!(c)) return fail(FLUID_ERR_INVALID, "null context");                                         \
		FL_TRY(ctx_bind(c));                                                                              \
	} while (0)

// ---------------------------------------------------------------------------
// device-level primitives (all asynchronous on c->stream)
// ---------------------------------------------------------------------------
// fluid_step.cu
int dev_add_source(fluid_ctx* c, float* x, const float* s, float dt);
int dev_add_source2(fluid_ctx* c, float* x, const float* s, float* y, const float* t, float dt);
int dev_set_bnd(fluid_ctx* c, int b, float* x);
int dev_zero_bnd(fluid_ctx* c, float* x);
int dev_advect(fluid_ctx* c, int b, float* d, const float* d0, const float* u, const float* v, float dt);
int dev_advect_velocity(fluid_ctx* c, float* u, float* v, const float* u0, const float* v0, float dt);
int dev_project(fluid_ctx* c, float* u, float* v, float*& p, float* div);
int dev_vel_step(fluid_ctx* c, float* u, float* v, float* u0, float* v0, float visc, float dt, bool sources_added = false);
int dev_dens_step(fluid_ctx* c, float* x, float* x0, const float* u, const float* v, float diff, float dt);
int dev_vort_confinement(fluid_ctx* c, float* u, float* v, float vorticity);
int dev_damp(fluid_ctx* c, float* vel, float amount, float dt);
int sync_sources(fluid_ctx* c);
int dev_update(fluid_ctx* c, float dt);
int dev_update_hooks(fluid_ctx* c, float dt, cudaEvent_t dens_ready, const std::function<int()>& after_velocity);
int dev_halo_refresh(fluid_ctx* c, int n, const int* fields);
int dev_render(fluid_ctx* c, int sim, const float* color_rgb, float* host_pixels, float* host_lines);
// Diffuse1 (FluidSystem.cpp:810-814): a = dt*diff*N*N left to right in float.
static inline float diffusion_a(const fluid_ctx* c, float coef, float dt) { return dt * coef * c->g.N * c->g.N; }
static inline float fluid_diff(const fluid_ctx* c, int fl) { return c->fl_set[fl] ? c->fl_diff[fl] : c->diff; }
// visc_avg, FluidSystem.cpp:26-30: float sum over the fluids, divided by the count
static inline float visc_avg(const fluid_ctx* c)
{
	if (c->n_fluids == 1) return c->visc;
	float v = 0;
	for (int fl = 0; fl < c->n_fluids; fl++) v += c->fl_set[fl] ? c->fl_visc[fl] : c->visc;
	v /= (float) c->n_fluids;
	return v;
}
// fluid_relax.cu
int dev_lin_solve(fluid_ctx* c, int b, float*& x, const float* x0, float a, float cc, bool x_is_zero = false);
int dev_blur(fluid_ctx* c, int b, float* x, float* x0, float a, float cc);
int division_for(fluid_ctx* c, float cc, double cell_updates, DivC* out);
bool solve_accepts_zero_x(const fluid_ctx* c, float a, float cc, int limit);
bool use_small_kernels(const fluid_ctx* c);
int choose_rb_rx(int N);
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
				  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
				  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int get_encode_fn(EncodeTiledFn* out);
// fluid_bordered.cu (FSType::ORIGINAL_BORDERED: cellInfo-masked primitives)
int bordered_alloc(fluid_ctx* c);
int bordered_reset(fluid_ctx* c);
int bordered_calc(fluid_ctx* c);
int bordered_disable_edges(fluid_ctx* c, int n, const int* edges);
int dev_lin_solve_b(fluid_ctx* c, int b, float* x, const float* x0, float a, float cc);
int dev_blur_b(fluid_ctx* c, int b, float* x, float* x0, float a, float cc);
int dev_project_b(fluid_ctx* c, float* u, float* v, float* p, float* div);
int dev_mf_coeffs(fluid_ctx* c, float dt);
int dev_lin_solve_bso(fluid_ctx* c, int b, float* x, const float* x0);
int dev_project_bo(fluid_ctx* c, float* u, float* v, float* div);
constexpr int kBndNone = -2; // advect: no SetBorder at all (Advect1B, FluidSystem.cpp:864-874)
// fluid_peer.cu
int peer_halo_exchange(fluid_ctx* c, int n, const int* ids);
// fluid_core.cu
int check_field_n(int f, int n);
int check_distinct_n(int n_fields, std::initializer_list<int> ids);
#define check_field(f) check_field_n((f), c->n_fields)
