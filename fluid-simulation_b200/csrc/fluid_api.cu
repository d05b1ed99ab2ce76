// libfluid_cuda.so -- the C ABI declared in include/fluid_cuda.h.
//
// Host side of the drop-in boundary: a resident context that plays the role of
// the reference's FluidSystem (FSType::ORIGINAL, one fluid; FluidSystem.cpp:18-128)
// and the Stam-signature host-pointer entry points layered on top of it.
// There is no CPU fallback anywhere in this file: every arithmetic step is a
// kernel launch, and without a CUDA device every entry point fails.
#include "../../include/fluid_cuda.h"

#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <vector>

#include "grid.cuh"
#include "kernels_basic.cuh"
#include "kernels_relax.cuh"
#include "kernels_rb_stream.cuh"
#include "kernels_rb_tile.cuh"
#include "kernels_blur_stream.cuh"
#include "kernels_small.cuh"

#include <functional>
#include <map>

using namespace fluid;

// ---------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------
static thread_local char t_err[512] = "";
static thread_local unsigned long long t_launches = 0;

static int fail(int code, const char* fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(t_err, sizeof t_err, fmt, ap);
	va_end(ap);
	return code;
}

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
	int* d_oob = nullptr;
	float* d_pixels = nullptr; // headless display buffers, allocated on first fluid_ctx_render
	float* d_lines = nullptr;      // set by advect when a backtrace leaves the rows stored on this slab
	int wavefront_tile = 0;    // 0 = automatic, else 32 or 64
	bool small_kernels = true; // N <= kSmallMaxN: whole solve / whole blur in one single-CTA launch per member
	int rb_tile_sweeps = 2;
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
		int holdoff = 0;
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
	int rb_rx = 256; // tile window width, see choose_rb_rx
	// divisors whose short division sequence has been verified exhaustively (grid.cuh: DivC)
	int fast_div_mode = 1; // 0 never, 1 when the solve is large enough to repay the 1.5 ms check, 2 always
	std::map<uint32_t, DivC> divs;
	std::map<uint32_t, double> div_work; // cell updates seen per not-yet-verified divisor
	unsigned int* d_divbins = nullptr;
	float* rb_scratch = nullptr;
	std::map<std::pair<const float*, int>, CUtensorMap> tmaps; // (array, window width) -> descriptor
	// optional per-kernel-class device timing (CUDA event pairs on c->stream)
	bool prof_on = false;
	struct ProfSeg { cudaEvent_t a, b; int cls; unsigned long long launches; };
	std::vector<ProfSeg> prof_segs;
	std::vector<cudaEvent_t> prof_pool;
};

// RAII segment: brackets the launches of one primitive with two events when
// profiling is on; costs nothing otherwise.
struct ProfScope {
	fluid_ctx* c;
	int idx = -1;
	unsigned long long l0 = 0;
	ProfScope(fluid_ctx* ctx, int cls) : c(ctx)
	{
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
		if (idx < 0) return;
		cudaEventRecord(c->prof_segs[idx].b, c->stream);
		c->prof_segs[idx].launches = t_launches - l0;
	}
};

static int ctx_bind(fluid_ctx* c)
{
	CU_TRY(cudaSetDevice(c->device));
	return FLUID_OK;
}

static inline long long n4_of(const fluid_ctx* c) { return c->g.plane * c->batch / 4; }

static inline dim3 rowgroup_grid(const fluid_ctx* c, int rows) { return dim3(cdiv(c->g.pitch / 4, 32), cdiv(rows, 8), c->batch); }
static inline dim3 cell_grid(const fluid_ctx* c) { return dim3(cdiv(c->g.N, 32), cdiv(c->g.NY, 8), c->batch); }
static const dim3 kBlk2D(32, 8);

static int flat_blocks(const fluid_ctx* c)
{
	long long need = (n4_of(c) + 255) / 256;
	return (int) std::min<long long>(need, 148LL * 16);
}

// ---------------------------------------------------------------------------
// device-level primitives (all asynchronous on c->stream)
// ---------------------------------------------------------------------------
static int dev_add_source(fluid_ctx* c, float* x, const float* s, float dt)
{
	ProfScope ps(c, FLUID_PROF_ADD_SOURCE);
	LAUNCH(k_add_source, flat_blocks(c), 256, 0, c->stream, (float4*) x, (const float4*) s, dt, n4_of(c));
	return FLUID_OK;
}

static int dev_add_source2(fluid_ctx* c, float* x, const float* s, float* y, const float* t, float dt)
{
	ProfScope ps(c, FLUID_PROF_ADD_SOURCE);
	LAUNCH(k_add_source2, flat_blocks(c), 256, 0, c->stream, (float4*) x, (const float4*) s, (float4*) y,
	       (const float4*) t, dt, n4_of(c));
	return FLUID_OK;
}

static int dev_set_bnd(fluid_ctx* c, int b, float* x)
{
	const int n = std::max(c->g.N, c->g.NY);
	LAUNCH(k_set_bnd, dim3(cdiv(n, 256), 1, c->batch), 256, 0, c->stream, c->g, b, x);
	return FLUID_OK;
}

static int dev_zero_bnd(fluid_ctx* c, float* x)
{
	const int n = std::max(c->g.N, c->g.NY) + 2;
	LAUNCH(k_zero_bnd, dim3(cdiv(n, 256), 1, c->batch), 256, 0, c->stream, c->g, x);
	return FLUID_OK;
}

// ---- red-black tile path ---------------------------------------------------
constexpr int kRbRY = 64, kRbRX = 256, kRbNT = 512;
constexpr size_t kRbSmem = (2 * (size_t) (kRbRY + 2) + kRbRY) * kRbRX * sizeof(float); // two x windows + x0 staging

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
				  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
				  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled is fetched through the runtime so that the library
// has no link-time dependency on libcuda (it must load on a machine without a driver).
static int get_encode_fn(EncodeTiledFn* out)
{
	static EncodeTiledFn fn = nullptr;
	if (!fn) {
		void* p = nullptr;
		cudaDriverEntryPointQueryResult qres;
		CU_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
		if (qres != cudaDriverEntryPointSuccess || !p) return fail(FLUID_ERR_CUDA, "driver does not provide cuTensorMapEncodeTiled");
		fn = (EncodeTiledFn) p;
	}
	*out = fn;
	return FLUID_OK;
}

// TMA descriptor of one field: 3-D tensor (pitch, NY+2, batch) of floats, box = one tile window.
static int field_tmap(fluid_ctx* c, const float* base, int rx, const CUtensorMap** out)
{
	auto it = c->tmaps.find({ base, rx });
	if (it == c->tmaps.end()) {
		EncodeTiledFn enc = nullptr;
		FL_TRY(get_encode_fn(&enc));
		CUtensorMap m;
		const cuuint64_t dims[3] = { (cuuint64_t) c->g.pitch, (cuuint64_t) (c->g.NY + 2), (cuuint64_t) c->batch };
		const cuuint64_t strides[2] = { (cuuint64_t) c->g.pitch * sizeof(float), (cuuint64_t) c->g.plane * sizeof(float) };
		const cuuint32_t box[3] = { (cuuint32_t) rx, (cuuint32_t) kRbRY, 1 };
		const cuuint32_t estr[3] = { 1, 1, 1 };
		CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*) base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
				 CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
		if (r != CUDA_SUCCESS) return fail(FLUID_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int) r);
		it = c->tmaps.emplace(std::make_pair(base, rx), m).first;
	}
	*out = &it->second;
	return FLUID_OK;
}

// DivC for divisor cc: verified window if known / worth establishing, else range = 0 (callers use div.rn).
static int division_for(fluid_ctx* c, float cc, double cell_updates, DivC* out)
{
	out->c = cc;
	out->y = 1.0f / cc;
	out->lo = 0;
	out->range = 0;
	if (!(cc > 0.0f) || !isfinite(cc) || !isfinite(out->y) || c->fast_div_mode == 0) return FLUID_OK;
	uint32_t key;
	memcpy(&key, &cc, sizeof key);
	auto it = c->divs.find(key);
	if (it != c->divs.end()) {
		*out = it->second;
		return FLUID_OK;
	}
	// 2^31 trial divisions cost about as much as 2.5e8 cell updates: run the proof once that much work
	// has been (or is about to be) done with this divisor -- at once for a large grid, after a few
	// steps for a medium one, never for a divisor that changes every frame on a small grid
	if (c->fast_div_mode == 1) {
		double& seen = c->div_work[key];
		seen += cell_updates;
		if (seen < 2.5e8) return FLUID_OK;
	}
	if (!c->d_divbins) CU_TRY(cudaMalloc(&c->d_divbins, 256 * sizeof(unsigned int)));
	CU_TRY(cudaMemsetAsync(c->d_divbins, 0, 256 * sizeof(unsigned int), c->stream));
	LAUNCH(k_verify_division, c->num_sms * 8, 256, 0, c->stream, *out, c->d_divbins);
	unsigned int bins[256];
	CU_TRY(cudaMemcpyAsync(bins, c->d_divbins, sizeof bins, cudaMemcpyDeviceToHost, c->stream));
	CU_TRY(cudaStreamSynchronize(c->stream));
	if (bins[127] == 0) {
		int elo = 127, ehi = 127;
		while (elo > 1 && bins[elo - 1] == 0) elo--;
		while (ehi < 254 && bins[ehi + 1] == 0) ehi++;
		out->lo = (uint32_t) elo << 23;
		out->range = (uint32_t) (ehi - elo + 1) << 23;
	}
	c->divs[key] = *out;
	return FLUID_OK;
}

constexpr int kStreamHoldoff = 64;

// May this solve use the streaming kernel?  Consumes the count of an earlier launch if it has arrived.
static bool stream_allowed(fluid_ctx* c, int which)
{
	auto& fb = c->feedback[which];
	if (fb.pending && cudaEventQuery(fb.arrived) == cudaSuccess) {
		fb.pending = false;
		if (getenv("FLUID_DEBUG_FEEDBACK")) fprintf(stderr, "feedback[%d]: %d of %lld segments redone\n", which, *fb.h_count, fb.segments);
		if (fb.segments > 0 && (long long) *fb.h_count * 12 > fb.segments) fb.holdoff = kStreamHoldoff;
	}
	if (fb.holdoff > 0) {
		fb.holdoff--;
		return false;
	}
	return c->rb_stream;
}

// After the last streaming launch of a solve: send its redo count home (asynchronously).
static int stream_report(fluid_ctx* c, int which, long long segments, int buffer = -1)
{
	if (buffer < 0) buffer = which;
	auto& fb = c->feedback[which];
	if (fb.pending) return FLUID_OK; // an earlier count is still in flight; one at a time
	if (!fb.h_count) {
		CU_TRY(cudaMallocHost(&fb.h_count, sizeof(int)));
		CU_TRY(cudaEventCreateWithFlags(&fb.arrived, cudaEventDisableTiming));
	}
	CU_TRY(cudaMemcpyAsync(fb.h_count, c->d_redo[buffer], sizeof(int), cudaMemcpyDeviceToHost, c->stream));
	CU_TRY(cudaEventRecord(fb.arrived, c->stream));
	fb.segments = segments;
	fb.pending = true;
	return FLUID_OK;
}

// The list of rectangles a streaming kernel hands to its exact fallback: [0] = count, [1..] = entries.
static int redo_reserve(fluid_ctx* c, int which, long long entries)
{
	if (entries > 0x7fffffffLL) return fail(FLUID_ERR_UNSUPPORTED, "too many streaming tasks");
	if (entries > c->redo_capacity[which]) {
		if (c->d_redo[which]) CU_TRY(cudaFree(c->d_redo[which]));
		c->d_redo[which] = nullptr;
		CU_TRY(cudaMalloc(&c->d_redo[which], sizeof(int) * (size_t) entries));
		c->redo_capacity[which] = (int) entries;
	}
	return FLUID_OK;
}

// Geometry of one pass of S sweeps: the tile grid, which tiles are left to the tile kernel (the frame)
// and the rectangle the streaming kernel produces.
struct PassGeom {
	int TX, TY, ntx, nty;
	int ftx0, ftx1, fty0, fty1; // streamed tiles: ftx0 <= tx < ftx1, fty0 <= ty < fty1; ftx1 = 0: none, the tile kernel does everything
	StreamRect rc;
};

template <int S>
static PassGeom pass_geometry(const fluid_ctx* c, bool stream_ok) // stream_ok: verified division at hand and streaming allowed
{
	PassGeom pg{};
	pg.TY = kRbRY - 4 * S;
	pg.TX = c->rb_rx - 2 * (((2 * S + 3) / 4) * 4);
	pg.ntx = cdiv(c->g.N, pg.TX);
	pg.nty = cdiv(c->g.NY, pg.TY);
	pg.ftx0 = pg.ftx1 = pg.fty0 = pg.fty1 = 0;
	if (!stream_ok) return pg; // the streaming kernel only knows the verified division
	constexpr int L = StreamGeom<S>::L, HALO = StreamGeom<S>::HALO;
	if (StreamGeom<S>::OUT > pg.TX) return pg;             // a redo tile must cover a strip's width
	// Rows: all of them; the streaming kernel does the bottom and top walls' set_bnd (or, for a slab that
	// ends in a neighbour's rows, treats the row beyond as fixed, like the tile kernel).
	if (c->g.NY < 8 * L) return pg;
	const int fy = pg.nty;
	// Columns: the streaming kernel handles the left and right walls itself when a row is a whole number of
	// float4 groups; otherwise it keeps one tile column away from them as well.
	int fx0, fx1;
	if (c->g.N % 4 == 0 && c->g.N >= 2 * StreamGeom<S>::OUT) {
		fx0 = 0;
		fx1 = pg.ntx;
	} else {
		fx0 = 1;
		fx1 = pg.ntx - 1;
		while (fx1 > 1 && pg.TX * fx1 + HALO > c->g.N + 1) fx1--;
		if (fx1 <= 1) return pg;
	}
	const long long width = (fx0 == 0) ? c->g.N : (long long) (fx1 - 1) * pg.TX;
	if (width * c->g.NY * c->batch < (1LL << 21)) return pg; // too little to bother
	pg.ftx0 = fx0;
	pg.ftx1 = fx1;
	pg.fty1 = fy;
	pg.rc.col0 = (fx0 == 0) ? XO + 1 : XO + 1 + pg.TX;
	pg.rc.col1 = (fx0 == 0) ? XO + 1 + c->g.N : XO + 1 + pg.TX * fx1;
	pg.rc.row0 = 1;
	pg.rc.row1 = c->g.NY + 1;
	pg.rc.nstrips = cdiv(pg.rc.col1 - pg.rc.col0, StreamGeom<S>::OUT);
	const int rows = pg.rc.row1 - pg.rc.row0;
	// chunks of about 256 rows, nudged so that the warp tasks fill whole waves of resident warps
	int nchunks = std::max(1, cdiv(rows, 256));
	// resident warps per SM of k_rb_stream<S> (register bound: 167 / 115 / 89 registers per thread)
	const long long slots = (long long) c->num_sms * (S >= 4 ? 12 : (S == 2 ? 16 : 20));
	constexpr int kWallSplit = 2; // wall strips: tasks of half the rows
	const bool walls = (fx0 == 0);
	const long long per_chunk = (long long) (pg.rc.nstrips + (walls ? 2 * (kWallSplit - 1) : 0)) * c->batch;
	if (per_chunk * nchunks > slots / 2) {
		const long long waves = std::max(1LL, (per_chunk * nchunks + slots / 2) / slots);
		nchunks = (int) std::max(1LL, std::min<long long>(rows / 32, (waves * slots) / per_chunk));
	}
	const int unit = 2 * kWallSplit; // chunk and wall task both an even number of rows: one colour phase per launch
	// the bottom and top chunks run ~3S slower row steps next to their wall: make them that much shorter
	const int shift = (nchunks > 2) ? 12 * S : 0;
	pg.rc.row_shift = shift;
	pg.rc.chunk_rows = (cdiv(rows + 2 * shift, nchunks) + unit - 1) / unit * unit;
	pg.rc.nchunks = cdiv(rows + shift, pg.rc.chunk_rows);
	pg.rc.wall_rows = walls ? pg.rc.chunk_rows / kWallSplit : 0;
	pg.rc.wall_chunks = walls ? cdiv(rows + shift, pg.rc.wall_rows) : 0;
	pg.rc.seg_rows = pg.TY;
	pg.rc.segs = cdiv(pg.rc.chunk_rows, pg.TY);
	pg.rc.per_sim = 2 * pg.rc.wall_chunks + (pg.rc.nstrips - (walls ? 2 : 0)) * pg.rc.nchunks;
	return pg;
}

template <int S, int RX>
static int rb_tile_pass_rx(fluid_ctx* c, int b, const float* in, float* out, const float* x0, float a, const DivC& dv,
			   const PassGeom& pg)
{
	static bool configured[64] = {}; // per template instance and per device: function attributes are per device
	constexpr size_t smem = (2 * (size_t) (kRbRY + 2) + kRbRY) * RX * sizeof(float); // two x windows + x0 staging
	auto kern = k_rb_tile<S, kRbRY, RX, kRbNT>;
	if (c->device >= 64 || !configured[c->device]) {
		CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
		if (c->device < 64) configured[c->device] = true;
	}
	const CUtensorMap *tm, *tm0;
	FL_TRY(field_tmap(c, in, RX, &tm));
	FL_TRY(field_tmap(c, x0, RX, &tm0));
	long long ntiles = (long long) pg.ntx * pg.nty;
	if (pg.ftx1 > 0) ntiles -= (long long) (pg.ftx1 - pg.ftx0) * (pg.fty1 - pg.fty0); // the frame only, possibly nothing
	ntiles *= c->batch;
	if (ntiles > 0x7fffffffLL) return fail(FLUID_ERR_UNSUPPORTED, "too many tiles");
	// persistent: one CTA per SM; behind the streaming kernel the redo list adds tiles the host cannot count
	const int grid = (pg.ftx1 > 0) ? c->num_sms : (int) std::min<long long>(ntiles, c->num_sms);
	LAUNCH(kern, grid, kRbNT, smem, c->stream, *tm, *tm0, c->g, b, out, a, dv, pg.ntx, pg.nty, (int) ntiles, pg.ftx0, pg.ftx1,
	       pg.fty0, pg.fty1, c->d_redo[c->lane ? 2 : 0], pg.rc, StreamGeom<S>::OUT, StreamGeom<S>::HALO);
	return FLUID_OK;
}

// One pass of S red-black sweeps, in -> out: the streaming kernel on the interior rectangle (if the grid
// is large enough to have one) and the tile kernel on the frame of tiles around it, or on everything.
// The window width of the tile kernel is chosen per context so that ceil(N / (RX-8)) windows waste the
// fewest columns (N=512: 3 x 192 instead of 3 x 256; N <= 120: one 128-wide window instead of 256).
template <int S>
static int rb_tile_pass(fluid_ctx* c, int b, const float* in, float* out, const float* x0, float a, const DivC& dv, bool streaming,
			bool pow2, long long* segments)
{
	const PassGeom pg = pass_geometry<S>(c, streaming);
	*segments = 0;
	const bool whole = pg.ftx1 > 0 && pg.ftx0 == 0 && pg.ftx1 == pg.ntx && pg.fty0 == 0 && pg.fty1 == pg.nty; // the streaming kernel covers the grid
	(void) pow2;
	if (pg.ftx1 > 0) {
		const long long ntasks_ll = (long long) pg.rc.per_sim * c->batch;
		*segments = ntasks_ll * pg.rc.segs;
		const int rbuf = c->lane ? 2 : 0;
		FL_TRY(redo_reserve(c, rbuf, ntasks_ll * pg.rc.segs + 1));
		const int ntasks = (int) ntasks_ll;
		CU_TRY(cudaMemsetAsync(c->d_redo[rbuf], 0, sizeof(int), c->stream));
		LAUNCH((k_rb_stream<S, false>), cdiv(ntasks, 4), 128, 0, c->stream, c->g, pg.rc, in, x0, out, a, dv, ntasks, c->d_redo[rbuf], b);
	}
	// nothing but listed rectangles to do: the narrowest window covers a strip (112 columns) and has the
	// shortest colour passes -- a redo launch lasts as long as one tile does
	if (whole && StreamGeom<S>::OUT <= 128 - 2 * StreamGeom<S>::HALO) {
		PassGeom narrow = pg;
		narrow.TX = 128 - 2 * StreamGeom<S>::HALO;
		narrow.ntx = narrow.ftx1 = cdiv(c->g.N, narrow.TX);
		return rb_tile_pass_rx<S, 128>(c, b, in, out, x0, a, dv, narrow);
	}
	switch (c->rb_rx) {
	case 128: return rb_tile_pass_rx<S, 128>(c, b, in, out, x0, a, dv, pg);
	case 192: return rb_tile_pass_rx<S, 192>(c, b, in, out, x0, a, dv, pg);
	default: return rb_tile_pass_rx<S, 256>(c, b, in, out, x0, a, dv, pg);
	}
}

static int choose_rb_rx(int N)
{
	int best = 256;
	long long best_cost = -1;
	for (int rx : { 256, 192, 128 }) {
		const long long cost = (long long) cdiv(N, rx - 8) * rx;
		if (best_cost < 0 || cost < best_cost) { best = rx; best_cost = cost; }
	}
	return best;
}

// The streaming kernel covers the whole grid (rows are float4 groups, the grid is large): no tile kernel involved.
template <int S>
static bool stream_covers(const fluid_ctx* c, StreamRect* rc)
{
	const PassGeom pg = pass_geometry<S>(c, c->rb_stream);
	*rc = pg.rc;
	return pg.ftx1 > 0 && pg.ftx0 == 0 && pg.ftx1 == pg.ntx && pg.fty0 == 0 && pg.fty1 == pg.nty;
}

// a == 1 and c a power of two (Project's pressure solve, c = 4): the division is a multiplication, exact for
// every operand -- no verified window, no redo list, and five instead of twelve instructions per update, which
// makes the pass HBM-bound.
static bool is_pow2_solve(float a, float cc)
{
	int e = 0;
	return a == 1.0f && cc > 0.0f && isfinite(cc) && frexpf(cc, &e) == 0.5f && (1.0f / cc) >= 1.17549435e-38f;
}

template <int S>
static int rb_stream_pass_pow2(fluid_ctx* c, int b, const float* in, float* out, const float* x0, float a, const DivC& dv, bool x_is_zero)
{
	StreamRect rc;
	if (!stream_covers<S>(c, &rc)) return fail(FLUID_ERR_CUDA, "internal: streaming pass on a grid it does not cover");
	const int ntasks = (int) ((long long) rc.per_sim * c->batch);
	if (x_is_zero) LAUNCH((k_rb_stream<S, true, true>), cdiv(ntasks, 4), 128, 0, c->stream, c->g, rc, in, x0, out, a, dv, ntasks, (int*) nullptr, b);
	else LAUNCH((k_rb_stream<S, true, false>), cdiv(ntasks, 4), 128, 0, c->stream, c->g, rc, in, x0, out, a, dv, ntasks, (int*) nullptr, b);
	return FLUID_OK;
}

// Will a solve with these coefficients start with a pass that can be told "x is zero"?  (dev_project asks before
// it decides not to write p = 0.)
static bool solve_accepts_zero_x(const fluid_ctx* c, float a, float cc, int limit)
{
	StreamRect rc;
	if (!(c->gs_order == FLUID_GS_REDBLACK && c->rb_tile_sweeps >= 2 && limit >= 4 && is_pow2_solve(a, cc))) return false;
	return stream_covers<4>(c, &rc);
}

// `limit` red-black sweeps as passes of 4, 2 or 1 sweeps, ping-ponging between x and the scratch field.
// After an odd number of passes the result sits in the scratch array: if x is the storage of a context
// field, that field and the scratch simply trade arrays (c->f is the only place that maps field ids to
// storage, and x is updated for the caller) -- 20 sweeps are then five 4-sweep passes instead of four and
// two 2-sweep passes.  For any other array the number of passes is made even.
static int rb_tile_solve(fluid_ctx* c, int b, float*& x, const float* x0, float a, float cc, int limit, bool x_is_zero)
{
	int field = -1;
	for (int f = 0; f < c->n_fields; f++)
		if (c->f[f] == x) field = f;
	float*& scratch = c->lane ? c->rb_scratch2 : c->rb_scratch; // each lane ping-pongs (and trades arrays) with its own
	if (!scratch) {
		CU_TRY(cudaMalloc(&scratch, c->field_bytes));
		CU_TRY(cudaMemsetAsync(scratch, 0, c->field_bytes, c->stream));
	}
	DivC dv;
	FL_TRY(division_for(c, cc, (double) c->g.N * c->g.NY * c->batch * limit, &dv));
	const bool pow2 = is_pow2_solve(a, cc);
	// otherwise the streaming kernel relies on the proven window reaching the largest finite float
	const bool streaming = dv.range != 0 && (dv.lo + dv.range) == (255u << 23) && stream_allowed(c, 0);
	// power-of-two solves first take as many whole-grid streaming passes of 5 and 4 sweeps as fit
	int p5 = 0, p4 = 0;
	StreamRect rc_unused;
	if (pow2 && c->rb_tile_sweeps >= 2) {
		// (5 sweeps per pass were tried: 260 us per pass against 148 us for 4 -- ten dependent levels per row step
		// leave three warps per scheduler too little to overlap, and the strips shrink to 104 columns)
		if (limit >= 4 && stream_covers<4>(c, &rc_unused)) p4 = limit / 4;
	}
	if (x_is_zero && p5 + p4 == 0) return fail(FLUID_ERR_CUDA, "internal: zero-x hint given to a solve that cannot use it");
	limit -= 5 * p5 + 4 * p4;
	int n4 = 0, n2 = 0, n1 = 0; // passes of 4, 2 and 1 sweeps; the total number of passes must be even
	if (limit == 0) {
		if (field < 0 && (p5 + p4) % 2) return fail(FLUID_ERR_CUDA, "internal: odd pass count on an array that is not a field");
	} else if (c->rb_tile_sweeps >= 2) {
		int rest = limit;
		if (pass_geometry<4>(c, streaming).ftx1 > 0) { // 4 sweeps per pass pay off where the streaming kernel runs
			n4 = rest / 4;
			rest -= 4 * n4;
		}
		n2 = rest / 2;
		n1 = rest % 2;
		if (field < 0 && (p5 + p4 + n4 + n2 + n1) % 2) {
			if (n4 > 0) { n4 -= 1; n2 += 2; }
			else { n2 -= 1; n1 += 2; } // limit >= 2, so n2 >= 1 here
		}
	} else {
		n2 = (field < 0) ? (limit + p5 + p4) % 2 : 0; // an odd count needs one 2-sweep pass to make the number of passes even
		n1 = limit - 2 * n2;
	}
	const float* in = x;
	float* out = scratch;
	long long segments = 0;
	for (int k = 0; k < p5 + p4; k++) {
		FL_TRY(rb_stream_pass_pow2<4>(c, b, in, out, x0, a, dv, x_is_zero && k == 0));
		float* t = const_cast<float*>(in);
		in = out;
		out = t;
	}
	for (int k = 0; k < n4 + n2 + n1; k++) {
		if (k < n4) FL_TRY(rb_tile_pass<4>(c, b, in, out, x0, a, dv, streaming, pow2, &segments));
		else if (k < n4 + n2) FL_TRY(rb_tile_pass<2>(c, b, in, out, x0, a, dv, streaming, pow2, &segments));
		else FL_TRY(rb_tile_pass<1>(c, b, in, out, x0, a, dv, streaming, pow2, &segments));
		float* t = const_cast<float*>(in);
		in = out;
		out = t;
	}
	if (segments > 0) FL_TRY(stream_report(c, 0, segments, c->lane ? 2 : 0)); // how much of the last pass had to be redone
	if (in != x) { // odd number of passes: the field's values are in the scratch array now
		c->f[field] = scratch;
		scratch = x;
		x = c->f[field];
	}
	return FLUID_OK;
}

static inline bool use_small_kernels(const fluid_ctx* c)
{
	return c->small_kernels && c->slab_count == 1 && c->g.N <= kSmallMaxN && c->g.NY == c->g.N;
}

static int small_configure(fluid_ctx* c)
{
	static bool configured[64] = {};
	if (c->device >= 64 || !configured[c->device]) {
		const int bytes = (int) small_smem_bytes(kSmallMaxN);
		CU_TRY(cudaFuncSetAttribute(k_small_lin_solve<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
		CU_TRY(cudaFuncSetAttribute(k_small_lin_solve<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
		CU_TRY(cudaFuncSetAttribute(k_small_blur, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
		if (c->device < 64) configured[c->device] = true;
	}
	return FLUID_OK;
}

// LinearSolve1 (FluidSystem.cpp:534-548) in the context's relaxation order.
// x may come back pointing at another array (rb_tile_solve); callers keep using x, not a copy made before.
static int dev_lin_solve(fluid_ctx* c, int b, float*& x, const float* x0, float a, float cc, bool x_is_zero = false)
{
	ProfScope ps(c, FLUID_PROF_LIN_SOLVE);
	const int limit = a > 0 ? c->gs_iters : 1;
	if (use_small_kernels(c)) {
		FL_TRY(small_configure(c));
		if (c->gs_order == FLUID_GS_REDBLACK)
			LAUNCH(k_small_lin_solve<true>, c->batch, kSmallThreads, small_smem_bytes(c->g.N), c->stream, c->g, b, x, x0, a, cc, limit);
		else
			LAUNCH(k_small_lin_solve<false>, c->batch, kSmallThreads, small_smem_bytes(c->g.N), c->stream, c->g, b, x, x0, a, cc, limit);
	} else if (c->gs_order == FLUID_GS_REDBLACK && c->rb_tile_sweeps > 0 && limit >= 2 && b >= 0 && b <= 2) {
		FL_TRY(rb_tile_solve(c, b, x, x0, a, cc, limit, x_is_zero));
	} else if (c->gs_order == FLUID_GS_REDBLACK) {
		for (int k = 0; k < limit; k++) {
			LAUNCH(k_rb_half, rowgroup_grid(c, c->g.NY), kBlk2D, 0, c->stream, c->g, b, x, x0, a, cc, 0);
			LAUNCH(k_rb_half, rowgroup_grid(c, c->g.NY), kBlk2D, 0, c->stream, c->g, b, x, x0, a, cc, 1);
		}
	} else {
		if (c->slab_count > 1)
			return fail(FLUID_ERR_UNSUPPORTED, "the wavefront (reference-order) relaxation is a pipeline across slabs and is offered on one GPU only; use FLUID_GS_REDBLACK for row slabs");
		// small tiles: a launch is bound by the latency of one tile (2T-1 dependent steps), not by throughput.
		// Measured ms/step (T=16/32/64): N=1024 8.6/10.0/13.8, N=2048 14.5/14.0/17.4, N=8192 77/57/68.
		const int T = (c->wavefront_tile > 0) ? c->wavefront_tile : (c->g.N < 2048 ? 16 : 32);
		const int ntx = cdiv(c->g.N, T), nty = cdiv(c->g.NY, T);
		const int last = (ntx - 1) + (nty - 1) + 2 * (limit - 1);
		for (int s = 0; s <= last; s++) {
			if (T == 16) LAUNCH(k_gs_wavefront<16>, dim3(ntx, limit, c->batch), 16, 0, c->stream, c->g, b, x, x0, a, cc, s, ntx, nty);
			else if (T == 32) LAUNCH(k_gs_wavefront<32>, dim3(ntx, limit, c->batch), 32, 0, c->stream, c->g, b, x, x0, a, cc, s, ntx, nty);
			else LAUNCH(k_gs_wavefront<64>, dim3(ntx, limit, c->batch), 64, 0, c->stream, c->g, b, x, x0, a, cc, s, ntx, nty);
		}
	}
	CU_TRY(cudaGetLastError());
	return FLUID_OK;
}

// Blur (FluidSystem.cpp:722-743).  x and x0 are the caller's two buffers; on
// return the result is in x and the penultimate iterate in x0.
// Blur, L passes per launch (kernels_blur_stream.cuh): the task geometry, or false when the grid does not
// qualify (rows must be whole float4 groups and at least two strips wide; small grids have their own kernel).
constexpr int kBlurL = 5, kBlurSeg = 32;

static bool blur_stream_geometry(const fluid_ctx* c, StreamRect* rc)
{
	constexpr int OUT = BlurGeom<kBlurL>::OUT;
	const GridDesc& g = c->g;
	if (!c->rb_stream || g.N % 4 != 0 || g.N < 2 * OUT || g.NY < 8 * kBlurL) return false;
	if ((long long) g.N * g.NY * c->batch < (1LL << 21)) return false;
	*rc = StreamRect{};
	rc->col0 = XO + 1;
	rc->col1 = XO + 1 + g.N;
	rc->row0 = 1;
	rc->row1 = g.NY + 1;
	rc->nstrips = cdiv(g.N, OUT);
	const int rows = g.NY;
	int nchunks = std::max(1, cdiv(rows, 256));
	const long long slots = (long long) c->num_sms * 16; // 4 CTAs of 4 warps per SM
	constexpr int kWallSplit = 2;
	const long long per_chunk = (long long) (rc->nstrips + 2 * (kWallSplit - 1)) * c->batch;
	if (per_chunk * nchunks > slots / 2) {
		const long long waves = std::max(1LL, (per_chunk * nchunks + slots / 2) / slots);
		nchunks = (int) std::max(1LL, std::min<long long>(rows / 32, (waves * slots) / per_chunk));
	}
	const int shift = (nchunks > 2) ? 6 * kBlurL : 0;
	rc->row_shift = shift;
	rc->chunk_rows = (cdiv(rows + 2 * shift, nchunks) + kWallSplit - 1) / kWallSplit * kWallSplit;
	rc->nchunks = cdiv(rows + shift, rc->chunk_rows);
	rc->wall_rows = rc->chunk_rows / kWallSplit;
	rc->wall_chunks = cdiv(rows + shift, rc->wall_rows);
	rc->seg_rows = kBlurSeg;
	rc->segs = cdiv(rc->chunk_rows, kBlurSeg);
	rc->per_sim = 2 * rc->wall_chunks + (rc->nstrips - 2) * rc->nchunks;
	return true;
}

static int dev_blur(fluid_ctx* c, int b, float* x, float* x0, float a, float cc)
{
	ProfScope ps(c, FLUID_PROF_BLUR);
	int limit = a > 0 ? c->gs_iters : 1;
	if (limit % 2 == 0) limit += 1;
	if (use_small_kernels(c)) {
		FL_TRY(small_configure(c));
		LAUNCH(k_small_blur, c->batch, kSmallThreads, small_smem_bytes(c->g.N), c->stream, c->g, b, x, x0, a, cc, limit);
		CU_TRY(cudaGetLastError());
		return FLUID_OK;
	}
	float *dst = x, *src = x0;
	DivC dv;
	FL_TRY(division_for(c, cc, (double) c->g.N * c->g.NY * c->batch * limit, &dv));
	FL_TRY(dev_set_bnd(c, b, src));
	// kBlurL passes per launch while an even number of such launches fits (the result must land in x with the
	// penultimate iterate in x0, as the pass-per-launch sequence leaves them); single passes for the rest
	int done = 0;
	StreamRect rc;
	const bool window_ok = dv.range != 0 && (dv.lo + dv.range) == (255u << 23);
	if (window_ok && b >= 0 && b <= 2 && limit > 2 * kBlurL && blur_stream_geometry(c, &rc) && stream_allowed(c, 1)) {
		const int launches = 2 * ((limit - 1) / (2 * kBlurL));
		const long long ntasks = (long long) rc.per_sim * c->batch;
		FL_TRY(redo_reserve(c, 1, ntasks * rc.segs + 1));
		for (int l = 0; l < launches; l++) {
			if (l > 0) std::swap(dst, src);
			CU_TRY(cudaMemsetAsync(c->d_redo[1], 0, sizeof(int), c->stream));
			LAUNCH(k_blur_stream<kBlurL>, cdiv((int) ntasks, 4), 128, 0, c->stream, c->g, rc, src, dst, a, dv, (int) ntasks, c->d_redo[1], b);
			// few entries, each a chain of 2L barrier-separated phases over ~5000 cells: wide CTAs keep the chain short
			LAUNCH((k_blur_redo<kBlurL, kBlurSeg>), 2 * c->num_sms, 1024, 0, c->stream, c->g, rc, src, dst, a, cc, c->d_redo[1], b);
		}
		done = launches * kBlurL;
		FL_TRY(stream_report(c, 1, ntasks * rc.segs));
	}
	for (int k = done; k < limit; k++) {
		if (k > 0) std::swap(dst, src);
		LAUNCH(k_blur_pass, rowgroup_grid(c, c->g.NY), kBlk2D, 0, c->stream, c->g, b, dst, src, a, dv,
		       k + 1 < limit ? 1 : 0);
	}
	FL_TRY(dev_zero_bnd(c, dst));
	CU_TRY(cudaGetLastError());
	return FLUID_OK;
}

static int dev_advect(fluid_ctx* c, int b, float* d, const float* d0, const float* u, const float* v, float dt)
{
	ProfScope ps(c, FLUID_PROF_ADVECT);
	const float dtN = dt * c->g.N;
	const bool euler = (c->advect_mode == FLUID_ADVECT_EULER);
	if (c->slab_count > 1) {
		if (euler) LAUNCH((k_advect<true, 1>), cell_grid(c), kBlk2D, 0, c->stream, c->g, b, d, d0, u, v, dtN, c->d_oob, c->own_lo, c->own_lo + c->slab_rows - 1);
		else LAUNCH((k_advect<true, 0>), cell_grid(c), kBlk2D, 0, c->stream, c->g, b, d, d0, u, v, dtN, c->d_oob, c->own_lo, c->own_lo + c->slab_rows - 1);
	} else {
		if (euler) LAUNCH((k_advect<false, 1>), cell_grid(c), kBlk2D, 0, c->stream, c->g, b, d, d0, u, v, dtN, c->d_oob, 1, 0);
		else LAUNCH((k_advect<false, 0>), cell_grid(c), kBlk2D, 0, c->stream, c->g, b, d, d0, u, v, dtN, c->d_oob, 1, 0);
	}
	return FLUID_OK;
}

static int dev_advect_velocity(fluid_ctx* c, float* u, float* v, const float* u0, const float* v0, float dt)
{
	ProfScope ps(c, FLUID_PROF_ADVECT);
	const float dtN = dt * c->g.N;
	const bool euler = (c->advect_mode == FLUID_ADVECT_EULER);
	if (c->slab_count > 1) {
		if (euler) LAUNCH((k_advect_velocity<true, 1>), cell_grid(c), kBlk2D, 0, c->stream, c->g, u, v, u0, v0, dtN, c->d_oob, c->own_lo, c->own_lo + c->slab_rows - 1);
		else LAUNCH((k_advect_velocity<true, 0>), cell_grid(c), kBlk2D, 0, c->stream, c->g, u, v, u0, v0, dtN, c->d_oob, c->own_lo, c->own_lo + c->slab_rows - 1);
	} else {
		if (euler) LAUNCH((k_advect_velocity<false, 1>), cell_grid(c), kBlk2D, 0, c->stream, c->g, u, v, u0, v0, dtN, c->d_oob, 1, 0);
		else LAUNCH((k_advect_velocity<false, 0>), cell_grid(c), kBlk2D, 0, c->stream, c->g, u, v, u0, v0, dtN, c->d_oob, 1, 0);
	}
	return FLUID_OK;
}

// Project (FluidSystem.cpp:982-1007).
static int dev_project(fluid_ctx* c, float* u, float* v, float*& p, float* div)
{
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

// Diffuse1 (FluidSystem.cpp:810-814): a = dt*diff*N*N left to right in float.
static inline float diffusion_a(const fluid_ctx* c, float coef, float dt) { return dt * coef * c->g.N * c->g.N; }

// Halo exchange between row slabs.  The slabs overlap by slab_halo (+1) rows and every
// kernel computes on all stored rows, so stale values creep in from the outer
// halo edge by one row per stencil application; the fields named here are made
// valid again over the full halo depth by asking the host's transport (NCCL
// send/recv, peer copies ...) to refill the halo rows from the neighbours' owned
// rows.  With a 48-row halo that is needed once per 20-sweep solve instead of
// once per colour pass: 9 exchanges of ~3 MB per step instead of ~200 of 64 KB.
static int halo_exchange(fluid_ctx* c, std::initializer_list<const float*> arrays)
{
	if (c->slab_count <= 1) return FLUID_OK;
	if (!c->exchange) return fail(FLUID_ERR_INVALID, "this context is a row slab but no exchange callback is set (fluid_ctx_set_exchange)");
	int ids[fluid_ctx::kMaxFields], n = 0;
	for (const float* a : arrays)
		for (int f = 0; f < c->n_fields; f++)
			if (c->f[f] == a) ids[n++] = f;
	ProfScope ps(c, FLUID_PROF_EXCHANGE);
	if (!c->exchange_stream_ordered) CU_TRY(cudaStreamSynchronize(c->stream)); // the transport reads the owned rows
	const int rc = c->exchange(c->exchange_user, n, ids);
	if (rc != 0) return fail(FLUID_ERR_CUDA, "halo exchange callback failed with code %d", rc);
	return FLUID_OK;
}

// VelStep1 (FluidSystem.cpp:354-376).  The four pointers are locals: the
// reference's two Swap() calls per FlipFlopArr become local pointer swaps, so
// the result lands in the arrays passed as u and v.
static int dev_vel_step(fluid_ctx* c, float* u, float* v, float* u0, float* v0, float visc, float dt)
{
	// slabs: u, v (and the source arrays, which every slab fills itself) are valid over the whole halo on entry
	FL_TRY(dev_add_source2(c, u, u0, v, v0, dt));
	std::swap(u, u0);
	const float a = diffusion_a(c, visc, dt);
	std::swap(v, v0);
	if (c->overlap_density) { // the two diffusions are independent: v on a second lane beside u
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
		const int rc = dev_lin_solve(c, 2, v, v0, a, 1 + 4 * a);
		c->lane = 0;
		c->stream = main_stream;
		FL_TRY(rc);
		CU_TRY(cudaEventRecord(c->ev_join2, c->lane_stream));
		FL_TRY(dev_lin_solve(c, 1, u, u0, a, 1 + 4 * a));
		CU_TRY(cudaStreamWaitEvent(c->stream, c->ev_join2, 0));
	} else {
		FL_TRY(dev_lin_solve(c, 1, u, u0, a, 1 + 4 * a));
		FL_TRY(dev_lin_solve(c, 2, v, v0, a, 1 + 4 * a));
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
static int dev_dens_diffuse(fluid_ctx* c, float* x, float* x0, float diff, float dt)
{
	FL_TRY(dev_add_source(c, x, x0, dt));
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

static int dev_dens_step(fluid_ctx* c, float* x, float* x0, const float* u, const float* v, float diff, float dt)
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
static int dev_vort_confinement(fluid_ctx* c, float* u, float* v, float vorticity)
{
	FL_TRY(ensure_vort_scratch(c));
	ProfScope ps(c, FLUID_PROF_OTHER);
	LAUNCH(k_vort_curl, cell_grid(c), kBlk2D, 0, c->stream, c->g, u, v, c->vscr[0], c->vscr[1]);
	LAUNCH(k_vort_dir, cell_grid(c), kBlk2D, 0, c->stream, c->g, c->vscr[1], c->vscr[2], c->vscr[3]);
	LAUNCH(k_vort_apply, cell_grid(c), kBlk2D, 0, c->stream, c->g, u, v, c->vscr[0], c->vscr[2], c->vscr[3], vorticity);
	return FLUID_OK;
}

static int dev_damp(fluid_ctx* c, float* vel, float amount, float dt)
{
	const float f = powf(1 - amount, dt); // host libm, as the reference (FluidSystem.cpp:1181)
	ProfScope ps(c, FLUID_PROF_OTHER);
	LAUNCH(k_scale, flat_blocks(c), 256, 0, c->stream, (float4*) vel, f, n4_of(c));
	return FLUID_OK;
}

static int sync_sources(fluid_ctx* c)
{
	if (!c->src_dirty) return FLUID_OK;
	std::vector<SourceDev> flat;
	std::vector<int> first(c->batch + 1, 0);
	for (int s = 0; s < c->batch; s++) {
		first[s] = (int) flat.size();
		flat.insert(flat.end(), c->src[s].begin(), c->src[s].end());
	}
	first[c->batch] = (int) flat.size();
	if (c->d_src) CU_TRY(cudaFree(c->d_src));
	c->d_src = nullptr;
	if (!c->d_first) CU_TRY(cudaMalloc(&c->d_first, sizeof(int) * (c->batch + 1)));
	if (!flat.empty()) {
		CU_TRY(cudaMalloc(&c->d_src, sizeof(SourceDev) * flat.size()));
		CU_TRY(cudaMemcpy(c->d_src, flat.data(), sizeof(SourceDev) * flat.size(), cudaMemcpyHostToDevice));
	}
	CU_TRY(cudaMemcpy(c->d_first, first.data(), sizeof(int) * first.size(), cudaMemcpyHostToDevice));
	c->n_src = (int) flat.size();
	c->src_dirty = false;
	return FLUID_OK;
}

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

// FluidSystem::Update (FluidSystem.cpp:101-128) for FSType::ORIGINAL.
// FluidSystem::Update (FluidSystem.cpp:101-128): forcing, [vorticity], VelStep, DensStep per fluid, [damping].
// The densities' add_source + Blur touch neither velocity array, so on a single-GPU context they are enqueued on
// a side stream (fork after the forcing, join before the advection) and share the GPU with VelStep.
// `dens_ready`: an event (of another stream) after which the current density arrays are valid, or nullptr;
// `after_velocity`: called once VelStep has been enqueued on c->stream, or empty.  fluid_update() uses both to
// slip its bus transfers in; everybody else calls dev_update.
static int dev_update_hooks(fluid_ctx* c, float dt, cudaEvent_t dens_ready, const std::function<int()>& after_velocity)
{
	{
		ProfScope ps(c, FLUID_PROF_OTHER);
		CU_TRY(cudaMemsetAsync(c->f[FLUID_U_PREV], 0, c->field_bytes, c->stream));
		CU_TRY(cudaMemsetAsync(c->f[FLUID_V_PREV], 0, c->field_bytes, c->stream));
		for (int fl = 0; fl < c->n_fluids; fl++) CU_TRY(cudaMemsetAsync(c->f[FLUID_DENS_PREV_OF(fl)], 0, c->field_bytes, c->stream));
		if (c->n_src > 0) {
			FieldPtrs fp = {};
			fp.cur[0] = c->f[FLUID_U];
			fp.prev[0] = c->f[FLUID_U_PREV];
			fp.cur[1] = c->f[FLUID_V];
			fp.prev[1] = c->f[FLUID_V_PREV];
			for (int fl = 0; fl < c->n_fluids; fl++) {
				fp.cur[2 + fl] = c->f[FLUID_DENS_OF(fl)];
				fp.prev[2 + fl] = c->f[FLUID_DENS_PREV_OF(fl)];
			}
			LAUNCH(k_apply_sources, cdiv(c->batch, 64), 64, 0, c->stream, c->g, fp, c->d_src, c->d_first, c->batch);
		}
	}
	const bool concurrent = c->overlap_density; // slabs too: the halo exchanges of the densities come after the join
	auto diffuse_all = [&]() -> int {
		for (int fl = 0; fl < c->n_fluids; fl++) // DensStep1's loop over fluid_count (FluidSystem.cpp:291)
			FL_TRY(dev_dens_diffuse(c, c->f[FLUID_DENS_OF(fl)], c->f[FLUID_DENS_PREV_OF(fl)], fluid_diff(c, fl), dt));
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
	}
	if (c->vorticity != 0.0f) {
		FL_TRY(dev_vort_confinement(c, c->f[FLUID_U], c->f[FLUID_V], c->vorticity));
		FL_TRY(halo_exchange(c, { c->f[FLUID_U], c->f[FLUID_V] })); // three stencil phases consumed 3 halo rows
	}
	FL_TRY(dev_vel_step(c, c->f[FLUID_U], c->f[FLUID_V], c->f[FLUID_U_PREV], c->f[FLUID_V_PREV], visc_avg(c), dt));
	if (after_velocity) FL_TRY(after_velocity());
	if (concurrent) {
		CU_TRY(cudaStreamWaitEvent(c->stream, c->ev_join, 0));
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

static int dev_update(fluid_ctx* c, float dt) { return dev_update_hooks(c, dt, nullptr, {}); }

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
	if (c->d_divbins) cudaFree(c->d_divbins);
	for (auto& p : c->d_redo) if (p) cudaFree(p);
	if (c->side_stream) cudaStreamDestroy(c->side_stream);
	if (c->ev_fork) cudaEventDestroy(c->ev_fork);
	if (c->ev_join) cudaEventDestroy(c->ev_join);
	if (c->d_pixels) cudaFree(c->d_pixels);
	if (c->d_lines) cudaFree(c->d_lines);
	if (c->d_src) cudaFree(c->d_src);
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

extern "C" int fluid_ctx_create(const fluid_params* p, fluid_ctx** out)
{
	if (!p || !out) return fail(FLUID_ERR_INVALID, "null argument");
	*out = nullptr;
	if (p->N < 1 || p->N > 32768) return fail(FLUID_ERR_INVALID, "N=%d out of range [1,32768]", p->N);
	if (p->batch < 1 || p->batch > 65535) return fail(FLUID_ERR_INVALID, "batch=%d out of range [1,65535]", p->batch);
	if (p->gs_iters < 1) return fail(FLUID_ERR_INVALID, "gs_iters=%d must be >= 1", p->gs_iters);
	if (p->gs_order != FLUID_GS_WAVEFRONT && p->gs_order != FLUID_GS_REDBLACK) return fail(FLUID_ERR_INVALID, "unknown gs_order %d", p->gs_order);
	if (p->advect_mode != FLUID_ADVECT_MIDPOINT && p->advect_mode != FLUID_ADVECT_EULER) return fail(FLUID_ERR_INVALID, "unknown advect_mode %d", p->advect_mode);
	for (int r : p->reserved) if (r != 0) return fail(FLUID_ERR_INVALID, "reserved fields must be 0");
	if (p->n_fluids < 0 || p->n_fluids > FLUID_MAX_FLUIDS) return fail(FLUID_ERR_INVALID, "n_fluids=%d out of range [1,%d]", p->n_fluids, FLUID_MAX_FLUIDS);
	if (p->slab_rows < 0 || p->slab_rows > 1048576) return fail(FLUID_ERR_INVALID, "slab_rows=%d out of range", p->slab_rows);
	if (p->slab_count < 0 || (p->slab_count > 1 && (p->slab_index < 0 || p->slab_index >= p->slab_count)))
		return fail(FLUID_ERR_INVALID, "slab_index=%d / slab_count=%d is not a valid slab", p->slab_index, p->slab_count);
	if (p->slab_count > 1) {
		const int halo = p->slab_halo > 0 ? p->slab_halo : 48;
		if (halo < 44) return fail(FLUID_ERR_INVALID, "slab_halo=%d is too small: a projection consumes 42 halo rows (20 red-black sweeps + 2)", halo);
		if (p->slab_rows < halo + 1) return fail(FLUID_ERR_INVALID, "slab_rows=%d must exceed the halo depth %d", p->slab_rows, halo);
		if (p->batch != 1) return fail(FLUID_ERR_UNSUPPORTED, "row slabs and ensembles cannot be combined");
		if (p->gs_order != FLUID_GS_REDBLACK) return fail(FLUID_ERR_UNSUPPORTED, "row slabs need FLUID_GS_REDBLACK (the reference order is a pipeline across slabs)");
		if (p->gs_iters > 20) return fail(FLUID_ERR_UNSUPPORTED, "row slabs support gs_iters <= 20 (the halo is consumed at 2 rows per sweep)");
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

#define CHECK_CTX(c)                                                                                              \
	do {                                                                                                      \
		if (!(c)) return fail(FLUID_ERR_INVALID, "null context");                                         \
		FL_TRY(ctx_bind(c));                                                                              \
	} while (0)

static int check_field_n(int f, int n) { return (f >= 0 && f < n) ? FLUID_OK : fail(FLUID_ERR_INVALID, "field id %d out of range [0,%d)", f, n); }
#define check_field(f) check_field_n((f), c->n_fields)

static int check_distinct_n(int n_fields, std::initializer_list<int> ids)
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
		v.push_back({ slot, s.i, s.j, s.mode, s.value });
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

// FluidRenderer::DoBuffer without a window: colourise + velocity lines of ensemble member `sim` into
// device buffers, then (optionally) to dense host arrays.
extern "C" int fluid_ctx_render(fluid_ctx* c, int sim, const float* color_rgb, float* host_pixels, float* host_lines)
{
	CHECK_CTX(c);
	if (!color_rgb || !host_pixels) return fail(FLUID_ERR_INVALID, "null pointer");
	if (sim < 0 || sim >= c->batch) return fail(FLUID_ERR_INVALID, "sim %d out of range [0,%d)", sim, c->batch);
	if (c->slab_count > 1) return fail(FLUID_ERR_UNSUPPORTED, "render a slab by downloading it; the display buffer is per whole grid");
	const size_t cells = (size_t) (c->g.N + 2) * (c->g.NY + 2);
	if (!c->d_pixels) CU_TRY(cudaMalloc(&c->d_pixels, cells * 3 * sizeof(float)));
	if (host_lines && !c->d_lines) CU_TRY(cudaMalloc(&c->d_lines, cells * 4 * sizeof(float)));
	const long long off = (long long) sim * c->g.plane;
	LAUNCH(k_colourise, dim3(cdiv(c->g.N + 2, 32), cdiv(c->g.NY + 2, 8), 1), kBlk2D, 0, c->stream, c->g, c->f[FLUID_DENS] + off,
	       c->f[FLUID_U] + off, c->f[FLUID_V] + off, color_rgb[0], color_rgb[1], color_rgb[2], c->d_pixels,
	       host_lines ? c->d_lines : nullptr);
	CU_TRY(cudaMemcpyAsync(host_pixels, c->d_pixels, cells * 3 * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
	if (host_lines) CU_TRY(cudaMemcpyAsync(host_lines, c->d_lines, cells * 4 * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
	CU_TRY(cudaStreamSynchronize(c->stream));
	return FLUID_OK;
}

extern "C" int fluid_ctx_set_stream_kernel(fluid_ctx* c, int on)
{
	CHECK_CTX(c);
	c->rb_stream = on != 0;
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
	for (int s = 0; s < nsteps; s++) FL_TRY(dev_update(c, dt));
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
	if (c->slab_count > 1) {
		int oob = 0;
		CU_TRY(cudaMemcpy(&oob, c->d_oob, sizeof(int), cudaMemcpyDeviceToHost));
		if (oob) {
			CU_TRY(cudaMemset(c->d_oob, 0, sizeof(int)));
			return fail(FLUID_ERR_UNSUPPORTED, "advect: a backtrace reached rows that are not stored on this slab "
							   "(dt*N*|v| exceeded the %d-row halo); results are invalid -- use a deeper slab_halo", c->slab_halo);
		}
	}
	return FLUID_OK;
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

// ---------------------------------------------------------------------------
// Red-black lin_solve for callers that are not allowed to allocate.
//
// Who this is for: code that runs after its memory budget has been fixed --
// a real-time loop whose frame must not block on the allocator, a plug-in that
// lives inside somebody else's pre-sized arena, a host that has handed the whole
// GPU to a pool allocator -- and therefore cannot let a library call cudaMalloc
// behind its back (an allocation can fail late, takes a device-wide lock, and
// costs milliseconds).  The regular red-black path needs one scratch field for
// its tile ping-pong and allocates it per context; here that scratch is a fixed
// array inside the library image instead, which is why the grid size is capped.
//
// The array below is part of the loaded module: every GPU that loads
// libfluid_cuda.so carries these 4.3 MiB until the process exits, used or not.
// ---------------------------------------------------------------------------
#ifndef FLUID_STATIC_MAX_N
#define FLUID_STATIC_MAX_N 1024
#endif
#define FLUID_STATIC_PITCH (((XO + FLUID_STATIC_MAX_N + 2 + 31) / 32) * 32)
static __device__ __align__(256) float g_static_scratch[(size_t) FLUID_STATIC_PITCH * (FLUID_STATIC_MAX_N + 2)];

// In host code `g_static_scratch` is only a handle for the runtime, not a device
// address (a kernel handed &g_static_scratch[0] from the host would fault, or
// quietly use host memory).  The real address is asked for once per device, with
// that device current, and remembered here.  No heap is involved: fixed table.
#define FLUID_STATIC_MAX_DEVICES 64
static struct StaticSetup {
	bool ready;
	float* scratch; // device address of g_static_scratch on this device
	int num_sms;
} s_static[FLUID_STATIC_MAX_DEVICES];

static int static_setup(int dev, const StaticSetup** out)
{
	StaticSetup& st = s_static[dev];
	if (!st.ready) {
		void* p = nullptr;
		CU_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
		if (!p) return fail(FLUID_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
		int sms = 0;
		CU_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
		CU_TRY(cudaFuncSetAttribute(k_rb_tile<2, kRbRY, kRbRX, kRbNT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kRbSmem));
		CU_TRY(cudaFuncSetAttribute(k_rb_tile<1, kRbRY, kRbRX, kRbNT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kRbSmem));
		st.scratch = (float*) p;
		st.num_sms = sms;
		st.ready = true;
	}
	*out = &st;
	return FLUID_OK;
}

// TMA descriptor of one single-member pitched field, written into caller storage (stack).
static int encode_static_tmap(const float* base, const GridDesc& g, CUtensorMap* m)
{
	EncodeTiledFn enc = nullptr;
	FL_TRY(get_encode_fn(&enc));
	const cuuint64_t dims[3] = { (cuuint64_t) g.pitch, (cuuint64_t) (g.NY + 2), 1 };
	const cuuint64_t strides[2] = { (cuuint64_t) g.pitch * sizeof(float), (cuuint64_t) g.plane * sizeof(float) };
	const cuuint32_t box[3] = { (cuuint32_t) kRbRX, (cuuint32_t) kRbRY, 1 };
	const cuuint32_t estr[3] = { 1, 1, 1 };
	CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*) base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
			 CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
	if (r != CUDA_SUCCESS) return fail(FLUID_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int) r);
	return FLUID_OK;
}

template <int S>
static void static_tile_pass(const StaticSetup* st, const GridDesc& g, int b, const CUtensorMap& tm_in, float* out,
			     const CUtensorMap& tm_x0, float a, float cc, cudaStream_t stream)
{
	constexpr int TY = kRbRY - 4 * S, TX = kRbRX - 2 * (((2 * S + 3) / 4) * 4);
	const int ntx = cdiv(g.N, TX), nty = cdiv(g.NY, TY), ntiles = ntx * nty;
	DivC dv; // plain IEEE division: establishing a verified window would need a device buffer
	dv.c = cc;
	dv.y = 0.0f;
	dv.lo = dv.range = 0;
	LAUNCH((k_rb_tile<S, kRbRY, kRbRX, kRbNT>), std::min(ntiles, st->num_sms), kRbNT, kRbSmem, stream, tm_in, tm_x0, g, b, out, a,
	       dv, ntx, nty, ntiles, 0, 0, 0, 0, nullptr, StreamRect{}, 0, 0);
}

extern "C" int fluid_lin_solve_rb_static(int N, int b, float* x_dev, const float* x0_dev, float a, float cc, int iters, void* stream_)
{
	if (N < 1 || N > FLUID_STATIC_MAX_N)
		return fail(FLUID_ERR_INVALID, "fluid_lin_solve_rb_static: N=%d is outside [1,%d], the size the built-in scratch array was "
					       "compiled for; use a context (fluid_ctx_lin_solve) for larger grids", N, FLUID_STATIC_MAX_N);
	if (iters < 1) return fail(FLUID_ERR_INVALID, "fluid_lin_solve_rb_static: iters=%d must be >= 1", iters);
	if (!x_dev || !x0_dev || x_dev == x0_dev) return fail(FLUID_ERR_INVALID, "fluid_lin_solve_rb_static: x and x0 must be two device arrays");
	if (((uintptr_t) x_dev & 15) || ((uintptr_t) x0_dev & 15))
		return fail(FLUID_ERR_INVALID, "fluid_lin_solve_rb_static: device arrays must be 16-byte aligned");
	int dev = -1;
	CU_TRY(cudaGetDevice(&dev));
	if (dev < 0 || dev >= FLUID_STATIC_MAX_DEVICES) return fail(FLUID_ERR_UNSUPPORTED, "device ordinal %d not supported", dev);
	const StaticSetup* st = nullptr;
	FL_TRY(static_setup(dev, &st));
	cudaStream_t stream = (cudaStream_t) stream_;

	GridDesc g;
	g.N = N;
	g.NY = N;
	g.pitch = ((XO + N + 2 + 31) / 32) * 32;
	g.plane = (long long) g.pitch * (N + 2);
	g.j0 = 0;
	g.NYG = N;
	g.walls = 3;
	const int limit = a > 0 ? iters : 1;

	if (limit >= 2 && b >= 0 && b <= 2) {
		// The array still holds whatever the previous call (any N) left there.  The passes overwrite
		// every cell they later read, but the pad columns would keep stale values; start from zeros
		// like a freshly created context does.
		CU_TRY(cudaMemsetAsync(st->scratch, 0, (size_t) g.plane * sizeof(float), stream));
		CUtensorMap tm_x, tm_s, tm_0;
		FL_TRY(encode_static_tmap(x_dev, g, &tm_x));
		FL_TRY(encode_static_tmap(st->scratch, g, &tm_s));
		FL_TRY(encode_static_tmap(x0_dev, g, &tm_0));
		int n2 = limit / 2, n1 = limit % 2; // passes of two sweeps and of one; their number must be even
		if ((n2 + n1) % 2) { n2 -= 1; n1 += 2; }
		bool in_is_x = true;
		for (int k = 0; k < n2 + n1; k++) {
			const CUtensorMap& tm = in_is_x ? tm_x : tm_s;
			float* out = in_is_x ? st->scratch : x_dev;
			if (k < n2) static_tile_pass<2>(st, g, b, tm, out, tm_0, a, cc, stream);
			else static_tile_pass<1>(st, g, b, tm, out, tm_0, a, cc, stream);
			in_is_x = !in_is_x;
		}
	} else {
		// a single sweep, or a border mode the tile kernel does not maintain: in-place colour kernels, no scratch at all
		const dim3 grid(cdiv(g.pitch / 4, 32), cdiv(g.NY, 8), 1);
		for (int k = 0; k < limit; k++) {
			LAUNCH(k_rb_half, grid, kBlk2D, 0, stream, g, b, x_dev, x0_dev, a, cc, 0);
			LAUNCH(k_rb_half, grid, kBlk2D, 0, stream, g, b, x_dev, x0_dev, a, cc, 1);
		}
	}
	CU_TRY(cudaGetLastError());
	return FLUID_OK;
}

// ---------------------------------------------------------------------------
// fluid_multi: several GPUs from ONE plain-C host process.
//
// The same row-slab contexts as the one-process-per-GPU (NCCL) arrangement, but
// the halo transport is native: peer copies (cudaMemcpyPeerAsync over NVLink)
// ordered purely by CUDA events.  Each slab is stepped by its own host thread
// (the step functions are blocking enqueue loops); at an exchange point a thread
//   1. records "my owned rows are final" on its stream,
//   2. meets the others at a host barrier (so that every such event exists),
//   3. makes its stream wait for its neighbours' events and pulls their boundary
//      rows into its halo rows,
//   4. records "I have read my neighbours' rows", meets the barrier again and
//      makes its stream wait for the neighbours' versions of that event before
//      its next kernel may overwrite the rows they were reading.
// No stream is ever drained and the host never waits for the device: the step
// stays one asynchronous stream of work per GPU.  No counterpart in the
// reference (single process, single thread, no GPU).
// ---------------------------------------------------------------------------
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>

namespace {
class HostBarrier {
public:
	explicit HostBarrier(int n) : n_(n) {}
	// returns false when the barrier has been broken by a failing participant
	bool wait()
	{
		std::unique_lock<std::mutex> lk(mu_);
		if (broken_) return false;
		const unsigned gen = gen_;
		if (++count_ == n_) {
			count_ = 0;
			gen_++;
			cv_.notify_all();
			return true;
		}
		cv_.wait(lk, [&] { return gen_ != gen || broken_; });
		return !broken_;
	}
	void abort()
	{
		std::lock_guard<std::mutex> lk(mu_);
		broken_ = true;
		cv_.notify_all();
	}
	void reset()
	{
		std::lock_guard<std::mutex> lk(mu_);
		broken_ = false;
		count_ = 0;
	}

private:
	std::mutex mu_;
	std::condition_variable cv_;
	int n_, count_ = 0;
	unsigned gen_ = 0;
	bool broken_ = false;
};
} // namespace

struct fluid_multi {
	std::vector<fluid_ctx*> slab;
	std::vector<cudaEvent_t> ready, done;
	HostBarrier* barrier = nullptr;
	struct Link { fluid_multi* m; int k; };
	std::vector<Link> links;
	int N = 0, rows = 0;
};

// the exchange callback of slab k (runs on slab k's host thread, inside its step call)
static int multi_exchange(void* user, int nfields, const int* fields)
{
	fluid_multi::Link* ln = (fluid_multi::Link*) user;
	fluid_multi* m = ln->m;
	const int k = ln->k, n = (int) m->slab.size();
	fluid_ctx* c = m->slab[k];
	if (cudaEventRecord(m->ready[k], c->stream) != cudaSuccess) return 10;
	if (!m->barrier->wait()) return 11;
	for (int nb : { k - 1, k + 1 })
		if (nb >= 0 && nb < n && cudaStreamWaitEvent(c->stream, m->ready[nb], 0) != cudaSuccess) return 12;
	for (int q = 0; q < nfields; q++) {
		float *sd, *rd, *su, *ru;
		size_t cnt;
		if (fluid_ctx_halo_regions(c, fields[q], &sd, &rd, &su, &ru, &cnt) != FLUID_OK) return 13;
		if (k > 0) { // my lower halo <- lower neighbour's top owned rows
			float *nsd, *nrd, *nsu, *nru;
			size_t ncnt;
			if (fluid_ctx_halo_regions(m->slab[k - 1], fields[q], &nsd, &nrd, &nsu, &nru, &ncnt) != FLUID_OK) return 13;
			if (cudaSetDevice(c->device) != cudaSuccess) return 14; // halo_regions bound the neighbour's device
			if (cudaMemcpyPeerAsync(rd, c->device, nsu, m->slab[k - 1]->device, cnt * sizeof(float), c->stream) != cudaSuccess) return 15;
		}
		if (k + 1 < n) { // my upper halo <- upper neighbour's bottom owned rows
			float *nsd, *nrd, *nsu, *nru;
			size_t ncnt;
			if (fluid_ctx_halo_regions(m->slab[k + 1], fields[q], &nsd, &nrd, &nsu, &nru, &ncnt) != FLUID_OK) return 13;
			if (cudaSetDevice(c->device) != cudaSuccess) return 14;
			if (cudaMemcpyPeerAsync(ru, c->device, nsd, m->slab[k + 1]->device, cnt * sizeof(float), c->stream) != cudaSuccess) return 15;
		}
	}
	if (cudaEventRecord(m->done[k], c->stream) != cudaSuccess) return 16;
	if (!m->barrier->wait()) return 17;
	for (int nb : { k - 1, k + 1 })
		if (nb >= 0 && nb < n && cudaStreamWaitEvent(c->stream, m->done[nb], 0) != cudaSuccess) return 18;
	return 0;
}

extern "C" int fluid_multi_destroy(fluid_multi* m)
{
	if (!m) return FLUID_OK;
	for (size_t k = 0; k < m->slab.size(); k++) {
		if (m->slab[k]) {
			cudaSetDevice(m->slab[k]->device);
			cudaStreamSynchronize(m->slab[k]->stream);
		}
	}
	for (size_t k = 0; k < m->ready.size(); k++) {
		if (k < m->slab.size() && m->slab[k]) cudaSetDevice(m->slab[k]->device);
		if (m->ready[k]) cudaEventDestroy(m->ready[k]);
		if (m->done[k]) cudaEventDestroy(m->done[k]);
	}
	for (auto* c : m->slab) fluid_ctx_destroy(c);
	delete m->barrier;
	delete m;
	return FLUID_OK;
}

extern "C" int fluid_multi_create(const fluid_params* p, int ndev, const int* devices, fluid_multi** out)
{
	if (!p || !out || !devices || ndev < 1) return fail(FLUID_ERR_INVALID, "bad arguments");
	*out = nullptr;
	if (p->slab_rows < 1) return fail(FLUID_ERR_INVALID, "fluid_multi_create: set slab_rows (rows per device); the global grid is N x slab_rows*ndev");
	fluid_multi* m = new (std::nothrow) fluid_multi();
	if (!m) return fail(FLUID_ERR_NOMEM, "host allocation failed");
	m->N = p->N;
	m->rows = p->slab_rows;
	m->barrier = new HostBarrier(ndev);
	m->slab.assign(ndev, nullptr);
	m->ready.assign(ndev, nullptr);
	m->done.assign(ndev, nullptr);
	m->links.resize(ndev);
	int rc = FLUID_OK;
	for (int k = 0; k < ndev && rc == FLUID_OK; k++) {
		fluid_params q = *p;
		q.device = devices[k];
		q.slab_index = k;
		q.slab_count = ndev;
		q.gs_order = FLUID_GS_REDBLACK;
		if (ndev == 1) { // a single device: one whole grid of N x slab_rows
			q.slab_count = 0;
			q.slab_index = q.slab_halo = 0;
		}
		rc = fluid_ctx_create(&q, &m->slab[k]);
		if (rc != FLUID_OK) break;
		if (cudaEventCreateWithFlags(&m->ready[k], cudaEventDisableTiming) != cudaSuccess ||
		    cudaEventCreateWithFlags(&m->done[k], cudaEventDisableTiming) != cudaSuccess) {
			rc = fail(FLUID_ERR_CUDA, "cudaEventCreate failed: %s", cudaGetErrorString(cudaGetLastError()));
			break;
		}
		m->links[k] = { m, k };
		m->slab[k]->exchange = multi_exchange;
		m->slab[k]->exchange_user = &m->links[k];
		m->slab[k]->exchange_stream_ordered = true;
	}
	// direct peer access between neighbours (NVLink); without it the copies are staged by the driver
	for (int k = 0; k + 1 < ndev && rc == FLUID_OK; k++) {
		const int a = devices[k], b = devices[k + 1];
		if (a == b) continue;
		int can = 0;
		if (cudaDeviceCanAccessPeer(&can, a, b) == cudaSuccess && can) {
			cudaSetDevice(a);
			cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
			if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) rc = fail(FLUID_ERR_CUDA, "enabling peer access %d->%d: %s", a, b, cudaGetErrorString(e));
			cudaSetDevice(b);
			e = cudaDeviceEnablePeerAccess(a, 0);
			if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) rc = fail(FLUID_ERR_CUDA, "enabling peer access %d->%d: %s", b, a, cudaGetErrorString(e));
			cudaGetLastError();
		}
	}
	if (rc != FLUID_OK) {
		char keep[sizeof t_err];
		memcpy(keep, t_err, sizeof keep);
		fluid_multi_destroy(m);
		memcpy(t_err, keep, sizeof keep);
		return rc;
	}
	*out = m;
	return FLUID_OK;
}

#define CHECK_MULTI(m)                                                                                            \
	do {                                                                                                      \
		if (!(m)) return fail(FLUID_ERR_INVALID, "null multi-GPU handle");                                \
	} while (0)

extern "C" int fluid_multi_count(fluid_multi* m) { return m ? (int) m->slab.size() : 0; }

extern "C" int fluid_multi_set_sources(fluid_multi* m, int n, const fluid_source* src)
{
	CHECK_MULTI(m);
	for (auto* c : m->slab) FL_TRY(fluid_ctx_set_sources(c, -1, n, src)); // global rows; each slab applies what it stores
	return FLUID_OK;
}

extern "C" int fluid_multi_upload(fluid_multi* m, int field, const float* host_global)
{
	CHECK_MULTI(m);
	for (auto* c : m->slab) FL_TRY(fluid_ctx_upload(c, field, 0, host_global));
	return FLUID_OK;
}

extern "C" int fluid_multi_download(fluid_multi* m, int field, float* host_global)
{
	CHECK_MULTI(m);
	for (auto* c : m->slab) FL_TRY(fluid_ctx_download(c, field, 0, host_global));
	return FLUID_OK;
}

// fluid_ctx_step on every slab, one host thread per slab; returns when all work is ENQUEUED.
extern "C" int fluid_multi_step(fluid_multi* m, float dt, int nsteps)
{
	CHECK_MULTI(m);
	const int n = (int) m->slab.size();
	std::vector<int> rcs(n, FLUID_OK);
	std::vector<std::string> errs(n);
	m->barrier->reset();
	auto work = [&](int k) {
		rcs[k] = fluid_ctx_step(m->slab[k], dt, nsteps);
		if (rcs[k] != FLUID_OK) {
			errs[k] = t_err; // thread-local message of the worker
			m->barrier->abort();
		}
	};
	std::vector<std::thread> th;
	for (int k = 1; k < n; k++) th.emplace_back(work, k);
	work(0);
	for (auto& t : th) t.join();
	for (int k = 0; k < n; k++)
		if (rcs[k] != FLUID_OK && errs[k].find("callback failed with code 11") == std::string::npos &&
		    errs[k].find("callback failed with code 17") == std::string::npos)
			return fail(rcs[k], "slab %d: %s", k, errs[k].c_str());
	for (int k = 0; k < n; k++)
		if (rcs[k] != FLUID_OK) return fail(rcs[k], "slab %d: %s", k, errs[k].c_str());
	return FLUID_OK;
}

extern "C" int fluid_multi_sync(fluid_multi* m)
{
	CHECK_MULTI(m);
	for (auto* c : m->slab) FL_TRY(fluid_ctx_sync(c));
	return FLUID_OK;
}

// Device time of the last fluid_multi_step: the slowest slab (max over devices), in ms.
extern "C" int fluid_multi_last_step_ms(fluid_multi* m, float* ms)
{
	CHECK_MULTI(m);
	if (!ms) return fail(FLUID_ERR_INVALID, "null pointer");
	float worst = 0;
	for (auto* c : m->slab) {
		float t = 0;
		FL_TRY(fluid_ctx_last_step_ms(c, &t));
		worst = std::max(worst, t);
	}
	*ms = worst;
	return FLUID_OK;
}
