// Relaxation: LinearSolve1 (FluidSystem.cpp:534-548) and Blur (FluidSystem.cpp:722-743) on resident
// fields -- which kernel family runs a solve (small-grid, streaming, tile, one colour per launch,
// reference-order wavefront), the pass schedule, the verified division and the redo lists.
#include "ctx.h"
#include <vector>

#include "kernels_relax.cuh"
#include "kernels_rowwave.cuh"
#include "kernels_rb_stream.cuh"
#include "kernels_rb_tile.cuh"
#include "kernels_blur_stream.cuh"
#include "kernels_small.cuh"

// ---- red-black tile path ---------------------------------------------------
constexpr int kRbRY = 64, kRbNT = 512;

// cuTensorMapEncodeTiled is fetched through the runtime so that the library
// has no link-time dependency on libcuda (it must load on a machine without a driver).
int get_encode_fn(EncodeTiledFn* out)
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

// The mismatch-free exponent window around 1.0 from the 256 per-exponent mismatch counts of k_verify_division.
static void window_from_bins(const unsigned int* bins, DivC* dv)
{
	dv->lo = dv->range = 0;
	dv->full = 1; // bins[256 ..]: k_verify_division_full
	for (int e = 256; e < 512; e++)
		if (bins[e] != 0) dv->full = 0;
	if (bins[127] != 0) return;
	int elo = 127, ehi = 127;
	while (elo > 1 && bins[elo - 1] == 0) elo--;
	while (ehi < 254 && bins[ehi + 1] == 0) ehi++;
	dv->lo = (uint32_t) elo << 23;
	dv->range = (uint32_t) (ehi - elo + 1) << 23;
}

constexpr size_t kMaxProvenDivisors = 64, kMaxSeenDivisors = 256;

static void remember_division(fluid_ctx* c, uint32_t key, const DivC& dv)
{
	if (c->divs.size() >= kMaxProvenDivisors) { // the least recently used entry leaves
		auto oldest = c->divs.begin();
		for (auto it = c->divs.begin(); it != c->divs.end(); ++it)
			if (it->second.last_use < oldest->second.last_use) oldest = it;
		c->divs.erase(oldest);
	}
	c->divs[key] = { dv, c->epoch };
	c->div_seen.erase(key);
}

// DivC for divisor cc: verified window if known, else range = 0 (callers use the exact division).
//
// The exhaustive proof costs 2^31 trial divisions for the window and 1.6e9 for the scaled sequence (about 3 ms of the whole
// GPU).  The reference's main loop feeds the solver the measured frame time (SystemManager.cpp:49-52), so an integration
// may see a new c = 1 + 4*dt*coef*N*N every frame.  Automatic mode therefore uses the QUICK proof (one binade of dividends
// instead of all of them, see grid.cuh), which is cheap enough to run for every new divisor and wait for.  Divisors the
// quick proof does not cover (c < 1 or >= 2^15) are proven exhaustively, in the background, once they have COME BACK --
// seen in an earlier frame or API call -- with at least 2.5e8 cell updates behind them; until that proof has landed, and for
// such divisors that never recur, solves use the exact division (tile / pass kernels).  Both tables are bounded.
// Mode 2 proves exhaustively at once and waits (tests, one-off solves).
int division_for(fluid_ctx* c, float cc, double cell_updates, DivC* out)
{
	out->c = cc;
	out->y = 1.0f / cc;
	out->rc = 1.0 / (double) cc;
	out->lo = 0;
	out->range = 0;
	out->full = 0;
	if (!(cc > 0.0f) || !isfinite(cc) || !isfinite(out->y) || c->fast_div_mode == 0) return FLUID_OK;
	uint32_t key;
	memcpy(&key, &cc, sizeof key);
	auto& pd = c->div_pending;
	if (pd.active && (c->fast_div_mode == 2 ? cudaEventSynchronize(pd.done) : cudaEventQuery(pd.done)) == cudaSuccess) {
		window_from_bins(pd.h_bins, &pd.dv); // a background proof has landed
		if (getenv("FLUID_DEBUG_FEEDBACK"))
			fprintf(stderr, "division proof for c = %.9g: window exponents [%u, %u), scaled sequence %s\n", pd.dv.c, pd.dv.lo >> 23,
				(pd.dv.lo + pd.dv.range) >> 23, pd.dv.full ? "proven" : "not proven");
		remember_division(c, pd.key, pd.dv);
		pd.active = false;
	}
	auto it = c->divs.find(key);
	if (it != c->divs.end()) {
		it->second.last_use = c->epoch;
		*out = it->second.dv;
		return FLUID_OK;
	}
	if (!c->d_divbins) CU_TRY(cudaMalloc(&c->d_divbins, 512 * sizeof(unsigned int)));
	if (c->fast_div_mode == 2) {
		CU_TRY(cudaMemsetAsync(c->d_divbins, 0, 512 * sizeof(unsigned int), c->stream));
		LAUNCH(k_verify_division, c->num_sms * 8, 256, 0, c->stream, *out, c->d_divbins);
		LAUNCH(k_verify_division_full, c->num_sms * 8, 256, 0, c->stream, *out, c->d_divbins + 256);
		unsigned int bins[512];
		CU_TRY(cudaMemcpyAsync(bins, c->d_divbins, sizeof bins, cudaMemcpyDeviceToHost, c->stream));
		CU_TRY(cudaStreamSynchronize(c->stream));
		window_from_bins(bins, out);
		remember_division(c, key, *out);
		return FLUID_OK;
	}
	// Automatic mode: the quick proof (grid.cuh: k_verify_division_quick), on the side stream, waited for -- ~0.1 ms of GPU and
	// one event wait per NEW divisor, so a run whose time step changes every frame (two new divisors per frame) keeps the
	// fast kernels.  Divisors it does not cover (c < 1, c >= 2^15) go the slow way below.
	if (cc >= 1.0f && cc < 32768.0f) {
		if (!pd.stream) {
			CU_TRY(cudaStreamCreateWithFlags(&pd.stream, cudaStreamNonBlocking));
			CU_TRY(cudaEventCreateWithFlags(&pd.done, cudaEventDisableTiming));
			CU_TRY(cudaMallocHost(&pd.h_bins, 512 * sizeof(unsigned int)));
		}
		if (pd.active) { // an exhaustive proof in flight uses the same buffers
			CU_TRY(cudaEventSynchronize(pd.done));
			pd.active = false;
		}
		const int ec = ilogbf(cc);
		const unsigned int lo_biased = (unsigned int) (27 + ec); // 2^(ec-100)
		CU_TRY(cudaMemsetAsync(c->d_divbins, 0, 2 * sizeof(unsigned int), pd.stream));
		LAUNCH(k_verify_division_quick, c->num_sms * 8, 256, 0, pd.stream, *out, (unsigned int) (ec + 3) << 23, lo_biased, c->d_divbins);
		CU_TRY(cudaMemcpyAsync(pd.h_bins, c->d_divbins, 2 * sizeof(unsigned int), cudaMemcpyDeviceToHost, pd.stream));
		CU_TRY(cudaEventRecord(pd.done, pd.stream));
		CU_TRY(cudaEventSynchronize(pd.done));
		if (pd.h_bins[0] == 0) {
			out->lo = lo_biased << 23;
			out->range = (255u - lo_biased) << 23;
		}
		out->full = (pd.h_bins[1] == 0) ? 1u : 0u;
		if (getenv("FLUID_DEBUG_FEEDBACK"))
			fprintf(stderr, "quick division proof for c = %.9g: window exponents [%u, %u), scaled sequence %s\n", cc, out->lo >> 23,
				(out->lo + out->range) >> 23, out->full ? "proven" : "not proven");
		remember_division(c, key, *out);
		return FLUID_OK;
	}
	if (c->div_seen.size() >= kMaxSeenDivisors && !c->div_seen.count(key)) c->div_seen.clear(); // divisors that never came back
	auto seen = c->div_seen.emplace(key, fluid_ctx::DivSeen{ 0.0, c->epoch }).first;
	seen->second.work += cell_updates;
	if (seen->second.first_epoch == c->epoch || seen->second.work < 2.5e8 || pd.active) return FLUID_OK;
	if (!pd.stream) {
		CU_TRY(cudaStreamCreateWithFlags(&pd.stream, cudaStreamNonBlocking));
		CU_TRY(cudaEventCreateWithFlags(&pd.done, cudaEventDisableTiming));
		CU_TRY(cudaMallocHost(&pd.h_bins, 512 * sizeof(unsigned int)));
	}
	pd.key = key;
	pd.dv = *out;
	CU_TRY(cudaMemsetAsync(c->d_divbins, 0, 512 * sizeof(unsigned int), pd.stream));
	LAUNCH(k_verify_division, c->num_sms * 8, 256, 0, pd.stream, *out, c->d_divbins);
	LAUNCH(k_verify_division_full, c->num_sms * 8, 256, 0, pd.stream, *out, c->d_divbins + 256);
	CU_TRY(cudaMemcpyAsync(pd.h_bins, c->d_divbins, 512 * sizeof(unsigned int), cudaMemcpyDeviceToHost, pd.stream));
	CU_TRY(cudaEventRecord(pd.done, pd.stream));
	pd.active = true;
	return FLUID_OK;
}

constexpr int kStreamHoldoff = 64, kStreamHoldoffMax = 8192;

// How may this solve use the streaming kernel?  0 = not at all (tile / pass kernels), 1 = windowed sequence with the redo
// list (segments that met a dividend outside the proven window are listed and recomputed exactly behind the launch),
// 2 = scaled sequence (exact on the fringe as well, two instructions more per update, nothing listed but NaN rows).
// Consumes the count of an earlier launch if it has arrived: when more than a twelfth of the segments came back (an
// ensemble of small grids is mostly fringe) the next kStreamHoldoff solves (twice as many after every trial that comes
// back the same way) use the scaled sequence if it is proven for this divisor (full_ok), else the tile / pass kernels.  Measured on 512 members of N=512, relaxation per frame: 25 ms
// on the tile kernel, 40 ms with an exact FP64 recomputation per flagged colour pass (round 2's first inline variant:
// nearly every pass of every row has some lane on the fringe, and the conversion pipe is narrow).
static int stream_allowed(fluid_ctx* c, int which, bool full_ok)
{
	auto& fb = c->feedback[which];
	if (fb.pending && cudaEventQuery(fb.arrived) == cudaSuccess) {
		fb.pending = false;
		if (getenv("FLUID_DEBUG_FEEDBACK")) fprintf(stderr, "feedback[%d]: %d of %lld segments redone\n", which, *fb.h_count, fb.segments);
		if (fb.segments > 0 && (long long) *fb.h_count * 12 > fb.segments) {
			fb.holdoff_len = std::min(std::max(fb.holdoff_len * 2, kStreamHoldoff), kStreamHoldoffMax);
			fb.holdoff = fb.holdoff_len;
		} else {
			fb.holdoff_len = fb.holdoff = 0; // the fringe has gone: back to the windowed sequence
		}
	}
	if (!c->rb_stream) return 0;
	if (c->stream_force_inline) return full_ok ? 2 : 1;
	if (fb.holdoff > 0) {
		fb.holdoff--;
		return full_ok ? 2 : 0;
	}
	// a trial at the end of a hold-off is ONE solve: its count takes a few solves to come home, and those stay as they were
	if (fb.holdoff_len > 0 && !fb.pending) fb.holdoff = 8;
	return 1;
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
	} else {
		// a thin grid (a row slab of a strongly scaled run: 1100 rows gave 380 tasks for 1776 warp slots): shorter
		// chunks, down to 64 rows (25 % more rows computed per task), until the tasks fill the resident warps
		nchunks = (int) std::max<long long>(nchunks, std::min<long long>(rows / 64, slots / per_chunk));
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
	pg.rc.nsims = c->batch;
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
		if (c->stream_inline) {
			LAUNCH((k_rb_stream<S, false, false, true>), cdiv(ntasks, 4), 128, 0, c->stream, c->g, pg.rc, in, x0, out, a, dv, ntasks, c->d_redo[rbuf], b);
			*segments = 0; // what is listed is not fringe: nothing to report
		} else {
			LAUNCH((k_rb_stream<S, false>), cdiv(ntasks, 4), 128, 0, c->stream, c->g, pg.rc, in, x0, out, a, dv, ntasks, c->d_redo[rbuf], b);
		}
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

int choose_rb_rx(int N)
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
	if (c->batch > 1 && S >= 4) { // ensembles: shorter unrolled blocks (see StreamRect::nsims)
		if (x_is_zero) LAUNCH((k_rb_stream<S, true, true, false, 2>), cdiv(ntasks, 4), 128, 0, c->stream, c->g, rc, in, x0, out, a, dv, ntasks, (int*) nullptr, b);
		else LAUNCH((k_rb_stream<S, true, false, false, 2>), cdiv(ntasks, 4), 128, 0, c->stream, c->g, rc, in, x0, out, a, dv, ntasks, (int*) nullptr, b);
		return FLUID_OK;
	}
	if (x_is_zero) LAUNCH((k_rb_stream<S, true, true>), cdiv(ntasks, 4), 128, 0, c->stream, c->g, rc, in, x0, out, a, dv, ntasks, (int*) nullptr, b);
	else LAUNCH((k_rb_stream<S, true, false>), cdiv(ntasks, 4), 128, 0, c->stream, c->g, rc, in, x0, out, a, dv, ntasks, (int*) nullptr, b);
	return FLUID_OK;
}

// Will a solve with these coefficients start with a pass that can be told "x is zero"?  (dev_project asks before
// it decides not to write p = 0.)
bool solve_accepts_zero_x(const fluid_ctx* c, float a, float cc, int limit)
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
	const int stream_mode = (dv.range != 0 && (dv.lo + dv.range) == (255u << 23)) ? stream_allowed(c, 0, dv.full != 0) : 0;
	const bool streaming = stream_mode != 0;
	c->stream_inline = (stream_mode == 2);
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
			// one pass more: split a 4-sweep pass into 2 + 2, else a 2-sweep pass into 1 + 1; with neither (one
			// single sweep left over after the power-of-two passes) a power-of-two pass gives way to 2 + 2
			if (n4 > 0) { n4 -= 1; n2 += 2; }
			else if (n2 > 0) { n2 -= 1; n1 += 2; }
			else if (p4 > 0) { p4 -= 1; n2 += 2; }
			else return fail(FLUID_ERR_CUDA, "internal: cannot make the number of passes even");
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

bool use_small_kernels(const fluid_ctx* c)
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
int dev_lin_solve(fluid_ctx* c, int b, float*& x, const float* x0, float a, float cc, bool x_is_zero)
{
	if (c->fs_type != FLUID_TYPE_ORIGINAL) return dev_lin_solve_b(c, b, x, x0, a, cc); // LinearSolve1B: cellInfo-masked
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
		const long long tickets = (long long) c->batch * limit * ntx;
		// One launch per solve (CTAs take (sweep, tile column) tickets and order themselves by progress counters) pays
		// where launches dominate: measured ms/step, persistent vs per-hyperplane launches: N=1024 6.4 vs 7.0, N=2048 12.8 vs
		// 11.4, N=8192 81 vs 47 -- a tile's waits, fences and L2-only loads cost more than the launch gaps they replace once
		// every launch carries thousands of tiles.  Automatic choice: grids up to 1024 cells wide; mode 2 forces it.
		// Row pipeline (kernels_rowwave.cuh): warps own 32-row bands per sweep and stream along the rows; one launch per solve.
		// Measured ms per step, row pipeline / tile tickets / launch per hyperplane: N=256 2.05 / 2.09 / 2.15, N=512 2.86 / 4.19 /
		// 4.65, N=1024 4.32 / 6.20 / 6.86, N=1536 6.1 / 8.49 / 8.77, N=2048 8.1-8.5 / 13.2 / 10.8, N=3072 14.9 / - / 14.6, N=4096
		// 23.4 / - / 19.0, N=8192 66.4 / - / 46.8 (profiles/r02_wavefront_modes.jsonl): the pipeline's duration is (bands +
		// sweeps) hand-offs of ~15 us plus one task, and a task runs at half its speed once every SM holds twelve of them.
		// Automatic: the pipeline up to N = 2560.
		const bool rowwave = c->wavefront_persistent == 3 || (c->wavefront_persistent == 1 && c->g.N >= 256 && c->g.N <= 2560 && c->g.NY >= 256);
		if (rowwave) {
			DivC dv;
			FL_TRY(division_for(c, cc, (double) c->g.N * c->g.NY * c->batch * limit, &dv));
			const int nb = rowwave_bands(c->g.NY);
			const long long tasks = (long long) c->batch * nb * limit;
			if (tasks < (1LL << 28)) {
				int*& wsync = c->d_wf_sync[c->lane ? 1 : 0];
				long long& wcap = c->wf_sync_capacity[c->lane ? 1 : 0];
				if (tasks + 2 > wcap) {
					if (wsync) CU_TRY(cudaFree(wsync));
					wsync = nullptr;
					CU_TRY(cudaMalloc(&wsync, sizeof(int) * (size_t) (tasks + 2)));
					wcap = tasks + 2;
				}
				CU_TRY(cudaMemsetAsync(wsync, 0, sizeof(int) * (size_t) (tasks + 2), c->stream));
				int e = 0;
				const bool pow2 = cc > 0.0f && isfinite(cc) && frexpf(cc, &e) == 0.5f && (1.0f / cc) >= 1.17549435e-38f;
				// diagnostics: FLUID_RW_STAMPS=<file> dumps four timestamps per task of every solve (blocking; see kernels_rowwave.cuh)
				static const char* stamp_path = getenv("FLUID_RW_STAMPS");
				unsigned long long* stamps = nullptr;
				if (stamp_path) {
					CU_TRY(cudaMalloc(&stamps, sizeof(unsigned long long) * 4 * (size_t) tasks));
					CU_TRY(cudaMemsetAsync(stamps, 0, sizeof(unsigned long long) * 4 * (size_t) tasks, c->stream));
				}
				static const int per_sm_env = getenv("FLUID_RW_CTAS_PER_SM") ? atoi(getenv("FLUID_RW_CTAS_PER_SM")) : 0; // (measurements only)
				// 12 CTAs fit an SM's shared memory, but a task runs faster the fewer share its SM, and the front of tasks that can
				// run at one time is narrow: ms per solve at 12 / 6 / 3 per SM: N=1024 1.56 / 1.55 / 1.47, N=2048 2.81 / 2.62 / 2.51,
				// N=4096 5.69 / 5.29 / 6.38, N=8192 12.0 / 12.9 / 21.5 (and 8 per SM: 5.35, 11.5); three per SM also leave room for
				// the u and v solves of a step to run side by side.
				const int per_sm = per_sm_env > 0 ? per_sm_env : (c->g.N <= 2560 ? 3 : 8);
				const int grid = (int) std::min<long long>(tasks, (long long) c->num_sms * per_sm);
				if (pow2) LAUNCH(k_gs_rowwave<2>, grid, 64, 0, c->stream, c->g, b, x, x0, a, dv, limit, nb, c->batch, wsync, c->d_oob, stamps);
				else if (dv.full) LAUNCH(k_gs_rowwave<1>, grid, 64, 0, c->stream, c->g, b, x, x0, a, dv, limit, nb, c->batch, wsync, c->d_oob, stamps);
				else LAUNCH(k_gs_rowwave<0>, grid, 64, 0, c->stream, c->g, b, x, x0, a, dv, limit, nb, c->batch, wsync, c->d_oob, stamps);
				c->rowwave_used = true;
				CU_TRY(cudaGetLastError());
				if (stamps) {
					std::vector<unsigned long long> h(4 * (size_t) tasks);
					CU_TRY(cudaMemcpyAsync(h.data(), stamps, sizeof(unsigned long long) * h.size(), cudaMemcpyDeviceToHost, c->stream));
					CU_TRY(cudaStreamSynchronize(c->stream));
					CU_TRY(cudaFree(stamps));
					if (FILE* f = fopen(stamp_path, "a")) {
						fprintf(f, "solve N=%d NY=%d nb=%d limit=%d batch=%d\n", c->g.N, c->g.NY, nb, limit, c->batch);
						for (long long t = 0; t < tasks; t++)
							fprintf(f, "%lld %llu %llu %llu %llu\n", t, h[4 * t], h[4 * t + 1], h[4 * t + 2], h[4 * t + 3]);
						fclose(f);
					}
				}
				return FLUID_OK;
			}
		}
		const bool persistent = c->wavefront_persistent == 2 || (c->wavefront_persistent == 1 && c->g.N <= 1024);
		if (persistent && tickets < (1LL << 28)) {
			int*& wsync = c->d_wf_sync[c->lane ? 1 : 0];
			long long& wcap = c->wf_sync_capacity[c->lane ? 1 : 0];
			if (tickets + 1 > wcap) {
				if (wsync) CU_TRY(cudaFree(wsync));
				wsync = nullptr;
				CU_TRY(cudaMalloc(&wsync, sizeof(int) * (size_t) (tickets + 1)));
				wcap = tickets + 1;
			}
			CU_TRY(cudaMemsetAsync(wsync, 0, sizeof(int) * (size_t) (tickets + 1), c->stream));
			const int grid = (int) std::min<long long>(tickets, (long long) c->num_sms * 32);
			if (T == 16) LAUNCH(k_gs_wavefront_persistent<16>, grid, 16, 0, c->stream, c->g, b, x, x0, a, cc, limit, ntx, nty, c->batch, wsync);
			else if (T == 32) LAUNCH(k_gs_wavefront_persistent<32>, grid, 32, 0, c->stream, c->g, b, x, x0, a, cc, limit, ntx, nty, c->batch, wsync);
			else LAUNCH(k_gs_wavefront_persistent<64>, grid, 64, 0, c->stream, c->g, b, x, x0, a, cc, limit, ntx, nty, c->batch, wsync);
			CU_TRY(cudaGetLastError());
			return FLUID_OK;
		}
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
	} else {
		nchunks = (int) std::max<long long>(nchunks, std::min<long long>(rows / 48, slots / per_chunk)); // thin grids, see pass_geometry
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
	rc->nsims = c->batch;
	return true;
}

int dev_blur(fluid_ctx* c, int b, float* x, float* x0, float a, float cc)
{
	if (c->fs_type != FLUID_TYPE_ORIGINAL) return dev_blur_b(c, b, x, x0, a, cc); // BlurB
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
	const int blur_mode = (window_ok && b >= 0 && b <= 2 && limit > 2 * kBlurL && blur_stream_geometry(c, &rc)) ? stream_allowed(c, 1, dv.full != 0) : 0;
	if (blur_mode != 0) {
		const int launches = 2 * ((limit - 1) / (2 * kBlurL));
		const long long ntasks = (long long) rc.per_sim * c->batch;
		FL_TRY(redo_reserve(c, 1, ntasks * rc.segs + 1));
		for (int l = 0; l < launches; l++) {
			if (l > 0) std::swap(dst, src);
			CU_TRY(cudaMemsetAsync(c->d_redo[1], 0, sizeof(int), c->stream));
			if (blur_mode == 2) // mostly fringe: the scaled sequence
				LAUNCH((k_blur_stream<kBlurL, true>), cdiv((int) ntasks, 4), 128, 0, c->stream, c->g, rc, src, dst, a, dv, (int) ntasks, c->d_redo[1], b);
			else
				LAUNCH(k_blur_stream<kBlurL>, cdiv((int) ntasks, 4), 128, 0, c->stream, c->g, rc, src, dst, a, dv, (int) ntasks, c->d_redo[1], b);
			// few entries, each a chain of 2L barrier-separated phases over ~5000 cells: wide CTAs keep the chain short
			LAUNCH((k_blur_redo<kBlurL, kBlurSeg>), 2 * c->num_sms, 1024, 0, c->stream, c->g, rc, src, dst, a, cc, c->d_redo[1], b);
		}
		done = launches * kBlurL;
		if (blur_mode == 1) FL_TRY(stream_report(c, 1, ntasks * rc.segs));
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
