// fluid_lin_solve_rb_static: the red-black solve for callers that must not allocate.  The scratch array,
// the set-up that asks the runtime for its device address and the entry point live in this one file.
#include "ctx.h"
#include "kernels_relax.cuh"
#include "kernels_rb_tile.cuh"

constexpr int kRbRY = 64, kRbRX = 256, kRbNT = 512;
constexpr size_t kRbSmem = (2 * (size_t) (kRbRY + 2) + kRbRY) * kRbRX * sizeof(float); // two x windows + x0 staging

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
	dv.rc = 1.0 / (double) cc;
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
