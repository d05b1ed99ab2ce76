/*
 * Headless host program in plain C: the part of the reference's main loop that
 * matters to the solver (SystemManager::Run -> Update -> FluidSystem::Update,
 * SystemManager.cpp:23-87,122-125) with the window, input and renderer left
 * out.  It owns the loop and the forcing and calls only the extern "C" layer
 * of include/fluid_cuda.h -- this is what "host code stays in C" looks like.
 *
 * usage: fluid_host_demo [N=64] [steps=100] [visc=1e-4] [diff=1e-4] [order=0|1] [batch=1] [gpus=1]
 *        gpus > 1: N rows per GPU (global grid N x N*gpus), red-black, peer-copy halos (fluid_multi_*)
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "fluid_cuda.h"

static uint64_t fnv1a64(const void *p, size_t n)
{
	const unsigned char *b = (const unsigned char *) p;
	uint64_t h = 0xcbf29ce484222325ULL;
	for (size_t k = 0; k < n; k++) {
		h ^= b[k];
		h *= 0x100000001b3ULL;
	}
	return h;
}

#define CHECK(call)                                                                        \
	do {                                                                               \
		if ((call) != FLUID_OK) {                                                  \
			fprintf(stderr, "%s failed: %s\n", #call, fluid_last_error());     \
			return 1;                                                          \
		}                                                                          \
	} while (0)

/* several GPUs from this one C process: row slabs, halos by peer copies */
static int run_multi(fluid_params p, int steps, int gpus)
{
	int dev[64];
	if (gpus > 64) gpus = 64;
	for (int k = 0; k < gpus; k++) dev[k] = k;
	p.slab_rows = p.N;
	p.batch = 1;
	fluid_multi *m;
	CHECK(fluid_multi_create(&p, gpus, dev, &m));
	const int cx = p.N / 2, cy = p.N * gpus / 2;
	fluid_source src[3] = {
		{ FLUID_DENS, cx, cy, FLUID_SRC_PREV_SET, 100.0f },
		{ FLUID_U, cx, cy, FLUID_SRC_PREV_SET, 5.0f },
		{ FLUID_V, cx, cy, FLUID_SRC_PREV_SET, 5.0f },
	};
	CHECK(fluid_multi_set_sources(m, 3, src));
	CHECK(fluid_multi_step(m, 0.1f, 3)); /* warm-up */
	CHECK(fluid_multi_sync(m));
	CHECK(fluid_multi_step(m, 0.1f, steps));
	CHECK(fluid_multi_sync(m));
	float ms = 0;
	CHECK(fluid_multi_last_step_ms(m, &ms));
	printf("N=%d x %d rows on %d GPUs, steps=%d red-black: %.3f ms/step, %.4g cell-updates/s\n", p.N, p.N * gpus, gpus, steps,
	       ms / (steps > 0 ? steps : 1), steps > 0 ? (double) p.N * p.N * gpus * steps / (ms * 1e-3) : 0.0);
	CHECK(fluid_multi_destroy(m));
	return 0;
}

int main(int argc, char **argv)
{
	fluid_params p;
	fluid_params_default(&p);
	p.N = argc > 1 ? atoi(argv[1]) : 64;
	int steps = argc > 2 ? atoi(argv[2]) : 100;
	p.visc = argc > 3 ? (float) atof(argv[3]) : 1e-4f;
	p.diff = argc > 4 ? (float) atof(argv[4]) : 1e-4f;
	p.gs_order = argc > 5 ? atoi(argv[5]) : FLUID_GS_WAVEFRONT;
	p.batch = argc > 6 ? atoi(argv[6]) : 1;
	const int gpus = argc > 7 ? atoi(argv[7]) : 1;
	if (gpus > 1) return run_multi(p, steps, gpus);

	fluid_ctx *ctx;
	CHECK(fluid_ctx_create(&p, &ctx));

	/* source S1: every step dens_prev = 100, u_prev = v_prev = 5 at the centre cell */
	const int c = p.N / 2;
	fluid_source src[3] = {
		{ FLUID_DENS, c, c, FLUID_SRC_PREV_SET, 100.0f },
		{ FLUID_U, c, c, FLUID_SRC_PREV_SET, 5.0f },
		{ FLUID_V, c, c, FLUID_SRC_PREV_SET, 5.0f },
	};
	CHECK(fluid_ctx_set_sources(ctx, -1, 3, src));

	CHECK(fluid_ctx_step(ctx, 0.1f, steps));
	float ms = 0;
	CHECK(fluid_ctx_last_step_ms(ctx, &ms));

	const size_t n = (size_t) (p.N + 2) * (p.N + 2);
	float *h = (float *) malloc(n * sizeof(float));
	const char *names[3] = { "u", "v", "dens" };
	printf("N=%d steps=%d order=%s batch=%d: %.3f ms/step, %.4g cell-updates/s\n", p.N, steps,
	       p.gs_order ? "red-black" : "wavefront", p.batch, ms / (steps > 0 ? steps : 1),
	       steps > 0 ? (double) p.N * p.N * p.batch * steps / (ms * 1e-3) : 0.0);
	for (int f = 0; f < 3; f++) {
		CHECK(fluid_ctx_download(ctx, f, 0, h));
		double mass = 0;
		for (int j = 1; j <= p.N; j++)
			for (int i = 1; i <= p.N; i++)
				mass += h[i + (size_t) (p.N + 2) * j];
		printf("  %-4s fnv1a64=%016llx interior sum=%.6g\n", names[f], (unsigned long long) fnv1a64(h, n * sizeof(float)), mass);
	}
	free(h);
	CHECK(fluid_ctx_destroy(ctx));
	return 0;
}
