/* CPU restatement of div_scaled (fluid-simulation_b200/csrc/grid.cuh), operation for operation, checked against the
 * machine's own IEEE single division.  Test infrastructure (tests/test_division_cpu.py compiles and runs it).
 * usage: div_scaled_check <first bit pattern> <last bit pattern, exclusive> <stride> <divisor>...
 *        div_scaled_check seq <divisor>...
 * prints one line per divisor: "<divisor> <dividends tried> <mismatches> <ties adjusted>" */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

static inline float div_seq(float t, float c, float y)
{
	const float q0 = t * y;
	const float rn = fmaf(c, q0, -t); /* the residual with the opposite sign: zeros keep their sign */
	return fmaf(rn, -y, q0);
}

static inline float div_scaled(float ts, float c, float y, unsigned long long* adjusted)
{
	const float qs = div_seq(ts, c, y);
	float q = qs * 0x1p-64f;
	const float dl = fmaf(q, -0x1p64f, qs);
	const float r2 = fmaf(-c, qs, ts);
	const uint32_t dsign = f2u(dl) & 0x80000000u;
	const int away = u2f(f2u(r2) ^ dsign) > 0.0f;
	if (fabsf(dl) == 0x1p-86f && away) {
		q += u2f(dsign | 1u);
		++*adjusted;
	}
	return q;
}

/* "seq <divisor>...": the claim behind the quick proof (k_verify_division_quick) -- div_seq commutes with scaling the
 * dividend by a power of two from 2^(ec-100) up, so one binade decides all.  Per divisor: mismatches against t / c on five
 * whole binades, and mantissas whose quotient significand differs between a binade and [1,2). */
static int seq_mode(int argc, char** argv)
{
	for (int k = 2; k < argc; k++) {
		const float c = strtof(argv[k], 0), y = 1.0f / c;
		const int ec = ilogbf(c);
		const uint32_t exps[5] = { (uint32_t) (27 + ec), 60u, 127u, 200u, 254u };
		unsigned long long bad = 0, moved = 0;
		for (uint32_t m = 0; m < (1u << 23); m++) {
			const float t1 = u2f((127u << 23) | m);
			const float q1 = div_seq(t1, c, y);
			for (int e = 0; e < 5; e++) {
				const float t = u2f((exps[e] << 23) | m);
				volatile float want = t / c;
				const float q = div_seq(t, c, y);
				bad += f2u(q) != f2u(want);
				moved += (f2u(q) & 0x7fffffu) != (f2u(q1) & 0x7fffffu) || (int) (f2u(q) >> 23) - (int) (f2u(q1) >> 23) != (int) exps[e] - 127;
			}
		}
		printf("%.9g %d %llu %llu\n", c, ec, bad, moved);
	}
	return 0;
}

int main(int argc, char** argv)
{
	if (argc >= 3 && strcmp(argv[1], "seq") == 0) return seq_mode(argc, argv);
	if (argc < 5) return 2;
	const unsigned long long first = strtoull(argv[1], 0, 0), last = strtoull(argv[2], 0, 0), stride = strtoull(argv[3], 0, 0);
	for (int k = 4; k < argc; k++) {
		const float c = strtof(argv[k], 0), y = 1.0f / c;
		unsigned long long bad = 0, adjusted = 0, tried = 0;
		for (unsigned long long b = first; b < last; b += stride) {
			for (int neg = 0; neg < 2; neg++) {
				const float t = u2f((uint32_t) b | (neg ? 0x80000000u : 0u));
				volatile float want = t / c;
				const float got = div_scaled(t * 0x1p64f, c, y, &adjusted);
				bad += f2u(got) != f2u(want);
				tried++;
			}
		}
		printf("%.9g %llu %llu %llu\n", c, tried, bad, adjusted);
	}
	return 0;
}
