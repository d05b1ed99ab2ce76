/* CPU restatement of div_scaled (fluid-simulation_b200/csrc/grid.cuh), operation for operation, checked against the
 * machine's own IEEE single division.  Test infrastructure (tests/test_division_cpu.py compiles and runs it).
 * usage: div_scaled_check <first bit pattern> <last bit pattern, exclusive> <stride> <divisor>...
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

int main(int argc, char** argv)
{
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
