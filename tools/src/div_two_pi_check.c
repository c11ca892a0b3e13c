/* Exhaustive proof by enumeration that  float(double(a) * C)  with C = RN53(1 / double(2 pi as float)) returns the bits of
 * the IEEE single-precision quotient  a / (float)(2 pi)  for EVERY float a (shade.cl:120,143 divide a throughput by
 * 2*M_PI as floats; mcpt_device.cuh::fdiv_two_pi uses the product).  Both sides are IEEE basic operations, so what holds
 * on the host's SSE2 unit holds for F2F / DMUL / F2F on the device (-ftz=false, -prec-div=true).
 *   gcc -O2 -fopenmp -o div_two_pi_check div_two_pi_check.c && ./div_two_pi_check [stride]      (stride 1 = all 2^32)
 * prints the number of mismatches (NaN inputs compare as "both NaN"). */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

int main(int argc, char** argv) {
	const uint64_t stride = argc > 1 ? strtoull(argv[1], 0, 10) : 1;
	const float two_pi = 6.28318530717958647692f;
	const volatile double b = (double)two_pi;
	const double C = 1.0 / b;
	uint64_t bad = 0, n = 0;
	uint32_t first_bad = 0;
#pragma omp parallel for reduction(+ : bad, n) schedule(static)
	for (uint64_t chunk = 0; chunk < 4096; ++chunk) {
		for (uint64_t u = chunk << 20; u < (chunk + 1) << 20; u += stride) {
			uint32_t bits = (uint32_t)u;
			float a;
			memcpy(&a, &bits, 4);
			volatile float q = a / two_pi;
			volatile float p = (float)((double)a * C);
			float qq = q, pp = p;
			uint32_t qb, pb;
			memcpy(&qb, &qq, 4);
			memcpy(&pb, &pp, 4);
			++n;
			if (qb != pb && !(isnan(qq) && isnan(pp))) {
				++bad;
				first_bad = bits;
			}
		}
	}
	/* variable divisor (operator/(f4, float), the triangle test): float(double(a) * (1 / double(s))) against a / s on
	 * random bit patterns of every magnitude, denormals and extreme ratios included */
	uint64_t bad2 = 0, n2 = 0;
	const uint64_t pairs = argc > 2 ? strtoull(argv[2], 0, 10) : 2000000000ull;
#pragma omp parallel for reduction(+ : bad2, n2) schedule(static)
	for (uint64_t chunk = 0; chunk < 4096; ++chunk) {
		uint64_t x = 0x9E3779B97F4A7C15ull * (chunk + 1);
		for (uint64_t k = 0; k < pairs / 4096; ++k) {
			x ^= x << 13; x ^= x >> 7; x ^= x << 17;
			uint32_t ab = (uint32_t)x, sb = (uint32_t)(x >> 32);
			if (k & 1) sb = (sb & 0x807fffffu) | (((ab >> 23) & 0xffu) + (k % 7) - 3u) << 23 & 0x7f800000u;   /* similar exponents half of the time */
			float a, s;
			memcpy(&a, &ab, 4);
			memcpy(&s, &sb, 4);
			volatile float q = a / s;
			volatile double r = 1.0 / (double)s;
			volatile float p = (float)((double)a * r);
			float qq = q, pp = p;
			uint32_t qb, pb;
			memcpy(&qb, &qq, 4);
			memcpy(&pb, &pp, 4);
			++n2;
			if (qb != pb && !(isnan(qq) && isnan(pp))) ++bad2;
		}
	}
	printf("{\"pairs_checked\": %llu, \"pair_mismatches\": %llu}\n", (unsigned long long)n2, (unsigned long long)bad2);
	bad += bad2;
	printf("{\"checked\": %llu, \"mismatches\": %llu, \"reciprocal\": \"%a\", \"example\": \"0x%08x\"}\n", (unsigned long long)n,
	       (unsigned long long)bad, C, first_bad);
	return bad != 0;
}
