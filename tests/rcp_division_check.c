/* Test helper (CPU): the corrected-reciprocal quotient of gs_device.cuh div_hh<AR_RCP> against the IEEE fp64 divide,
 * for every mesh width h = 1/(n+1) a grid of up to NMAX points per axis can have and many numerators per h.
 * Build: gcc -O2 -ffp-contract=off rcp_division_check.c -lm.  Prints the number of mismatches. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static uint64_t rng = 0x9E3779B97F4A7C15ull;
static uint64_t next(void)
{
    rng ^= rng << 13;
    rng ^= rng >> 7;
    rng ^= rng << 17;
    return rng;
}

static double rcp_div(double s, double hh, double y)
{
    double q = s * y;
    double r = fma(-hh, q, s);
    q = fma(r, y, q);
    r = fma(-hh, q, s);
    return fma(r, y, q);
}

int main(int argc, char** argv)
{
    const int nmax = argc > 1 ? atoi(argv[1]) : 4096;
    const int per = argc > 2 ? atoi(argv[2]) : 20000;
    long bad = 0, total = 0;
    for (int n = 1; n <= nmax; n++) {
        const double h = 1.0 / (double)(n + 1); /* src/cpu/CpuGridData.cpp:41 */
        const double hh = h * h;
        const double y = 1.0 / hh;
        for (int i = 0; i < per; i++) {
            /* numerators: random sign and mantissa, exponents over the range tap sums take (1e-30 .. 1e+8),
             * plus values with short mantissas (sums of a few small integers times 2^k) */
            uint64_t u = next();
            double m = 1.0 + (double)(u >> 12) * (1.0 / 4503599627370496.0);
            if ((i & 7) == 0) m = 1.0 + (double)((u >> 12) & 0xff) / 256.0;
            int e = (int)((u >> 3) % 128) - 100;
            double s = ldexp(m, e);
            if (u & 1) s = -s;
            const double a = s / hh, b = rcp_div(s, hh, y);
            total++;
            if (memcmp(&a, &b, 8) != 0) bad++;
        }
        const double z = rcp_div(0.0, hh, y);
        if (z != 0.0) bad++;
    }
    printf("%ld %ld\n", bad, total);
    return bad != 0;
}
