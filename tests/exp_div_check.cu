// Test harness (CPU only, built by tests/test_exp_div.py): gs_exp and gs_div of gs_device.cuh on the host — the same
// explicit-fma sequences the device runs (the host fma() is correctly rounded like __fma_rn) — against glibc's exp and the
// IEEE divide.  Nothing here is linked into the library.
#include "../gpu-solve_b200/csrc/gs_device.cuh"

#include <math.h>
#include <stdint.h>

static inline uint64_t next(uint64_t& s)
{
    s ^= s << 13; s ^= s >> 7; s ^= s << 17;
    return s;
}
static inline double unit(uint64_t& s) { return (double)(next(s) >> 11) / 9007199254740992.0; }

static long long ulps(double a, double b)
{
    long long x = gs_double_to_bits(a), y = gs_double_to_bits(b);
    return x > y ? x - y : y - x;
}

// exp: n arguments uniform in [lo, hi]; returns the largest distance to glibc's exp in units of the last place and the
// number of arguments where the two differ at all
extern "C" long long exp_check(double lo, double hi, long long n, long long* differing)
{
    uint64_t s = 88172645463325252ull;
    long long worst = 0, diff = 0;
    for (long long i = 0; i < n; i++) {
        const double x = lo + (hi - lo) * unit(s);
        const long long d = ulps(gs_exp(x), exp(x));
        if (d) diff++;
        if (d > worst) worst = d;
    }
    *differing = diff;
    return worst;
}

// divide: n quotients a / b with b of magnitude [blo, bhi] and a of magnitude up to amax, both signs; returns the largest
// distance to the IEEE quotient in ulps and the number of quotients that differ
extern "C" long long div_check(double blo, double bhi, double amax, long long n, long long* differing)
{
    uint64_t s = 1234567890123456789ull;
    long long worst = 0, diff = 0;
    for (long long i = 0; i < n; i++) {
        double b = blo * pow(bhi / blo, unit(s));
        double a = amax * (2.0 * unit(s) - 1.0) * pow(10.0, -12.0 * unit(s));
        if (next(s) & 1) b = -b;
        const long long d = ulps(gs_div(a, b), a / b);
        if (d) diff++;
        if (d > worst) worst = d;
    }
    *differing = diff;
    return worst;
}

extern "C" double exp_one(double x) { return gs_exp(x); }
extern "C" double div_one(double a, double b) { return gs_div(a, b); }
