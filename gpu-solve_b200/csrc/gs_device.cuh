// gs_device.cuh — device-side building blocks shared by kernels_simple.cu and the tuned
// kernels: the per-point arithmetic (one definition, so every kernel variant produces the same
// bits), the deterministic sum-of-squares reduction, and the mbarrier / TMA PTX wrappers.
//
// Arithmetic contract (DESIGN.md "Parity"): the library is compiled with -fmad=false, so every
// expression below is evaluated with separate IEEE fp64 multiplies and adds in the order written,
// which is the order of the reference (cited per function, paths relative to the reference root).
#pragma once
#include <cuda.h>

#include <string.h>
#include <math.h>

#include <atomic>
#include <utility>
#include <mutex>

#include "gs_internal.h"

// ---- per-device one-time kernel setup ----
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) and the occupancy of a kernel belong to a DEVICE, and the C ABI takes
// a device per handle: every launcher keeps its "prepared" flag and occupancy per device (slot = cudaGetDevice()).
constexpr int GS_MAX_DEVICES = 64;
struct DevOnce {
    std::atomic<int> v[GS_MAX_DEVICES];  // 0 = not set up on that device yet, else the value setup() reported (>= 1)
};
inline int gs_cur_device()
{
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= GS_MAX_DEVICES) d = 0;
    return d;
}
// runs setup(int* value) once per device (thread safe); *value_out = what it reported for the current device
template <class Setup>
inline cudaError_t gs_dev_once(DevOnce& st, int* value_out, Setup&& setup)
{
    static std::mutex mu;
    const int d = gs_cur_device();
    int o = st.v[d].load(std::memory_order_acquire);
    if (o == 0) {
        std::lock_guard<std::mutex> lock(mu);
        o = st.v[d].load(std::memory_order_acquire);
        if (o == 0) {
            int q = 1;
            const cudaError_t e = setup(&q);
            if (e != cudaSuccess) return e;
            o = q > 0 ? q : 1;
            st.v[d].store(o, std::memory_order_release);
        }
    }
    if (value_out) *value_out = o;
    return cudaSuccess;
}

// ---- host helpers shared by the tuned translation units (kernels_fast.cu) ----
// TMA descriptor of a level array viewed as [nzl][Y][pitch] doubles with a bw x bh x 1 box (cached)
bool gs_tensor_map(const double* base, const Geom& g, int bw, int bh, CUtensorMap* out);
// planes per CTA of a z-marching kernel: long chunks (each re-reads its rim planes), but short
// enough that small levels still put several CTAs on every SM
int gs_pick_zchunk(const LaunchCtx& cx, int tiles, int planes, int max_chunk);
// the same with wave quantisation taken into account (tune.wave): see kernels_fast.cu
int gs_pick_zchunk_waves(const LaunchCtx& cx, int tiles, int planes, int ctas_per_sm, int overhead, int min_chunk, int max_chunk,
                         bool even);

// ---------------------------------------------------------------------------------
// point arithmetic
// ---------------------------------------------------------------------------------
// s = sum_i c_i v(p+o_i) (already accumulated, i ascending: src/cpu/CpuSolver.cpp:56-60)
// -> r = f - (s/(h*h) + NL), and the Jacobi denominator (modes 1..3)
// src/cpu/CpuSolver.cpp:61-74 (residual) and :161-171 (denominator)
// AR = arithmetic flags of a level, chosen on the host (gs_api.cu create_impl); every value gives the same bits:
//   AR_POW2  h*h is a power of two: s * (1/hh) == s / hh exactly
//   AR_UNIT  the six off-centre stencil values are exactly -1: s + (-1.0)*v == s - v exactly (IEEE: a - b is
//            a + (-b), and -1.0*v is v with the sign flipped), so the six multiplies are dropped
//   AR_RCP   any other h*h: the quotient is formed from y = RN(1/hh) (host division, correctly rounded) by
//            q0 = RN(s*y), then twice r = fma(-hh, q, s) [exact], q = fma(r, y, q).  q0 is within 2 ulp of s/hh, the first
//            correction brings it within 1 ulp, and by Markstein's theorem (y correctly rounded, q within 1 ulp, exact
//            remainder) the second yields RN(s/hh) — the bits of the fp64 divide for 5 instructions instead of ~25
//            (tests/test_rcp_division.py checks the sequence against the divide for every h a grid up to 4096 can have).
// Instantiated combinations: 0 (general: true divide), AR_POW2, AR_POW2|AR_UNIT, AR_RCP, AR_RCP|AR_UNIT.
enum { AR_POW2 = 1, AR_UNIT = 2, AR_RCP = 4 };
#define AR_SWITCH(ar, STMT)                                                     \
    do {                                                                        \
        const int ar_ = (ar);                                                   \
        if (ar_ == (AR_POW2 | AR_UNIT)) {                                       \
            constexpr int ARV = AR_POW2 | AR_UNIT;                              \
            STMT;                                                               \
        } else if (ar_ & AR_POW2) {                                             \
            constexpr int ARV = AR_POW2;                                        \
            STMT;                                                               \
        } else if (ar_ == (AR_RCP | AR_UNIT)) {                                 \
            constexpr int ARV = AR_RCP | AR_UNIT;                               \
            STMT;                                                               \
        } else if (ar_ & AR_RCP) {                                              \
            constexpr int ARV = AR_RCP;                                         \
            STMT;                                                               \
        } else {                                                                \
            constexpr int ARV = 0;                                              \
            STMT;                                                               \
        }                                                                       \
    } while (0)

// The point functions are __host__ __device__ only so that a CPU test can run them over host arrays
// (tests/pc_host_check.cu); the library itself never calls them on the host.
#define GS_HD __host__ __device__ __forceinline__
#ifdef __CUDA_ARCH__
#define GS_FMA(a, b, c) __fma_rn(a, b, c)
#else
#define GS_FMA(a, b, c) fma(a, b, c)
#endif

// ---------------------------------------------------------------------------------
// programmatic dependent launch (sm_90+)
// ---------------------------------------------------------------------------------
// A V-cycle is a chain of ~50 dependent kernels; on the levels below 48 points per axis they take a few microseconds each
// and the chain is bound by the launch-to-launch latency of the graph.  A kernel launched with the attribute may be made
// resident while its predecessor still runs; it first tells its own successor to come (launch_dependents) and then waits for
// the predecessor's completion and memory flush (wait) before it touches anything, so the dependency itself is unchanged.
// Used by the one-thread-per-point kernels (kernels_simple.cu): graph V-cycle 0.125 -> 0.111 ms at 63^3, 0.177 -> 0.163 at
// 127^3, 0.471 -> 0.445 at 255^3 (profiles/r02r_pdl.txt).  On the large-shared-memory TMA kernels it gained nothing at 513^3
// and lost at the small sizes (early-resident CTAs hold 108 KB each), so those are launched plainly.  Without the attribute
// (option pdl = 0) both instructions are no-ops.
__device__ __forceinline__ void gs_pdl_sync()
{
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

template <typename... KArgs, typename... Args>
cudaError_t gs_launch(const LaunchCtx& cx, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, Args&&... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = cx.stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = cx.tune.pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// ---------------------------------------------------------------------------------
// exp() and the per-point divide of the non-linear modes
// ---------------------------------------------------------------------------------
// The exp() modes are instruction-bound (profiles/r02a_ncu_modes_summary.txt: 68 % of the issue slots, DRAM 27 %), and a
// third of what they issue is the library exp / divide moving immediates into registers and testing special cases.
// gs_exp: k = rint(x / ln 2), r = x - k ln 2 (two-part ln 2), Taylor polynomial of degree 13 on |r| <= 0.347 (truncation
// 4e-18, below 1 ulp in total; tests/test_exp_div.py measures it against glibc), scaled by 2^k through the exponent
// field; arguments outside +-700 (and NaN) go to the library.  North_star's tolerance for these modes is 1e-12 relative;
// CUDA's own exp differs from glibc's by up to an ulp as well.
GS_HD double gs_bits_to_double(long long b)
{
#ifdef __CUDA_ARCH__
    return __longlong_as_double(b);
#else
    double d;
    memcpy(&d, &b, 8);
    return d;
#endif
}
GS_HD long long gs_double_to_bits(double d)
{
#ifdef __CUDA_ARCH__
    return __double_as_longlong(d);
#else
    long long b;
    memcpy(&b, &d, 8);
    return b;
#endif
}

// the special cases stay out of line: one copy per kernel, not one per call site of an unrolled loop
__host__ __device__ __noinline__ inline double gs_exp_library(double x) { return exp(x); }
__host__ __device__ __noinline__ inline double gs_div_library(double a, double b) { return a / b; }

// coefficients as constant-bank operands of the fma (an immediate double costs two moves into uniform registers each time)
#define GS_EXP_TABLE                                                                                     \
    {1.6059043836821613e-10, /* 1/13! */ 2.0876756987868100e-09, /* 1/12! */ 2.5052108385441720e-08,     \
     2.7557319223985890e-07, /* 1/10! */ 2.7557319223985893e-06, /* 1/9! */ 2.4801587301587302e-05,      \
     1.9841269841269841e-04, /* 1/7! */ 1.3888888888888889e-03, /* 1/6! */ 8.3333333333333332e-03,       \
     4.1666666666666664e-02, /* 1/4! */ 1.6666666666666666e-01, /* 1/3! */ 0.5, 1.0,                     \
     1.4426950408889634074,        /* [13] log2(e) */                                                    \
     6755399441055744.0,           /* [14] 2^52 + 2^51: the integer lands in the low mantissa bits */    \
     -6.93147180369123816490e-01,  /* [15] -ln2, high part */                                            \
     -1.90821492927058770002e-10}  /* [16] -ln2, low part */
static __constant__ double gs_exp_c[17] = GS_EXP_TABLE;
static const double gs_exp_c_host[17] = GS_EXP_TABLE;
#ifdef __CUDA_ARCH__
#define GS_EXP_C(i) gs_exp_c[i]
#else
#define GS_EXP_C(i) gs_exp_c_host[i]
#endif

GS_HD double gs_exp(double x)
{
    // |x| >= 700, infinities and NaN: one integer compare on the high word (an fp64 compare costs two issue cycles)
    if ((unsigned)((gs_double_to_bits(x) >> 32) & 0x7fffffff) >= 0x4085E000u) return gs_exp_library(x);
    const double t = GS_FMA(x, GS_EXP_C(13), GS_EXP_C(14));
    const double kd = t - GS_EXP_C(14);
    const long long k = (long long)(int)(unsigned)gs_double_to_bits(t);
    double r = GS_FMA(kd, GS_EXP_C(15), x);
    r = GS_FMA(kd, GS_EXP_C(16), r);
    double p = GS_EXP_C(0);
#pragma unroll
    for (int i = 1; i <= 12; i++) p = GS_FMA(p, r, GS_EXP_C(i));
    p = GS_FMA(p, r, 1.0);
    return p * gs_bits_to_double((k + 1023) << 52);  // |x| < 700: k + 1023 in [13, 2033], a normal exponent
}

// a / b for the Jacobi update of the exp() modes (b = c0/h^2 + gamma (1 + w) e^w, a normal number in any run that has not
// already diverged): reciprocal seed refined by two Newton steps, quotient corrected once by its exact remainder — the
// sequence of the library's divide without its scaling of extreme exponents; operands outside the safe exponent range go
// to the library.  The result is the correctly rounded quotient except for rare last-bit cases (tests/test_exp_div.py).
GS_HD double gs_div(double a, double b)
{
    // exponents: b within 2^-930 .. 2^930, a below 2^930 (zeros, infinities and NaN fail the test); integer compares
    const unsigned eb = (unsigned)(gs_double_to_bits(b) >> 52) & 0x7ffu, ea = (unsigned)(gs_double_to_bits(a) >> 52) & 0x7ffu;
    if (eb - 93u > 1860u || ea > 1953u) return gs_div_library(a, b);
#ifdef __CUDA_ARCH__
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(b));
#else
    double y = gs_bits_to_double(gs_double_to_bits(1.0 / b) & ~0xffffffffll);  // the seed's precision: upper mantissa bits only
#endif
    double e = GS_FMA(-b, y, 1.0);
    y = GS_FMA(y, e, y);
    e = GS_FMA(-b, y, 1.0);
    y = GS_FMA(y, e, y);
    double q = a * y;
    const double rem = GS_FMA(-b, q, a);
    return GS_FMA(rem, y, q);
}

// s / (h*h) under the level's arithmetic flags; every branch returns the same bits
template <int AR>
GS_HD double div_hh(double s, const PointK& k)
{
    if (AR & AR_POW2) return s * k.inv_hh;
    if (AR & AR_RCP) {
        double q = s * k.inv_hh;
        double r = GS_FMA(-k.hh, q, s);
        q = GS_FMA(r, k.inv_hh, q);
        r = GS_FMA(-k.hh, q, s);
        return GS_FMA(r, k.inv_hh, q);
    }
    return s / k.hh;
}

template <int KM, int AR>
GS_HD double finish_residual(double s, double vc, double fc, double aux, const PointK& k, double& den)
{
    s = div_hh<AR>(s, k);
    den = 0.0;
    if (KM == KM_NONLINEAR) {
        double ex = gs_exp(vc);
        double nl = (k.gamma * vc) * ex;
        s += nl;
        den = k.preFac + (k.gamma * (1 + vc)) * ex;
    } else if (KM == KM_NEWTON_EXP) {
        double ex = gs_exp(aux);
        double g1 = k.gamma * (1 + aux);
        s += (g1 * vc) * ex;
        den = k.preFac + g1 * ex;
    } else if (KM == KM_NEWTON_COEF) {
        // aux = (gamma*(1+nv))*exp(nv) precomputed once per Newton step (ks_newton_coef);
        // (g1*ex)*v instead of (g1*v)*ex: one rounding apart, inside the exp() tolerance
        s += aux * vc;
        den = k.preFac + aux;
    }
    return fc - s;
}

template <int KM>
GS_HD double jacobi_update(double vc, double r, double den, const PointK& k)
{
    // src/cpu/CpuSolver.cpp:158 resp. :164,:171
    if (KM == KM_LINEAR) return vc + k.omega * (k.alpha * r);
    return vc + k.omega * gs_div(r, den);
}

// ---------------------------------------------------------------------------------
// z-marched trilinear interpolation of one x column for the fused prolongation + smoother pass (kernels_fused.cu PC)
// ---------------------------------------------------------------------------------
// CpuSolver::interpolate in closed form (kernels_simple.cu prolong_point): e(X,Y,Z) = avg_z(avg_y(avg_x(C))) with
// avg = 0.5*a + 0.5*b on odd indices and the identity on even ones, nesting x -> y -> z.  For a thread that owns the fine
// rows Y (even) and Y+1 at column X and marches along z, everything it needs on a coarse plane k are the x-interpolants
// ax of coarse rows j = Y/2 and j+1; they are carried from plane to plane, so only ODD fine planes load new coarse
// values.  `ld(k, row)` returns ax of coarse row j+row (row in {-1, 0, 1}) on GLOBAL coarse plane k.  Results per plane:
//   eA  row Y   (even)            eB  row Y+1 (odd)
//   eJ  row Y+2 (even: the rim row above the last row pair of a tile)
//   eM  row Y-1 (odd:  the rim row below the first row pair; only evaluated when want_m)
// Same operations in the same order as prolong_point, hence the same bits (tests/test_pc_host.py checks every point).
struct PcInterp {
    double a, j;  // ax(row j), ax(row j+1) on coarse plane zg >> 1 of the plane that comes next
    template <class LD>
    GS_HD void init(int zg0, LD&& ld)
    {
        a = ld(zg0 >> 1, 0);
        j = ld(zg0 >> 1, 1);
    }
    template <class LD>
    GS_HD void step(int zg, bool want_m, LD&& ld, double& eA, double& eB, double& eJ, double& eM)
    {
        eM = 0.0;
        if (zg & 1) {
            const int k1 = (zg >> 1) + 1;
            const double nA = ld(k1, 0), nJ = ld(k1, 1);
            eA = 0.5 * a + 0.5 * nA;
            eJ = 0.5 * j + 0.5 * nJ;
            eB = 0.5 * (0.5 * a + 0.5 * j) + 0.5 * (0.5 * nA + 0.5 * nJ);
            if (want_m) {
                const double m0 = 0.5 * ld(zg >> 1, -1) + 0.5 * a, m1 = 0.5 * ld(k1, -1) + 0.5 * nA;
                eM = 0.5 * m0 + 0.5 * m1;
            }
            a = nA;
            j = nJ;
        } else {
            eA = a;
            eJ = j;
            eB = 0.5 * a + 0.5 * j;
            if (want_m) eM = 0.5 * ld(zg >> 1, -1) + 0.5 * a;
        }
    }
};

// canonical tap order C,+x,-x,+y,-y,+z,-z (config order of the axis-aligned stencil)
template <int AR>
__device__ __forceinline__ double canon_tap_sum(double C, double E, double W, double N, double S, double U, double D, const PointK& k)
{
    double s = 0.0;
    s += k.c[0] * C;
    if (AR & AR_UNIT) {
        s -= E;
        s -= W;
        s -= N;
        s -= S;
        s -= U;
        s -= D;
        return s;
    }
    s += k.c[1] * E;
    s += k.c[2] * W;
    s += k.c[3] * N;
    s += k.c[4] * S;
    s += k.c[5] * U;
    s += k.c[6] * D;
    return s;
}

template <int KM, int AR>
__device__ __forceinline__ double canon_residual(double C, double E, double W, double N, double S, double U, double D, double fv,
                                                 double aux, const PointK& k, double& den)
{
    return finish_residual<KM, AR>(canon_tap_sum<AR>(C, E, W, N, S, U, D, k), C, fv, aux, k, den);
}

// ---------------------------------------------------------------------------------
// deterministic sum of squares: CTA tree, then the last CTA to finish adds the per-CTA
// partials in index order with the same tree (no floating-point atomics; the result does
// not depend on which CTA happens to be last)
// ---------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

// sum over the first NT threads (tid 0..NT-1, NT a multiple of 32) of the CTA; the other threads
// of the CTA (e.g. a TMA producer warp) must not call it.  Uses named barrier 1.
template <int NT>
__device__ __forceinline__ double block_sum(double v, double* sh, int tid)
{
    v = warp_sum(v);
    if ((tid & 31) == 0) sh[tid >> 5] = v;
    asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory");
    double t = 0.0;
    if (tid < 32) {
        t = tid < NT / 32 ? sh[tid] : 0.0;
        t = warp_sum(t);
    }
    asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory");
    return t;  // valid in thread 0
}

template <int NT>
__device__ __forceinline__ void grid_sum_finish(double block_total, NormBuf nb, double* sh, int tid)
{
    const unsigned nblk = gridDim.x * gridDim.y * gridDim.z;
    const unsigned bid = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
    __shared__ int last;
    if (tid == 0) {
        nb.partials[bid] = block_total;
        __threadfence();
        unsigned t = atomicAdd(nb.ticket, 1u);
        last = (t == nblk - 1);
    }
    asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory");
    if (!last) return;
    __threadfence();
    double acc = 0.0;
    for (unsigned i = tid; i < nblk; i += NT) acc += __ldcg(&nb.partials[i]);
    double tot = block_sum<NT>(acc, sh, tid);
    if (tid == 0) {
        *nb.result = tot;
        *nb.ticket = 0u;
    }
}

// ---------------------------------------------------------------------------------
// mbarrier / TMA (cp.async.bulk.tensor, SASS UTMALDG)
// ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.b32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    while (!mbar_try_wait(bar, parity)) {
    }
}
// 3-D tiled load of one box; the box must START on a 16-byte boundary in global memory
// (tools/tma_probe.cu: an 8-byte-aligned start raises "illegal instruction" on B200)
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// the same on 32-bit shared-window addresses (no generic->shared conversion in inner loops)
__device__ __forceinline__ void mbar_wait_a(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.b32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ void mbar_arrive_a(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ double lds_f64(uint32_t a)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts_f64(uint32_t a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory"); }
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
// consumer-only barrier (named barrier 2) for kernels whose producer warp runs ahead
template <int NT>
__device__ __forceinline__ void consumer_sync()
{
    asm volatile("bar.sync 2, %0;" ::"n"(NT) : "memory");
}
