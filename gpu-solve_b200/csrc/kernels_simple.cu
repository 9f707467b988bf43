// kernels_simple.cu — one thread per lattice point, any seven in-range stencil
// offsets in config order.  These are the general-purpose kernels (arbitrary
// offsets, every mode) and the in-library cross-check of the tuned kernels in
// kernels_fast.cu.  They are CUDA kernels for sm_100a, not a CPU fallback.
//
// Arithmetic contract (DESIGN.md "Parity"): the library is compiled with
// -fmad=false, so every expression below is evaluated with separate IEEE fp64
// multiplies and adds in the order written, which is the order of the reference
// (cited per function, paths relative to the reference root).  Linear-mode
// fields are therefore bit-identical to GpuSolve-cpu; modes with exp() differ by
// CUDA-vs-glibc exp rounding only.
#include <string.h>

#include <mutex>

#include "gs_device.cuh"

namespace {



constexpr int BX = 64, BY = 4;  // 256 threads, one x-row of 64 points per 2 warps

struct SmoothP {
    Geom g;
    PointK k;
    i64 toff[7];
    int zb;
    const double* vin;
    double* vout;
    const double* f;
    const double* aux;
};

GS_HD double tap_sum(const double* __restrict__ a, i64 c, const double (&cv)[7], const i64 (&toff)[7])
{
    // src/cpu/CpuSolver.cpp:56-60: stencilsum = 0.0; stencilsum += values[i] * v(p + o_i), i ascending
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 7; i++) s += cv[i] * a[c + toff[i]];
    return s;
}

// The per-point bodies are functions of (x, y, local plane), shared by the grid kernels below (and by the CPU check of the
// interpolation state machine, tests/pc_host_check.cu).
template <int KM, bool POW2>
GS_HD void smooth_point(const Geom& g, const PointK& k, const i64 (&toff)[7], const double* vin, double* vout,
                                             const double* f, const double* auxp, int x, int y, int zl)
{
    const i64 c = gs_idx(g, x, y, zl);
    const double vc = vin[c];
    double s = tap_sum(vin, c, k.c, toff);
    const double aux = (KM >= KM_NEWTON_EXP) ? auxp[c] : 0.0;
    double den;
    const double r = finish_residual<KM, POW2>(s, vc, f[c], aux, k, den);
    vout[c] = jacobi_update<KM>(vc, r, den, k);
}

template <int KM, bool POW2>
__global__ void __launch_bounds__(BX* BY) smooth_simple(SmoothP p)
{
    gs_pdl_sync();
    const int x = blockIdx.x * BX + threadIdx.x + 1;
    const int y = blockIdx.y * BY + threadIdx.y + 1;
    const int zl = blockIdx.z + p.zb;
    if (x > p.g.nx || y > p.g.ny) return;
    smooth_point<KM, POW2>(p.g, p.k, p.toff, p.vin, p.vout, p.f, p.aux, x, y, zl);
}

struct ResidP {
    Geom g;
    PointK k;
    i64 toff[7];
    int zb;
    const double* v;
    const double* f;
    const double* aux;
    double* r;
    NormBuf nb;
};

// CpuSolver::compResidual, src/cpu/CpuSolver.cpp:45-83
template <int KM, bool POW2>
GS_HD double residual_point(const Geom& g, const PointK& k, const i64 (&toff)[7], const double* v, const double* f,
                                                 const double* auxp, double* rout, int x, int y, int zl)
{
    const i64 c = gs_idx(g, x, y, zl);
    const double vc = v[c];
    double s = tap_sum(v, c, k.c, toff);
    const double aux = (KM >= KM_NEWTON_EXP) ? auxp[c] : 0.0;
    double den;
    const double r = finish_residual<KM, POW2>(s, vc, f[c], aux, k, den);
    if (rout) rout[c] = r;
    return r;
}

template <int KM, bool POW2>
__global__ void __launch_bounds__(BX* BY) residual_simple(ResidP p)
{
    gs_pdl_sync();
    __shared__ double sh[32];
    const int x = blockIdx.x * BX + threadIdx.x + 1;
    const int y = blockIdx.y * BY + threadIdx.y + 1;
    const int zl = blockIdx.z + p.zb;
    double sq = 0.0;
    if (x <= p.g.nx && y <= p.g.ny) {
        const double r = residual_point<KM, POW2>(p.g, p.k, p.toff, p.v, p.f, p.aux, p.r, x, y, zl);
        sq = r * r;
    }
    const int tid = threadIdx.x + threadIdx.y * BX;
    double tot = block_sum<BX * BY>(sq, sh, tid);
    grid_sum_finish<BX * BY>(tot, p.nb, sh, tid);
}

struct RestrictP {
    Geom gf, gc;
    int zb;
    const double* fine;
    double* coarse;
    double* coarse2;
};

// CpuSolver::restrict, src/cpu/CpuSolver.cpp:211-238: running sum over the 27
// neighbours, x offset outermost, z offset innermost, weights 0.125*w*w*w
GS_HD void restrict_point(const Geom& gf, const Geom& gc, const double* fine, double* coarse, double* coarse2,
                                               int x, int y, int zc)
{
    const int zf = 2 * (zc + gc.z0) - gf.z0;  // local fine plane of the centre
    const i64 fc = gs_idx(gf, 2 * x, 2 * y, zf);
    const double w[3] = {0.5, 1.0, 0.5};
    double acc = 0.0;
#pragma unroll
    for (int ii = -1; ii < 2; ii++)
#pragma unroll
        for (int jj = -1; jj < 2; jj++)
#pragma unroll
            for (int kk = -1; kk < 2; kk++) {
                const double fac = 0.125 * w[ii + 1] * w[jj + 1] * w[kk + 1];
                acc += fac * fine[fc + ii + jj * gf.pitch + kk * gf.plane];
            }
    const i64 cc = gs_idx(gc, x, y, zc);
    coarse[cc] = acc;
    if (coarse2) coarse2[cc] = acc;
}

__global__ void __launch_bounds__(BX* BY) restrict_simple(RestrictP p)
{
    gs_pdl_sync();
    const int x = blockIdx.x * BX + threadIdx.x + 1;
    const int y = blockIdx.y * BY + threadIdx.y + 1;
    const int zc = blockIdx.z + p.zb;  // local coarse plane
    if (x > p.gc.nx || y > p.gc.ny) return;
    restrict_point(p.gf, p.gc, p.fine, p.coarse, p.coarse2, x, y, zc);
}

struct ProlongP {
    Geom gf, gc;
    int zb;
    const double* vin;
    double* vout;
    const double* coarse;
    double* e_out;
};

// CpuSolver::interpolate (src/cpu/CpuSolver.cpp:240-290) in closed form, then
// v += e (:131-132).  The four dependent passes of the reference (inject, x, y, z
// averages) nest as  e = avg_z(avg_y(avg_x(C)))  with avg = 0.5*a + 0.5*b on odd
// indices and identity on even ones; same operations, same order => same bits.
GS_HD double prolong_point(const double* __restrict__ cv, const Geom& gc, int X, int Y, int Zg)
{
    const int i = X >> 1, j = Y >> 1, k = (Zg >> 1) - gc.z0;
    const bool ox = X & 1, oy = Y & 1, oz = Zg & 1;
    const i64 b = gs_idx(gc, i, j, k);
    auto ax = [&](i64 q) { return ox ? 0.5 * cv[q] + 0.5 * cv[q + 1] : cv[q]; };
    auto ay = [&](i64 q) { return oy ? 0.5 * ax(q) + 0.5 * ax(q + gc.pitch) : ax(q); };
    return oz ? 0.5 * ay(b) + 0.5 * ay(b + gc.plane) : ay(b);
}

GS_HD void prolong_correct_point(const Geom& gf, const Geom& gc, const double* vin, double* vout, const double* coarse,
                                                      double* e_out, int x, int y, int zl)
{
    const double e = prolong_point(coarse, gc, x, y, zl + gf.z0);
    const i64 c = gs_idx(gf, x, y, zl);
    if (e_out) e_out[c] = e;
    if (vout) vout[c] = vin[c] + e;
}

__global__ void __launch_bounds__(BX* BY) prolong_correct_simple(ProlongP p)
{
    gs_pdl_sync();
    const int x = blockIdx.x * BX + threadIdx.x + 1;
    const int y = blockIdx.y * BY + threadIdx.y + 1;
    const int zl = blockIdx.z + p.zb;
    if (x > p.gf.nx || y > p.gf.ny) return;
    prolong_correct_point(p.gf, p.gc, p.vin, p.vout, p.coarse, p.e_out, x, y, zl);
}

struct FasP {
    Geom g;
    PointK k;
    i64 toff[7];
    int zb;
    const double* restv;
    double* f;
    double* r_out;
};

// CpuSolver::applyStencil (src/cpu/CpuSolver.cpp:182-209) on restV, then f += r (:112)
template <bool POW2>
GS_HD void fas_rhs_point(const Geom& g, const PointK& k, const i64 (&toff)[7], const double* restv, double* f,
                                              double* r_out, int x, int y, int zl)
{
    const i64 c = gs_idx(g, x, y, zl);
    const double vc = restv[c];
    double s = tap_sum(restv, c, k.c, toff);
    s = POW2 ? s * k.inv_hh : s / k.hh;
    const double nl = (k.gamma * vc) * gs_exp(vc);
    s += nl;
    if (r_out) r_out[c] = s;
    f[c] = f[c] + s;
}

template <bool POW2>
__global__ void __launch_bounds__(BX* BY) fas_rhs_simple(FasP p)
{
    gs_pdl_sync();
    const int x = blockIdx.x * BX + threadIdx.x + 1;
    const int y = blockIdx.y * BY + threadIdx.y + 1;
    const int zl = blockIdx.z + p.zb;
    if (x > p.g.nx || y > p.g.ny) return;
    fas_rhs_point<POW2>(p.g, p.k, p.toff, p.restv, p.f, p.r_out, x, y, zl);
}

struct NewtonFP {
    Geom g;
    PointK k;
    i64 toff[7];
    int zb;
    const double* nv;
    const double* newtonF;
    double* f;
    NormBuf nb;
};

// NewtonSolver::compF, src/cpu/NewtonSolver.cpp:48-81
template <bool POW2>
__global__ void __launch_bounds__(BX* BY) newton_F_simple(NewtonFP p)
{
    gs_pdl_sync();
    __shared__ double sh[32];
    const int x = blockIdx.x * BX + threadIdx.x + 1;
    const int y = blockIdx.y * BY + threadIdx.y + 1;
    const int zl = blockIdx.z + p.zb;
    double sq = 0.0;
    if (x <= p.g.nx && y <= p.g.ny) {
        const i64 c = gs_idx(p.g, x, y, zl);
        const double vc = p.nv[c];
        double s = tap_sum(p.nv, c, p.k.c, p.toff);
        s = POW2 ? s * p.k.inv_hh : s / p.k.hh;
        const double ex = gs_exp(vc);
        const double nl = (p.k.gamma * vc) * ex;
        s += nl;
        const double fv = p.newtonF[c] - s;
        p.f[c] = fv;
        sq = fv * fv;
    }
    const int tid = threadIdx.x + threadIdx.y * BX;
    double tot = block_sum<BX * BY>(sq, sh, tid);
    grid_sum_finish<BX * BY>(tot, p.nb, sh, tid);
}

GS_HD void newton_coef_point(const Geom& g, double gamma, const double* nv, double* coef, int x, int y, int zl)
{
    const i64 c = gs_idx(g, x, y, zl);
    const double n = nv[c];
    coef[c] = (gamma * (1 + n)) * gs_exp(n);
}

__global__ void __launch_bounds__(BX* BY) newton_coef_simple(Geom g, int zb, double gamma, const double* __restrict__ nv,
                                                              double* __restrict__ coef)
{
    gs_pdl_sync();
    const int x = blockIdx.x * BX + threadIdx.x + 1;
    const int y = blockIdx.y * BY + threadIdx.y + 1;
    const int zl = blockIdx.z + zb;
    if (x > g.nx || y > g.ny) return;
    newton_coef_point(g, gamma, nv, coef, x, y, zl);
}

// Vector3::operator+= / -= (src/cpu/Vector3.cpp:34-54); the halo of every src
// operand the V-cycle uses is zero, so the interior is the whole effect
GS_HD void axpy_point(const Geom& g, double* dst, const double* src, int sign, int x, int y, int zl)
{
    const i64 c = gs_idx(g, x, y, zl);
    dst[c] = sign >= 0 ? dst[c] + src[c] : dst[c] - src[c];
}

__global__ void __launch_bounds__(BX* BY) axpy_simple(Geom g, int zb, double* __restrict__ dst, const double* __restrict__ src,
                                                       int sign)
{
    gs_pdl_sync();
    const int x = blockIdx.x * BX + threadIdx.x + 1;
    const int y = blockIdx.y * BY + threadIdx.y + 1;
    const int zl = blockIdx.z + zb;
    if (x > g.nx || y > g.ny) return;
    axpy_point(g, dst, src, sign, x, y, zl);
}

// src/cpu/CpuGridData.cpp:7-12
__device__ __forceinline__ double rhs_f0(double x)
{
    return (100 * x * (x - 1.0) * x * (x - 1.0) * x * (x - 1.0) * x * (x - 1.0));
}
__device__ __forceinline__ double rhs_f2(double x)
{
    return (100.0 * 4.0 * (x - 1.0) * (x - 1.0) * x * x * (14.0 * x * x - 14.0 * x + 3));
}

// src/cpu/CpuGridData.cpp:45-78.  Linear: interior only, sampled at the zero-based
// interior index (:47-58).  Otherwise: whole array, sampled at the array index (:62-76).
__global__ void __launch_bounds__(BX* BY) init_rhs_simple(Geom g, int mode, double h, double gamma, double* __restrict__ f)
{
    gs_pdl_sync();
    const int x = blockIdx.x * BX + threadIdx.x;
    const int y = blockIdx.y * BY + threadIdx.y;
    const int zl = blockIdx.z;
    const int zg = zl + g.z0;
    if (x >= g.X || y >= g.Y) return;
    if (mode == GS_MODE_LINEAR) {
        if (x < 1 || x > g.nx || y < 1 || y > g.ny || zg < 1 || zg > g.nz) return;
        const double px = (x - 1) * h, py = (y - 1) * h, pz = (zg - 1) * h;
        const double val = -(rhs_f2(px) * rhs_f0(py) * rhs_f0(pz) + rhs_f0(px) * rhs_f2(py) * rhs_f0(pz) +
                             rhs_f0(px) * rhs_f0(py) * rhs_f2(pz));
        f[gs_idx(g, x, y, zl)] = val;
    } else {
        if (zg < 0 || zg >= g.Zg) return;
        const double px = x * h, py = y * h, pz = zg * h;
        const double val = 2.0 * ((py - py * py) * (pz - pz * pz) + (px - px * px) * (pz - pz * pz) + (px - px * px) * (py - py * py)) +
                           gamma * (px - px * px) * (py - py * py) * (pz - pz * pz) *
                               gs_exp((px - px * px) * (py - py * py) * (pz - pz * pz));
        f[gs_idx(g, x, y, zl)] = val;
    }
}


inline dim3 grid_for(int nx, int ny, int nz) { return dim3((nx + BX - 1) / BX, (ny + BY - 1) / BY, nz); }

inline void fill_toff(const Geom& g, const StencilOff& so, i64 (&toff)[7])
{
    for (int i = 0; i < 7; i++) toff[i] = so.dx[i] + so.dy[i] * g.pitch + so.dz[i] * g.plane;
}

inline cudaError_t done(const LaunchCtx& cx)
{
    if (cx.launches) ++*cx.launches;
    return cudaGetLastError();
}

}  // namespace

cudaError_t ks_zero(const LaunchCtx& cx, double* a, size_t count)
{
    return cudaMemsetAsync(a, 0, count * sizeof(double), cx.stream);
}

cudaError_t ks_copy(const LaunchCtx& cx, double* dst, const double* src, size_t count)
{
    return cudaMemcpyAsync(dst, src, count * sizeof(double), cudaMemcpyDeviceToDevice, cx.stream);
}

#define KM_DISPATCH(KERNEL, km, pow2, P)                                                                    \
    do {                                                                                                    \
        cudaError_t le__ = cudaSuccess;                                                                     \
        if (pow2) {                                                                                         \
            switch (km) {                                                                                   \
            case KM_LINEAR: le__ = gs_launch(cx, KERNEL<KM_LINEAR, true>, grid, block, 0, P); break;       \
            case KM_NONLINEAR: le__ = gs_launch(cx, KERNEL<KM_NONLINEAR, true>, grid, block, 0, P); break; \
            case KM_NEWTON_EXP: le__ = gs_launch(cx, KERNEL<KM_NEWTON_EXP, true>, grid, block, 0, P); break; \
            default: le__ = gs_launch(cx, KERNEL<KM_NEWTON_COEF, true>, grid, block, 0, P); break;         \
            }                                                                                               \
        } else {                                                                                            \
            switch (km) {                                                                                   \
            case KM_LINEAR: le__ = gs_launch(cx, KERNEL<KM_LINEAR, false>, grid, block, 0, P); break;      \
            case KM_NONLINEAR: le__ = gs_launch(cx, KERNEL<KM_NONLINEAR, false>, grid, block, 0, P); break; \
            case KM_NEWTON_EXP: le__ = gs_launch(cx, KERNEL<KM_NEWTON_EXP, false>, grid, block, 0, P); break; \
            default: le__ = gs_launch(cx, KERNEL<KM_NEWTON_COEF, false>, grid, block, 0, P); break;        \
            }                                                                                               \
        }                                                                                                   \
        if (le__ != cudaSuccess) return le__;                                                               \
    } while (0)

cudaError_t ks_smooth(const LaunchCtx& cx, int km, bool pow2, const Geom& g, int zb, int ze, const PointK& k,
                      const StencilOff& so, const double* vin, double* vout, const double* f, const double* aux)
{
    if (ze <= zb || g.nx <= 0 || g.ny <= 0) return cudaSuccess;
    SmoothP p{g, k, {}, zb, vin, vout, f, aux};
    fill_toff(g, so, p.toff);
    dim3 grid = grid_for(g.nx, g.ny, ze - zb), block(BX, BY);
    KM_DISPATCH(smooth_simple, km, pow2, p);
    return done(cx);
}

cudaError_t ks_residual(const LaunchCtx& cx, int km, bool pow2, const Geom& g, int zb, int ze, const PointK& k,
                        const StencilOff& so, const double* v, const double* f, const double* aux, double* r)
{
    if (ze <= zb || g.nx <= 0 || g.ny <= 0) return cudaErrorInvalidValue;
    ResidP p{g, k, {}, zb, v, f, aux, r, cx.norm};
    fill_toff(g, so, p.toff);
    dim3 grid = grid_for(g.nx, g.ny, ze - zb), block(BX, BY);
    if ((size_t)grid.x * grid.y * grid.z > cx.norm.cap) return cudaErrorInvalidValue;
    KM_DISPATCH(residual_simple, km, pow2, p);
    return done(cx);
}

// ---------------------------------------------------------------------------------
// residual + norm with the reduction scratch in static device memory (gs_internal.h ks_residual_static)
// ---------------------------------------------------------------------------------
// Layout of the array: [0, MAX) one partial sum per CTA, [MAX] the result, [MAX + 1] the ticket (its first four bytes).
// In HOST code the name g_static_norm_scratch is only a stand-in for the symbol, not a device address: the address is asked
// for once per device, with that device current, by cudaGetSymbolAddress below.
static __device__ __align__(256) double g_static_norm_scratch[GS_STATIC_SCRATCH_MAX_CTAS + 2];

size_t ks_residual_ctas(const Geom& g, int planes)
{
    if (planes <= 0 || g.nx <= 0 || g.ny <= 0) return 0;
    const dim3 grid = grid_for(g.nx, g.ny, planes);
    return (size_t)grid.x * grid.y * grid.z;
}

cudaError_t ks_residual_static(const LaunchCtx& cx_in, int km, bool pow2, const Geom& g, int zb, int ze, const PointK& k,
                               const StencilOff& so, const double* v, const double* f, const double* aux, double* r,
                               double** result_dev)
{
    constexpr int MAXDEV = 64;
    static std::mutex mu;
    static double* base_of[MAXDEV] = {};  // device address of the array, per device
    if (ze <= zb || g.nx <= 0 || g.ny <= 0) return cudaErrorInvalidValue;
    const size_t ctas = ks_residual_ctas(g, ze - zb);
    if (ctas > (size_t)GS_STATIC_SCRATCH_MAX_CTAS) return cudaErrorInvalidValue;
    int dev = -1;
    if (cudaError_t e = cudaGetDevice(&dev)) return e;
    if (dev < 0 || dev >= MAXDEV) return cudaErrorInvalidDevice;
    double* base = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!base_of[dev]) {
            void* ptr = nullptr;
            if (cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_norm_scratch)) return e;
            base_of[dev] = static_cast<double*>(ptr);
        }
        base = base_of[dev];
    }
    // the array keeps the last call's contents: result and ticket start from zero (the partial sums need not: every CTA
    // writes its slot before the last one reads them)
    if (cudaError_t e = cudaMemsetAsync(base + GS_STATIC_SCRATCH_MAX_CTAS, 0, 2 * sizeof(double), cx_in.stream)) return e;
    LaunchCtx cx = cx_in;
    cx.tune.pdl = 0;  // a plain launch behind the memset
    cx.norm.partials = base;
    cx.norm.result = base + GS_STATIC_SCRATCH_MAX_CTAS;
    cx.norm.ticket = reinterpret_cast<unsigned int*>(base + GS_STATIC_SCRATCH_MAX_CTAS + 1);
    cx.norm.cap = GS_STATIC_SCRATCH_MAX_CTAS;
    ResidP p{g, k, {}, zb, v, f, aux, r, cx.norm};
    fill_toff(g, so, p.toff);
    dim3 grid = grid_for(g.nx, g.ny, ze - zb), block(BX, BY);
    KM_DISPATCH(residual_simple, km, pow2, p);
    if (result_dev) *result_dev = cx.norm.result;
    return done(cx);
}

cudaError_t ks_restrict(const LaunchCtx& cx, const Geom& gf, const Geom& gc, int zb, int ze, const double* fine,
                        double* coarse, double* coarse2)
{
    if (ze <= zb || gc.nx <= 0 || gc.ny <= 0) return cudaSuccess;
    RestrictP p{gf, gc, zb, fine, coarse, coarse2};
    if (cudaError_t le = gs_launch(cx, restrict_simple, grid_for(gc.nx, gc.ny, ze - zb), dim3(BX, BY), 0, p)) return le;
    return done(cx);
}

cudaError_t ks_prolong_correct(const LaunchCtx& cx, const Geom& gf, const Geom& gc, int zb, int ze, const double* vin,
                               double* vout, const double* coarse, double* e_out)
{
    if (ze <= zb) return cudaSuccess;
    ProlongP p{gf, gc, zb, vin, vout, coarse, e_out};
    if (cudaError_t le = gs_launch(cx, prolong_correct_simple, grid_for(gf.nx, gf.ny, ze - zb), dim3(BX, BY), 0, p)) return le;
    return done(cx);
}

cudaError_t ks_fas_rhs(const LaunchCtx& cx, bool pow2, const Geom& g, int zb, int ze, const PointK& k,
                       const StencilOff& so, const double* restv, double* f, double* r_out)
{
    if (ze <= zb) return cudaSuccess;
    FasP p{g, k, {}, zb, restv, f, r_out};
    fill_toff(g, so, p.toff);
    dim3 grid = grid_for(g.nx, g.ny, ze - zb), block(BX, BY);
    if (cudaError_t le = pow2 ? gs_launch(cx, fas_rhs_simple<true>, grid, block, 0, p) : gs_launch(cx, fas_rhs_simple<false>, grid, block, 0, p)) return le;
    return done(cx);
}

cudaError_t ks_newton_F(const LaunchCtx& cx, bool pow2, const Geom& g, int zb, int ze, const PointK& k,
                        const StencilOff& so, const double* nv, const double* newtonF, double* f)
{
    if (ze <= zb) return cudaErrorInvalidValue;
    NewtonFP p{g, k, {}, zb, nv, newtonF, f, cx.norm};
    fill_toff(g, so, p.toff);
    dim3 grid = grid_for(g.nx, g.ny, ze - zb), block(BX, BY);
    if ((size_t)grid.x * grid.y * grid.z > cx.norm.cap) return cudaErrorInvalidValue;
    if (cudaError_t le = pow2 ? gs_launch(cx, newton_F_simple<true>, grid, block, 0, p) : gs_launch(cx, newton_F_simple<false>, grid, block, 0, p)) return le;
    return done(cx);
}

cudaError_t ks_newton_coef(const LaunchCtx& cx, const Geom& g, int zb, int ze, double gamma, const double* nv, double* coef)
{
    if (ze <= zb) return cudaSuccess;
    if (cudaError_t le = gs_launch(cx, newton_coef_simple, grid_for(g.nx, g.ny, ze - zb), dim3(BX, BY), 0, g, zb, gamma, nv, coef)) return le;
    return done(cx);
}

cudaError_t ks_axpy(const LaunchCtx& cx, const Geom& g, int zb, int ze, double* dst, const double* src, int sign)
{
    if (ze <= zb) return cudaSuccess;
    if (cudaError_t le = gs_launch(cx, axpy_simple, grid_for(g.nx, g.ny, ze - zb), dim3(BX, BY), 0, g, zb, dst, src, sign)) return le;
    return done(cx);
}

cudaError_t ks_init_rhs(const LaunchCtx& cx, const Geom& g, int mode, double h, double gamma, double* f)
{
    if (cudaError_t le = gs_launch(cx, init_rhs_simple, grid_for(g.X, g.Y, g.nzl), dim3(BX, BY), 0, g, mode, h, gamma, f)) return le;
    return done(cx);
}
