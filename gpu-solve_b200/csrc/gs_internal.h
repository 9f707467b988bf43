// gs_internal.h — data structures shared by the C-ABI layer (gs_api.cu) and the
// kernel translation units.  Nothing here crosses the C ABI.
//
// Device layout of every level array (DESIGN.md "Data layout in HBM"):
//   x fastest, row pitch padded to a multiple of 16 doubles with logical x = 1
//   (first interior column) on a 128-byte boundary:  elem(x,y,zl) =
//   GS_XOFF + x + pitch*y + plane*zl,  plane = pitch*Y.  The one-cell Dirichlet
//   halo of the reference (src/cpu/CpuGridData.cpp:32-36) is materialised and
//   stays zero.  `zl` is a LOCAL plane: a rank holds global planes
//   [z0, z0 + nzl) of the (nz+2) logical planes (all of them on one GPU).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/gpusolve_cuda.h"
#include "../../include/gpusolve_cuda_tuning.h"

#define GS_XOFF 15  // doubles in front of logical x = 0, so that x = 1 sits at element 16

typedef long long i64;

struct Geom {
    int X, Y, Zg;    // logical array dims (n + 2) of the whole level
    int nx, ny, nz;  // interior points of the whole level
    int z0;          // global plane index of local plane 0
    int nzl;         // local planes held
    i64 pitch;       // doubles per row
    i64 plane;       // doubles per plane (= pitch * Y)
};

__host__ __device__ __forceinline__ i64 gs_idx(const Geom& g, int x, int y, int zl)
{
    return (i64)GS_XOFF + x + g.pitch * y + g.plane * (i64)zl;
}

// per-level constants of the point update, computed on the host exactly as the
// reference does (src/cpu/CpuSolver.cpp:144-145)
struct PointK {
    double c[7];     // stencil values, config order
    double hh;       // level.h * level.h
    double inv_hh;   // RN(1/hh): s*inv_hh == s/hh bit for bit when hh is a power of two, else the seed of the corrected quotient (AR_RCP)
    double alpha;    // (h*h)/c[0]
    double preFac;   // c[0]/(h*h)
    double omega, gamma;
};

struct StencilOff {
    int dx[7], dy[7], dz[7];
};

// kernel "mode" template values
enum { KM_LINEAR = 0, KM_NONLINEAR = 1, KM_NEWTON_EXP = 2, KM_NEWTON_COEF = 3 };

// device-side scratch for deterministic norms (fixed-order two-stage reduction)
struct NormBuf {
    double* partials;     // one slot per CTA of the reducing kernel
    unsigned int* ticket; // last-CTA-done counter (self-resetting)
    double* result;       // device scalar: sum of squares
    size_t cap;           // slots in partials
};

// run-time tuning knobs (gs_set_option / GPUSOLVE_* environment)
struct Tune {
    int smooth_variant;  // 0 simple, 1x z-march through L1, 2x TMA-staged z-march (kernels_fast.cu)
    int zchunk;          // planes per CTA of the z-marching kernels
    int fast_min_n;      // levels with fewer interior points per axis use the simple kernels
    int rr_variant;      // fused residual+restriction kernel: 0 = five-stage ring, 1 = six-stage (default)
    int fuse2;           // two sweeps per pass (kernels_fused.cu): 0 off, 1 = five-stage ring, 2 = four-stage
    int fuse2_min_n;     // levels with fewer interior points per axis sweep one at a time
    int pc_min_n;        // levels with fewer interior points per axis keep prolongation + correction as its own pass
    int pdl;             // one-thread-per-point kernels are launched with programmatic stream serialization (kernels_simple.cu)
    int s2_ty;           // tile height of the two-sweep kernel: 0 = per mode (12 linear, 10 otherwise), 10 / 12 forced (A/B)
    int zchunk2;         // planes per CTA of the two-sweep kernel
    int wave;            // 1: size z chunks so the grid fills whole waves of resident CTAs (gs_pick_zchunk_waves)
};

struct LaunchCtx {
    cudaStream_t stream;
    Tune tune;
    NormBuf norm;
    uint64_t* launches;   // host counter of kernels launched
    int sm_count;
};

// ---- kernels_ctl.cu: device-side stopping test of the solve loop (conditional CUDA graph) ----
constexpr int GS_STOP_CAP = 64;  // most cycles one device-side loop can run (its residual history is a fixed array)
struct StopCtl {
    double initial;               // norm before the first cycle
    unsigned count, pad;          // cycles run
    double hist[GS_STOP_CAP + 1]; // [0] initial, [i] after cycle i
};
// sqrt(cx.norm.result) -> initial / hist[0], loop condition = maxiter > 0
cudaError_t kc_stop_init(const LaunchCtx& cx, StopCtl* c, unsigned long long cond, unsigned maxiter);
// sqrt(cx.norm.result) -> hist[++count]; loop condition = !(res <= initial/(1/tol)) && count < maxiter
cudaError_t kc_stop_check(const LaunchCtx& cx, StopCtl* c, unsigned long long cond, unsigned maxiter, double tol);

// ---- kernels_simple.cu: one thread per lattice point, any 7 in-range offsets ----
// All take LOCAL plane ranges [zb, ze) of points to update.
cudaError_t ks_smooth(const LaunchCtx& cx, int km, bool pow2, const Geom& g, int zb, int ze, const PointK& k,
                      const StencilOff& so, const double* vin, double* vout, const double* f, const double* aux);
// r = f - A(v) (stored if r != nullptr) and sum of squares -> cx.norm.result
cudaError_t ks_residual(const LaunchCtx& cx, int km, bool pow2, const Geom& g, int zb, int ze, const PointK& k,
                        const StencilOff& so, const double* v, const double* f, const double* aux, double* r);
// No-allocation variant of ks_residual (gs_residual_static, no counterpart in
// the reference): the per-CTA partial sums, the ticket and the result of the norm reduction live in a file-scope static
// __device__ array of kernels_simple.cu instead of the handle's cudaMalloc'ed NormBuf.  Calls nothing that allocates.
// *result_dev receives the DEVICE address of the sum of squares (valid once the stream has run the kernel).  The launch
// needs ks_residual_ctas(g, ze - zb) <= GS_STATIC_SCRATCH_MAX_CTAS reduction slots (else cudaErrorInvalidValue).
// Every call on a device shares the one array: calls that overlap on the GPU race.
constexpr int GS_STATIC_SCRATCH_MAX_CTAS = 65536;  // 512 KB of partial sums per GPU that loads the library
size_t ks_residual_ctas(const Geom& g, int planes);
cudaError_t ks_residual_static(const LaunchCtx& cx, int km, bool pow2, const Geom& g, int zb, int ze, const PointK& k,
                               const StencilOff& so, const double* v, const double* f, const double* aux, double* r,
                               double** result_dev);
// coarse(interior) = full weighting of fine; coarse local planes [zb, ze)
cudaError_t ks_restrict(const LaunchCtx& cx, const Geom& gf, const Geom& gc, int zb, int ze, const double* fine,
                        double* coarse, double* coarse2);
// vout = vin + P(coarse v [- coarse restV]) on fine local planes [zb, ze); e_out (optional) receives e
cudaError_t ks_prolong_correct(const LaunchCtx& cx, const Geom& gf, const Geom& gc, int zb, int ze, const double* vin,
                               double* vout, const double* coarse, double* e_out);
// FAS: f += A(restV)/hh + gamma*restV*exp(restV)   (r_out optional: receives A(restV) like the reference's r)
cudaError_t ks_fas_rhs(const LaunchCtx& cx, bool pow2, const Geom& g, int zb, int ze, const PointK& k,
                       const StencilOff& so, const double* restv, double* f, double* r_out);
// Newton F: f = newtonF - (A(nv)/hh + gamma*nv*exp(nv)), sum of squares -> cx.norm.result
cudaError_t ks_newton_F(const LaunchCtx& cx, bool pow2, const Geom& g, int zb, int ze, const PointK& k,
                        const StencilOff& so, const double* nv, const double* newtonF, double* f);
// coef = (gamma*(1+nv))*exp(nv) over interior
cudaError_t ks_newton_coef(const LaunchCtx& cx, const Geom& g, int zb, int ze, double gamma, const double* nv,
                           double* coef);
// dst = dst + sign*src over interior points of local planes [zb, ze)
cudaError_t ks_axpy(const LaunchCtx& cx, const Geom& g, int zb, int ze, double* dst, const double* src, int sign);
// analytic right-hand side, src/cpu/CpuGridData.cpp:45-78 (local planes [0, nzl))
cudaError_t ks_init_rhs(const LaunchCtx& cx, const Geom& g, int mode, double h, double gamma, double* f);

// whole-array helpers (cudaMemsetAsync / cudaMemcpyAsync on the stream)
cudaError_t ks_zero(const LaunchCtx& cx, double* a, size_t count);
cudaError_t ks_copy(const LaunchCtx& cx, double* dst, const double* src, size_t count);

// ---- kernels_fast.cu: tuned sm_100a kernels for the canonical axis-aligned stencil ----
bool kf_available();
// one-time per-process setup (dynamic shared memory attributes); call outside stream capture
cudaError_t kf_prepare();
cudaError_t kf_smooth(const LaunchCtx& cx, int km, int ar, const Geom& g, int zb, int ze, const PointK& k,
                      const double* vin, double* vout, const double* f, const double* aux, bool* handled, const StencilOff* gen = nullptr);
// r = f - A(v) (stored if r != nullptr) and sum of squares -> cx.norm.result
cudaError_t kf_residual(const LaunchCtx& cx, int km, int ar, const Geom& g, int zb, int ze, const PointK& k,
                        const double* v, const double* f, const double* aux, double* r, bool* handled, const StencilOff* gen = nullptr);
// vout = sweep(vin) and sum of squares of the residual of vin -> cx.norm.result
cudaError_t kf_smooth_norm(const LaunchCtx& cx, int km, int ar, const Geom& g, int zb, int ze, const PointK& k,
                           const double* vin, double* vout, const double* f, const double* aux, bool* handled,
                           const StencilOff* gen = nullptr);
// `gen` != nullptr in the three calls above: the config's seven offsets (any taps inside the one-cell rim) instead of the
// canonical axis-aligned stencil — the GEN instantiations of stencil_tma

// ---- kernels_fused.cu: two Jacobi sweeps per pass (temporal blocking, bit-identical to two single sweeps) ----
cudaError_t kx_prepare();
// vout = S(S(vin)) on local planes [zb, ze); needs two valid planes of vin outside the range.  with_norm: also
// sum of squares of the residual of vin -> cx.norm.result
// coarse != nullptr: the sweeps start from vin + P(coarse) (prolongation + correction folded into the pass; gc = geometry of
// the coarse level; needs the coarse planes of fine planes zb-2 .. ze+1; no norm, no vzero)
cudaError_t kx_smooth2(const LaunchCtx& cx, int km, int ar, const Geom& g, int zb, int ze, const PointK& k,
                       const double* vin, double* vout, const double* f, const double* aux, bool with_norm, bool vzero,
                       bool* handled, const double* coarse = nullptr, const Geom* gc = nullptr, bool prepare_pc = false);

// ---- kernels_transfer.cu: tuned grid-transfer kernels (canonical stencil) ----
cudaError_t kt_prepare();
cudaError_t kt_prolong_correct(const LaunchCtx& cx, const Geom& gf, const Geom& gc, int zb, int ze, const double* vin,
                               double* vout, const double* coarse, bool* handled);
cudaError_t kt_restrict(const LaunchCtx& cx, const Geom& gf, const Geom& gc, int zcb, int zce, const double* fine, double* coarse,
                        double* coarse2, bool* handled);
cudaError_t kt_resid_restrict(const LaunchCtx& cx, int km, int ar, const Geom& gf, const Geom& gc, int zcb, int zce,
                              const PointK& k, const double* v, const double* f, const double* aux, double* coarse, bool* handled,
                              double* rv1 = nullptr, double* rv2 = nullptr);  // rv1/rv2 (mode 1): also R(v) -> both, within rounding
