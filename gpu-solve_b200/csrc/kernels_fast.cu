// kernels_fast.cu — tuned sm_100a smoother kernels for the canonical axis-aligned 7-point
// stencil (config order C,+x,-x,+y,-y,+z,-z; any coefficient values).
//
// Both variants are 2.5-D: a CTA owns an xy tile and marches along z (the slowest
// axis) keeping the z-1 / z / z+1 centre values in a register queue, so every v plane
// is fetched from L2/HBM once per sweep (24 B per lattice update: read v, read f,
// write v').
//   variant 2  "tma":    xy tiles (with their one-cell rim) and the f tile are staged in a
//                        shared-memory ring by TMA (cp.async.bulk.tensor.3d, SASS UTMALDG)
//                        issued by one producer warp; consumer warps wait on mbarriers —
//                        no __syncthreads in the steady state, prefetch depth NSTAGE-2 planes.
//   variant 1  "zmarch": no shared memory; the four in-plane neighbours are re-read
//                        through L1 (they were fetched one step earlier as some thread's
//                        z+1 value).
// Per-point arithmetic is the same expression sequence as kernels_simple.cu (and the
// reference), compiled with -fmad=false: results are bit-identical to the simple kernels.
#include <cuda.h>

#include <map>
#include <mutex>
#include <tuple>

#include "gs_internal.h"

namespace {

// ---------------------------------------------------------------------------------
// shared point arithmetic (identical to kernels_simple.cu: src/cpu/CpuSolver.cpp:56-74,158-171)
// ---------------------------------------------------------------------------------
template <int KM, bool POW2>
__device__ __forceinline__ double point_update(double C, double E, double W, double N, double S, double U, double D, double fv,
                                               double aux, const PointK& k)
{
    double s = 0.0;
    s += k.c[0] * C;
    s += k.c[1] * E;
    s += k.c[2] * W;
    s += k.c[3] * N;
    s += k.c[4] * S;
    s += k.c[5] * U;
    s += k.c[6] * D;
    s = POW2 ? s * k.inv_hh : s / k.hh;
    double den = 0.0;
    if (KM == KM_NONLINEAR) {
        double ex = exp(C);
        s += (k.gamma * C) * ex;
        den = k.preFac + (k.gamma * (1 + C)) * ex;
    } else if (KM == KM_NEWTON_EXP) {
        double ex = exp(aux);
        double g1 = k.gamma * (1 + aux);
        s += (g1 * C) * ex;
        den = k.preFac + g1 * ex;
    } else if (KM == KM_NEWTON_COEF) {
        s += aux * C;
        den = k.preFac + aux;
    }
    const double r = fv - s;
    if (KM == KM_LINEAR) return C + k.omega * (k.alpha * r);
    return C + k.omega * (r / den);
}

struct FastP {
    Geom g;
    PointK k;
    int zb, ze, zchunk;
    const double* vin;
    double* vout;
    const double* f;
    const double* aux;
};

// ---------------------------------------------------------------------------------
// variant 1: register-queue z-march, in-plane neighbours through L1
// ---------------------------------------------------------------------------------
template <int KM, bool POW2, int BX, int BY>
__global__ void __launch_bounds__(BX* BY) smooth_zmarch(FastP p)
{
    const int x = blockIdx.x * BX + threadIdx.x + 1;
    const int y = blockIdx.y * BY + threadIdx.y + 1;
    if (x > p.g.nx || y > p.g.ny) return;
    const int z0 = p.zb + blockIdx.z * p.zchunk;
    const int z1 = min(z0 + p.zchunk, p.ze);
    i64 c = gs_idx(p.g, x, y, z0);
    const i64 pitch = p.g.pitch, plane = p.g.plane;
    const double* __restrict__ v = p.vin;
    double D = v[c - plane], C = v[c];
#pragma unroll 2
    for (int z = z0; z < z1; z++) {
        const double U = v[c + plane];
        const double fv = p.f[c];
        const double aux = (KM >= KM_NEWTON_EXP) ? p.aux[c] : 0.0;
        const double E = v[c + 1], W = v[c - 1], N = v[c + pitch], S = v[c - pitch];
        p.vout[c] = point_update<KM, POW2>(C, E, W, N, S, U, D, fv, aux, p.k);
        D = C;
        C = U;
        c += plane;
    }
}

// ---------------------------------------------------------------------------------
// variant 2: TMA-staged z-march
// ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
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
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}

template <int TX, int TY, int NSTAGE>
struct TmaCfg {
    // v box with its rim.  A TMA box must START on a 16-byte boundary in global memory (measured on this
    // B200/driver: an 8-byte-aligned start raises "illegal instruction", tools/tma_probe.cu), and x0 sits on a
    // 128-byte boundary, so the box takes two columns on either side instead of one.
    static constexpr int VW = TX + 4, VH = TY + 2;
    static constexpr int V_BYTES = VW * VH * 8;
    static constexpr int V_STRIDE = (V_BYTES + 127) / 128 * 128;  // every TMA destination 128-B aligned
    static constexpr int F_BYTES = TX * TY * 8;
    static constexpr int STAGE = V_STRIDE + F_BYTES;
    static constexpr int SMEM = NSTAGE * STAGE + 2 * NSTAGE * 8 + 128;
    static constexpr int CONSUMERS = TX * TY;
    static constexpr int THREADS = CONSUMERS + 32;
};

template <int KM, bool POW2, int TX, int TY, int NSTAGE>
__global__ void __launch_bounds__(TX* TY + 32) smooth_tma(const __grid_constant__ CUtensorMap map_v,
                                                            const __grid_constant__ CUtensorMap map_f, FastP p)
{
    using Cfg = TmaCfg<TX, TY, NSTAGE>;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);  // stays a shared-space pointer (LDS, not LD)
    uint64_t* full = (uint64_t*)(smem + NSTAGE * Cfg::STAGE);
    uint64_t* empty = full + NSTAGE;

    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * TX + 1, y0 = blockIdx.y * TY + 1;  // first interior point of the tile
    const int z0 = p.zb + blockIdx.z * p.zchunk;
    const int z1 = min(z0 + p.zchunk, p.ze);
    const int nz = z1 - z0;         // planes this CTA updates
    const int nplanes = nz + 2;     // v planes z0-1 .. z1 it consumes

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], Cfg::CONSUMERS / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (tid >= Cfg::CONSUMERS) {
        // ---- producer warp: one lane streams planes into the ring ----
        if (tid == Cfg::CONSUMERS) {
            for (int q = 0; q < nplanes; q++) {
                const int s = q % NSTAGE;
                if (q >= NSTAGE) mbar_wait(&empty[s], ((q / NSTAGE) - 1) & 1);
                const bool has_f = q >= 1 && q <= nz;
                mbar_expect_tx(&full[s], Cfg::V_BYTES + (has_f ? Cfg::F_BYTES : 0));
                unsigned char* st = smem + s * Cfg::STAGE;
                tma_load_3d(st, &map_v, &full[s], GS_XOFF + x0 - 2, y0 - 1, z0 - 1 + q);
                if (has_f) tma_load_3d(st + Cfg::V_STRIDE, &map_f, &full[s], GS_XOFF + x0, y0, z0 - 1 + q);
            }
        }
        return;
    }

    // ---- consumers: one lattice column per thread ----
    const int tx = tid % TX, ty = tid / TX;
    const int x = x0 + tx, y = y0 + ty;
    const bool live = x <= p.g.nx && y <= p.g.ny;
    const int lane0 = (tid & 31) == 0;
    const int ctr = (ty + 1) * Cfg::VW + tx + 2;  // centre inside a v box
    i64 c = gs_idx(p.g, min(x, p.g.nx), min(y, p.g.ny), z0);

    auto vbox = [&](int q) { return (const double*)(smem + (q % NSTAGE) * Cfg::STAGE); };
    auto fbox = [&](int q) { return (const double*)(smem + (q % NSTAGE) * Cfg::STAGE + Cfg::V_STRIDE); };

    mbar_wait(&full[0], 0);
    double D = vbox(0)[ctr];
    __syncwarp();
    if (lane0) mbar_arrive(&empty[0]);
    mbar_wait(&full[1 % NSTAGE], (1 / NSTAGE) & 1);
    double C = vbox(1)[ctr];

    for (int q = 1; q <= nz; q++) {
        const double aux = (KM >= KM_NEWTON_EXP) ? p.aux[c] : 0.0;
        mbar_wait(&full[(q + 1) % NSTAGE], ((q + 1) / NSTAGE) & 1);
        const double* vb = vbox(q);
        const double U = vbox(q + 1)[ctr];
        const double E = vb[ctr + 1], W = vb[ctr - 1], N = vb[ctr + Cfg::VW], S = vb[ctr - Cfg::VW];
        const double fv = fbox(q)[ty * TX + tx];
        __syncwarp();
        if (lane0) mbar_arrive(&empty[q % NSTAGE]);
        const double nv = point_update<KM, POW2>(C, E, W, N, S, U, D, fv, aux, p.k);
        if (live) p.vout[c] = nv;
        D = C;
        C = U;
        c += p.g.plane;
    }
    // the last plane (q = nz+1) was only read for its centre; nobody refills it
}

// ---------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn()
{
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (tried) return fn;
    tried = true;
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
        fn = (EncodeTiledFn)sym;
    return fn;
}

// one descriptor per (array, geometry, box); arrays live as long as the handle, so cache them
bool tensor_map(const double* base, const Geom& g, int bw, int bh, CUtensorMap* out)
{
    typedef std::tuple<const void*, i64, int, int, int, int> Key;
    static std::map<Key, CUtensorMap> cache;
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    Key key(base, g.pitch, g.Y, g.nzl, bw, bh);
    auto it = cache.find(key);
    if (it != cache.end()) {
        *out = it->second;
        return true;
    }
    EncodeTiledFn enc = encode_fn();
    if (!enc) return false;
    cuuint64_t dims[3] = {(cuuint64_t)g.pitch, (cuuint64_t)g.Y, (cuuint64_t)g.nzl};
    cuuint64_t strides[2] = {(cuuint64_t)g.pitch * 8, (cuuint64_t)g.plane * 8};
    cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUtensorMap m;
    CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    if (cache.size() > 4096) cache.clear();
    cache[key] = m;
    *out = m;
    return true;
}

// planes per CTA: as long as tune.zchunk allows (each chunk re-reads two planes), but short
// enough that small levels still put several CTAs on every SM
int pick_zchunk(const LaunchCtx& cx, int tiles, int planes, int max_chunk)
{
    const int want = (cx.sm_count * 8 + tiles - 1) / tiles;  // chunks wanted along z
    int chunk = (planes + want - 1) / want;
    if (chunk < 8) chunk = 8;
    if (chunk > max_chunk) chunk = max_chunk;
    if (chunk > planes) chunk = planes;
    return chunk;
}

template <int KM, bool POW2, int TX, int TY, int NSTAGE>
cudaError_t launch_tma(const LaunchCtx& cx, FastP p, bool* handled)
{
    using Cfg = TmaCfg<TX, TY, NSTAGE>;
    CUtensorMap mv, mf;
    if (!tensor_map(p.vin, p.g, Cfg::VW, Cfg::VH, &mv) || !tensor_map(p.f, p.g, TX, TY, &mf)) return cudaSuccess;  // not handled
    p.zchunk = pick_zchunk(cx, ((p.g.nx + TX - 1) / TX) * ((p.g.ny + TY - 1) / TY), p.ze - p.zb, p.zchunk);
    const int nzc = (p.ze - p.zb + p.zchunk - 1) / p.zchunk;
    dim3 grid((p.g.nx + TX - 1) / TX, (p.g.ny + TY - 1) / TY, nzc);
    smooth_tma<KM, POW2, TX, TY, NSTAGE><<<grid, Cfg::THREADS, Cfg::SMEM, cx.stream>>>(mv, mf, p);
    *handled = true;
    return cudaGetLastError();
}

template <int KM, bool POW2, int TX, int TY, int NSTAGE>
cudaError_t prepare_tma()
{
    return cudaFuncSetAttribute(smooth_tma<KM, POW2, TX, TY, NSTAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                TmaCfg<TX, TY, NSTAGE>::SMEM);
}

template <int KM, bool POW2, int BX, int BY>
cudaError_t launch_zmarch(const LaunchCtx& cx, FastP p, bool* handled)
{
    p.zchunk = pick_zchunk(cx, ((p.g.nx + BX - 1) / BX) * ((p.g.ny + BY - 1) / BY), p.ze - p.zb, p.zchunk);
    const int nzc = (p.ze - p.zb + p.zchunk - 1) / p.zchunk;
    dim3 grid((p.g.nx + BX - 1) / BX, (p.g.ny + BY - 1) / BY, nzc);
    smooth_zmarch<KM, POW2, BX, BY><<<grid, dim3(BX, BY), 0, cx.stream>>>(p);
    *handled = true;
    return cudaGetLastError();
}

// id, TX, TY, NSTAGE
#define TMA_VARIANTS(X) X(2, 64, 4, 4) X(21, 64, 8, 4) X(22, 128, 4, 4) X(23, 64, 4, 6) X(24, 32, 8, 4) X(25, 128, 2, 6)
// id, BX, BY
#define ZMARCH_VARIANTS(X) X(1, 64, 4) X(11, 32, 8) X(12, 128, 2) X(13, 64, 8)

template <int KM, bool POW2>
cudaError_t dispatch(const LaunchCtx& cx, const FastP& p, bool* handled)
{
    switch (cx.tune.smooth_variant) {
#define X(id, bx, by) \
    case id: return launch_zmarch<KM, POW2, bx, by>(cx, p, handled);
        ZMARCH_VARIANTS(X)
#undef X
#define X(id, tx, ty, ns) \
    case id: return launch_tma<KM, POW2, tx, ty, ns>(cx, p, handled);
        TMA_VARIANTS(X)
#undef X
    default: return cudaSuccess;  // not handled: caller uses the simple kernel
    }
}

template <int KM, bool POW2>
cudaError_t prepare_all()
{
    cudaError_t e = cudaSuccess;
#define X(id, tx, ty, ns) \
    if (e == cudaSuccess) e = prepare_tma<KM, POW2, tx, ty, ns>();
    TMA_VARIANTS(X)
#undef X
    return e;
}

}  // namespace

bool kf_available() { return encode_fn() != nullptr; }

cudaError_t kf_prepare()
{
    cudaError_t e = cudaSuccess;
#define P(KMV)                                                \
    if (e == cudaSuccess) e = prepare_all<KMV, true>();       \
    if (e == cudaSuccess) e = prepare_all<KMV, false>();
    P(KM_LINEAR) P(KM_NONLINEAR) P(KM_NEWTON_EXP) P(KM_NEWTON_COEF)
#undef P
    return e;
}

cudaError_t kf_smooth(const LaunchCtx& cx, int km, bool pow2, const Geom& g, int zb, int ze, const PointK& k, const double* vin,
                      double* vout, const double* f, const double* aux, bool* handled)
{
    *handled = false;
    if (ze <= zb) return cudaSuccess;
    // small levels are latency-bound: the one-thread-per-point kernel has the most parallelism there
    if (g.nx < cx.tune.fast_min_n || g.ny < 8 || ze - zb < 8) return cudaSuccess;
    FastP p{g, k, zb, ze, 0, vin, vout, f, aux};
    p.zchunk = cx.tune.zchunk > 0 ? cx.tune.zchunk : 64;
    if (p.zchunk > ze - zb) p.zchunk = ze - zb;
    cudaError_t e;
#define GO(KMV)                                                                       \
    e = pow2 ? dispatch<KMV, true>(cx, p, handled) : dispatch<KMV, false>(cx, p, handled); \
    break
    switch (km) {
    case KM_LINEAR: GO(KM_LINEAR);
    case KM_NONLINEAR: GO(KM_NONLINEAR);
    case KM_NEWTON_EXP: GO(KM_NEWTON_EXP);
    default: GO(KM_NEWTON_COEF);
    }
#undef GO
    if (e == cudaSuccess && *handled && cx.launches) ++*cx.launches;
    return e;
}
