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

#include "gs_device.cuh"

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn()
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

// one descriptor per (array, geometry, box); arrays live as long as the handle, so cache them (gs_destroy drops them)
typedef std::tuple<const void*, i64, int, int, int, int> TmKey;
static std::map<TmKey, CUtensorMap> g_tm_cache;
static std::mutex g_tm_mu;

bool gs_tensor_map(const double* base, const Geom& g, int bw, int bh, CUtensorMap* out)
{
    typedef TmKey Key;
    std::map<Key, CUtensorMap>& cache = g_tm_cache;
    std::lock_guard<std::mutex> lock(g_tm_mu);
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

// forget the descriptors of an array that is about to be freed (gs_destroy): a later allocation may reuse the address
void gs_tensor_map_drop(const void* base)
{
    if (!base) return;
    std::lock_guard<std::mutex> lock(g_tm_mu);
    for (auto it = g_tm_cache.begin(); it != g_tm_cache.end();) {
        if (std::get<0>(it->first) == base) it = g_tm_cache.erase(it);
        else ++it;
    }
}

// planes per CTA: as long as tune.zchunk allows (each chunk re-reads two planes), but short
// enough that small levels still put several CTAs on every SM
int gs_pick_zchunk(const LaunchCtx& cx, int tiles, int planes, int max_chunk)
{
    const int want = (cx.sm_count * 8 + tiles - 1) / tiles;  // chunks wanted along z
    int chunk = (planes + want - 1) / want;
    if (chunk < 8) chunk = 8;
    if (chunk > max_chunk) chunk = max_chunk;
    if (chunk > planes) chunk = planes;
    return chunk;
}


// Wave-aware variant: a z-marching CTA costs about (chunk + overhead) plane-times — `overhead` = rim planes it
// re-reads / recomputes plus the pipeline fill — and the grid runs in ceil(CTAs / resident slots) waves, the
// last of which leaves SMs idle when it is partial.  Choose the chunk count that minimises waves x CTA cost, so the
// grid is (close to) a whole multiple of sm_count x resident CTAs per SM.
int gs_pick_zchunk_waves(const LaunchCtx& cx, int tiles, int planes, int ctas_per_sm, int overhead, int min_chunk, int max_chunk,
                         bool even)
{
    if (planes <= 0) return 1;
    const long slots = (long)cx.sm_count * (ctas_per_sm > 0 ? ctas_per_sm : 1);
    if (max_chunk > planes) max_chunk = planes;
    if (min_chunk > max_chunk) min_chunk = max_chunk;
    long best_cost = -1;
    int best_chunk = max_chunk;
    for (int n = 1; n <= planes; n++) {
        int chunk = (planes + n - 1) / n;
        if (even) chunk += chunk & 1;
        if (chunk > max_chunk) continue;
        if (chunk < min_chunk) break;
        const long nch = (planes + chunk - 1) / chunk;
        const long waves = (tiles * nch + slots - 1) / slots;
        const long cost = waves * (chunk + overhead);
        if (best_cost < 0 || cost < best_cost) {
            best_cost = cost;
            best_chunk = chunk;
        }
    }
    return best_chunk;
}

bool kf_available() { return encode_fn() != nullptr; }

namespace {

enum { OP_SMOOTH = 0, OP_RESID = 1, OP_SMOOTH_NORM = 2 };  // 2: the sweep's new iterate AND sum r^2 of the old one

struct FastP {
    Geom g;
    PointK k;
    int zb, ze, zchunk;
    const double* vin;
    double* vout;  // OP_SMOOTH: the new iterate; OP_RESID: r (may be null: norm only)
    const double* f;
    const double* aux;
    NormBuf nb;    // OP_RESID
    bool gen;                    // config-driven offsets: take the GEN kernels
    int gdx[7], gdy[7], gdz[7];  // GEN kernels: the config's seven offsets (src/gridParams.h:7-26), each in {-1,0,1}
};

template <int KM, int AR>
__device__ __forceinline__ double point_update(double C, double E, double W, double N, double S, double U, double D, double fv,
                                               double aux, const PointK& k)
{
    double den;
    const double r = canon_residual<KM, AR>(C, E, W, N, S, U, D, fv, aux, k, den);
    return jacobi_update<KM>(C, r, den, k);
}

// ---------------------------------------------------------------------------------
// variant 1: register-queue z-march, in-plane neighbours through L1
// ---------------------------------------------------------------------------------
template <int KM, int AR, int BX, int BY>
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
        p.vout[c] = point_update<KM, AR>(C, E, W, N, S, U, D, fv, aux, p.k);
        D = C;
        C = U;
        c += plane;
    }
}

// ---------------------------------------------------------------------------------
// variant 2: TMA-staged z-march
// ---------------------------------------------------------------------------------
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

// GEN: config-driven offsets (any seven taps inside the one-cell rim, src/gridParams.h:7-26, src/main.cpp:63-82).  The v
// box already carries the full xy rim and the ring holds planes z-1, z, z+1, so tap i is one shared-memory read at the
// CTA-uniform displacement (dx_i, dy_i) in the box of plane z+dz_i; the sum runs over i = 0..6 in config order like the
// reference (src/cpu/CpuSolver.cpp:56-60), hence the same bits as kernels_simple.cu.  The only structural difference
// from the canonical kernel: a stage is released one step later (plane q-1 is still read at step q).
template <int KM, int AR, int OP, int TX, int TY, int NSTAGE, bool GEN = false>
__global__ void __launch_bounds__(TX* TY + 32) stencil_tma(const __grid_constant__ CUtensorMap map_v,
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
        mbar_fence_init();
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

    double sq = 0.0;
    mbar_wait(&full[0], 0);
    double D = vbox(0)[ctr];  // (stage 0 is released at the end of the first iteration, once D has been used)
    mbar_wait(&full[1 % NSTAGE], (1 / NSTAGE) & 1);
    double C = vbox(1)[ctr];
    int goff[7];  // GEN: displacement of tap i inside a v box
    if (GEN) {
#pragma unroll
        for (int i = 0; i < 7; i++) goff[i] = ctr + p.gdy[i] * Cfg::VW + p.gdx[i];
    }

    // the Newton coefficient / iterate is the one operand outside the TMA ring: fetched one plane ahead into a register and
    // three planes ahead into L2, so no step waits a DRAM round trip for it (70 % of the stalls before, profiles/r02a_ncu_modes_summary.txt)
    double aux_next = (KM >= KM_NEWTON_EXP) ? p.aux[c] : 0.0;
    for (int q = 1; q <= nz; q++) {
        const double aux = aux_next;
        if (KM >= KM_NEWTON_EXP) {
            if (q < nz) aux_next = p.aux[c + p.g.plane];
            if (q + 3 <= nz) prefetch_l2(p.aux + c + 3 * p.g.plane);
        }
        mbar_wait(&full[(q + 1) % NSTAGE], ((q + 1) / NSTAGE) & 1);
        const double* vb = vbox(q);
        const double U = vbox(q + 1)[ctr];
        const double fv = fbox(q)[ty * TX + tx];
        if (GEN) {
            const double* vd = vbox(q - 1);
            const double* vu = vbox(q + 1);
            double s = 0.0;
#pragma unroll
            for (int i = 0; i < 7; i++) {
                const double* b = p.gdz[i] < 0 ? vd : (p.gdz[i] > 0 ? vu : vb);
                s += p.k.c[i] * b[goff[i]];
            }
            double den;
            const double r = finish_residual<KM, AR>(s, C, fv, aux, p.k, den);
            if (live) {
                if (OP == OP_RESID) {
                    if (p.vout) p.vout[c] = r;
                } else {
                    p.vout[c] = jacobi_update<KM>(C, r, den, p.k);
                }
                if (OP != OP_SMOOTH) sq += r * r;
            }
            __syncwarp();
            if (lane0) mbar_arrive(&empty[(q - 1) % NSTAGE]);  // plane q-1 was read in this step for the last time
            D = C;
            C = U;
            c += p.g.plane;
            continue;
        }
        const double E = vb[ctr + 1], W = vb[ctr - 1], N = vb[ctr + Cfg::VW], S = vb[ctr - Cfg::VW];
        if (OP == OP_SMOOTH) {
            const double nv = point_update<KM, AR>(C, E, W, N, S, U, D, fv, aux, p.k);
            if (live) p.vout[c] = nv;
        } else if (OP == OP_SMOOTH_NORM) {
            double den;
            const double r = canon_residual<KM, AR>(C, E, W, N, S, U, D, fv, aux, p.k, den);
            if (live) {
                p.vout[c] = jacobi_update<KM>(C, r, den, p.k);
                sq += r * r;
            }
        } else {
            double den;
            const double r = canon_residual<KM, AR>(C, E, W, N, S, U, D, fv, aux, p.k, den);
            if (live) {
                if (p.vout) p.vout[c] = r;
                sq += r * r;  // per-thread running sum in plane order: deterministic
            }
        }
        // release the stage of plane q only now, after the arithmetic above has consumed everything that was loaded from
        // it: the arrival is then ordered behind the completion of those shared-memory loads by data dependence, not just
        // by issue order (a ring refilled under a slow warp is what broke the producer-warp residual+restriction kernel)
        __syncwarp();
        if (lane0) {
            if (q == 1) mbar_arrive(&empty[0]);
            mbar_arrive(&empty[q % NSTAGE]);
        }
        D = C;
        C = U;
        c += p.g.plane;
    }
    // the last plane (q = nz+1) was only read for its centre; nobody refills it
    if (OP != OP_SMOOTH) {
        __shared__ double sh[32];
        const double tot = block_sum<Cfg::CONSUMERS>(sq, sh, tid);
        grid_sum_finish<Cfg::CONSUMERS>(tot, p.nb, sh, tid);
    }
}

// ---------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------
template <int KM, int AR, int OP, int TX, int TY, int NSTAGE, bool GEN = false>
cudaError_t launch_tma(const LaunchCtx& cx, FastP p, bool* handled)
{
    using Cfg = TmaCfg<TX, TY, NSTAGE>;
    // raise the dynamic shared memory limit once per instantiation and device (outside stream capture: kf_prepare()
    // instantiates every kernel the dispatchers can reach)
    static DevOnce once;
    int occ = 1;
    cudaError_t pe = gs_dev_once(once, &occ, [](int* o) {
        cudaError_t e = cudaFuncSetAttribute(stencil_tma<KM, AR, OP, TX, TY, NSTAGE, GEN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::SMEM);
        if (e == cudaSuccess)
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(o, stencil_tma<KM, AR, OP, TX, TY, NSTAGE, GEN>, Cfg::THREADS, Cfg::SMEM);
        return e;
    });
    if (pe != cudaSuccess) return pe;
    if (!p.vin) return cudaSuccess;  // preparation call
    CUtensorMap mv, mf;
    if (!gs_tensor_map(p.vin, p.g, Cfg::VW, Cfg::VH, &mv) || !gs_tensor_map(p.f, p.g, TX, TY, &mf)) return cudaSuccess;  // not handled
    const int tiles = ((p.g.nx + TX - 1) / TX) * ((p.g.ny + TY - 1) / TY);
    if (cx.tune.wave) p.zchunk = gs_pick_zchunk_waves(cx, tiles, p.ze - p.zb, occ, 2 + NSTAGE / 2, 8, p.zchunk, false);
    else p.zchunk = gs_pick_zchunk(cx, tiles, p.ze - p.zb, p.zchunk);
    const int nzc = (p.ze - p.zb + p.zchunk - 1) / p.zchunk;
    dim3 grid((p.g.nx + TX - 1) / TX, (p.g.ny + TY - 1) / TY, nzc);
    if (OP != OP_SMOOTH && (size_t)grid.x * grid.y * grid.z > cx.norm.cap) return cudaSuccess;  // not handled
    stencil_tma<KM, AR, OP, TX, TY, NSTAGE, GEN><<<grid, Cfg::THREADS, Cfg::SMEM, cx.stream>>>(mv, mf, p);
    *handled = true;
    return cudaGetLastError();
}

template <int KM, int AR, int BX, int BY>
cudaError_t launch_zmarch(const LaunchCtx& cx, FastP p, bool* handled)
{
    if (!p.vin) return cudaSuccess;
    p.zchunk = gs_pick_zchunk(cx, ((p.g.nx + BX - 1) / BX) * ((p.g.ny + BY - 1) / BY), p.ze - p.zb, p.zchunk);
    const int nzc = (p.ze - p.zb + p.zchunk - 1) / p.zchunk;
    dim3 grid((p.g.nx + BX - 1) / BX, (p.g.ny + BY - 1) / BY, nzc);
    smooth_zmarch<KM, AR, BX, BY><<<grid, dim3(BX, BY), 0, cx.stream>>>(p);
    *handled = true;
    return cudaGetLastError();
}

// id, TX, TY, NSTAGE
#define TMA_VARIANTS(X) X(2, 64, 4, 4) X(21, 64, 8, 4) X(22, 128, 4, 4)
// id, BX, BY
#define ZMARCH_VARIANTS(X) X(1, 64, 4) X(13, 64, 8)

template <int KM, int AR, int OP>
cudaError_t dispatch_gen(const LaunchCtx& cx, const FastP& p, bool* handled)
{
    // one more stage than the canonical kernels: three planes are being read, two are in flight
    if (p.g.nx >= 96) return launch_tma<KM, AR, OP, 128, 4, 5, true>(cx, p, handled);
    return launch_tma<KM, AR, OP, 64, 4, 5, true>(cx, p, handled);
}

template <int KM, int AR>
cudaError_t dispatch_smooth(const LaunchCtx& cx, const FastP& p, int variant, bool* handled)
{
    if (p.gen) return dispatch_gen<KM, AR, OP_SMOOTH>(cx, p, handled);
    switch (variant) {
#define X(id, bx, by) \
    case id: return launch_zmarch<KM, AR, bx, by>(cx, p, handled);
        ZMARCH_VARIANTS(X)
#undef X
#define X(id, tx, ty, ns) \
    case id: return launch_tma<KM, AR, OP_SMOOTH, tx, ty, ns>(cx, p, handled);
        TMA_VARIANTS(X)
#undef X
    default: return cudaSuccess;  // not handled: caller uses the simple kernel
    }
}

template <int KM, int AR, int OP>
cudaError_t dispatch_resid(const LaunchCtx& cx, const FastP& p, bool* handled)
{
    if (p.gen) return dispatch_gen<KM, AR, OP>(cx, p, handled);
    // wide tiles on big levels, the narrow tile where 128 columns would leave most of a CTA idle
    if (p.g.nx >= 96) return launch_tma<KM, AR, OP, 128, 4, 4>(cx, p, handled);
    return launch_tma<KM, AR, OP, 64, 4, 4>(cx, p, handled);
}

// run `what` for every (mode, pow2) instantiation
#define FOR_ALL_KM(CALL)                                           \
    do {                                                           \
        switch (km) {                                              \
        case KM_LINEAR: CALL(KM_LINEAR); break;                    \
        case KM_NONLINEAR: CALL(KM_NONLINEAR); break;              \
        case KM_NEWTON_EXP: CALL(KM_NEWTON_EXP); break;            \
        default: CALL(KM_NEWTON_COEF); break;                      \
        }                                                          \
    } while (0)

}  // namespace

static void set_gen(FastP& p, const StencilOff* so)
{
    p.gen = so != nullptr;
    for (int i = 0; i < 7; i++) {
        p.gdx[i] = so ? so->dx[i] : 0;
        p.gdy[i] = so ? so->dy[i] : 0;
        p.gdz[i] = so ? so->dz[i] : 0;
    }
}

cudaError_t kf_smooth(const LaunchCtx& cx, int km, int ar, const Geom& g, int zb, int ze, const PointK& k, const double* vin,
                      double* vout, const double* f, const double* aux, bool* handled, const StencilOff* gen)
{
    *handled = false;
    if (vin) {
        if (ze <= zb) return cudaSuccess;
        // small levels are latency-bound: the one-thread-per-point kernel has the most parallelism there
        if (g.nx < cx.tune.fast_min_n || g.ny < 8 || ze - zb < 8) return cudaSuccess;
    }
    FastP p{g, k, zb, ze, 0, vin, vout, f, aux, cx.norm};
    set_gen(p, gen);
    p.zchunk = cx.tune.zchunk > 0 ? cx.tune.zchunk : 64;
    if (vin && p.zchunk > ze - zb) p.zchunk = ze - zb;
    int variant = cx.tune.smooth_variant;
    if (vin && variant == 22 && g.nx < 96) variant = 2;
    cudaError_t e = cudaSuccess;
#define CALL(KMV) AR_SWITCH(ar, e = (dispatch_smooth<KMV, ARV>(cx, p, variant, handled)))
    FOR_ALL_KM(CALL);
#undef CALL
    if (e == cudaSuccess && *handled && cx.launches) ++*cx.launches;
    return e;
}

cudaError_t kf_residual(const LaunchCtx& cx, int km, int ar, const Geom& g, int zb, int ze, const PointK& k, const double* v,
                        const double* f, const double* aux, double* r, bool* handled, const StencilOff* gen)
{
    *handled = false;
    if (v) {
        if (ze <= zb) return cudaSuccess;
        if (g.nx < cx.tune.fast_min_n || g.ny < 8 || ze - zb < 8) return cudaSuccess;
    }
    FastP p{g, k, zb, ze, 0, v, r, f, aux, cx.norm};
    set_gen(p, gen);
    p.zchunk = cx.tune.zchunk > 0 ? cx.tune.zchunk : 64;
    if (v && p.zchunk > ze - zb) p.zchunk = ze - zb;
    cudaError_t e = cudaSuccess;
#define CALL(KMV) AR_SWITCH(ar, e = (dispatch_resid<KMV, ARV, OP_RESID>(cx, p, handled)))
    FOR_ALL_KM(CALL);
#undef CALL
    if (e == cudaSuccess && *handled && cx.launches) ++*cx.launches;
    return e;
}

// vout = one Jacobi sweep of vin, and sum of squares of the residual of vin -> cx.norm.result: the residual the
// V-cycle returns (src/cpu/CpuSolver.cpp:138) is the one the next cycle's first pre-smoothing sweep starts from (:88)
cudaError_t kf_smooth_norm(const LaunchCtx& cx, int km, int ar, const Geom& g, int zb, int ze, const PointK& k,
                           const double* vin, double* vout, const double* f, const double* aux, bool* handled, const StencilOff* gen)
{
    *handled = false;
    if (vin) {
        if (ze <= zb) return cudaSuccess;
        if (g.nx < cx.tune.fast_min_n || g.ny < 8 || ze - zb < 8) return cudaSuccess;
    }
    FastP p{g, k, zb, ze, 0, vin, vout, f, aux, cx.norm};
    set_gen(p, gen);
    p.zchunk = cx.tune.zchunk > 0 ? cx.tune.zchunk : 64;
    if (vin && p.zchunk > ze - zb) p.zchunk = ze - zb;
    cudaError_t e = cudaSuccess;
#define CALL(KMV) AR_SWITCH(ar, e = (dispatch_resid<KMV, ARV, OP_SMOOTH_NORM>(cx, p, handled)))
    FOR_ALL_KM(CALL);
#undef CALL
    if (e == cudaSuccess && *handled && cx.launches) ++*cx.launches;
    return e;
}

// one-time per-device setup: every launcher raises its kernel's dynamic shared memory limit on a
// null-pointer call, so that nothing but launches happens later under stream capture
cudaError_t kf_prepare()
{
    LaunchCtx cx{};
    Geom g{};
    PointK k{};
    bool handled;
    cudaError_t e = cudaSuccess;
    const int variants[] = {1, 13, 2, 21, 22};
    for (int km = 0; km < 4 && e == cudaSuccess; km++)
        for (int pw = 0; pw < 8 && e == cudaSuccess; pw++) {
            for (int var : variants) {
                cx.tune.smooth_variant = var;
                if (e == cudaSuccess) e = kf_smooth(cx, km, pw, g, 0, 0, k, nullptr, nullptr, nullptr, nullptr, &handled, nullptr);
            }
            StencilOff so{};
            for (int nx : {64, 128}) {
                g.nx = nx;
                for (const StencilOff* gen : {(const StencilOff*)nullptr, (const StencilOff*)&so}) {
                    if (e == cudaSuccess) e = kf_residual(cx, km, pw, g, 0, 0, k, nullptr, nullptr, nullptr, nullptr, &handled, gen);
                    if (e == cudaSuccess) e = kf_smooth_norm(cx, km, pw, g, 0, 0, k, nullptr, nullptr, nullptr, nullptr, &handled, gen);
                    if (e == cudaSuccess && gen) e = kf_smooth(cx, km, pw, g, 0, 0, k, nullptr, nullptr, nullptr, nullptr, &handled, gen);
                }
            }
            g.nx = 0;
        }
    return e;
}
