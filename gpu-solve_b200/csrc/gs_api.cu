// gs_api.cu — the C ABI of include/gpusolve_cuda.h: level hierarchy and device
// memory, V-cycle / solve / Newton orchestration (host control that the reference
// keeps in CpuSolver::solve/vcycle and NewtonSolver::solve/findError), CUDA-graph
// capture of the V-cycle, field I/O, config parsing.  All numerical work happens
// in the kernels of kernels_simple.cu / kernels_fast.cu; there is no CPU
// arithmetic path here beyond sqrt() of the reduced norm and level constants.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <string>
#include <vector>

#include "gs_comm.h"
#include "gs_internal.h"

int gs_pick_zchunk_waves(const LaunchCtx& cx, int tiles, int planes, int ctas_per_sm, int overhead, int min_chunk, int max_chunk,
                         bool even);
void gs_tensor_map_drop(const void* base);

// ---------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

#define CU(expr)                                                                                         \
    do {                                                                                                 \
        cudaError_t e__ = (expr);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return fail(GS_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

#define GS(expr)                  \
    do {                          \
        int rc__ = (expr);        \
        if (rc__ != GS_OK) return rc__; \
    } while (0)

// ---------------------------------------------------------------------------------
// handle
// ---------------------------------------------------------------------------------
// slab boundaries: see gs_partition
struct Partition {
    int nlev;
    int replicated_from;              // first level held whole by every rank
    std::vector<std::vector<int>> b;  // b[level][rank] .. b[level][world]
};

constexpr int GS_LEVEL_ARRAYS = 8;  // v buffer A, v buffer B, f, r, restV, newtonV, coef, (spare)

struct Level {
    Geom g;
    double h;
    PointK k;
    bool pow2, unit;
    int ar;  // arithmetic flags of the tuned kernels (gs_device.cuh AR_*), same bits either way
    int zlo, zhi;     // owned GLOBAL planes [zlo, zhi) (both physical halo planes belong to the end ranks)
    int ib, ie;       // LOCAL plane range [ib, ie) of owned interior planes
    int min_interior; // fewest owned interior planes of any rank on this level (= ie - ib on one GPU / replicated levels)
    bool replicated;  // whole level held (and computed) by every rank
    size_t count;     // doubles per array
    double *v, *vtmp, *f, *r, *restV, *newtonV, *coef;
    double* phys[GS_LEVEL_ARRAYS];  // the arrays as allocated (v and vtmp trade places, these do not): index = slot registered with the peer transport
    bool coef_valid;
    int gv;           // slabs: ghost planes of v that are current on each side (0..2); refreshed lazily by need_v_halo
    bool vzero_next;  // enqueue_vcycle: v is to be zeroed before the pre-smoothing that follows (done inside its first pass if possible)
    int pre_valid;    // vtmp == this many Jacobi sweeps of v for the current f/aux (written by the fused sweep+norm kernels)
};

// a whole solve loop as one graph: initial norm, WHILE (conditional node) { V-cycle; stopping test }, history copy
struct LoopGraph {
    cudaGraphExec_t exec = nullptr;  // null: this configuration cannot loop on the device (cycle not parity neutral)
    uint64_t launches_head = 0, launches_cycle = 0;
    int pre_valid_after = 0;
};

struct EventPair {
    cudaEvent_t a, b;
    int op, level;
};

struct gs_handle {
    gs_params p;
    int device, rank, world;
    std::vector<Level> lev;
    StencilOff so;
    bool canonical;
    double* newtonF;
    bool newton_begun;
    cudaStream_t stream;
    LaunchCtx cx;
    double* h_norm;  // pinned
    uint64_t launches;
    int opt_kernels, opt_graph, opt_timing, opt_fuse_norm, opt_graph_multi, opt_vzero, opt_timers, opt_unit, opt_rcp_div;
    std::map<uint64_t, cudaGraphExec_t> graphs;
    std::map<uint64_t, uint64_t> graphs_launches;  // kernels per replay of each graph
    std::map<uint64_t, int> graphs_pre_valid;      // lev[0].pre_valid after a replay of each graph
    std::map<uint64_t, std::vector<int>> graphs_gv;  // Level::gv after a replay of each graph
    std::map<uint64_t, uint64_t> graphs_swapped;   // bit l: level l ends a replay with v and vtmp exchanged
    std::vector<double> history;
    std::vector<EventPair> ev_used;
    std::vector<cudaEvent_t> ev_free;
    double op_ms[GS_OP_COUNT];
    uint64_t op_launches[GS_OP_COUNT];
    std::vector<double> lvl_ms;          // [level][op]
    std::vector<uint64_t> lvl_launches;  // [level][op]
    std::vector<uint64_t> lvl_sweeps;    // [level] Jacobi sweeps run by the GS_OP_SMOOTH launches
    GsComm* comm;
    Partition part;
    int replicated_from;
    double* stage;        // device staging buffer of field I/O (flat PCIe transfers, pitch conversion in HBM), two halves
    size_t stage_count;   // doubles per half
    cudaStream_t io_stream;        // PCIe copies of field I/O; the pitch conversion runs on `stream`, one chunk behind
    cudaEvent_t io_copied[2], io_converted[2];
    int opt_fuse_rv;      // FAS: R(v) -> restV, v computed by the fused residual+restriction pass (kernels_transfer.cu RV)
    int opt_fuse_pc;      // prolongation + correction folded into the first post-smoothing pass (kernels_fused.cu PC)
    int opt_p2p;          // slabs: ghost planes, coarse gather and norm through peer-mapped memory (kernels_comm.cu) instead of NCCL
    std::vector<long long> arena_offs;  // byte offset of every registered array in the arena (index level * GS_LEVEL_ARRAYS + slot; last = newtonF)
    char* arena;          // slabs: every level array lives in ONE allocation, shared with the peers by CUDA IPC
    size_t arena_bytes;
    int opt_device_stop;  // solve loops without printing run as ONE conditional graph, stopping test on the device (kernels_ctl.cu)
    StopCtl* d_stop;      // device-side loop state
    StopCtl* h_stop;      // pinned copy, filled by the graph's last node
    std::map<uint64_t, struct LoopGraph> loops;
};

namespace {

struct OpScope {  // CUDA-event bracket around one operator when "timing" is on
    gs_handle* h;
    int op, level;
    cudaEvent_t a, b;
    uint64_t l0;
    OpScope(gs_handle* h_, int op_, int level_) : h(h_), op(op_), level(level_), a(nullptr), b(nullptr), l0(h_->launches)
    {
        if (!h->opt_timing) return;
        a = get();
        b = get();
        cudaEventRecord(a, h->stream);
    }
    ~OpScope()
    {
        h->op_launches[op] += h->launches - l0;
        h->lvl_launches[(size_t)level * GS_OP_COUNT + op] += h->launches - l0;
        if (!h->opt_timing) return;
        cudaEventRecord(b, h->stream);
        h->ev_used.push_back({a, b, op, level});
    }
    cudaEvent_t get()
    {
        cudaEvent_t e;
        if (!h->ev_free.empty()) {
            e = h->ev_free.back();
            h->ev_free.pop_back();
        } else {
            cudaEventCreate(&e);
        }
        return e;
    }
};

// arithmetic flags of a level for the tuned kernels (gs_device.cuh AR_*)
int level_ar(const Level& L, bool unit_ok, bool rcp_ok)
{
    const int u = (L.unit && unit_ok) ? 2 : 0;
    if (L.pow2) return 1 | u;
    return rcp_ok ? (4 | u) : 0;
}

int num_levels_for(const gs_params& p)
{
    // src/cpu/CpuGridData.cpp:19
    uint64_t m = std::min(std::min(p.dim[0], p.dim[1]), p.dim[2]);
    return (int)floor(log((double)m) / log(2.0)) + 1;
}

void level_dims(const gs_params& p, int level, uint64_t n[3])
{
    // src/cpu/CpuGridData.cpp:23-29: integer halving per axis
    for (int d = 0; d < 3; d++) n[d] = p.dim[d];
    for (int l = 0; l < level; l++)
        for (int d = 0; d < 3; d++) n[d] /= 2;
}

bool is_pow2(double x)
{
    int e;
    return x > 0 && std::isfinite(x) && frexp(x, &e) == 0.5;
}

int validate(const gs_params& p)
{
    if (p.mode < 0 || p.mode > 2) return fail(GS_EINVAL, "Invalid mode %d", p.mode);
    for (int d = 0; d < 3; d++)
        if (p.dim[d] < 1 || p.dim[d] > 4000) return fail(GS_EINVAL, "grid dimension %d out of range: %llu", d, (unsigned long long)p.dim[d]);
    for (int i = 0; i < 7; i++)
        for (int d = 0; d < 3; d++)
            if (p.stencil_off[i][d] < -1 || p.stencil_off[i][d] > 1)
                return fail(GS_EINVAL, "stencil offset [%d][%d]=%d outside {-1,0,1} (halo width is 1)", i, d, p.stencil_off[i][d]);
    if (p.stencil_val[0] == 0.0) return fail(GS_EINVAL, "stencil centre value (first entry) must be non-zero");
    return GS_OK;
}

bool canonical_offsets(const gs_params& p)
{
    static const int c[7][3] = {{0, 0, 0}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
    for (int i = 0; i < 7; i++)
        for (int d = 0; d < 3; d++)
            if (p.stencil_off[i][d] != c[i][d]) return false;
    return true;
}

Partition make_partition(const gs_params& p, int world)
{
    Partition P;
    P.nlev = num_levels_for(p);
    P.b.resize(P.nlev);
    int min_planes = 4;
    if (const char* e = getenv("GPUSOLVE_SLAB_MIN")) min_planes = std::max(1, atoi(e));
    double min_points = 3.0e6;
    if (const char* e = getenv("GPUSOLVE_SLAB_MIN_POINTS")) min_points = atof(e);
    // deepest level that still gives every rank >= min_planes planes; never the last two levels
    int deepest = -1;
    if (world > 1) {
        for (int l = 0; l + 2 < P.nlev; l++) {
            uint64_t n[3];
            level_dims(p, l, n);
            int Zm1 = (int)n[2] + 1;
            int blk = (Zm1 + world - 1) / world;
            // small levels are latency-bound: every rank computes them whole (no halo exchange) once the
            // level has fewer than min_points interior points
            const double pts = (double)n[0] * (double)n[1] * (double)n[2];
            if (blk >= min_planes && (world - 1) * blk < Zm1 && pts >= min_points) deepest = l;
            else break;
        }
    }
    P.replicated_from = deepest + 1;
    for (int l = 0; l < P.nlev; l++) {
        uint64_t n[3];
        level_dims(p, l, n);
        const int Z = (int)n[2] + 2;
        P.b[l].assign(world + 1, 0);
        if (l >= P.replicated_from) {
            for (int r = 0; r <= world; r++) P.b[l][r] = r == 0 ? 0 : Z;  // everyone holds [0, Z)
            continue;
        }
        uint64_t nd[3];
        level_dims(p, deepest, nd);
        const int blk_deep = ((int)nd[2] + 1 + world - 1) / world;
        for (int r = 0; r < world; r++) P.b[l][r] = std::min(Z, (r * blk_deep) << (deepest - l));
        P.b[l][world] = Z;
    }
    return P;
}

int alloc_zero(double** p, size_t count)
{
    cudaError_t e = cudaMalloc((void**)p, count * sizeof(double));
    if (e != cudaSuccess) return fail(GS_ENOMEM, "cudaMalloc of %zu bytes failed: %s", count * sizeof(double), cudaGetErrorString(e));
    CU(cudaMemset(*p, 0, count * sizeof(double)));
    return GS_OK;
}

double** field_ptr(gs_handle* h, int level, int field)
{
    if (field == GS_FIELD_NEWTONF) return level == 0 ? &h->newtonF : nullptr;
    if (level < 0 || level >= (int)h->lev.size()) return nullptr;
    Level& L = h->lev[level];
    switch (field) {
    case GS_FIELD_V: return &L.v;
    case GS_FIELD_F: return &L.f;
    case GS_FIELD_R: return &L.r;
    case GS_FIELD_RESTV: return &L.restV;
    case GS_FIELD_NEWTONV: return &L.newtonV;
    }
    return nullptr;
}

int check_level(const gs_handle* h, int level)
{
    if (!h) return fail(GS_EINVAL, "null handle");
    if (level < 0 || level >= (int)h->lev.size()) return fail(GS_EINVAL, "level %d out of range [0,%d)", level, (int)h->lev.size());
    return GS_OK;
}

int smoother_km(const gs_handle* h)
{
    switch (h->p.mode) {
    case GS_MODE_LINEAR: return KM_LINEAR;
    case GS_MODE_NONLINEAR: return KM_NONLINEAR;
    default: return h->opt_kernels ? KM_NEWTON_COEF : KM_NEWTON_EXP;
    }
}

bool partitioned(const gs_handle* h, int level) { return h->world > 1 && !h->lev[level].replicated; }

// Slabs: which kernel runs decides how many ghost refreshes an operator needs, and the refreshes are a handshake between
// neighbours — so every rank must take the same decision.  The z-marching kernels decline short plane ranges; ask for the
// fewest planes any rank owns instead of this rank's own count.
bool slab_planes_ok(const gs_handle* h, int level, int need) { return !partitioned(h, level) || h->lev[level].min_interior >= need; }

// offsets argument of the single-sweep tuned kernels (kernels_fast.cu): null = canonical axis-aligned stencil, else the
// config's seven offsets (GEN instantiations)
const StencilOff* gen_offsets(const gs_handle* h) { return h->canonical ? nullptr : &h->so; }

// ---- multi-GPU helpers (no-ops on one GPU) ----------------------------------------
// refresh `depth` ghost planes on both sides of a partitioned level array
// registered index (peer transport) of a level array given by pointer, or -1
int array_slot(const gs_handle* h, int level, const double* a)
{
    const Level& L = h->lev[level];
    for (int k = 0; k < GS_LEVEL_ARRAYS; k++)
        if (L.phys[k] == a) return level * GS_LEVEL_ARRAYS + k;
    return -1;
}

// global plane of LOCAL plane 0 of rank r's slab on a partitioned level (create_impl: two ghost planes per side)
int slab_z0(const gs_handle* h, int level, int r)
{
    if (r < 0 || r >= h->world) return 0;
    return std::max(0, h->part.b[level][r] - 2);
}

bool use_p2p(const gs_handle* h) { return h->comm && h->opt_p2p && gs_comm_p2p(h->comm); }

// refresh `depth` ghost planes on both sides of a partitioned level array: one kernel on the peer transport (push +
// arrival, kernels_comm.cu), a grouped ncclSend/ncclRecv otherwise
int exchange_halo(gs_handle* h, int level, double* a, int depth)
{
    if (!partitioned(h, level)) return GS_OK;
    Level& L = h->lev[level];
    const int lo = L.zlo - L.g.z0, hi = L.zhi - L.g.z0;  // local owned range
    int rc;
    const int k = use_p2p(h) ? array_slot(h, level, a) : -1;
    if (k >= 0)
        rc = gs_comm_halo_p2p(h->comm, k, a, L.g.plane, lo, hi, depth, L.zlo, L.zhi, L.g.Zg, L.g.z0, slab_z0(h, level, h->rank - 1),
                              slab_z0(h, level, h->rank + 1), h->stream);
    else rc = gs_comm_halo(h->comm, a, L.g.plane, lo, hi, depth, L.zlo, L.zhi, L.g.Zg, h->stream);
    if (rc != 0) return fail(GS_ECOMM, "halo exchange failed on level %d: %s", level, gs_comm_error(h->comm));
    h->launches += k >= 0 ? 1 : 0;
    return GS_OK;
}

// make `depth` ghost planes of the level's v current (slabs only).  Operators that write v just mark the ghosts
// stale (Level::gv = 0); the exchange happens once, two planes deep, right before the next operator that reads them.
int need_v_halo(gs_handle* h, int level, int depth)
{
    if (!partitioned(h, level)) return GS_OK;
    Level& L = h->lev[level];
    if (L.gv >= depth) return GS_OK;
    const int owned = L.zhi - L.zlo;
    const int d = owned >= 2 ? 2 : depth;
    GS(exchange_halo(h, level, L.v, d));
    L.gv = d;
    return GS_OK;
}

// ---- operators ------------------------------------------------------------------
int ensure_coef(gs_handle* h, int level)
{
    Level& L = h->lev[level];
    if (L.coef_valid) return GS_OK;
    OpScope t(h, GS_OP_OTHER, level);
    CU(ks_newton_coef(h->cx, L.g, 0, L.g.nzl, h->p.gamma, L.newtonV, L.coef));
    L.coef_valid = true;
    return GS_OK;
}

const double* smoother_aux(gs_handle* h, int level, int km)
{
    Level& L = h->lev[level];
    if (km == KM_NEWTON_EXP) return L.newtonV;
    if (km == KM_NEWTON_COEF) return L.coef;
    return nullptr;
}

// CpuSolver::jacobi (src/cpu/CpuSolver.cpp:141-179) as fused ping-pong sweeps
int do_smooth(gs_handle* h, int level, uint64_t iters)
{
    Level& L = h->lev[level];
    const int km = smoother_km(h);
    if (km == KM_NEWTON_COEF && iters > 0) GS(ensure_coef(h, level));
    OpScope t(h, GS_OP_SMOOTH, level);
    if (L.pre_valid > 0 && iters >= (uint64_t)L.pre_valid) {
        // the first sweep(s) were already produced together with the last residual norm (do_norm_residual)
        iters -= L.pre_valid;
        L.pre_valid = 0;
        std::swap(L.v, L.vtmp);
        L.gv = 0;
    }
    L.pre_valid = 0;
    const double* aux = smoother_aux(h, level, km);
    const bool tuned1 = h->opt_kernels != 0 && slab_planes_ok(h, level, 8);
    const bool tuned = tuned1 && h->canonical && slab_planes_ok(h, level, 16);  // two sweeps per pass: canonical stencil only
    if (L.vzero_next) {
        // CpuSolver.cpp:100-101 v.fill(0) followed by the pre-smoothing: the two-sweep kernel takes the zero iterate as a
        // constant (no memset, no read of v); its output buffer becomes v, the stale one is only ever written again
        L.vzero_next = false;
        bool handled = false;
        // (slabs: the zero iterate needs no ghost refresh either; f carries its one ghost plane from finish_coarse)
        if (iters >= 2 && tuned)
            CU(kx_smooth2(h->cx, km, L.ar, L.g, L.ib, L.ie, L.k, L.v, L.vtmp, L.f, aux, false, true, &handled));
        if (handled) {
            iters -= 2;
            h->lvl_sweeps[level] += 2;
            std::swap(L.v, L.vtmp);
            L.gv = 0;
        } else {
            CU(ks_zero(h->cx, L.v, L.count));
            L.gv = 2;
        }
    }
    while (iters > 0) {
        bool handled = false;
        GS(need_v_halo(h, level, (iters >= 2 && tuned && h->cx.tune.fuse2) ? 2 : 1));
        if (iters >= 2 && tuned) {
            // two sweeps per pass (kernels_fused.cu); a slab needs two current ghost planes of v for it
            if (!partitioned(h, level) || L.gv >= 2)
                CU(kx_smooth2(h->cx, km, L.ar, L.g, L.ib, L.ie, L.k, L.v, L.vtmp, L.f, aux, false, false, &handled));
            if (handled) {
                iters -= 2;
                h->lvl_sweeps[level] += 2;
            }
        }
        if (!handled) {
            if (tuned1) CU(kf_smooth(h->cx, km, L.ar, L.g, L.ib, L.ie, L.k, L.v, L.vtmp, L.f, aux, &handled, gen_offsets(h)));
            if (!handled) CU(ks_smooth(h->cx, km, L.pow2, L.g, L.ib, L.ie, L.k, h->so, L.v, L.vtmp, L.f, aux));
            iters--;
            h->lvl_sweeps[level]++;
        }
        std::swap(L.v, L.vtmp);
        L.gv = 0;
    }
    return GS_OK;
}

// sum over ranks of the device scalar, copy to the pinned host slot (async)
int norm_to_host(gs_handle* h, int level)
{
    if (partitioned(h, level)) {
        // peer path: every rank adds the partial sums in rank order (bit-identical everywhere); else ncclAllReduce
        const int rc = use_p2p(h) ? gs_comm_norm_p2p(h->comm, h->cx.norm.result, h->stream)
                                  : gs_comm_allreduce_sum(h->comm, h->cx.norm.result, 1, h->stream);
        if (rc != 0) return fail(GS_ECOMM, "norm reduction failed: %s", gs_comm_error(h->comm));
        if (use_p2p(h)) h->launches += 1;
    }
    CU(cudaMemcpyAsync(h->h_norm, h->cx.norm.result, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    return GS_OK;
}

// CpuSolver::compResidual (src/cpu/CpuSolver.cpp:45-83); the square root is taken by the caller
int do_residual(gs_handle* h, int level, bool store_r, int zb, int ze)
{
    Level& L = h->lev[level];
    int km = smoother_km(h);
    if (km == KM_NEWTON_COEF) GS(ensure_coef(h, level));
    OpScope t(h, GS_OP_RESIDUAL, level);
    GS(need_v_halo(h, level, 1));
    bool handled = false;
    if (h->opt_kernels)
        CU(kf_residual(h->cx, km, L.ar, L.g, zb, ze, L.k, L.v, L.f, smoother_aux(h, level, km), store_r ? L.r : nullptr, &handled,
                       gen_offsets(h)));
    if (!handled)
        CU(ks_residual(h->cx, km, L.pow2, L.g, zb, ze, L.k, h->so, L.v, L.f, smoother_aux(h, level, km), store_r ? L.r : nullptr));
    return GS_OK;
}

// level-0 residual norm when nothing needs r itself (end of a V-cycle, start of a solve).  The next thing that
// happens to level 0 is then a pre-smoothing sweep, which starts from exactly this residual, so the tuned path
// produces that sweep in the same pass (vtmp = S(v), v untouched, Level::pre_valid) instead of reading v and f twice.
int do_norm_residual(gs_handle* h)
{
    Level& L = h->lev[0];
    const int km = smoother_km(h);
    if (h->opt_fuse_norm && h->opt_kernels && h->p.pre > 0 && !L.pre_valid) {
        if (km == KM_NEWTON_COEF) GS(ensure_coef(h, 0));
        OpScope t(h, GS_OP_RESIDUAL, 0);
        bool handled = false;
        const double* aux = smoother_aux(h, 0, km);
        const bool two = h->p.pre >= 2 && h->canonical && slab_planes_ok(h, 0, 16);
        GS(need_v_halo(h, 0, (two && h->cx.tune.fuse2) ? 2 : 1));
        if (two && (!partitioned(h, 0) || L.gv >= 2)) {
            CU(kx_smooth2(h->cx, km, L.ar, L.g, L.ib, L.ie, L.k, L.v, L.vtmp, L.f, aux, true, false, &handled));
            if (handled) {
                L.pre_valid = 2;
                return GS_OK;
            }
        }
        if (slab_planes_ok(h, 0, 8))
            CU(kf_smooth_norm(h->cx, km, L.ar, L.g, L.ib, L.ie, L.k, L.v, L.vtmp, L.f, aux, &handled, gen_offsets(h)));
        if (handled) {
            L.pre_valid = 1;
            return GS_OK;
        }
    }
    return do_residual(h, 0, false, L.ib, L.ie);
}

// restrict a level array onto the next level (src/cpu/CpuSolver.cpp:211-238);
// on the first replicated level the slabs are gathered so every rank holds it whole
// LOCAL coarse planes of level+1 this rank computes when restricting from `level`
void coarse_range(gs_handle* h, int level, int* zb, int* ze)
{
    Level& F = h->lev[level];
    Level& C = h->lev[level + 1];
    *zb = C.ib;
    *ze = C.ie;
    if (partitioned(h, level) && C.replicated) {
        // fine level in slabs, coarse level held whole: the coarse planes whose centre fine plane this rank owns
        *zb = std::max(1, (F.zlo + 1) / 2);
        *ze = std::min(C.g.nz + 1, (F.zhi + 1) / 2);
    }
}

// make a freshly restricted coarse array consistent across ranks: gather the planes onto every rank on the
// first replicated level, else refresh one ghost plane per side
int finish_coarse(gs_handle* h, int level, double* coarse)
{
    Level& C = h->lev[level + 1];
    if (partitioned(h, level) && C.replicated) {
        const int k = use_p2p(h) ? array_slot(h, level + 1, coarse) : -1;
        const int rc = k >= 0 ? gs_comm_gather_planes_p2p(h->comm, k, coarse, C.g.plane, level, h->stream)
                              : gs_comm_gather_planes(h->comm, coarse, C.g.plane, level, h->stream);
        if (rc != 0) return fail(GS_ECOMM, "coarse gather failed: %s", gs_comm_error(h->comm));
        if (k >= 0) h->launches += 1;
        return GS_OK;
    }
    return exchange_halo(h, level + 1, coarse, 1);
}

int do_restrict(gs_handle* h, int level, const double* fine, double* coarse, double* coarse2)
{
    Level& F = h->lev[level];
    Level& C = h->lev[level + 1];
    int zb, ze;
    coarse_range(h, level, &zb, &ze);
    bool handled = false;
    if (h->opt_kernels) CU(kt_restrict(h->cx, F.g, C.g, zb, ze, fine, coarse, coarse2, &handled));
    if (!handled) CU(ks_restrict(h->cx, F.g, C.g, zb, ze, fine, coarse, coarse2));
    GS(finish_coarse(h, level, coarse));
    if (coarse2) GS(finish_coarse(h, level, coarse2));
    return GS_OK;
}

// compResidual(level) ; restrict(r -> f_{level+1})     src/cpu/CpuSolver.cpp:93-98
// with_rv (FAS cycles, survey K5): the fused kernel also restricts v itself into restV and v of the coarse level in the same
// pass; *rv_done tells the caller (do_fas_coarse_rhs) that this part of its work has been done
int do_restrict_residual(gs_handle* h, int level, bool with_rv = false, bool* rv_done = nullptr)
{
    if (rv_done) *rv_done = false;
    Level& L = h->lev[level];
    Level& C = h->lev[level + 1];
    OpScope t(h, GS_OP_RESTRICT, level);
    int zb = L.ib, ze = L.ie;
    int km = smoother_km(h);
    if (km == KM_NEWTON_COEF) GS(ensure_coef(h, level));
    if (partitioned(h, level)) {
        // the lowest owned coarse plane needs r one fine plane below the slab, which needs v two below
        GS(need_v_halo(h, level, 2));
        if (L.zlo > 0) zb = L.ib - 1;
    }
    if (h->opt_kernels && h->canonical) {
        // fused: the residual never goes to memory (kernels_transfer.cu)
        int czb, cze;
        coarse_range(h, level, &czb, &cze);
        bool handled = false;
        // One-GPU handles only: on slabs a rank with fewer than four coarse planes has to take the separate (bit-exact)
        // restriction, and the two orders of summation differ in the last bit — which would cost the property that slab
        // fields equal the one-GPU fields bit for bit (seen at world 8, n = 255).  Slab handles therefore always form R(v)
        // with the separate restriction; compare them with a one-GPU handle that has fuse_rv = 0.
        if (with_rv && rv_done && km == KM_NONLINEAR && h->opt_fuse_rv && h->world == 1) {
            CU(kt_resid_restrict(h->cx, km, L.ar, L.g, C.g, czb, cze, L.k, L.v, L.f, nullptr, C.f, &handled, C.restV, C.v));
            if (handled) {
                C.pre_valid = 0;
                GS(finish_coarse(h, level, C.f));
                GS(finish_coarse(h, level, C.restV));
                GS(finish_coarse(h, level, C.v));
                C.gv = 1;  // finish_coarse refreshed one ghost plane
                *rv_done = true;
                return GS_OK;
            }
        }
        CU(kt_resid_restrict(h->cx, km, L.ar, L.g, C.g, czb, cze, L.k, L.v, L.f, smoother_aux(h, level, km), C.f, &handled));
        if (handled) return finish_coarse(h, level, C.f);
    }
    {
        // unfused: the residual goes to memory (config-driven offsets: GEN kernel; small levels: one thread per point)
        bool handled = false;
        if (h->opt_kernels && !h->canonical)
            CU(kf_residual(h->cx, km, L.ar, L.g, zb, ze, L.k, L.v, L.f, smoother_aux(h, level, km), L.r, &handled, gen_offsets(h)));
        if (!handled) CU(ks_residual(h->cx, km, L.pow2, L.g, zb, ze, L.k, h->so, L.v, L.f, smoother_aux(h, level, km), L.r));
    }
    GS(do_restrict(h, level, L.r, C.f, nullptr));
    return GS_OK;
}

// FAS coarse right-hand side, src/cpu/CpuSolver.cpp:106-112
int do_fas_coarse_rhs(gs_handle* h, int level, bool rv_done = false)
{
    Level& C = h->lev[level + 1];
    OpScope t(h, GS_OP_RESTRICT, level);
    if (!rv_done) {  // else restV = v_{l+1} = R(v_l) came out of the fused residual+restriction pass
        GS(need_v_halo(h, level, 1));
        GS(do_restrict(h, level, h->lev[level].v, C.restV, C.v));
        C.gv = 1;  // finish_coarse refreshed one ghost plane of both outputs
    }
    CU(ks_fas_rhs(h->cx, C.pow2, C.g, C.ib, C.ie, C.k, h->so, C.restV, C.f, C.r));
    GS(exchange_halo(h, level + 1, C.f, 1));  // f changed on the owned planes only
    return GS_OK;
}

// mode 1: v_{l+1} -= restV_{l+1}   src/cpu/CpuSolver.cpp:121-125
int fas_coarse_delta(gs_handle* h, int level)
{
    Level& C = h->lev[level + 1];
    if (h->p.mode != GS_MODE_NONLINEAR) return GS_OK;
    OpScope t(h, GS_OP_PROLONG, level);
    CU(ks_axpy(h->cx, C.g, 0, C.g.nzl, C.v, C.restV, -1));
    C.gv = std::min(C.gv, 1);  // restV carries one current ghost plane
    return GS_OK;
}

// e = P v_{l+1}; v_l += e   src/cpu/CpuSolver.cpp:128-132 (after fas_coarse_delta)
int do_prolong_correct(gs_handle* h, int level)
{
    Level& F = h->lev[level];
    Level& C = h->lev[level + 1];
    OpScope t(h, GS_OP_PROLONG, level);
    F.pre_valid = 0;
    GS(need_v_halo(h, level + 1, 1));  // the interpolation reads one coarse plane outside the slab
    // an odd number of sweeps per cycle would leave v/vtmp swapped; writing the corrected
    // field out of place in that case keeps the buffer parity (and the CUDA graph) fixed
    const bool flip = ((h->p.pre + h->p.post) & 1) != 0;
    bool handled = false;
    if (h->opt_kernels) CU(kt_prolong_correct(h->cx, F.g, C.g, F.ib, F.ie, F.v, flip ? F.vtmp : F.v, C.v, &handled));
    if (!handled) CU(ks_prolong_correct(h->cx, F.g, C.g, F.ib, F.ie, F.v, flip ? F.vtmp : F.v, C.v, nullptr));
    if (flip) std::swap(F.v, F.vtmp);
    F.gv = 0;
    return GS_OK;
}

// Prolongation + correction folded into the first post-smoothing pass (north_star fusion list; src/cpu/CpuSolver.cpp:128-134):
// vtmp = S(S(v + P v_{l+1})), the corrected iterate never goes to memory (kernels_fused.cu smooth2u_tma<.., PC>).  *done =
// false: not applicable here (small level, odd sweep total, slab, option off) and nothing was launched.
int try_prolong_smooth2(gs_handle* h, int level, bool* done)
{
    *done = false;
    Level& F = h->lev[level];
    Level& C = h->lev[level + 1];
    const int km = smoother_km(h);
    const bool flip = ((h->p.pre + h->p.post) & 1) != 0;
    if (!h->opt_fuse_pc || !h->opt_kernels || !h->canonical || !h->cx.tune.fuse2 || h->p.post < 2 || flip ||
        km == KM_NEWTON_EXP)
        return GS_OK;
    // Only the linear mode gains: with exp() in the sweeps (mode 1) or the per-point coefficient (mode 2) the fused pass runs
    // out of registers and loses to the two separate passes at 513^3 — mode 1: 2.31 ms against 1.64 + 0.44; mode 2, V(3,3)
    // cycle 5.61 against 5.46 ms, V(2,2) 4.04 against 3.82 (profiles/r02k_levels_mode1_fuse_pc.txt, r02ad_fuse_pc_mode2.txt).
    // fuse_pc = 2 forces it in every mode for A/B and for the parity test.
    if (km != KM_LINEAR && h->opt_fuse_pc < 2) return GS_OK;
    // ... and only the large levels: at 255^3 and below the fused pass (latency-bound, one wave or two of CTAs) is no faster
    // than the two it replaces (0.206 against 0.143 + 0.051 ms for the level-1 smoother launches of a 513^3 cycle)
    if (F.g.nx < h->cx.tune.pc_min_n && h->opt_fuse_pc < 2) return GS_OK;
    if (F.g.nx < h->cx.tune.fuse2_min_n || F.g.ny < 16 || F.ie - F.ib < 16 || !slab_planes_ok(h, level, 16))
        return GS_OK;  // kx_smooth2 would decline (on this rank or on another)
    if (km == KM_NEWTON_COEF) GS(ensure_coef(h, level));
    OpScope t(h, GS_OP_SMOOTH, level);
    // Slabs: the pass needs v + e on two planes outside the slab.  The uncorrected v ghosts are normally still current (the
    // residual+restriction on the way down refreshed them and v has not changed since); e on the ghost planes comes from
    // the coarse iterate one plane below and two planes above the coarse slab.  So the only exchange of the whole
    // prolongation + first post-smoothing pass is a two-plane refresh of the COARSE level (a quarter of the fine volume).
    GS(need_v_halo(h, level, 2));
    GS(need_v_halo(h, level + 1, 2));
    bool handled = false;
    CU(kx_smooth2(h->cx, km, F.ar, F.g, F.ib, F.ie, F.k, F.v, F.vtmp, F.f, smoother_aux(h, level, km), false, false, &handled, C.v, &C.g));
    if (!handled) return GS_OK;
    F.pre_valid = 0;
    h->lvl_sweeps[level] += 2;
    std::swap(F.v, F.vtmp);
    F.gv = 0;
    *done = true;
    return GS_OK;
}

int zero_field(gs_handle* h, Level& L, double* a)
{
    L.pre_valid = false;
    OpScope t(h, GS_OP_OTHER, (int)(&L - h->lev.data()));
    CU(ks_zero(h->cx, a, L.count));
    if (a == L.v) L.gv = 2;  // ghost planes included
    return GS_OK;
}

// everything CpuSolver::vcycle (src/cpu/CpuSolver.cpp:85-139) does, queued on the stream
int enqueue_vcycle(gs_handle* h, bool host_norm = true)
{
    const int last = (int)h->lev.size() - 1;
    for (int i = 0; i < last; i++) {
        GS(do_smooth(h, i, h->p.pre));
        bool rv_done = false;
        GS(do_restrict_residual(h, i, h->p.mode == GS_MODE_NONLINEAR, &rv_done));
        if (h->p.mode != GS_MODE_NONLINEAR) {
            if (h->opt_vzero) {
                h->lev[i + 1].vzero_next = true;  // consumed by the do_smooth that follows
                h->lev[i + 1].pre_valid = 0;
            } else GS(zero_field(h, h->lev[i + 1], h->lev[i + 1].v));
        } else GS(do_fas_coarse_rhs(h, i, rv_done));
    }
    GS(do_smooth(h, last, h->p.pre + h->p.post));
    if ((h->p.pre + h->p.post) & 1) {  // restore buffer parity on the (tiny) coarsest level
        Level& L = h->lev[last];
        CU(ks_copy(h->cx, L.vtmp, L.v, L.count));
        std::swap(L.v, L.vtmp);
    }
    for (int i = last; i > 0; i--) {
        GS(fas_coarse_delta(h, i - 1));
        bool fused = false;
        GS(try_prolong_smooth2(h, i - 1, &fused));
        if (!fused) GS(do_prolong_correct(h, i - 1));
        GS(do_smooth(h, i - 1, h->p.post - (fused ? 2 : 0)));
    }
    GS(do_norm_residual(h));
    if (host_norm) GS(norm_to_host(h, 0));
    return GS_OK;
}

// Key of a captured V-cycle: only the state that changes BETWEEN cycles of one handle — which of the two buffers is `v` on
// each level, whether level 0 already holds pre-smoothing sweeps from the fused norm pass, and (slabs) which ghost planes
// are current, which decides the exchanges the captured cycle contains.  Everything else a capture bakes in (kernel
// variants, tuning, options) is covered by drop_graphs(): gs_set_option destroys the cache, so an option changed after
// the first gs_vcycle can never replay a stale configuration.
uint64_t parity_key(const gs_handle* h)
{
    uint64_t key = 0xcbf29ce484222325ull;
    auto mix = [&](uint64_t v) { key = (key ^ v) * 0x100000001b3ull; };
    for (size_t l = 0; l < h->lev.size(); l++) {
        const Level& L = h->lev[l];
        mix((L.v > L.vtmp ? 1u : 0u) | (uint64_t)(partitioned(h, (int)l) ? L.gv : 0) << 1 | (uint64_t)(L.vzero_next ? 1 : 0) << 3);
    }
    mix((uint64_t)h->lev[0].pre_valid);
    return key;
}

void drop_graphs(gs_handle* h)
{
    for (auto& kv : h->graphs) cudaGraphExecDestroy(kv.second);
    h->graphs.clear();
    h->graphs_launches.clear();
    h->graphs_pre_valid.clear();
    h->graphs_gv.clear();
    h->graphs_swapped.clear();
    for (auto& kv : h->loops)
        if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    h->loops.clear();
}

int run_vcycle(gs_handle* h, double* norm)
{
    const bool need_coef = smoother_km(h) == KM_NEWTON_COEF;
    if (need_coef)
        for (size_t l = 0; l < h->lev.size(); l++) GS(ensure_coef(h, (int)l));
    const bool graph_ok = h->opt_graph && !h->opt_timing && (h->world == 1 || h->opt_graph_multi);
    if (!graph_ok) {
        GS(enqueue_vcycle(h));
    } else {
        const uint64_t key = parity_key(h);
        auto it = h->graphs.find(key);
        if (it == h->graphs.end()) {
            std::vector<double*> v0, t0;
            for (auto& L : h->lev) {
                v0.push_back(L.v);
                t0.push_back(L.vtmp);
            }
            const uint64_t l0 = h->launches;
            const int pre0 = h->lev[0].pre_valid;
            std::vector<int> gv0, gv1;
            for (auto& L : h->lev) gv0.push_back(L.gv);
            cudaGraph_t g = nullptr;
            CU(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
            int rc = enqueue_vcycle(h);
            cudaError_t ce = cudaStreamEndCapture(h->stream, &g);
            // capture ran the host-side pointer swaps once; undo them, the launch below redoes them
            uint64_t swapped = 0;
            for (size_t l = 0; l < h->lev.size(); l++) {
                if (h->lev[l].v != v0[l]) swapped |= 1ull << l;
                h->lev[l].v = v0[l];
                h->lev[l].vtmp = t0[l];
            }
            uint64_t per_cycle = h->launches - l0;
            h->launches = l0;
            const int pre1 = h->lev[0].pre_valid;
            h->lev[0].pre_valid = pre0;
            for (size_t l = 0; l < h->lev.size(); l++) {
                gv1.push_back(h->lev[l].gv);
                h->lev[l].gv = gv0[l];
            }
            if (rc != GS_OK) {
                if (g) cudaGraphDestroy(g);
                return rc;
            }
            CU(ce);
            cudaGraphExec_t ex = nullptr;
            CU(cudaGraphInstantiate(&ex, g, 0));
            cudaGraphDestroy(g);
            it = h->graphs.emplace(key, ex).first;
            h->graphs_launches[key] = per_cycle;
            h->graphs_pre_valid[key] = pre1;
            h->graphs_swapped[key] = swapped;
            h->graphs_gv[key] = gv1;
        }
        CU(cudaGraphLaunch(it->second, h->stream));
        h->launches += h->graphs_launches[key];
        h->lev[0].pre_valid = h->graphs_pre_valid[key];
        {
            const std::vector<int>& gv1 = h->graphs_gv[key];
            for (size_t l = 0; l < h->lev.size(); l++) h->lev[l].gv = gv1[l];
        }
        // host-side view of the buffer parity after the cycle (the graph is keyed by the parity it starts from)
        const uint64_t swapped = h->graphs_swapped[key];
        for (size_t l = 0; l < h->lev.size(); l++)
            if (swapped >> l & 1) std::swap(h->lev[l].v, h->lev[l].vtmp);
    }
    CU(cudaStreamSynchronize(h->stream));
    if (h->comm && gs_comm_p2p_failed(h->comm)) return fail(GS_ECOMM, "a peer did not arrive at a ghost-plane exchange within the time limit");
    *norm = sqrt(*h->h_norm);
    return GS_OK;
}


long long ms_since(std::chrono::high_resolution_clock::time_point t0)
{
    return std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::high_resolution_clock::now() - t0).count();
}

int residual_norm_sync(gs_handle* h, int level, bool store_r, double* norm)
{
    Level& L = h->lev[level];
    GS(do_residual(h, level, store_r, L.ib, L.ie));
    GS(norm_to_host(h, level));
    CU(cudaStreamSynchronize(h->stream));
    *norm = sqrt(*h->h_norm);
    return GS_OK;
}

int drain_events_impl(gs_handle* h)
{
    CU(cudaStreamSynchronize(h->stream));
    for (auto& ep : h->ev_used) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, ep.a, ep.b) == cudaSuccess) {
            h->op_ms[ep.op] += t;
            h->lvl_ms[(size_t)ep.level * GS_OP_COUNT + ep.op] += t;
        }
        h->ev_free.push_back(ep.a);
        h->ev_free.push_back(ep.b);
    }
    h->ev_used.clear();
    return GS_OK;
}

// Timer::stop's partial timers (src/Timer.cpp:17-24): ", name: Nms (Cx) name: Nms (Cx) " after "Took Nms".  With
// GPUSOLVE_TIMERS=1 the operators of the V-cycle are reported that way (CUDA-event time, plain launches instead of
// the graph); "sumBuffer" — the SYCL backend's norm reduction that runExperiments.py:81-85 sums up — is part of the
// residual kernels here and reported as 0ms.
void print_partial_timers(gs_handle* h)
{
    static const char* names[GS_OP_COUNT] = {"smooth", "residual", "restrict", "prolong", "other"};
    if (drain_events_impl(h) != GS_OK) return;
    std::cout << ", ";
    for (int op = 0; op < GS_OP_COUNT; op++) {
        if (h->op_launches[op] == 0 && h->op_ms[op] == 0.0) continue;
        std::cout << names[op] << ": " << (long long)(h->op_ms[op] + 0.5) << "ms (" << h->op_launches[op] << "x) ";
    }
    std::cout << "sumBuffer: 0ms (1x) ";
    memset(h->op_ms, 0, sizeof h->op_ms);
    memset(h->op_launches, 0, sizeof h->op_launches);
}

// the early-exit test exactly as spelled in src/cpu/CpuSolver.cpp:39 / src/cpu/NewtonSolver.cpp:37
bool reached(double res, double initial, double tol) { return res <= initial / (1.0 / tol); }

// ---- the solve loop on the device ---------------------------------------------------------------
// One graph per (cycle state, maxiter, tol): [initial residual norm (fused with the first pre-smoothing pass) ;
// stop_init] -> WHILE conditional node { captured V-cycle ; stop_check } -> copy of the loop state to pinned memory.
// The body must leave the handle in the state it found it (buffer parity, fused pre-sweeps); it does for every
// configuration the V-cycle graph itself is parity neutral for, otherwise the caller keeps the host loop.
struct HostState {
    std::vector<double*> v, t;
    std::vector<int> gv;
    int pre_valid;
    uint64_t launches;
};

HostState save_state(const gs_handle* h)
{
    HostState st;
    for (auto& L : h->lev) {
        st.v.push_back(L.v);
        st.t.push_back(L.vtmp);
        st.gv.push_back(L.gv);
    }
    st.pre_valid = h->lev[0].pre_valid;
    st.launches = h->launches;
    return st;
}

void restore_state(gs_handle* h, const HostState& st)
{
    for (size_t l = 0; l < h->lev.size(); l++) {
        h->lev[l].v = st.v[l];
        h->lev[l].vtmp = st.t[l];
        h->lev[l].gv = st.gv[l];
        h->lev[l].vzero_next = false;
    }
    h->lev[0].pre_valid = st.pre_valid;
    h->launches = st.launches;
}

bool same_buffers(const HostState& a, const HostState& b) { return a.v == b.v && a.t == b.t && a.pre_valid == b.pre_valid; }

int build_loop_graph(gs_handle* h, uint64_t maxiter, double tol, LoopGraph* out)
{
    *out = LoopGraph();
    if (!h->d_stop) {
        CU(cudaMalloc((void**)&h->d_stop, sizeof(StopCtl)));
        CU(cudaMallocHost((void**)&h->h_stop, sizeof(StopCtl)));
    }
    const HostState s0 = save_state(h);
    cudaGraph_t g = nullptr;
    CU(cudaGraphCreate(&g, 0));
    cudaGraphConditionalHandle cond;
    cudaError_t e = cudaGraphConditionalHandleCreate(&cond, g, 0, 0);
    int rc = GS_OK;
    cudaGraph_t body = nullptr, dummy = nullptr;
    HostState sA = s0, sB = s0;
    bool capturing = false;
    do {
        if (e != cudaSuccess) break;
        // ---- head: initial norm + stop_init, the WHILE node, the history copy ----
        if ((e = cudaStreamBeginCaptureToGraph(h->stream, g, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal)) != cudaSuccess) break;
        capturing = true;
        if ((rc = do_norm_residual(h)) != GS_OK) break;
        if ((e = kc_stop_init(h->cx, h->d_stop, (unsigned long long)cond, (unsigned)maxiter)) != cudaSuccess) break;
        cudaStreamCaptureStatus cs;
        const cudaGraphNode_t* deps = nullptr;
        size_t ndeps = 0;
        if ((e = cudaStreamGetCaptureInfo_v2(h->stream, &cs, nullptr, nullptr, &deps, &ndeps)) != cudaSuccess) break;
        cudaGraphNodeParams cp = {cudaGraphNodeTypeConditional};
        cp.conditional.handle = cond;
        cp.conditional.type = cudaGraphCondTypeWhile;
        cp.conditional.size = 1;
        cudaGraphNode_t cnode;
        if ((e = cudaGraphAddNode(&cnode, g, deps, ndeps, &cp)) != cudaSuccess) break;
        body = cp.conditional.phGraph_out[0];
        if ((e = cudaStreamUpdateCaptureDependencies(h->stream, &cnode, 1, cudaStreamSetCaptureDependencies)) != cudaSuccess) break;
        if ((e = cudaMemcpyAsync(h->h_stop, h->d_stop, sizeof(StopCtl), cudaMemcpyDeviceToHost, h->stream)) != cudaSuccess) break;
        e = cudaStreamEndCapture(h->stream, &dummy);
        capturing = false;
        if (e != cudaSuccess) break;
        sA = save_state(h);
        // ---- body: one V-cycle (no host copy of the norm) + stop_check ----
        if ((e = cudaStreamBeginCaptureToGraph(h->stream, body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal)) != cudaSuccess) break;
        capturing = true;
        if ((rc = enqueue_vcycle(h, false)) != GS_OK) break;
        if ((e = kc_stop_check(h->cx, h->d_stop, (unsigned long long)cond, (unsigned)maxiter, tol)) != cudaSuccess) break;
        e = cudaStreamEndCapture(h->stream, &dummy);
        capturing = false;
        if (e != cudaSuccess) break;
        sB = save_state(h);
    } while (0);
    if (capturing) {
        cudaStreamEndCapture(h->stream, &dummy);
        cudaGetLastError();
    }
    restore_state(h, s0);
    if (rc == GS_OK && e == cudaSuccess && same_buffers(sA, sB)) {
        cudaGraphExec_t ex = nullptr;
        e = cudaGraphInstantiate(&ex, g, 0);
        if (e == cudaSuccess) {
            out->exec = ex;
            out->launches_head = sA.launches - s0.launches;
            out->launches_cycle = sB.launches - sA.launches;
            out->pre_valid_after = sA.pre_valid;
        }
    }
    cudaGraphDestroy(g);
    if (rc != GS_OK) return rc;
    if (e != cudaSuccess) {
        // conditional nodes unavailable (old driver) or a body node they do not admit: not an error, the host loop runs
        cudaGetLastError();
        out->exec = nullptr;
    }
    return GS_OK;
}

// returns GS_OK with *used = false when this solve has to take the host loop
int solve_mg_device(gs_handle* h, uint64_t maxiter, double tol, std::vector<double>* hist, uint64_t* cycles, bool* converged, bool* used)
{
    *used = false;
    if (!h->opt_device_stop || !h->opt_graph || h->opt_timing || h->world != 1 || maxiter > (uint64_t)GS_STOP_CAP) return GS_OK;
    if (smoother_km(h) == KM_NEWTON_COEF)
        for (size_t l = 0; l < h->lev.size(); l++) GS(ensure_coef(h, (int)l));
    uint64_t key = parity_key(h);
    uint64_t tb;
    memcpy(&tb, &tol, 8);
    key = (key ^ maxiter) * 0x100000001b3ull;
    key = (key ^ tb) * 0x100000001b3ull;
    auto it = h->loops.find(key);
    if (it == h->loops.end()) {
        LoopGraph lg;
        GS(build_loop_graph(h, maxiter, tol, &lg));
        it = h->loops.emplace(key, lg).first;
    }
    const LoopGraph& lg = it->second;
    if (!lg.exec) return GS_OK;
    CU(cudaGraphLaunch(lg.exec, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    const StopCtl& c = *h->h_stop;
    h->lev[0].pre_valid = lg.pre_valid_after;
    h->launches += lg.launches_head + (uint64_t)c.count * lg.launches_cycle;
    *cycles = c.count;
    if (hist)
        for (unsigned i = 0; i <= c.count; i++) hist->push_back(c.hist[i]);
    *converged = c.count > 0 && reached(c.hist[c.count], c.initial, tol);
    *h->h_norm = c.count > 0 ? c.hist[c.count] * c.hist[c.count] : c.initial * c.initial;
    *used = true;
    return GS_OK;
}

// CpuSolver::solve, src/cpu/CpuSolver.cpp:12-43
int solve_mg(gs_handle* h, uint64_t maxiter, double tol, bool print, std::vector<double>* hist, uint64_t* cycles,
             bool* converged)
{
    double initial;
    if (!print) {
        bool used = false;
        *cycles = 0;
        *converged = false;
        GS(solve_mg_device(h, maxiter, tol, hist, cycles, converged, &used));
        if (used) return GS_OK;
    }
    GS(do_norm_residual(h));
    GS(norm_to_host(h, 0));
    CU(cudaStreamSynchronize(h->stream));
    initial = sqrt(*h->h_norm);
    if (print) std::cout << "Inital residual: " << initial << '\n';
    if (hist) hist->push_back(initial);
    *cycles = 0;
    *converged = false;
    const bool timers = print && h->opt_timers;
    const int timing0 = h->opt_timing;
    if (timers) {
        h->opt_timing = 1;
        drain_events_impl(h);
        memset(h->op_ms, 0, sizeof h->op_ms);
        memset(h->op_launches, 0, sizeof h->op_launches);
    }
    for (uint64_t i = 0; i < maxiter; i++) {
        auto t0 = std::chrono::high_resolution_clock::now();
        double res;
        int rc = run_vcycle(h, &res);
        if (rc != GS_OK) {
            h->opt_timing = timing0;
            return rc;
        }
        ++*cycles;
        if (hist) hist->push_back(res);
        if (print) {
            std::cout << "iter: " << i << " residual: " << res << ' ';
            std::cout << "Took " << ms_since(t0) << "ms";  // src/Timer.cpp:13-27
            if (timers) print_partial_timers(h);
            std::cout << '\n';
        }
        if (reached(res, initial, tol)) {
            *converged = true;
            break;
        }
    }
    h->opt_timing = timing0;
    return GS_OK;
}

int newton_begin(gs_handle* h)
{
    // NewtonSolver::solve line 12: newtonF = level0.f, "never gets changed"
    Level& L = h->lev[0];
    CU(cudaMemcpyAsync(h->newtonF, L.f, L.count * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    h->newton_begun = true;
    return GS_OK;
}

// NewtonSolver::compF, src/cpu/NewtonSolver.cpp:48-81
int newton_compF(gs_handle* h, double* norm)
{
    Level& L = h->lev[0];
    L.pre_valid = false;  // f changes
    {
        OpScope t(h, GS_OP_RESIDUAL, 0);
        // F(newtonV) = newtonF - (A newtonV / h^2 + gamma newtonV e^newtonV) is the mode-1 residual of the pair (newtonV,
        // newtonF), expression for expression (gs_device.cuh finish_residual<KM_NONLINEAR> vs kernels_simple.cu
        // newton_F_simple): the tuned residual kernel computes it, stores it as f and reduces its norm
        bool handled = false;
        if (h->opt_kernels)
            CU(kf_residual(h->cx, KM_NONLINEAR, L.ar, L.g, L.ib, L.ie, L.k, L.newtonV, h->newtonF, nullptr, L.f, &handled, gen_offsets(h)));
        if (!handled) CU(ks_newton_F(h->cx, L.pow2, L.g, L.ib, L.ie, L.k, h->so, L.newtonV, h->newtonF, L.f));
        GS(exchange_halo(h, 0, L.f, 1));  // the residual+restriction of a slab reads f one plane outside it
    }
    GS(norm_to_host(h, 0));
    CU(cudaStreamSynchronize(h->stream));
    *norm = sqrt(*h->h_norm);
    return GS_OK;
}

// NewtonSolver::findError, src/cpu/NewtonSolver.cpp:83-108
int newton_find_error(gs_handle* h, uint64_t* inner)
{
    const int nlev = (int)h->lev.size();
    {
        OpScope t(h, GS_OP_RESTRICT, 0);
        // :88-92 — newtonV is NOT restricted to the coarsest level (stays zero there)
        for (int i = 1; i + 1 < nlev; i++) GS(do_restrict(h, i - 1, h->lev[i - 1].newtonV, h->lev[i].newtonV, nullptr));
    }
    for (auto& L : h->lev) L.coef_valid = false;
    uint64_t cycles = 0;
    bool conv = false;
    GS(solve_mg(h, 10, 0.1, false, nullptr, &cycles, &conv));  // :94-100
    *inner += cycles;
    Level& L0 = h->lev[0];
    {
        OpScope t(h, GS_OP_OTHER, 0);
        GS(need_v_halo(h, 0, 2));  // the update runs over the ghost planes too: newtonV's ghosts stay current
        CU(ks_axpy(h->cx, L0.g, 0, L0.g.nzl, L0.newtonV, L0.v, +1));  // :106-107
    }
    L0.coef_valid = false;
    L0.pre_valid = false;
    return GS_OK;
}

int newton_step(gs_handle* h, double* norm, uint64_t* inner)
{
    if (!h->newton_begun) GS(newton_begin(h));
    double tmp;
    GS(newton_compF(h, &tmp));                    // :21
    GS(zero_field(h, h->lev[0], h->lev[0].v));    // :22
    GS(newton_find_error(h, inner));              // :24
    GS(newton_compF(h, norm));                    // :26
    return GS_OK;
}

// NewtonSolver::solve, src/cpu/NewtonSolver.cpp:10-44
int solve_newton(gs_handle* h, gs_solve_info* info)
{
    GS(newton_begin(h));
    double initial;
    GS(newton_compF(h, &initial));
    std::cout << "Inital newton residual: " << initial << '\n';  // printed unconditionally, :16
    h->history.push_back(initial);
    info->initial_residual = info->final_residual = initial;
    for (uint64_t i = 0; i < h->p.maxiter; i++) {
        auto t0 = std::chrono::high_resolution_clock::now();
        double res;
        GS(newton_step(h, &res, &info->inner_cycles));
        info->iterations++;
        info->final_residual = res;
        h->history.push_back(res);
        std::cout << "newton iter: " << i << " residual: " << res << ' ';
        std::cout << "Took " << ms_since(t0) << "ms\n";
        if (reached(res, initial, h->p.tol)) {
            info->converged = 1;
            break;
        }
    }
    return GS_OK;
}

int create_impl(const gs_params* p, int device, int rank, int world, const void* nccl_id, gs_handle** out)
{
    if (!p || !out) return fail(GS_EINVAL, "null argument");
    GS(validate(*p));
    if (world < 1 || rank < 0 || rank >= world) return fail(GS_EINVAL, "bad rank/world %d/%d", rank, world);
    int ndev = 0;
    CU(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(GS_EINVAL, "device %d not in [0,%d)", device, ndev);
    CU(cudaSetDevice(device));

    gs_handle* h = new gs_handle();
    h->p = *p;
    h->p.h = 1.0 / (double)(p->dim[1] + 1);  // src/main.cpp:84
    h->device = device;
    h->rank = rank;
    h->world = world;
    h->newtonF = nullptr;
    h->newton_begun = false;
    h->launches = 0;
    h->opt_kernels = 1;
    h->opt_graph = 1;
    h->opt_timing = 0;
    h->opt_fuse_norm = 1;
    h->opt_vzero = 1;
    h->opt_unit = 1;
    h->opt_rcp_div = 1;
    if (const char* e = getenv("GPUSOLVE_RCP_DIV")) h->opt_rcp_div = atoi(e) != 0;
    h->opt_timers = 0;
    if (const char* e = getenv("GPUSOLVE_TIMERS")) h->opt_timers = atoi(e) != 0;
    h->opt_graph_multi = 1;  // capture the NCCL halo exchanges of a slab V-cycle into the graph as well
    h->comm = nullptr;
    h->h_norm = nullptr;
    h->stage = nullptr;
    h->stage_count = 0;
    h->io_stream = nullptr;
    for (int b = 0; b < 2; b++) h->io_copied[b] = h->io_converted[b] = nullptr;
    h->arena = nullptr;
    h->arena_bytes = 0;
    h->opt_p2p = 1;
    if (const char* e = getenv("GPUSOLVE_P2P")) h->opt_p2p = atoi(e) != 0;
    h->opt_fuse_rv = 1;
    if (const char* e = getenv("GPUSOLVE_FUSE_RV")) h->opt_fuse_rv = atoi(e) != 0;
    h->opt_fuse_pc = 1;  // validated on the GPU (profiles/r02a_*): same bits, one pass per level and cycle fewer
    if (const char* e = getenv("GPUSOLVE_FUSE_PC")) h->opt_fuse_pc = std::max(0, std::min(2, atoi(e)));
    h->opt_device_stop = 1;
    if (const char* e = getenv("GPUSOLVE_DEVICE_STOP")) h->opt_device_stop = atoi(e) != 0;
    h->d_stop = nullptr;
    h->h_stop = nullptr;
    memset(h->op_ms, 0, sizeof h->op_ms);
    memset(h->op_launches, 0, sizeof h->op_launches);
    for (int i = 0; i < 7; i++) {
        h->so.dx[i] = p->stencil_off[i][0];
        h->so.dy[i] = p->stencil_off[i][1];
        h->so.dz[i] = p->stencil_off[i][2];
    }
    h->canonical = canonical_offsets(*p);
    if (const char* e = getenv("GPUSOLVE_KERNELS")) h->opt_kernels = atoi(e) != 0;
    if (const char* e = getenv("GPUSOLVE_GRAPH")) h->opt_graph = atoi(e) != 0;
    if (const char* e = getenv("GPUSOLVE_FUSE_NORM")) h->opt_fuse_norm = atoi(e) != 0;
    if (const char* e = getenv("GPUSOLVE_GRAPH_MULTI")) h->opt_graph_multi = atoi(e) != 0;

    Partition part = make_partition(*p, world);
    h->part = part;
    h->replicated_from = part.replicated_from;
    const int nlev = part.nlev;
    h->lev.resize(nlev);
    h->lvl_ms.assign((size_t)nlev * GS_OP_COUNT, 0.0);
    h->lvl_launches.assign((size_t)nlev * GS_OP_COUNT, 0);
    h->lvl_sweeps.assign((size_t)nlev, 0);
    const int ghost = 2;
    size_t max_blocks = 1;
    int rc = GS_OK;
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    for (int l = 0; l < nlev && rc == GS_OK; l++) {
        Level& L = h->lev[l];
        memset(&L, 0, sizeof L);
        uint64_t n[3];
        level_dims(*p, l, n);
        Geom& g = L.g;
        g.nx = (int)n[0];
        g.ny = (int)n[1];
        g.nz = (int)n[2];
        g.X = g.nx + 2;
        g.Y = g.ny + 2;
        g.Zg = g.nz + 2;
        L.replicated = l >= part.replicated_from;
        L.zlo = L.replicated ? 0 : part.b[l][rank];
        L.zhi = L.replicated ? g.Zg : part.b[l][rank + 1];
        if (world == 1 || L.replicated) {
            g.z0 = 0;
            g.nzl = g.Zg;
        } else {
            g.z0 = std::max(0, L.zlo - ghost);
            g.nzl = std::min(g.Zg, L.zhi + ghost) - g.z0;
        }
        g.pitch = ((GS_XOFF + g.X + 15) / 16) * 16;
        g.plane = g.pitch * g.Y;
        L.count = (size_t)g.plane * g.nzl;
        // owned interior planes, local indices
        L.ib = std::max(1, L.zlo) - g.z0;
        L.ie = std::min(g.nz + 1, L.zhi) - g.z0;
        L.min_interior = L.ie - L.ib;
        if (world > 1 && !L.replicated)
            for (int r = 0; r < world; r++)
                L.min_interior = std::min(L.min_interior, std::min(g.nz + 1, part.b[l][r + 1]) - std::max(1, part.b[l][r]));
        L.h = 1.0 / (double)(n[1] + 1);  // src/cpu/CpuGridData.cpp:41 (Y dimension)
        PointK& k = L.k;
        for (int i = 0; i < 7; i++) k.c[i] = p->stencil_val[i];
        k.hh = L.h * L.h;
        L.pow2 = is_pow2(k.hh);
        k.inv_hh = 1.0 / k.hh;
        bool unit = true;
        for (int i = 1; i < 7; i++) unit = unit && p->stencil_val[i] == -1.0;
        L.unit = unit;
        L.ar = level_ar(L, true, h->opt_rcp_div != 0);
        k.alpha = (L.h * L.h) / p->stencil_val[0];   // src/cpu/CpuSolver.cpp:145
        k.preFac = p->stencil_val[0] / (L.h * L.h);  // :144
        k.omega = p->omega;
        k.gamma = p->gamma;
        size_t blocks = (size_t)((g.nx + 63) / 64) * ((g.ny + 3) / 4) * (size_t)g.nzl;
        max_blocks = std::max(max_blocks, blocks);
    }
    // ---- device memory.  One GPU: one allocation per array.  Slabs: ONE arena per rank (control block of the peer
    // transport first, then every level array at a 2 MB-aligned offset), so that a single CUDA IPC handle maps all of a
    // rank's arrays into its peers (gs_comm_enable_p2p).
    {
        const bool arena = world > 1;
        std::vector<long long> offs((size_t)nlev * GS_LEVEL_ARRAYS + 1, -1);
        size_t total = 1 << 16;  // P2PBlock
        auto place = [&](size_t count) {
            const size_t bytes = count * sizeof(double);
            const size_t al = bytes >= ((size_t)2 << 20) ? ((size_t)2 << 20) : 4096;
            total = (total + al - 1) / al * al;
            const size_t at = total;
            total += bytes;
            return at;
        };
        for (int l = 0; l < nlev; l++) {
            const bool want[GS_LEVEL_ARRAYS] = {true, true, true, true, p->mode == GS_MODE_NONLINEAR, p->mode == GS_MODE_NEWTON,
                                                p->mode == GS_MODE_NEWTON, false};
            for (int k = 0; k < GS_LEVEL_ARRAYS; k++)
                if (want[k]) offs[(size_t)l * GS_LEVEL_ARRAYS + k] = (long long)place(h->lev[l].count);
        }
        if (p->mode == GS_MODE_NEWTON) offs[(size_t)nlev * GS_LEVEL_ARRAYS] = (long long)place(h->lev[0].count);
        if (arena) {
            cudaError_t ae = cudaMalloc((void**)&h->arena, total);
            if (ae != cudaSuccess) rc = fail(GS_ENOMEM, "cudaMalloc of the %zu-byte slab arena failed: %s", total, cudaGetErrorString(ae));
            else if ((ae = cudaMemset(h->arena, 0, total)) != cudaSuccess) rc = fail(GS_ECUDA, "arena memset: %s", cudaGetErrorString(ae));
            h->arena_bytes = total;
        }
        for (int l = 0; l < nlev && rc == GS_OK; l++) {
            Level& L = h->lev[l];
            for (int k = 0; k < GS_LEVEL_ARRAYS && rc == GS_OK; k++) {
                const long long o = offs[(size_t)l * GS_LEVEL_ARRAYS + k];
                if (o < 0) continue;
                if (arena) L.phys[k] = (double*)(h->arena + o);
                else rc = alloc_zero(&L.phys[k], L.count);
            }
            L.v = L.phys[0];
            L.vtmp = L.phys[1];
            L.f = L.phys[2];
            L.r = L.phys[3];
            L.restV = L.phys[4];
            L.newtonV = L.phys[5];
            L.coef = L.phys[6];
        }
        if (rc == GS_OK && p->mode == GS_MODE_NEWTON) {
            if (arena) h->newtonF = (double*)(h->arena + offs[(size_t)nlev * GS_LEVEL_ARRAYS]);
            else rc = alloc_zero(&h->newtonF, h->lev[0].count);
        }
        h->arena_offs = offs;
    }
    if (rc != GS_OK) {
        gs_destroy(h);
        return rc;
    }
    cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMallocHost((void**)&h->h_norm, sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc((void**)&h->cx.norm.partials, max_blocks * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc((void**)&h->cx.norm.ticket, sizeof(unsigned));
    if (e == cudaSuccess) e = cudaMalloc((void**)&h->cx.norm.result, sizeof(double));
    if (e == cudaSuccess) e = cudaMemset(h->cx.norm.ticket, 0, sizeof(unsigned));
    if (e == cudaSuccess) e = cudaMemset(h->cx.norm.result, 0, sizeof(double));
    if (e != cudaSuccess) {
        gs_destroy(h);
        return fail(GS_ECUDA, "handle setup failed: %s", cudaGetErrorString(e));
    }
    h->cx.norm.cap = max_blocks;
    h->cx.stream = h->stream;
    h->cx.launches = &h->launches;
    h->cx.sm_count = prop.multiProcessorCount;
    h->cx.tune.smooth_variant = 22;
    h->cx.tune.zchunk = 64;
    h->cx.tune.fast_min_n = 48;
    if (const char* ev = getenv("GPUSOLVE_SMOOTH_VARIANT")) h->cx.tune.smooth_variant = atoi(ev);
    if (const char* ev = getenv("GPUSOLVE_ZCHUNK")) h->cx.tune.zchunk = std::max(1, atoi(ev));
    if (const char* ev = getenv("GPUSOLVE_FAST_MIN_N")) h->cx.tune.fast_min_n = atoi(ev);
    h->cx.tune.rr_variant = 1;  // six-stage ring
    h->cx.tune.fuse2 = 1;
    h->cx.tune.fuse2_min_n = 96;
    h->cx.tune.pc_min_n = 384;
    if (const char* ev = getenv("GPUSOLVE_PC_MIN_N")) h->cx.tune.pc_min_n = atoi(ev);
    h->cx.tune.pdl = 1;
    if (const char* ev = getenv("GPUSOLVE_PDL")) h->cx.tune.pdl = atoi(ev) != 0;
    h->cx.tune.s2_ty = 0;
    if (const char* ev = getenv("GPUSOLVE_S2_TY")) h->cx.tune.s2_ty = atoi(ev);
    h->cx.tune.zchunk2 = 128;
    h->cx.tune.wave = 1;
    if (const char* ev = getenv("GPUSOLVE_WAVE")) h->cx.tune.wave = atoi(ev) != 0;
    if (const char* ev = getenv("GPUSOLVE_FUSE2")) h->cx.tune.fuse2 = std::max(0, std::min(1, atoi(ev)));
    if (const char* ev = getenv("GPUSOLVE_RR_VARIANT")) h->cx.tune.rr_variant = std::max(0, std::min(1, atoi(ev)));
    e = kf_prepare();
    if (e == cudaSuccess) e = kt_prepare();
    if (e == cudaSuccess) e = kx_prepare();
    if (e != cudaSuccess) {
        gs_destroy(h);
        return fail(GS_ECUDA, "kernel setup failed: %s", cudaGetErrorString(e));
    }

    if (world > 1) {
        std::vector<int> bounds;
        for (int l = 0; l < nlev; l++)
            for (int r = 0; r <= world; r++) bounds.push_back(part.b[l][r]);
        h->comm = gs_comm_create(rank, world, nccl_id, nlev, bounds.data(), part.replicated_from);
        if (!h->comm || gs_comm_error(h->comm)[0]) {
            int code = fail(GS_ECOMM, "communicator setup failed: %s", h->comm ? gs_comm_error(h->comm) : "out of memory");
            gs_destroy(h);
            return code;
        }
        // peer-mapped transport: every rank maps every other rank's arena (CUDA IPC).  All ranks decide together; if any
        // of them cannot (no peer access between the devices), all keep NCCL send/recv and say so once.
        if (h->opt_p2p && cudaDeviceSynchronize() == cudaSuccess &&
            gs_comm_enable_p2p(h->comm, h->arena, h->arena_bytes, h->arena_offs.data(), (int)h->arena_offs.size(), h->cx.sm_count,
                               h->stream) != 0) {
            if (rank == 0) fprintf(stderr, "gpusolve: %s; ghost planes travel by NCCL send/recv\n", gs_comm_error(h->comm));
        }
    }
    // the zero fills above ran on the legacy stream, which the handle's non-blocking stream does not wait for
    e = cudaDeviceSynchronize();
    // analytic right-hand side on level 0, src/cpu/CpuGridData.cpp:45-78
    if (e == cudaSuccess) e = ks_init_rhs(h->cx, h->lev[0].g, p->mode, h->p.h, p->gamma, h->lev[0].f);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) {
        gs_destroy(h);
        return fail(GS_ECUDA, "right-hand-side initialisation failed: %s", cudaGetErrorString(e));
    }
    *out = h;
    return GS_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------
extern "C" {

const char* gs_last_error(void) { return g_err; }
int gs_version(void) { return 100; }

int gs_device_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return fail(GS_ECUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
    return n;
}

// src/main.cpp:33-85 — same token order, same operator>> parsing
int gs_parse_config(const char* path, gs_params* out)
{
    if (!path || !out) return fail(GS_EINVAL, "null argument");
    std::ifstream in(path);
    if (!in) return fail(GS_EIO, "\"%s\" does not exist or is not a file", path);
    gs_params p;
    memset(&p, 0, sizeof p);
    in >> p.maxiter >> p.tol >> p.dim[0] >> p.dim[1] >> p.dim[2];
    int mode = -1;
    in >> mode;
    p.mode = mode;
    in >> p.pre >> p.post >> p.omega >> p.gamma;
    for (int i = 0; i < 7; i++) in >> p.stencil_val[i];
    for (int d = 0; d < 3; d++)
        for (int i = 0; i < 7; i++) in >> p.stencil_off[i][d];
    if (!in) return fail(GS_EINVAL, "config file \"%s\" is truncated or malformed (14 lines expected, README.md:19-33)", path);
    p.h = 1.0 / (double)(p.dim[1] + 1);
    p.print_progress = 1;
    GS(validate(p));
    *out = p;
    return GS_OK;
}

int gs_create(const gs_params* p, int device, gs_handle** out) { return create_impl(p, device, 0, 1, nullptr, out); }

int gs_comm_unique_id(void* buf128)
{
    if (!buf128) return fail(GS_EINVAL, "null argument");
    const char* err = gs_comm_make_id(buf128);
    if (err) return fail(GS_ECOMM, "%s", err);
    return GS_OK;
}

int gs_create_slab(const gs_params* p, int device, int rank, int world, const void* nccl_id, gs_handle** out)
{
    if (world > 1 && !nccl_id) return fail(GS_EINVAL, "nccl_id is required when world > 1");
    return create_impl(p, device, rank, world, nccl_id, out);
}

int gs_destroy(gs_handle* h)
{
    if (!h) return GS_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    drop_graphs(h);
    if (h->comm && h->stream) gs_comm_disable_p2p(h->comm, h->stream);  // collective: peers unmap this arena before it is freed
    for (auto& L : h->lev)
        for (double* a : L.phys)
            if (a) {
                gs_tensor_map_drop(a);  // the address may come back from a later cudaMalloc with another geometry
                if (!h->arena) cudaFree(a);
            }
    if (h->newtonF && !h->arena) cudaFree(h->newtonF);
    if (h->arena) cudaFree(h->arena);
    if (h->io_stream) cudaStreamDestroy(h->io_stream);
    for (int b = 0; b < 2; b++) {
        if (h->io_copied[b]) cudaEventDestroy(h->io_copied[b]);
        if (h->io_converted[b]) cudaEventDestroy(h->io_converted[b]);
    }
    if (h->stage) cudaFree(h->stage);
    if (h->cx.norm.partials) cudaFree(h->cx.norm.partials);
    if (h->cx.norm.ticket) cudaFree(h->cx.norm.ticket);
    if (h->cx.norm.result) cudaFree(h->cx.norm.result);
    if (h->h_norm) cudaFreeHost(h->h_norm);
    if (h->d_stop) cudaFree(h->d_stop);
    if (h->h_stop) cudaFreeHost(h->h_stop);
    for (auto& ep : h->ev_used) {
        cudaEventDestroy(ep.a);
        cudaEventDestroy(ep.b);
    }
    for (auto e : h->ev_free) cudaEventDestroy(e);
    if (h->comm) gs_comm_destroy(h->comm);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return GS_OK;
}

int gs_partition(const gs_params* p, int world, int rank, int level, uint64_t* z_begin, uint64_t* z_end, int* replicated_from)
{
    if (!p) return fail(GS_EINVAL, "null argument");
    GS(validate(*p));
    if (world < 1 || rank < 0 || rank >= world) return fail(GS_EINVAL, "bad rank/world %d/%d", rank, world);
    Partition P = make_partition(*p, world);
    if (level < 0 || level >= P.nlev) return fail(GS_EINVAL, "level %d out of range [0,%d)", level, P.nlev);
    if (z_begin) *z_begin = (uint64_t)P.b[level][rank];
    if (z_end) *z_end = (uint64_t)P.b[level][rank + 1];
    if (level >= P.replicated_from) {
        if (z_begin) *z_begin = 0;
        if (z_end) *z_end = (uint64_t)P.b[level][world];
    }
    if (replicated_from) *replicated_from = P.replicated_from;
    return GS_OK;
}

int gs_plan_zchunk(int tiles, int planes, int sm_count, int ctas_per_sm, int overhead, int min_chunk, int max_chunk, int even)
{
    if (tiles < 1 || planes < 1 || sm_count < 1 || min_chunk < 1 || max_chunk < min_chunk) return fail(GS_EINVAL, "bad chunk-plan arguments");
    LaunchCtx cx{};
    cx.sm_count = sm_count;
    return gs_pick_zchunk_waves(cx, tiles, planes, ctas_per_sm, overhead, min_chunk, max_chunk, even != 0);
}

// measurement hook (gpusolve_cuda_tuning.h): `reps` ghost-plane refreshes of the level's v, `depth` planes deep, back to
// back on the handle's stream; *ms_per_exchange = CUDA-event time / reps.  Collective on slabs; GS_EINVAL if the level is
// not partitioned.  The ghost planes receive what they already hold, so the fields do not change.
int gs_time_halo_exchange(gs_handle* h, int level, int depth, int reps, double* ms_per_exchange)
{
    GS(check_level(h, level));
    if (!partitioned(h, level) || depth < 1 || depth > 2 || reps < 1) return fail(GS_EINVAL, "level %d is not partitioned (or bad depth / reps)", level);
    CU(cudaSetDevice(h->device));
    Level& L = h->lev[level];
    cudaEvent_t a, b;
    CU(cudaEventCreate(&a));
    CU(cudaEventCreate(&b));
    GS(exchange_halo(h, level, L.v, depth));  // warm-up
    CU(cudaEventRecord(a, h->stream));
    for (int i = 0; i < reps; i++) GS(exchange_halo(h, level, L.v, depth));
    CU(cudaEventRecord(b, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, a, b));
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    if (h->comm && gs_comm_p2p_failed(h->comm)) return fail(GS_ECOMM, "a peer did not arrive at a ghost-plane exchange within the time limit");
    if (ms_per_exchange) *ms_per_exchange = ms / reps;
    return GS_OK;
}

int gs_num_levels(const gs_handle* h) { return h ? (int)h->lev.size() : fail(GS_EINVAL, "null handle"); }

int gs_get_level_info(const gs_handle* h, int level, uint64_t dim[3], double* hval)
{
    GS(check_level(h, level));
    const Level& L = h->lev[level];
    if (dim) {
        dim[0] = L.g.nx;
        dim[1] = L.g.ny;
        dim[2] = L.g.nz;
    }
    if (hval) *hval = L.h;
    return GS_OK;
}

// dense host planes <-> padded device planes through a flat device staging buffer: the PCIe transfer is one contiguous
// copy per chunk (full link speed) on its own stream, the pitch conversion is a device-to-device 2-D copy in HBM on the
// handle's stream.  The staging buffer has two halves, so the link never waits for a conversion: chunk k+1 crosses PCIe
// while chunk k is being converted.
static int staged_copy(gs_handle* h, const Geom& g, double* dev, double* dense, size_t nplanes, bool to_host)
{
    const size_t XY = (size_t)g.X * g.Y;
    size_t want = std::max<size_t>(XY, (size_t)4 << 20);  // 32 MB of doubles per half, or one plane if that is larger
    want = std::min(want, XY * nplanes);
    if (!h->io_stream) {
        CU(cudaStreamCreateWithFlags(&h->io_stream, cudaStreamNonBlocking));
        for (int b = 0; b < 2; b++) {
            CU(cudaEventCreateWithFlags(&h->io_copied[b], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&h->io_converted[b], cudaEventDisableTiming));
        }
    }
    if (h->stage_count < want) {
        if (h->stage) cudaFree(h->stage);
        h->stage = nullptr;
        h->stage_count = 0;
        if (cudaMalloc((void**)&h->stage, 2 * want * sizeof(double)) != cudaSuccess)
            return fail(GS_ENOMEM, "staging buffer of %zu bytes", 2 * want * sizeof(double));
        h->stage_count = want;
    }
    const size_t ppc = std::max<size_t>(1, h->stage_count / XY);
    int k = 0;
    for (size_t p0 = 0; p0 < nplanes; p0 += ppc, k++) {
        const size_t n = std::min(ppc, nplanes - p0);
        const int b = k & 1;
        double* st = h->stage + (size_t)b * h->stage_count;
        double* d = dev + GS_XOFF + p0 * (size_t)g.plane;
        double* hp = dense + p0 * XY;
        if (to_host) {
            if (k >= 2) CU(cudaStreamWaitEvent(h->stream, h->io_copied[b], 0));  // this half has left for the host
            CU(cudaMemcpy2DAsync(st, g.X * sizeof(double), d, g.pitch * sizeof(double), g.X * sizeof(double), g.Y * n,
                                 cudaMemcpyDeviceToDevice, h->stream));
            CU(cudaEventRecord(h->io_converted[b], h->stream));
            CU(cudaStreamWaitEvent(h->io_stream, h->io_converted[b], 0));
            CU(cudaMemcpyAsync(hp, st, n * XY * sizeof(double), cudaMemcpyDeviceToHost, h->io_stream));
            CU(cudaEventRecord(h->io_copied[b], h->io_stream));
        } else {
            if (k >= 2) CU(cudaStreamWaitEvent(h->io_stream, h->io_converted[b], 0));  // this half has been converted
            CU(cudaMemcpyAsync(st, hp, n * XY * sizeof(double), cudaMemcpyHostToDevice, h->io_stream));
            CU(cudaEventRecord(h->io_copied[b], h->io_stream));
            CU(cudaStreamWaitEvent(h->stream, h->io_copied[b], 0));
            CU(cudaMemcpy2DAsync(d, g.pitch * sizeof(double), st, g.X * sizeof(double), g.X * sizeof(double), g.Y * n,
                                 cudaMemcpyDeviceToDevice, h->stream));
            CU(cudaEventRecord(h->io_converted[b], h->stream));
        }
    }
    // the caller synchronises h->stream; make that cover the copies on the I/O stream too
    if (to_host && k > 0) {
        CU(cudaStreamWaitEvent(h->stream, h->io_copied[(k - 1) & 1], 0));
        if (k > 1) CU(cudaStreamWaitEvent(h->stream, h->io_copied[k & 1], 0));
    }
    return GS_OK;
}

// dense host array <-> padded device slab; z-fastest is transposed on the host
static int field_io(gs_handle* h, int level, int field, double* host, int layout, bool to_host)
{
    GS(check_level(h, level));
    if (!host) return fail(GS_EINVAL, "null host buffer");
    if (layout != GS_LAYOUT_X_FASTEST && layout != GS_LAYOUT_Z_FASTEST) return fail(GS_EINVAL, "bad layout %d", layout);
    if (field == GS_FIELD_E) return fail(GS_EUNSUPPORTED, "e is never materialised by the fused kernels; use gs_prolong");
    double** slot = field_ptr(h, level, field);
    if (!slot || !*slot) return fail(GS_EINVAL, "field %d does not exist on level %d in mode %d", field, level, h->p.mode);
    CU(cudaSetDevice(h->device));
    Level& L = h->lev[level];
    const Geom& g = L.g;
    const size_t X = g.X, Y = g.Y, Z = g.Zg;
    std::vector<double> tmp;
    double* dense = host;
    if (layout == GS_LAYOUT_Z_FASTEST) {
        tmp.resize(X * Y * Z);
        dense = tmp.data();
        if (!to_host)
            for (size_t x = 0; x < X; x++)
                for (size_t y = 0; y < Y; y++)
                    for (size_t z = 0; z < Z; z++) dense[x + X * (y + Y * z)] = host[z + Z * (y + Y * x)];
    }
    CU(cudaStreamSynchronize(h->stream));
    if (to_host) {
        if (partitioned(h, level)) {
            // every rank returns the whole logical array: gather owned planes through a device staging buffer
            int rc = gs_comm_gather_to_host(h->comm, *slot, g.plane, g.pitch, g.z0, level, dense, X, Y, Z, h->stream);
            if (rc != 0) return fail(GS_ECOMM, "field gather failed: %s", gs_comm_error(h->comm));
        } else {
            GS(staged_copy(h, g, *slot, dense, Z, true));
            CU(cudaStreamSynchronize(h->stream));
        }
        if (layout == GS_LAYOUT_Z_FASTEST)
            for (size_t x = 0; x < X; x++)
                for (size_t y = 0; y < Y; y++)
                    for (size_t z = 0; z < Z; z++) host[z + Z * (y + Y * x)] = dense[x + X * (y + Y * z)];
    } else {
        // local planes [z0, z0+nzl) of the dense array (ghost planes included, so they start consistent)
        // everything on the handle's own (non-blocking) stream: a legacy-stream cudaMemcpy from pageable memory
        // returns once the data is staged, and the DMA would race with the next kernel queued on h->stream
        GS(staged_copy(h, g, *slot, dense + X * Y * (size_t)g.z0, (size_t)g.nzl, false));
        if (field == GS_FIELD_V)  // keep the halo of both ping-pong buffers identical
            CU(cudaMemcpyAsync(L.vtmp, L.v, L.count * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        L.pre_valid = 0;
        if (field == GS_FIELD_V) L.gv = 2;  // every rank received its ghost planes from the dense array
        if (field == GS_FIELD_NEWTONV) L.coef_valid = false;
        if (field == GS_FIELD_NEWTONF) h->newton_begun = true;
    }
    return GS_OK;
}

int gs_get_field(gs_handle* h, int level, int field, double* host, int layout)
{
    return field_io(h, level, field, host, layout, true);
}

int gs_set_field(gs_handle* h, int level, int field, const double* host, int layout)
{
    return field_io(h, level, field, const_cast<double*>(host), layout, false);
}

// Vector3::dump, src/cpu/Vector3.cpp:56-78
int gs_dump_field(gs_handle* h, int level, int field, const char* path)
{
    GS(check_level(h, level));
    if (!path) return fail(GS_EINVAL, "null path");
    const Geom& g = h->lev[level].g;
    const size_t X = g.X, Y = g.Y, Z = g.Zg;
    std::vector<double> a(X * Y * Z);
    GS(gs_get_field(h, level, field, a.data(), GS_LAYOUT_X_FASTEST));
    std::ofstream out(path);
    if (!out) return fail(GS_EIO, "cannot open \"%s\" for writing", path);
    out << X << ' ' << Y << ' ' << Z << '\n';
    for (size_t x = 0; x < X; x++)
        for (size_t y = 0; y < Y; y++)
            for (size_t z = 0; z < Z; z++) out << x << ' ' << y << ' ' << z << ' ' << a[x + X * (y + Y * z)] << '\n';
    return out ? GS_OK : fail(GS_EIO, "write to \"%s\" failed", path);
}

int gs_smooth(gs_handle* h, int level, uint64_t iters)
{
    GS(check_level(h, level));
    CU(cudaSetDevice(h->device));
    return do_smooth(h, level, iters);
}

int gs_residual(gs_handle* h, int level, double* norm)
{
    GS(check_level(h, level));
    CU(cudaSetDevice(h->device));
    double n;
    GS(residual_norm_sync(h, level, true, &n));
    if (norm) *norm = n;
    return GS_OK;
}

// gs_residual without any allocation on the call path (include/gpusolve_cuda.h): the one-thread-per-point residual kernel
// with its reduction scratch in a static __device__ array (kernels_simple.cu ks_residual_static), the result read back
// through the pinned word the handle already owns.  No cudaMalloc / cudaFree / allocating library call from here down.
int gs_residual_static(gs_handle* h, int level, double* norm)
{
    GS(check_level(h, level));
    CU(cudaSetDevice(h->device));
    if (partitioned(h, level))
        return fail(GS_EUNSUPPORTED, "gs_residual_static: level %d is split into slabs (the norm would need the peer exchange); use gs_residual", level);
    Level& L = h->lev[level];
    const size_t ctas = ks_residual_ctas(L.g, L.ie - L.ib);
    if (ctas > (size_t)GS_STATIC_SCRATCH_MAX_CTAS)
        return fail(GS_EINVAL,
                    "gs_residual_static: level %d (%d x %d x %d) needs %zu reduction slots, the static scratch holds %d "
                    "(ceil(nx/64) * ceil(ny/4) * nz <= %d, e.g. up to 255^3); use gs_residual",
                    level, L.g.nx, L.g.ny, L.g.nz, ctas, GS_STATIC_SCRATCH_MAX_CTAS, GS_STATIC_SCRATCH_MAX_CTAS);
    const int km = smoother_km(h);
    if (km == KM_NEWTON_COEF) GS(ensure_coef(h, level));  // a kernel into an array allocated by gs_create
    double* res_dev = nullptr;
    CU(ks_residual_static(h->cx, km, L.pow2, L.g, L.ib, L.ie, L.k, h->so, L.v, L.f, smoother_aux(h, level, km), L.r, &res_dev));
    CU(cudaMemcpyAsync(h->h_norm, res_dev, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    if (norm) *norm = sqrt(*h->h_norm);
    return GS_OK;
}

int gs_restrict(gs_handle* h, int level, int src_field, int dst_field)
{
    GS(check_level(h, level));
    GS(check_level(h, level + 1));
    CU(cudaSetDevice(h->device));
    double** s = field_ptr(h, level, src_field);
    double** d = field_ptr(h, level + 1, dst_field);
    if (!s || !*s || !d || !*d) return fail(GS_EINVAL, "field %d/%d not available in mode %d", src_field, dst_field, h->p.mode);
    OpScope t(h, GS_OP_RESTRICT, level);
    if (partitioned(h, level)) {
        if (*s == h->lev[level].v) GS(need_v_halo(h, level, 1));
        else GS(exchange_halo(h, level, *s, 1));
    }
    GS(do_restrict(h, level, *s, *d, nullptr));
    if (*d == h->lev[level + 1].v) h->lev[level + 1].gv = 1;
    if (dst_field == GS_FIELD_NEWTONV) h->lev[level + 1].coef_valid = false;
    return GS_OK;
}

int gs_restrict_residual(gs_handle* h, int level)
{
    GS(check_level(h, level));
    GS(check_level(h, level + 1));
    CU(cudaSetDevice(h->device));
    return do_restrict_residual(h, level);
}

int gs_fas_coarse_rhs(gs_handle* h, int level)
{
    GS(check_level(h, level));
    GS(check_level(h, level + 1));
    if (h->p.mode != GS_MODE_NONLINEAR) return fail(GS_EINVAL, "gs_fas_coarse_rhs needs mode 1 (CpuSolver.cpp:184 asserts the same)");
    CU(cudaSetDevice(h->device));
    return do_fas_coarse_rhs(h, level);
}

int gs_prolong(gs_handle* h, int level, double* host_e, int layout)
{
    GS(check_level(h, level));
    GS(check_level(h, level + 1));
    CU(cudaSetDevice(h->device));
    Level& F = h->lev[level];
    Level& C = h->lev[level + 1];
    {
        OpScope t(h, GS_OP_PROLONG, level);
        CU(cudaMemsetAsync(F.r, 0, F.count * sizeof(double), h->stream));
        CU(ks_prolong_correct(h->cx, F.g, C.g, F.ib, F.ie, nullptr, nullptr, C.v, F.r));
    }
    return gs_get_field(h, level, GS_FIELD_R, host_e, layout);
}

int gs_prolong_correct(gs_handle* h, int level)
{
    GS(check_level(h, level));
    GS(check_level(h, level + 1));
    CU(cudaSetDevice(h->device));
    GS(fas_coarse_delta(h, level));
    return do_prolong_correct(h, level);
}

int gs_vcycle(gs_handle* h, double* norm)
{
    if (!h) return fail(GS_EINVAL, "null handle");
    CU(cudaSetDevice(h->device));
    double n;
    GS(run_vcycle(h, &n));
    if (norm) *norm = n;
    return GS_OK;
}

int gs_newton_compF(gs_handle* h, double* norm)
{
    if (!h) return fail(GS_EINVAL, "null handle");
    if (h->p.mode != GS_MODE_NEWTON) return fail(GS_EINVAL, "gs_newton_compF needs mode 2");
    CU(cudaSetDevice(h->device));
    if (!h->newton_begun) GS(newton_begin(h));
    double n;
    GS(newton_compF(h, &n));
    if (norm) *norm = n;
    return GS_OK;
}

int gs_newton_step(gs_handle* h, double* norm, uint64_t* inner_cycles)
{
    if (!h) return fail(GS_EINVAL, "null handle");
    if (h->p.mode != GS_MODE_NEWTON) return fail(GS_EINVAL, "gs_newton_step needs mode 2");
    CU(cudaSetDevice(h->device));
    double n;
    uint64_t inner = 0;
    GS(newton_step(h, &n, &inner));
    if (norm) *norm = n;
    if (inner_cycles) *inner_cycles = inner;
    return GS_OK;
}

int gs_solve(gs_handle* h, gs_solve_info* info)
{
    if (!h) return fail(GS_EINVAL, "null handle");
    CU(cudaSetDevice(h->device));
    gs_solve_info local;
    memset(&local, 0, sizeof local);
    h->history.clear();
    int rc;
    if (h->p.mode == GS_MODE_NEWTON) {
        rc = solve_newton(h, &local);
    } else {
        bool conv = false;
        rc = solve_mg(h, h->p.maxiter, h->p.tol, h->p.print_progress != 0, &h->history, &local.iterations, &conv);
        local.converged = conv;
        if (!h->history.empty()) {
            local.initial_residual = h->history.front();
            local.final_residual = h->history.back();
        }
    }
    std::cout.flush();
    if (info) *info = local;
    return rc;
}

int gs_get_history(const gs_handle* h, double* norms, int cap, int* count)
{
    if (!h) return fail(GS_EINVAL, "null handle");
    int n = (int)h->history.size();
    if (count) *count = n;
    if (norms)
        for (int i = 0; i < n && i < cap; i++) norms[i] = h->history[i];
    return GS_OK;
}

int gs_reset_timing(gs_handle* h)
{
    if (!h) return fail(GS_EINVAL, "null handle");
    CU(cudaStreamSynchronize(h->stream));
    for (auto& ep : h->ev_used) {
        h->ev_free.push_back(ep.a);
        h->ev_free.push_back(ep.b);
    }
    h->ev_used.clear();
    memset(h->op_ms, 0, sizeof h->op_ms);
    memset(h->op_launches, 0, sizeof h->op_launches);
    std::fill(h->lvl_ms.begin(), h->lvl_ms.end(), 0.0);
    std::fill(h->lvl_launches.begin(), h->lvl_launches.end(), 0);
    std::fill(h->lvl_sweeps.begin(), h->lvl_sweeps.end(), 0);
    return GS_OK;
}

static int drain_events(gs_handle* h) { return drain_events_impl(h); }

int gs_get_level_timing(gs_handle* h, int level, double ms[GS_OP_COUNT], uint64_t launches[GS_OP_COUNT])
{
    GS(check_level(h, level));
    GS(drain_events(h));
    for (int i = 0; i < GS_OP_COUNT; i++) {
        if (ms) ms[i] = h->lvl_ms[(size_t)level * GS_OP_COUNT + i];
        if (launches) launches[i] = h->lvl_launches[(size_t)level * GS_OP_COUNT + i];
    }
    return GS_OK;
}

int gs_get_level_sweeps(gs_handle* h, int level, uint64_t* sweeps)
{
    GS(check_level(h, level));
    if (sweeps) *sweeps = h->lvl_sweeps[level];
    return GS_OK;
}

int gs_get_timing(gs_handle* h, double ms[GS_OP_COUNT], uint64_t launches[GS_OP_COUNT])
{
    if (!h) return fail(GS_EINVAL, "null handle");
    GS(drain_events(h));
    for (int i = 0; i < GS_OP_COUNT; i++) {
        if (ms) ms[i] = h->op_ms[i];
        if (launches) launches[i] = h->op_launches[i];
    }
    return GS_OK;
}

uint64_t gs_launch_count(const gs_handle* h) { return h ? h->launches : 0; }

int gs_set_option(gs_handle* h, const char* name, int64_t value)
{
    if (!h || !name) return fail(GS_EINVAL, "null argument");
    std::string n(name);
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    drop_graphs(h);  // a captured V-cycle bakes in every option: never replay one captured under other settings
    if (n == "kernels") {
        h->opt_kernels = value != 0;
        for (auto& L : h->lev) L.coef_valid = false;
    } else if (n == "graph") h->opt_graph = value != 0;
    else if (n == "fuse_norm") {
        h->opt_fuse_norm = value != 0;
        h->lev[0].pre_valid = false;
    }
    else if (n == "rcp_div") {  // 0: true fp64 divide by h*h on levels where it is not a power of two (same bits, slower)
        h->opt_rcp_div = value != 0;
        for (auto& L : h->lev) L.ar = level_ar(L, h->opt_unit != 0, h->opt_rcp_div != 0);
    } else if (n == "vzero") h->opt_vzero = value != 0;
    else if (n == "timers") h->opt_timers = value != 0;
    else if (n == "timing") h->opt_timing = value != 0;
    else if (n == "smooth_variant") h->cx.tune.smooth_variant = (int)value;
    else if (n == "zchunk") h->cx.tune.zchunk = std::max<int64_t>(1, value);
    else if (n == "fast_min_n") h->cx.tune.fast_min_n = (int)value;
    else if (n == "wave") h->cx.tune.wave = value != 0;
    else if (n == "device_stop") h->opt_device_stop = value != 0;
    else if (n == "p2p") h->opt_p2p = value != 0;
    else if (n == "fuse_rv") h->opt_fuse_rv = value != 0;
    else if (n == "fuse_pc") h->opt_fuse_pc = (int)std::max<int64_t>(0, std::min<int64_t>(2, value));
    else if (n == "unit_taps") {  // 0: keep the six multiplies by -1.0 (A/B check; the bits do not change)
        h->opt_unit = value != 0;
        for (auto& L : h->lev) L.ar = level_ar(L, h->opt_unit != 0, h->opt_rcp_div != 0);
    } else if (n == "rr_variant") h->cx.tune.rr_variant = (int)std::max<int64_t>(0, std::min<int64_t>(1, value));
    else if (n == "fuse2") {
        h->cx.tune.fuse2 = (int)std::max<int64_t>(0, std::min<int64_t>(2, value));
        h->lev[0].pre_valid = 0;
    } else if (n == "fuse2_min_n") h->cx.tune.fuse2_min_n = (int)value;
    else if (n == "pc_min_n") h->cx.tune.pc_min_n = (int)value;
    else if (n == "pdl") h->cx.tune.pdl = value != 0;
    else if (n == "s2_ty") h->cx.tune.s2_ty = (int)value;
    else if (n == "zchunk2") h->cx.tune.zchunk2 = (int)std::max<int64_t>(8, value);
    else return fail(GS_EINVAL, "unknown option \"%s\"", name);
    return GS_OK;
}

void* gs_stream(gs_handle* h) { return h ? (void*)h->stream : nullptr; }

int gs_synchronize(gs_handle* h)
{
    if (!h) return fail(GS_EINVAL, "null handle");
    CU(cudaStreamSynchronize(h->stream));
    return GS_OK;
}

}  // extern "C"
