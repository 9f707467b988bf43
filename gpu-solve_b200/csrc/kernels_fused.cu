// kernels_fused.cu — temporally blocked smoother for sm_100a (canonical 7-point stencil):
// TWO damped-Jacobi sweeps per pass over the level.
//
//   smooth2u_tma  v'' = S(S(v))   (src/cpu/CpuSolver.cpp:141-179 applied twice), plane loop unrolled over the period
//                 of its rings (the kernel the V-cycle uses); smooth2_tma is the same algorithm with a rolled loop.
//                 24 B per lattice point for two sweeps (read v, read f, write v''), i.e.
//                 12 B per lattice update instead of 24.
//
// Jacobi is order-independent (the reference computes r of the whole level before it touches
// v, src/cpu/CpuSolver.cpp:149-177), so the second sweep of a point may run as soon as the
// first sweep of its seven neighbours is known.  A CTA owns a 60 x TY output tile, marches
// along z and keeps a two-deep wavefront:
//   * the v tile with a two-cell rim (64 x (TY+4) box) and the f tile with a one-cell rim
//     (64 x (TY+2) box) of every plane arrive in a shared-memory ring by TMA
//     (cp.async.bulk.tensor.3d, one "full" mbarrier per stage).  There is no producer warp:
//     thread 0 issues the load of plane q+NSTAGE-1 once the split-phase barrier below has told it
//     that every warp is done with plane q-1;
//   * a thread owns two y-adjacent columns of the 62 x (TY+2) first-sweep region with
//     register queues along z for the input and for the first-sweep values of its own columns;
//     first-sweep planes go through a two-slot shared-memory ring so that the in-plane
//     neighbours of the second sweep can be read one plane later.  The ring is guarded by ONE
//     split-phase mbarrier instead of a CTA barrier per plane: a warp arrives after storing its
//     first-sweep plane and waits only after the next plane's first sweep, so slow warps have
//     most of a plane of slack;
//   * first-sweep values outside the interior are forced to zero — what the unfused path reads
//     from the halo of the ping-pong buffer.
// Per-point arithmetic is point_update of kernels_fast.cu (canon_residual + jacobi_update,
// -fmad=false): every intermediate and final value is bit-identical to two single sweeps.
// With NORM the kernel also returns sum r^2 of the INPUT iterate (the first sweep's residual),
// each point counted by the CTA whose output tile holds it.
#include <cuda.h>

#include <type_traits>

#include "gs_device.cuh"

namespace {

struct Fused2P {
    Geom g;
    PointK k;
    int zb, ze, zchunk;  // LOCAL planes [zb, ze) to update, planes per CTA
    const double* vin;
    double* vout;
    const double* f;
    const double* aux;
    NormBuf nb;
    // PC kernels: the iterate the sweeps start from is vin + P(coarse) (prolongation + correction folded into the pass)
    const double* coarse;
    Geom gc;
};

template <int TY, int NSTAGE>
struct S2Cfg {
    static constexpr int BW = 64;    // box width: columns x0-2 .. x0+61
    static constexpr int TXO = 60;   // output columns x0 .. x0+59
    static constexpr int VH = TY + 4, FH = TY + 2;
    static constexpr int V_BYTES = BW * VH * 8, F_BYTES = BW * FH * 8;  // multiples of 512
    static constexpr int STAGE = V_BYTES + F_BYTES;
    static constexpr int S1_BYTES = BW * FH * 8;
    static constexpr int NCONS = BW * FH / 2;  // two first-sweep rows per thread
    static constexpr int THREADS = NCONS;      // no producer warp: thread 0 issues the TMA loads (512 threads -> 64 registers at 2 CTAs/SM)
    static constexpr int SMEM = NSTAGE * STAGE + 2 * S1_BYTES + (NSTAGE + 1) * 8 + 128;
    static_assert(TY % 2 == 0, "two rows per thread");
    static_assert(NCONS % 32 == 0 && NCONS <= 992, "consumer threads");
};

template <int KM, int AR, bool NORM, int TY, int NSTAGE>
__global__ void __launch_bounds__(S2Cfg<TY, NSTAGE>::THREADS, 2) smooth2_tma(const __grid_constant__ CUtensorMap map_v,
                                                                           const __grid_constant__ CUtensorMap map_f, Fused2P p)
{
    using Cfg = S2Cfg<TY, NSTAGE>;
    constexpr int BW = Cfg::BW;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    double* s1ring = (double*)(smem + NSTAGE * Cfg::STAGE);
    uint64_t* full = (uint64_t*)(smem + NSTAGE * Cfg::STAGE + 2 * Cfg::S1_BYTES);
    uint64_t* s1bar = full + NSTAGE;  // phase p completes when every warp has stored its first-sweep plane p+1

    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * Cfg::TXO + 1, y0 = blockIdx.y * TY + 1;  // first output point of the tile
    const int z0 = p.zb + blockIdx.z * p.zchunk;
    const int z1 = min(z0 + p.zchunk, p.ze);
    const int nzo = z1 - z0;      // planes this CTA writes
    // plane index q: local plane z0-2+q.  v planes q = 0 .. nzo+3, f planes q = 1 .. nzo+2,
    // first sweep on q = 1 .. nzo+2, second sweep on q = 2 .. nzo+1

    // plane P of the input -> ring stage P % NSTAGE (thread 0 only).  A stage is free for plane P once every warp has
    // finished the iteration that read plane P - NSTAGE, which the split-phase barrier s1bar below tells.
    auto issue = [&](int P) {
        const int s = P % NSTAGE;
        const bool has_f = P >= 1 && P <= nzo + 2;
        mbar_expect_tx(&full[s], Cfg::V_BYTES + (has_f ? Cfg::F_BYTES : 0));
        unsigned char* st = smem + s * Cfg::STAGE;
        // out-of-range rows / planes / columns (negative or past the array) are zero-filled by TMA
        tma_load_3d(st, &map_v, &full[s], GS_XOFF + x0 - 2, y0 - 2, z0 - 2 + P);
        if (has_f) tma_load_3d(st + Cfg::V_BYTES, &map_f, &full[s], GS_XOFF + x0 - 2, y0 - 1, z0 - 2 + P);
    };
    if (tid == 0) {
        for (int s = 0; s < NSTAGE; s++) mbar_init(&full[s], 1);
        mbar_init(s1bar, Cfg::NCONS / 32);
        mbar_fence_init();
        for (int P = 0; P < NSTAGE && P <= nzo + 3; P++) issue(P);
    }
    __syncthreads();

    // ---- consumers ----
    const int lx = tid & (BW - 1), tr = tid / BW;
    constexpr int LAST_TR = Cfg::FH / 2 - 1;
    const int x = x0 - 2 + lx;
    const int ya = y0 - 1 + 2 * tr, yb = ya + 1;
    const int iv = (2 * tr + 1) * BW + lx;  // (x, ya) inside a v box; (x, yb) = iv + BW
    const int is = (2 * tr) * BW + lx;      // (x, ya) inside an f box / first-sweep slot
    // first sweep defined (else the value is the zero of the Dirichlet halo)
    const bool x1ok = lx >= 1 && lx <= BW - 2 && x >= 1 && x <= p.g.nx;
    const bool a1 = x1ok && ya >= 1 && ya <= p.g.ny;
    const bool b1 = x1ok && yb <= p.g.ny;
    // points of the output tile
    const bool xo = lx >= 2 && lx <= BW - 3 && x <= p.g.nx;
    const bool oa = xo && tr > 0 && ya <= p.g.ny;
    const bool ob = xo && tr < LAST_TR && yb <= p.g.ny;
    const int isS = tr > 0 ? is - BW : is;                  // south of row a (unused when tr == 0)
    const int isN = tr < LAST_TR ? is + 2 * BW : is + BW;   // north of row b (unused when tr == LAST_TR)
    const int eo = lx < BW - 1 ? 1 : 0, wo = lx > 0 ? 1 : 0;  // the box-edge lanes own no first-sweep point: keep their reads inside
    const int lane0 = (tid & 31) == 0;
    const i64 plane = p.g.plane;
    i64 ca = gs_idx(p.g, min(max(x, 1), p.g.nx), min(max(ya, 1), p.g.ny), z0 - 2);  // plane q = 0
    const i64 ab = (yb <= p.g.ny && ya >= 1) ? p.g.pitch : 0;                    // offset of row b from row a
    const int zg0 = p.g.z0 + z0 - 2;  // global plane of q = 0

    auto vbox = [&](int q) { return (const double*)(smem + (q % NSTAGE) * Cfg::STAGE); };
    auto fbox = [&](int q) { return (const double*)(smem + (q % NSTAGE) * Cfg::STAGE + Cfg::V_BYTES); };
    auto s1slot = [&](int q) { return s1ring + (q & 1) * (Cfg::S1_BYTES / 8); };

    double sq = 0.0;
    mbar_wait(&full[0], 0);
    double Da = vbox(0)[iv], Db = vbox(0)[iv + BW];
    mbar_wait(&full[1 % NSTAGE], (1 / NSTAGE) & 1);
    double Ca = vbox(1)[iv], Cb = vbox(1)[iv + BW];
    double D1a = 0.0, D1b = 0.0, C1a = 0.0, C1b = 0.0;  // first-sweep values of the own columns
    double fpa = 0.0, fpb = 0.0, xpa = 0.0, xpb = 0.0;  // f / aux of the previous plane

    for (int q = 1; q <= nzo + 2; q++) {
        ca += plane;  // (x, ya) on plane q
        double xa = 0.0, xb = 0.0;
        const int zg = zg0 + q;
        const bool zok = zg >= 1 && zg <= p.g.nz;
        if (KM >= KM_NEWTON_EXP) {
            if (a1 && zok) xa = p.aux[ca];
            if (b1 && zok) xb = p.aux[ca + ab];
        }
        mbar_wait(&full[(q + 1) % NSTAGE], ((q + 1) / NSTAGE) & 1);
        const double* vb = vbox(q);
        const double* vu = vbox(q + 1);
        const double Ua = vu[iv], Ub = vu[iv + BW];
        const double Ea = vb[iv + eo], Wa = vb[iv - wo], Sa = vb[iv - BW];
        const double Eb = vb[iv + BW + eo], Wb = vb[iv + BW - wo], Nb = vb[iv + 2 * BW];
        const double fa = fbox(q)[is], fb = fbox(q)[is + BW];

        // ---- first sweep on plane q ----
        double dena, denb;
        const double ra = canon_residual<KM, AR>(Ca, Ea, Wa, Cb, Sa, Ua, Da, fa, xa, p.k, dena);
        const double rb = canon_residual<KM, AR>(Cb, Eb, Wb, Nb, Ca, Ub, Db, fb, xb, p.k, denb);
        double U1a = jacobi_update<KM>(Ca, ra, dena, p.k);
        double U1b = jacobi_update<KM>(Cb, rb, denb, p.k);
        U1a = (a1 && zok) ? U1a : 0.0;
        U1b = (b1 && zok) ? U1b : 0.0;
        if (NORM && q >= 2 && q <= nzo + 1) {
            if (oa) sq += ra * ra;
            if (ob) sq += rb * rb;
        }
        // ---- second sweep on plane q-1 (its first-sweep neighbours were stored one iteration ago) ----
        // Split-phase CTA barrier: a warp ARRIVES after storing its first-sweep plane (end of the iteration) and WAITS
        // here, after the next plane's first sweep, so stragglers have most of an iteration of slack.  Passing the
        // wait also means every warp has finished iteration q-1, i.e. its reads of slot (q-2)&1 = q&1, which is
        // the slot written below.
        if (q >= 2) {
            mbar_wait(s1bar, (q - 2) & 1);
            if (tid == 0) {  // every warp is past iteration q-1: the stages of planes <= q-1 are free
                if (q == 2 && NSTAGE <= nzo + 3) issue(NSTAGE);
                if (q - 1 + NSTAGE <= nzo + 3) issue(q - 1 + NSTAGE);
            }
        }
        if (q >= 3) {
            const double* sr = s1slot(q - 1);
            const double E1a = sr[is + eo], W1a = sr[is - wo], S1a = sr[isS];
            const double E1b = sr[is + BW + eo], W1b = sr[is + BW - wo], N1b = sr[isN];
            double d2;
            const double r2a = canon_residual<KM, AR>(C1a, E1a, W1a, C1b, S1a, U1a, D1a, fpa, xpa, p.k, d2);
            const double na = jacobi_update<KM>(C1a, r2a, d2, p.k);
            const double r2b = canon_residual<KM, AR>(C1b, E1b, W1b, N1b, C1a, U1b, D1b, fpb, xpb, p.k, d2);
            const double nb = jacobi_update<KM>(C1b, r2b, d2, p.k);
            if (oa) p.vout[ca - plane] = na;
            if (ob) p.vout[ca - plane + ab] = nb;
        }
        double* sw = s1slot(q);
        sw[is] = U1a;
        sw[is + BW] = U1b;
        __syncwarp();
        if (lane0) mbar_arrive(s1bar);
        Da = Ca; Ca = Ua; Db = Cb; Cb = Ub;
        D1a = C1a; C1a = U1a; D1b = C1b; C1b = U1b;
        fpa = fa; fpb = fb; xpa = xa; xpb = xb;
    }
    if (NORM) {
        __shared__ double sh[32];
        const double tot = block_sum<Cfg::NCONS>(sq, sh, tid);
        grid_sum_finish<Cfg::NCONS>(tot, p.nb, sh, tid);
    }
}

// ---------------------------------------------------------------------------------
// smooth2u_tma: the same algorithm with the plane loop unrolled over a full period of its
// rotating state, so nothing rotates at run time
// ---------------------------------------------------------------------------------
// The z-march of smooth2_tma carries three rotating things: the six-stage TMA ring, the two-slot
// first-sweep ring and the three-deep register queues of the own columns.  Unrolled by six (their
// common period) every ring stage, slot, mbarrier and queue register is a compile-time constant of
// the body copy: shared-memory accesses are [one per-thread base register + immediate], the queues
// need no register moves, and `q % NSTAGE` disappears.  (The rolled kernel issues ~210 instructions
// per plane for 52 fp64 operations and is issue-bound at 83 % of the issue slots: ncu,
// profiles/r01h_*.)  Six stages at two CTAs per SM need the smaller 60 x 12 tile.
template <int TY>
struct S2uCfg {
    static constexpr int NSTAGE = 6;
    static constexpr int BW = 64, TXO = 60;
    static constexpr int VH = TY + 4, FH = TY + 2;
    static constexpr int V_BYTES = BW * VH * 8, F_BYTES = BW * FH * 8;
    static constexpr int STAGE = V_BYTES + F_BYTES;
    static constexpr int S1_BYTES = BW * FH * 8;
    static constexpr int NCONS = BW * FH / 2;
    static constexpr int THREADS = NCONS;
    static constexpr int RING = NSTAGE * STAGE;
    static constexpr int BAR_OFF = RING + 2 * S1_BYTES + 1024;  // 1 KB pad: masked edge lanes read one row past a slot
    static constexpr int SMEM = BAR_OFF + (NSTAGE + 1) * 8 + 128;
    static_assert(TY % 2 == 0 && NCONS % 32 == 0, "two rows per thread, whole warps");
};

// VZ: the input iterate is identically zero (first pre-smoothing pass on a coarse level, src/cpu/CpuSolver.cpp:100-101):
// v is neither zeroed by a memset nor read — the taps are the constant 0.0, same arithmetic, same bits.
// PC: prolongation + correction folded into the first post-smoothing pass (src/cpu/CpuSolver.cpp:128-134: interpolate,
// v += e, jacobi).  The corrected iterate v + P v_c is never written to memory: as plane q+2 of v arrives in the ring,
// every thread adds e to the box entries it owns (its two rows, plus the outermost rim row for the first / last row pair)
// IN PLACE in shared memory — two planes ahead of the first sweep, so the split-phase barrier of step q+1 already orders
// these stores before any neighbour reads them at step q+2 (tests/ring_model.py smooth2_prolong; one plane ahead would
// race).  e is the closed form of the four interpolation passes (kernels_simple.cu prolong_point, nesting x -> y -> z),
// marched along z: the thread keeps ax of its coarse rows j and j+1 on the current coarse plane and loads four coarse
// values per ODD fine plane.  17 + 24 B per point become 25.
template <int KM, int AR, bool NORM, int TY, bool VZ, bool PC = false>
__global__ void __launch_bounds__(S2uCfg<TY>::THREADS, 2) smooth2u_tma(const __grid_constant__ CUtensorMap map_v,
                                                                      const __grid_constant__ CUtensorMap map_f, Fused2P p)
{
    static_assert(!(PC && VZ), "a zero iterate is not corrected");
    using Cfg = S2uCfg<TY>;
    constexpr int BW = Cfg::BW, NS = Cfg::NSTAGE;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    const uint32_t sbase = smem_u32(smem);
    const uint32_t bars = sbase + Cfg::BAR_OFF;  // full[0..NS), then s1bar
    const uint32_t s1bar = bars + NS * 8;

    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * Cfg::TXO + 1, y0 = blockIdx.y * TY + 1;
    const int z0 = p.zb + blockIdx.z * p.zchunk;
    const int z1 = min(z0 + p.zchunk, p.ze);
    const int nzo = z1 - z0;
    const int last = nzo + 2;  // first sweep on planes q = 1 .. last, second sweep on 2 .. last-1 (see smooth2_tma)

    // thread 0: plane P of the input -> stage S of the ring
    auto issue = [&](int P, int S) {
        const bool has_f = P >= 1 && P <= last;
        const uint32_t bar = bars + S * 8, dst = sbase + S * Cfg::STAGE;
        if (VZ && !has_f) {
            // nothing to load; the plain arrival completes the stage's phase so its parity stays in step with the plane count
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
            return;
        }
        // PC: the stage was last written through the generic proxy (the in-place correction): order those stores before
        // the bulk copy that overwrites them
        if (PC) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar),
                     "r"((VZ ? 0 : Cfg::V_BYTES) + (has_f ? Cfg::F_BYTES : 0))
                     : "memory");
        if (!VZ)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
                             dst),
                         "l"(&map_v), "r"(bar), "r"(GS_XOFF + x0 - 2), "r"(y0 - 2), "r"(z0 - 2 + P)
                         : "memory");
        if (has_f)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
                             dst + Cfg::V_BYTES),
                         "l"(&map_f), "r"(bar), "r"(GS_XOFF + x0 - 2), "r"(y0 - 1), "r"(z0 - 2 + P)
                         : "memory");
    };
    if (tid == 0) {
        for (int s = 0; s <= NS; s++) {
            const uint32_t cnt = s < NS ? 1 : Cfg::NCONS / 32;
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bars + s * 8), "r"(cnt));
        }
        mbar_fence_init();
#pragma unroll
        for (int P = 0; P < NS; P++)
            if (P <= nzo + 3) issue(P, P);
    }
    __syncthreads();

    const int lx = tid & (BW - 1), tr = tid / BW;
    constexpr int LAST_TR = Cfg::FH / 2 - 1;
    const int x = x0 - 2 + lx;
    const int ya = y0 - 1 + 2 * tr, yb = ya + 1;
    const bool x1ok = lx >= 1 && lx <= BW - 2 && x >= 1 && x <= p.g.nx;
    const bool a1 = x1ok && ya >= 1 && ya <= p.g.ny;   // first sweep defined (else the zero of the Dirichlet halo)
    const bool b1 = x1ok && yb <= p.g.ny;
    const bool xo = lx >= 2 && lx <= BW - 3 && x <= p.g.nx;
    const bool oa = xo && tr > 0 && ya <= p.g.ny;      // points of the output tile
    const bool ob = xo && tr < LAST_TR && yb <= p.g.ny;
    const bool lane0 = (tid & 31) == 0;
    // per-thread shared-memory base: element (x, ya) of a v box in stage 0; everything else is an immediate away.
    // Edge lanes (lx = 0, 63; tr = 0, LAST_TR) read neighbours that belong to another row, slot or the pad: those
    // values only enter results that are masked (a1/b1) or never stored (oa/ob).
    const uint32_t tb = sbase + ((2 * tr + 1) * BW + lx) * 8;
    constexpr int ROW = BW * 8;
    constexpr int FOFF = Cfg::V_BYTES - ROW;                  // (x, ya) of the f box of the same stage
    constexpr int S1OFF = Cfg::RING - ROW;                    // (x, ya) of first-sweep slot 0
    const i64 plane = p.g.plane;
    const i64 ca0 = gs_idx(p.g, min(max(x, 1), p.g.nx), min(max(ya, 1), p.g.ny), z0 - 2);  // plane q = 0
    const i64 ab = (yb <= p.g.ny && ya >= 1) ? p.g.pitch : 0;
    double* vo = p.vout + ca0;          // advanced one plane per iteration: (x, ya) on plane q - 1 when the second sweep stores
    const double* xp = KM >= KM_NEWTON_EXP ? p.aux + ca0 + plane : nullptr;  // (x, ya) on plane q
    const int qlo = 1 - (p.g.z0 + z0 - 2), qhi = p.g.nz - (p.g.z0 + z0 - 2);  // planes q inside the interior: qlo <= q <= qhi

    // ---- PC: per-thread state of the interpolation (coarse rows j = ya/2 = yb/2 and j+1, coarse column i = x/2) ----
    const bool pxodd = x & 1;
    const int pco = GS_XOFF + min(max(x >> 1, 0), p.gc.X - 2) + (int)p.gc.pitch * min(max(ya >> 1, 0), p.gc.Y - 2);
    const bool pinx = x >= 1 && x <= p.g.nx;
    const bool pmA = pinx && ya >= 1 && ya <= p.g.ny, pmB = pinx && yb >= 1 && yb <= p.g.ny;
    // the rim row nobody owns as a centre: row 0 of the box (y0-2, odd: coarse rows j-1, j) for the first row pair, row
    // TY+3 (y0+TY+1, even: coarse row j+1) for the last
    const bool pmR = pinx && (tr == 0 ? (y0 - 2 >= 1) : (tr == LAST_TR && y0 + TY + 1 <= p.g.ny));
    const int zgq0 = p.g.z0 + z0 - 2;  // global fine plane of q = 0
    PcInterp pci{0.0, 0.0};
    // ax of coarse row j + row on GLOBAL coarse plane k (clamped into the local array; clamped values are never used)
    auto pc_ld = [&](int k, int row) {
        const double* c = p.coarse + (i64)min(max(k - p.gc.z0, 0), p.gc.nzl - 1) * p.gc.plane + pco + row * (int)p.gc.pitch;
        return pxodd ? 0.5 * __ldg(c) + 0.5 * __ldg(c + 1) : __ldg(c);
    };
    // add e to the own entries of plane P, whose (x, ya) entry sits at shared address a
    auto pc_correct = [&](uint32_t a, int P) {
        const int zg = zgq0 + P;
        double eA, eB, eJ, eM;
        pci.step(zg, tr == 0 && pmR, pc_ld, eA, eB, eJ, eM);
        if (zg >= 1 && zg <= p.g.nz) {
            if (pmA) sts_f64(a, lds_f64(a) + eA);
            if (pmB) sts_f64(a + ROW, lds_f64(a + ROW) + eB);
            if (pmR) {
                const uint32_t r = tr == 0 ? a - ROW : a + 2 * ROW;
                sts_f64(r, lds_f64(r) + (tr == 0 ? eM : eJ));
            }
        }
    };

    double vq[3][2], s1q[3][2], fq[2][2], xq[2][2];
#pragma unroll
    for (int i = 0; i < 3; i++) s1q[i][0] = s1q[i][1] = 0.0;
#pragma unroll
    for (int i = 0; i < 2; i++) fq[i][0] = fq[i][1] = xq[i][0] = xq[i][1] = 0.0;
    double sq = 0.0;
#pragma unroll
    for (int i = 0; i < 3; i++) vq[i][0] = vq[i][1] = 0.0;
    if (PC) {
        // planes 0, 1, 2 are corrected up front (a CTA barrier orders them); step q then corrects plane q + 2
        pci.init(zgq0, pc_ld);
#pragma unroll
        for (int P = 0; P < 3; P++)
            if (P <= nzo + 3) {
                mbar_wait_a(bars + P * 8, 0);
                pc_correct(tb + P * Cfg::STAGE, P);
            }
        __syncthreads();
    }
    if (!VZ) {
        mbar_wait_a(bars + 0, 0);
        vq[0][0] = lds_f64(tb);
        vq[0][1] = lds_f64(tb + ROW);
        mbar_wait_a(bars + 8, 0);
        vq[1][0] = lds_f64(tb + Cfg::STAGE);
        vq[1][1] = lds_f64(tb + Cfg::STAGE + ROW);
    }

    // one plane: K = q mod 6 (ring stage, queue slots, parities), PH = 1 / 2 for the start-up planes q = 1 / 2, else 0
    auto step = [&](auto Kc, auto PHc, const int q, const uint32_t par) {
        constexpr int K = decltype(Kc)::value, PH = decltype(PHc)::value;
        constexpr int SC = K % NS, SU = (K + 1) % NS;                  // stages of planes q, q+1
        constexpr int iC = K % 3, iU = (K + 1) % 3, iD = (K + 2) % 3;   // queue slots of planes q, q+1, q-1 (q-2 for the first-sweep queue)
        constexpr int jC = K & 1, jP = (K + 1) & 1;                    // f / first-sweep slot of plane q, of plane q-1
        const bool zok = q >= qlo && q <= qhi;
        double xa = 0.0, xb = 0.0;
        if (KM >= KM_NEWTON_EXP) {
            if (a1 && zok) xa = xp[0];
            if (b1 && zok) xb = xp[ab];
            // the coefficient is the one operand that does not come through the TMA ring: without help every step waits a
            // DRAM round trip for it (52 % of the stalls, profiles/r02a_ncu_modes_summary.txt).  Next plane into L1, the
            // plane three ahead into L2; no register is held meanwhile.
            if (a1 && q + 3 <= qhi) {
                asm volatile("prefetch.global.L1 [%0];" ::"l"(xp + plane));
                asm volatile("prefetch.global.L1 [%0];" ::"l"(xp + plane + ab));
                prefetch_l2(xp + 3 * plane);
                prefetch_l2(xp + 3 * plane + ab);
            }
        }
        const uint32_t c = tb + SC * Cfg::STAGE, u = tb + SU * Cfg::STAGE;
        double Ea = 0.0, Wa = 0.0, Sa = 0.0, Eb = 0.0, Wb = 0.0, Nb = 0.0;
        if (PC) {
            // plane q+2 (two ahead of the first sweep): wait for it and add e to the own entries.  Plane q+1 was waited
            // for one step ago, so the wait below on its stage returns at once.
            if (q + 2 <= nzo + 3) {
                constexpr int S2 = (K + 2) % NS;
                mbar_wait_a(bars + S2 * 8, K >= NS - 2 ? par ^ 1u : par);
                pc_correct(tb + S2 * Cfg::STAGE, q + 2);
            }
        }
        if (VZ) {
            mbar_wait_a(bars + SC * 8, par);  // only the f box of plane q
        } else {
            mbar_wait_a(bars + SU * 8, SU == 0 ? par ^ 1u : par);
            vq[iU][0] = lds_f64(u);
            vq[iU][1] = lds_f64(u + ROW);
            Ea = lds_f64(c + 8), Wa = lds_f64(c - 8), Sa = lds_f64(c - ROW);
            Eb = lds_f64(c + ROW + 8), Wb = lds_f64(c + ROW - 8), Nb = lds_f64(c + 2 * ROW);
        }
        fq[jC][0] = lds_f64(c + FOFF);
        fq[jC][1] = lds_f64(c + FOFF + ROW);
        xq[jC][0] = xa;
        xq[jC][1] = xb;

        // ---- first sweep on plane q ----
        double dena, denb;
        const double ra = canon_residual<KM, AR>(vq[iC][0], Ea, Wa, vq[iC][1], Sa, vq[iU][0], vq[iD][0], fq[jC][0], xa, p.k, dena);
        const double rb = canon_residual<KM, AR>(vq[iC][1], Eb, Wb, Nb, vq[iC][0], vq[iU][1], vq[iD][1], fq[jC][1], xb, p.k, denb);
        const double u1a = jacobi_update<KM>(vq[iC][0], ra, dena, p.k);
        const double u1b = jacobi_update<KM>(vq[iC][1], rb, denb, p.k);
        // (s1q[iC] held plane q-3, last used by the second sweep of the previous step)
        const double D1a = s1q[iU][0], D1b = s1q[iU][1];  // plane q-2
        s1q[iC][0] = (a1 && zok) ? u1a : 0.0;
        s1q[iC][1] = (b1 && zok) ? u1b : 0.0;
        if (NORM && PH != 1) {
            if (q != last) {
                if (oa) sq += ra * ra;
                if (ob) sq += rb * rb;
            }
        }
        if (PH != 1) {
            // split-phase CTA barrier (see smooth2_tma): every warp has finished step q-1
            mbar_wait_a(s1bar, (uint32_t)jC);
            if (tid == 0) {
                if (PH == 2 && NS <= nzo + 3) issue(NS, 0);
                if (q + NS - 1 <= nzo + 3) issue(q + NS - 1, (K + NS - 1) % NS);
            }
        }
        if (PH == 0) {
            // ---- second sweep on plane q-1 ----
            const uint32_t s = tb + S1OFF + jP * Cfg::S1_BYTES;
            const double E1a = lds_f64(s + 8), W1a = lds_f64(s - 8), S1a = lds_f64(s - ROW);
            const double E1b = lds_f64(s + ROW + 8), W1b = lds_f64(s + ROW - 8), N1b = lds_f64(s + 2 * ROW);
            double d2;
            const double r2a = canon_residual<KM, AR>(s1q[iD][0], E1a, W1a, s1q[iD][1], S1a, s1q[iC][0], D1a, fq[jP][0], xq[jP][0], p.k, d2);
            const double na = jacobi_update<KM>(s1q[iD][0], r2a, d2, p.k);
            const double r2b = canon_residual<KM, AR>(s1q[iD][1], E1b, W1b, N1b, s1q[iD][0], s1q[iC][1], D1b, fq[jP][1], xq[jP][1], p.k, d2);
            const double nb = jacobi_update<KM>(s1q[iD][1], r2b, d2, p.k);
            if (oa) vo[0] = na;
            if (ob) vo[ab] = nb;
        }
        sts_f64(tb + S1OFF + jC * Cfg::S1_BYTES, s1q[iC][0]);
        sts_f64(tb + S1OFF + jC * Cfg::S1_BYTES + ROW, s1q[iC][1]);
        __syncwarp();
        if (lane0) mbar_arrive_a(s1bar);
        vo += plane;
        if (KM >= KM_NEWTON_EXP) xp += plane;
    };
#define GS_IC(v) std::integral_constant<int, v>()
    step(GS_IC(1), GS_IC(1), 1, 0u);
    step(GS_IC(2), GS_IC(2), 2, 0u);
    int q = 3;
    uint32_t par = 0;
    bool more = true;
#define GS_STEP(K)                       \
    if (q > last) {                      \
        more = false;                    \
        break;                           \
    }                                    \
    step(GS_IC(K), GS_IC(0), q, par);    \
    q++;
    do {
        GS_STEP(3) GS_STEP(4) GS_STEP(5)
    } while (0);
    while (more) {
        par ^= 1u;
        GS_STEP(0) GS_STEP(1) GS_STEP(2) GS_STEP(3) GS_STEP(4) GS_STEP(5)
    }
#undef GS_STEP
#undef GS_IC
    if (NORM) {
        __shared__ double sh[32];
        const double tot = block_sum<Cfg::NCONS>(sq, sh, tid);
        grid_sum_finish<Cfg::NCONS>(tot, p.nb, sh, tid);
    }
}
template <int KM, int AR, bool NORM, int TY, int NSTAGE>
cudaError_t launch_s2(const LaunchCtx& cx, Fused2P p, bool* handled)
{
    using Cfg = S2Cfg<TY, NSTAGE>;
    static DevOnce once;
    int occ = 1;
    cudaError_t pe = gs_dev_once(once, &occ, [](int* o) {
        cudaError_t e = cudaFuncSetAttribute(smooth2_tma<KM, AR, NORM, TY, NSTAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::SMEM);
        if (e == cudaSuccess)
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(o, smooth2_tma<KM, AR, NORM, TY, NSTAGE>, Cfg::THREADS, Cfg::SMEM);
        return e;
    });
    if (pe != cudaSuccess) return pe;
    if (!p.vin) return cudaSuccess;  // preparation call
    CUtensorMap mv, mf;
    if (!gs_tensor_map(p.vin, p.g, Cfg::BW, Cfg::VH, &mv) || !gs_tensor_map(p.f, p.g, Cfg::BW, Cfg::FH, &mf)) return cudaSuccess;
    const int tx = (p.g.nx + Cfg::TXO - 1) / Cfg::TXO, ty = (p.g.ny + TY - 1) / TY;
    // every chunk re-reads four planes and recomputes two first-sweep planes: keep chunks long
    const int planes = p.ze - p.zb;
    const int want = (cx.sm_count * 4 + tx * ty - 1) / (tx * ty);
    int chunk = (planes + want - 1) / want;
    chunk = max(chunk, min(planes, 32));
    chunk = min(chunk, p.zchunk);
    // four rim planes re-read, two first-sweep planes recomputed, pipeline fill
    if (cx.tune.wave) chunk = gs_pick_zchunk_waves(cx, tx * ty, planes, occ, 6, min(planes, 16), p.zchunk, false);
    p.zchunk = chunk;
    dim3 grid(tx, ty, (planes + chunk - 1) / chunk);
    if (NORM && (size_t)grid.x * grid.y * grid.z > cx.norm.cap) return cudaSuccess;
    smooth2_tma<KM, AR, NORM, TY, NSTAGE><<<grid, Cfg::THREADS, Cfg::SMEM, cx.stream>>>(mv, mf, p);
    *handled = true;
    return cudaGetLastError();
}

template <int KM, int AR, bool NORM, int TY, bool VZ, bool PC = false>
cudaError_t launch_s2u(const LaunchCtx& cx, Fused2P p, bool* handled)
{
    using Cfg = S2uCfg<TY>;
    static DevOnce once;
    int occ = 1;
    cudaError_t pe = gs_dev_once(once, &occ, [](int* o) {
        cudaError_t e = cudaFuncSetAttribute(smooth2u_tma<KM, AR, NORM, TY, VZ, PC>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM);
        if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(o, smooth2u_tma<KM, AR, NORM, TY, VZ, PC>, Cfg::THREADS, Cfg::SMEM);
        return e;
    });
    if (pe != cudaSuccess) return pe;
    if (!p.vin) return cudaSuccess;  // preparation call
    CUtensorMap mv, mf;
    if (!gs_tensor_map(p.f, p.g, Cfg::BW, Cfg::FH, &mf)) return cudaSuccess;
    if (VZ) mv = mf;  // never dereferenced
    else if (!gs_tensor_map(p.vin, p.g, Cfg::BW, Cfg::VH, &mv)) return cudaSuccess;
    const int tx = (p.g.nx + Cfg::TXO - 1) / Cfg::TXO, ty = (p.g.ny + TY - 1) / TY;
    const int planes = p.ze - p.zb;
    // every chunk re-reads four planes, recomputes two first-sweep planes and fills the ring
    if (cx.tune.wave) {
        p.zchunk = gs_pick_zchunk_waves(cx, tx * ty, planes, occ, 6, min(planes, 16), p.zchunk, false);
    } else {
        const int want = (cx.sm_count * 4 + tx * ty - 1) / (tx * ty);
        p.zchunk = min(p.zchunk, max((planes + want - 1) / want, min(planes, 32)));
    }
    dim3 grid(tx, ty, (planes + p.zchunk - 1) / p.zchunk);
    if (NORM && (size_t)grid.x * grid.y * grid.z > cx.norm.cap) return cudaSuccess;
    smooth2u_tma<KM, AR, NORM, TY, VZ, PC><<<grid, Cfg::THREADS, Cfg::SMEM, cx.stream>>>(mv, mf, p);
    *handled = true;
    return cudaGetLastError();
}

template <int KM, int AR, bool NORM>
cudaError_t dispatch_s2(const LaunchCtx& cx, const Fused2P& p, int variant, bool vzero, bool pc, bool* handled)
{
    // Tile height: 12 rows (448 threads, 72 registers at two CTAs per SM); with the per-point Newton coefficient the state
    // no longer fits and spills, so that mode takes 10 rows (384 threads, 80 registers): 3.01 against 3.14 ms for the three
    // level-0 smoother launches of a V(3,3) cycle at 513^3, while modes 0 and 1 lose 6 % and 2 % with the smaller tile
    // (profiles/r02m_tile_height.txt).  tune.s2_ty forces one or the other for A/B (preparation calls set up both).
    constexpr int TYD = KM == KM_NEWTON_COEF ? 10 : 12;
    const int ty = !p.vin ? 0 : (cx.tune.s2_ty == 10 || cx.tune.s2_ty == 12) ? cx.tune.s2_ty : TYD;
    cudaError_t e = cudaSuccess;
#define GS_S2U(NORMV, VZV, PCV)                                                                             \
    do {                                                                                                    \
        if (ty == 0 || ty == 12) e = launch_s2u<KM, AR, NORMV, 12, VZV, PCV>(cx, p, handled);               \
        if (e == cudaSuccess && (ty == 0 || ty == 10)) e = launch_s2u<KM, AR, NORMV, 10, VZV, PCV>(cx, p, handled); \
    } while (0)
    if (pc) {  // prolongation + correction folded in: the unrolled kernel only, no norm, a real input iterate
        if (variant == 1 || vzero) return cudaSuccess;
        if constexpr (KM == KM_NEWTON_EXP || NORM) return cudaSuccess;
        else {
            GS_S2U(false, false, true);
            return e;
        }
    }
    // the unrolled kernel spills heavily with exp() of the Newton iterate in the loop (KM_NEWTON_EXP, only reached
    // with the precomputed-coefficient path switched off): that mode keeps the rolled kernel
    if (variant == 1 || KM == KM_NEWTON_EXP) return vzero ? cudaSuccess : launch_s2<KM, AR, NORM, 14, 5>(cx, p, handled);
    if constexpr (KM != KM_NEWTON_EXP) {
        // the zero-input variant exists where the V-cycle zeroes the coarse iterate: modes 0 and 2, no norm wanted
        if (vzero) {
            if constexpr (!NORM && (KM == KM_LINEAR || KM == KM_NEWTON_COEF)) GS_S2U(false, true, false);
            return e;
        }
        GS_S2U(NORM, false, false);
    }
#undef GS_S2U
    return e;
}

}  // namespace

// vout = S(S(vin)); with_norm: also sum r(vin)^2 -> cx.norm.result.  The level must hold two valid
// planes of vin and one of f (and aux) outside [zb, ze).  vzero: vin is to be taken as identically zero and is not read
// (handled only by the unrolled kernel in modes 0 and 2 without norm).
cudaError_t kx_smooth2(const LaunchCtx& cx, int km, int ar, const Geom& g, int zb, int ze, const PointK& k, const double* vin,
                       double* vout, const double* f, const double* aux, bool with_norm, bool vzero, bool* handled, const double* coarse,
                       const Geom* gc, bool prepare_pc)
{
    *handled = false;
    if (vin) {
        if (ze <= zb || cx.tune.fuse2 == 0) return cudaSuccess;
        if (g.nx < cx.tune.fuse2_min_n || g.ny < 16 || ze - zb < 16) return cudaSuccess;  // small levels: single sweeps
    }
    Fused2P p{g, k, zb, ze, 0, vin, vout, f, aux, cx.norm, coarse, gc ? *gc : Geom{}};
    const bool pc = coarse != nullptr || prepare_pc;
    p.zchunk = cx.tune.zchunk2 > 0 ? cx.tune.zchunk2 : 128;
    const int lo = vin ? cx.tune.fuse2 - 1 : 0, hi = vin ? cx.tune.fuse2 - 1 : 1;
    cudaError_t e = cudaSuccess;
    for (int var = lo; var <= hi && e == cudaSuccess; var++) {
#define CALL(KMV)                                                                    \
    AR_SWITCH(ar, e = with_norm ? (dispatch_s2<KMV, ARV, true>(cx, p, var, vzero, pc, handled)) \
                                : (dispatch_s2<KMV, ARV, false>(cx, p, var, vzero, pc, handled)))
        switch (km) {
        case KM_LINEAR: CALL(KM_LINEAR); break;
        case KM_NONLINEAR: CALL(KM_NONLINEAR); break;
        case KM_NEWTON_EXP: CALL(KM_NEWTON_EXP); break;
        default: CALL(KM_NEWTON_COEF); break;
        }
#undef CALL
    }
    if (e == cudaSuccess && *handled && cx.launches) ++*cx.launches;
    return e;
}

cudaError_t kx_prepare()
{
    LaunchCtx cx{};
    Geom g{};
    PointK k{};
    bool handled;
    cudaError_t e = cudaSuccess;
    for (int km = 0; km < 4 && e == cudaSuccess; km++)
        for (int pw = 0; pw < 8 && e == cudaSuccess; pw++)
            for (int nm = 0; nm < 2 && e == cudaSuccess; nm++)
                for (int vz = 0; vz < 2 && e == cudaSuccess; vz++)
                    for (int pc = 0; pc < 2 && e == cudaSuccess; pc++)
                        e = kx_smooth2(cx, km, pw, g, 0, 0, k, nullptr, nullptr, nullptr, nullptr, nm != 0, vz != 0, &handled, nullptr, nullptr,
                                       pc != 0);
    return e;
}
