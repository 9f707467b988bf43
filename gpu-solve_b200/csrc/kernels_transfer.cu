// kernels_transfer.cu — tuned sm_100a grid-transfer kernels (canonical 7-point stencil):
//
//   prolong_correct_zmarch   v_l += P v_{l+1}           (src/cpu/CpuSolver.cpp:240-290 + :131-132)
//                            17 B per fine point: read v, write v, read 1/8 coarse point.
//   resid_restrict_tma       f_{l+1} = R (f_l - A v_l)  (src/cpu/CpuSolver.cpp:45-83 + :211-238)
//                            17 B per fine point: read v and f, write 1/8 coarse point; the
//                            residual is never stored.
//
// Both keep the reference's operation order per point (see the comments at each sum), so in
// linear mode their output is bit-identical to kernels_simple.cu and to GpuSolve-cpu.
#include "gs_device.cuh"

namespace {

// ---------------------------------------------------------------------------------
// prolongation + correction
// ---------------------------------------------------------------------------------
// One thread owns the 2x2 fine columns x in {2i+1, 2i+2}, y in {2j+1, 2j+2} and marches
// along z.  With C = coarse v, the closed form of the reference's four interpolation passes is
//   ax(j',k') = X odd ? 0.5*C(i,j',k') + 0.5*C(i+1,j',k') : C(X/2,j',k')
//   ay(k')    = Y odd ? 0.5*ax(j,k')  + 0.5*ax(j+1,k')   : ax(Y/2,k')
//   e(X,Y,Z)  = Z odd ? 0.5*ay(k)     + 0.5*ay(k+1)      : ay(Z/2)
// (kernels_simple.cu prolong_point; nesting x -> y -> z as in CpuSolver::interpolate).  The
// thread keeps ay of the current coarse plane in registers: four coarse loads per coarse
// plane serve eight fine points.  x pairs start at an even element, so v moves as double2.
struct ProlongFastP {
    Geom gf, gc;
    int zb, ze, zchunk;
    const double* vin;
    double* vout;
    const double* coarse;
};

struct Ay4 {
    double oo, eo, oe, ee;  // (x odd|even, y odd|even)
};

__device__ __forceinline__ Ay4 load_ay(const double* __restrict__ cv, i64 b, i64 pitch)
{
    const double c00 = __ldg(cv + b), c10 = __ldg(cv + b + 1), c01 = __ldg(cv + b + pitch), c11 = __ldg(cv + b + pitch + 1);
    const double axo_j = 0.5 * c00 + 0.5 * c10, axe_j = c10;
    const double axo_j1 = 0.5 * c01 + 0.5 * c11, axe_j1 = c11;
    Ay4 a;
    a.oo = 0.5 * axo_j + 0.5 * axo_j1;
    a.eo = 0.5 * axe_j + 0.5 * axe_j1;
    a.oe = axo_j1;
    a.ee = axe_j1;
    return a;
}

constexpr int PBX = 32, PBY = 8;
DevOnce g_prolong_once;  // resident CTAs per SM of prolong_correct_zmarch, per device
cudaError_t prolong_occ(int* occ);

__global__ void __launch_bounds__(PBX* PBY) prolong_correct_zmarch(ProlongFastP p)
{
    const int i = blockIdx.x * PBX + threadIdx.x;  // coarse column pair origin
    const int j = blockIdx.y * PBY + threadIdx.y;
    const int xo = 2 * i + 1, yo = 2 * j + 1;
    if (xo > p.gf.nx || yo > p.gf.ny) return;
    const bool has_xe = xo + 1 <= p.gf.nx, has_ye = yo + 1 <= p.gf.ny;
    const int z0 = p.zb + blockIdx.z * p.zchunk;
    const int z1 = min(z0 + p.zchunk, p.ze);
    int zg = z0 + p.gf.z0;                 // global fine plane
    int kc = (zg >> 1) - p.gc.z0;          // local coarse plane of zg/2
    i64 cb = gs_idx(p.gc, i, j, kc);
    Ay4 a = load_ay(p.coarse, cb, p.gc.pitch);
    i64 c = gs_idx(p.gf, xo, yo, z0);
    const i64 pitch = p.gf.pitch, plane = p.gf.plane;
    const double2 zero2 = make_double2(0.0, 0.0);
    auto mid = [](const Ay4& a, const Ay4& n) {
        Ay4 e;
        e.oo = 0.5 * a.oo + 0.5 * n.oo;
        e.eo = 0.5 * a.eo + 0.5 * n.eo;
        e.oe = 0.5 * a.oe + 0.5 * n.oe;
        e.ee = 0.5 * a.ee + 0.5 * n.ee;
        return e;
    };
    auto put = [&](i64 at, double2 v0, double2 v1, const Ay4& e) {
        v0.x = v0.x + e.oo;
        v0.y = v0.y + e.eo;
        if (has_xe) *reinterpret_cast<double2*>(p.vout + at) = v0;
        else p.vout[at] = v0.x;
        if (has_ye) {
            v1.x = v1.x + e.oe;
            v1.y = v1.y + e.ee;
            if (has_xe) *reinterpret_cast<double2*>(p.vout + at + pitch) = v1;
            else p.vout[at + pitch] = v1.x;
        }
    };
    // Two fine planes per iteration, all four v rows loaded before anything is stored: vin may alias vout (in-place
    // correction), so the compiler cannot hoist a load over a store by itself, and one 16-byte load in flight per
    // thread is not enough to cover the HBM latency.  Of two consecutive planes exactly one is odd (needs the next
    // coarse plane).
    for (int z = z0; z < z1; z += 2, zg += 2, c += 2 * plane) {
        const bool two = z + 1 < z1;
        const double2 v00 = *reinterpret_cast<const double2*>(p.vin + c);
        const double2 v01 = has_ye ? *reinterpret_cast<const double2*>(p.vin + c + pitch) : zero2;
        const double2 v10 = two ? *reinterpret_cast<const double2*>(p.vin + c + plane) : zero2;
        const double2 v11 = (two && has_ye) ? *reinterpret_cast<const double2*>(p.vin + c + plane + pitch) : zero2;
        Ay4 e0 = a, e1 = a;
        if (zg & 1) {
            cb += p.gc.plane;
            const Ay4 n = load_ay(p.coarse, cb, p.gc.pitch);
            e0 = mid(a, n);
            a = n;
            e1 = n;
        } else if (two) {
            cb += p.gc.plane;
            const Ay4 n = load_ay(p.coarse, cb, p.gc.pitch);
            e1 = mid(a, n);
            a = n;
        }
        put(c, v00, v01, e0);
        if (two) put(c + plane, v10, v11, e1);
    }
}

cudaError_t prolong_occ(int* occ)
{
    return gs_dev_once(g_prolong_once, occ,
                       [](int* o) { return cudaOccupancyMaxActiveBlocksPerMultiprocessor(o, prolong_correct_zmarch, PBX * PBY, 0); });
}

// ---------------------------------------------------------------------------------
// plain full-weighting restriction (of v in the FAS coarse right-hand side, of newtonV down the hierarchy, of a stored
// residual when the stencil offsets are config-driven)
// ---------------------------------------------------------------------------------
// src/cpu/CpuSolver.cpp:211-238.  One thread per coarse column (X,Y), marching along coarse z: the nine fine values of
// fine plane 2Z+1 are the plane 2(Z+1)-1 of the next coarse point, so every fine plane is loaded once per thread —
// rows as an aligned double2 (x = 2X-1, 2X) plus the scalar x = 2X+1, which the neighbouring thread also loads (L1 hit).
// The 27-term sum runs in the reference's order (x offset outermost, z innermost, weights 0.125*w*w*w evaluated left to
// right), so the result is bit-identical to kernels_simple.cu restrict_point.  Up to two outputs (FAS: restV and v).
struct RestrictFastP {
    Geom gf, gc;
    int zcb, zce, zchunk;  // LOCAL coarse planes [zcb, zce), planes per CTA
    const double* fine;
    double* coarse;
    double* coarse2;
};

constexpr int RBX = 32, RBY = 8;
DevOnce g_restrict_once;

struct Nine {
    double v[3][3];  // [x offset -1,0,+1][y offset -1,0,+1]
};

__device__ __forceinline__ Nine load_nine(const double* __restrict__ f, i64 c, i64 pitch)
{
    // c = element (2X-1, 2Y-1) of the plane: 16-byte aligned (x = 2X-1 is odd, GS_XOFF + x even)
    Nine n;
#pragma unroll
    for (int j = 0; j < 3; j++) {
        const double2 a = __ldg(reinterpret_cast<const double2*>(f + c + j * pitch));
        n.v[0][j] = a.x;
        n.v[1][j] = a.y;
        n.v[2][j] = __ldg(f + c + j * pitch + 2);
    }
    return n;
}

__global__ void __launch_bounds__(RBX* RBY) restrict_zmarch(RestrictFastP p)
{
    const int X = blockIdx.x * RBX + threadIdx.x + 1, Y = blockIdx.y * RBY + threadIdx.y + 1;
    if (X > p.gc.nx || Y > p.gc.ny) return;
    const int zc0 = p.zcb + blockIdx.z * p.zchunk;
    const int zc1 = min(zc0 + p.zchunk, p.zce);
    const int zf = 2 * (zc0 + p.gc.z0) - p.gf.z0;  // local fine plane of the first centre
    i64 c = gs_idx(p.gf, 2 * X - 1, 2 * Y - 1, zf - 1);
    i64 cc = gs_idx(p.gc, X, Y, zc0);
    const i64 pitch = p.gf.pitch, plane = p.gf.plane;
    Nine A = load_nine(p.fine, c, pitch);
    for (int zc = zc0; zc < zc1; zc++) {
        const Nine B = load_nine(p.fine, c + plane, pitch);
        const Nine C = load_nine(p.fine, c + 2 * plane, pitch);
        double acc = 0.0;
        const double w[3] = {0.5, 1.0, 0.5};
#pragma unroll
        for (int ii = 0; ii < 3; ii++)
#pragma unroll
            for (int jj = 0; jj < 3; jj++) {
                acc += (0.125 * w[ii] * w[jj] * 0.5) * A.v[ii][jj];
                acc += (0.125 * w[ii] * w[jj] * 1.0) * B.v[ii][jj];
                acc += (0.125 * w[ii] * w[jj] * 0.5) * C.v[ii][jj];
            }
        p.coarse[cc] = acc;
        if (p.coarse2) p.coarse2[cc] = acc;
        A = C;
        c += 2 * plane;
        cc += p.gc.plane;
    }
}

// ---------------------------------------------------------------------------------
// residual fused with full-weighting restriction
// ---------------------------------------------------------------------------------
// A CTA owns CX x CY coarse columns and marches along coarse z.  Coarse point (X,Y,Z) needs the
// residual at the 27 fine points (2X+ii, 2Y+jj, 2Z+kk); every fine residual needs its 7 v taps.
//   * fine v planes (xy box of the tile plus rim) arrive in a shared-memory ring by TMA, issued by
//     thread 0 after each coarse plane's CTA barrier; f (and the Newton coefficient) are read straight
//     from global memory as aligned double2, one plane ahead into registers and two planes ahead into L2.
//   * compute thread (X,Y) owns the 2x2 fine columns x in {2X-1, 2X}, y in {2Y-1, 2Y} with a
//     register queue along z; "strip" threads own the single columns of the extra fine column
//     x = 2(X0+CX)-1 and row y = 2(Y0+CY)-1 that the last coarse column/row of the tile needs
//     (the only redundant work: neighbouring tiles compute that strip as their own).
//   * residual planes go to a 5-slot shared ring (odd-x and even-x columns de-interleaved, so
//     coarse threads read them conflict-free); after two new planes one barrier, then thread (X,Y)
//     forms the 27-term running sum in the reference's order — x offset outermost, z innermost,
//     weights 0.125*w*w*w (src/cpu/CpuSolver.cpp:219-234) — and writes the coarse point.
// Residuals at fine halo points (even grids: 2X+1 = nx+1) are forced to zero, which is what the
// reference reads there (r is only ever written in the interior of a zero-initialised array).
struct ResidRestrictP {
    Geom gf, gc;
    PointK k;
    int zcb, zce, zchunk;  // LOCAL coarse planes [zcb, zce), planes per CTA
    const double* v;
    const double* f;
    const double* aux;
    double* coarse;
    // RV kernels (FAS, survey K5): the same pass also restricts v itself, R(v) -> rv1 and rv2 (restV and the coarse iterate)
    double* rv1;
    double* rv2;
};

// There is no producer warp: thread 0 issues the TMA loads right after the per-coarse-plane CTA barrier, which already
// tells that every warp is done with the planes before it (no empty barriers).  An earlier version with a producer warp
// and full/empty mbarriers lost parity when launched back to back (a stage was refilled while slow warps still read it;
// tools/debug_ops.py reproduces such races), and was slower.
template <int CX, int CY, int NSTAGE>
struct RRCfg {
    // the first coarse plane reads v planes 0..4 before the first CTA barrier, so the ring must hold five
    static_assert(NSTAGE >= 5, "resid_restrict_tma needs at least five stages");
    static constexpr int FX = 2 * CX, FY = 2 * CY;
    static constexpr int VW = FX + 4, VH = FY + 3;  // v box: x from 2*X0-3 (16-byte aligned start), y from 2*Y0-2
    static constexpr int V_BYTES = VW * VH * 8;
    static constexpr int V_STRIDE = (V_BYTES + 127) / 128 * 128;
    static constexpr int RH = FY + 1;               // residual rows y = 2*Y0-1 .. 2*(Y0+CY)-1
    static constexpr int ROW_O = CX + 1, ROW_E = CX;
    static constexpr int RSLOT = RH * (ROW_O + ROW_E);  // doubles per residual plane
    static constexpr int NSLOT = 5;
    static constexpr int NCOMP = CX * CY;
    static constexpr int NSTRIP_USED = RH + FX;
    static constexpr int NSTRIP = (NSTRIP_USED + 31) / 32 * 32;
    static constexpr int NCONS = NCOMP + NSTRIP;
    static constexpr int THREADS = NCONS;
    static constexpr int SMEM = NSTAGE * V_STRIDE + NSLOT * RSLOT * 8 + NSTAGE * 8 + 128;
};

__device__ __forceinline__ double2 lds2(const double* p) { return *reinterpret_cast<const double2*>(p); }

// RV: full weighting of v itself in the same pass (src/cpu/CpuSolver.cpp:106-107, the two restrict(v) calls of the FAS
// coarse right-hand side): the nine v values a coarse point needs on a fine plane are the thread's own 2 x 2, the east and
// north neighbours it loads for the residual anyway, and one corner.  The plane sums are combined as
// 0.125 (0.5 P(2Z-1) + P(2Z) + 0.5 P(2Z+1)): the same 27 exact products as the reference's running sum, added plane by
// plane instead of z-innermost, i.e. equal to it within rounding (~1e-16), not bit for bit — used only inside the V-cycle
// of the exp() mode, whose tolerance is 1e-12; the operator entry points keep restrict_zmarch.
template <int KM, int AR, int CX, int CY, int NSTAGE, bool RV = false>
__global__ void __launch_bounds__(RRCfg<CX, CY, NSTAGE>::THREADS, 2)
    resid_restrict_tma(const __grid_constant__ CUtensorMap map_v, ResidRestrictP p)
{
    using Cfg = RRCfg<CX, CY, NSTAGE>;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    double* rbuf = (double*)(smem + NSTAGE * Cfg::V_STRIDE);
    uint64_t* full = (uint64_t*)(rbuf + Cfg::NSLOT * Cfg::RSLOT);

    const int tid = threadIdx.x;
    const int X0 = 1 + blockIdx.x * CX, Y0 = 1 + blockIdx.y * CY;
    const int xf0 = 2 * X0, yf0 = 2 * Y0;
    const int zc0 = p.zcb + blockIdx.z * p.zchunk;
    const int zc1 = min(zc0 + p.zchunk, p.zce);
    const int nzc = zc1 - zc0;
    const int zgA = 2 * (zc0 + p.gc.z0) - 1;  // GLOBAL fine plane of residual plane t = 0
    const int zfA = zgA - p.gf.z0;            // the same, local
    const int nvp = 2 * nzc + 3;              // v planes zfA-1 .. zfA+2*nzc+1

    auto issue = [&](int q) {  // v plane q of this CTA -> ring stage q % NSTAGE
        const int s = q % NSTAGE;
        mbar_expect_tx(&full[s], Cfg::V_BYTES);
        tma_load_3d(smem + s * Cfg::V_STRIDE, &map_v, &full[s], GS_XOFF + xf0 - 3, yf0 - 2, zfA - 1 + q);
    };
    int next = 0;  // thread 0: first plane not yet requested
    if (tid == 0) {
        for (int s = 0; s < NSTAGE; s++) mbar_init(&full[s], 1);
        mbar_fence_init();
        for (; next < NSTAGE && next < nvp; next++) issue(next);
    }
    __syncthreads();
    // called right after the CTA barrier that ends coarse plane j — all warps have finished do_plane(2j+2), whose
    // neighbour reads were the last use of v plane 2j+3: the stages of planes <= 2j+3 are free
    auto refill = [&](int j) {
        if (tid == 0)
            for (; next < nvp && next <= 2 * j + 3 + NSTAGE; next++) issue(next);
    };

    auto vbox = [&](int q) { return (const double*)(smem + (q % NSTAGE) * Cfg::V_STRIDE); };
    auto slot = [&](int t) { return rbuf + (t % Cfg::NSLOT) * Cfg::RSLOT; };
    const i64 pitch = p.gf.pitch, plane = p.gf.plane;
    constexpr int VW = Cfg::VW;
    constexpr int EOFF = Cfg::RH * Cfg::ROW_O;  // even-x block of a residual slot

    if (tid < Cfg::NCOMP) {
        // ---- compute thread: fine columns (x1,x2) x (y1,y2) ----
        const int lx = tid % CX, ly = tid / CX;
        const int X = X0 + lx, Y = Y0 + ly;
        const int x1 = 2 * X - 1, y1 = 2 * Y - 1;
        const bool mx1 = x1 <= p.gf.nx, mx2 = x1 + 1 <= p.gf.nx, my1 = y1 <= p.gf.ny, my2 = y1 + 1 <= p.gf.ny;
        const bool ld1 = mx1 && my1, ld2 = mx1 && my2;  // rows whose f/aux pair may be read
        const int o11 = (2 * ly + 1) * VW + 2 * lx + 2;
        const i64 c0 = gs_idx(p.gf, min(x1, p.gf.nx), min(y1, p.gf.ny), zfA);
        const double2 zero2 = make_double2(0.0, 0.0);

        mbar_wait(&full[0], 0);
        double2 D1 = lds2(vbox(0) + o11), D2 = lds2(vbox(0) + o11 + VW);
        mbar_wait(&full[1 % NSTAGE], (1 / NSTAGE) & 1);
        double2 C1 = lds2(vbox(1) + o11), C2 = lds2(vbox(1) + o11 + VW);
        // f / aux of the first residual plane
        double2 f1 = ld1 ? __ldg(reinterpret_cast<const double2*>(p.f + c0)) : zero2;
        double2 f2 = ld2 ? __ldg(reinterpret_cast<const double2*>(p.f + c0 + pitch)) : zero2;
        double2 a1 = zero2, a2 = zero2;
        if (KM >= KM_NEWTON_EXP) {
            a1 = ld1 ? __ldg(reinterpret_cast<const double2*>(p.aux + c0)) : zero2;
            a2 = ld2 ? __ldg(reinterpret_cast<const double2*>(p.aux + c0 + pitch)) : zero2;
        }
        double rB11 = 0, rB21 = 0, rB12 = 0, rB22 = 0, r11 = 0, r21 = 0, r12 = 0, r22 = 0;
        double pv = 0, pvA = 0, pvB = 0;  // RV: weighted nine-point sums of v on the last plane, on planes 2j and 2j+1

        // one residual plane t (v centre plane q = t+1): own four points, stored to slot t
        auto do_plane = [&](int t) {
            const int q = t + 1;
            // prefetch f / aux of the next plane (there is one while t < 2*nzc): requested before the wait for the v
            // plane, so the wait is part of its lead time
            double2 nf1 = zero2, nf2 = zero2, na1 = zero2, na2 = zero2;
            if (t + 1 < 2 * nzc) {  // and the plane after it into L2: one plane of lead does not cover the HBM latency
                const i64 cp = c0 + (i64)(t + 2) * plane;
                if (ld1) prefetch_l2(p.f + cp);
                if (ld2) prefetch_l2(p.f + cp + pitch);
                if (KM >= KM_NEWTON_EXP) {
                    if (ld1) prefetch_l2(p.aux + cp);
                    if (ld2) prefetch_l2(p.aux + cp + pitch);
                }
            }
            if (t < 2 * nzc) {
                const i64 cn = c0 + (i64)(t + 1) * plane;
                if (ld1) nf1 = __ldg(reinterpret_cast<const double2*>(p.f + cn));
                if (ld2) nf2 = __ldg(reinterpret_cast<const double2*>(p.f + cn + pitch));
                if (KM >= KM_NEWTON_EXP) {
                    if (ld1) na1 = __ldg(reinterpret_cast<const double2*>(p.aux + cn));
                    if (ld2) na2 = __ldg(reinterpret_cast<const double2*>(p.aux + cn + pitch));
                }
            }
            mbar_wait(&full[(q + 1) % NSTAGE], ((q + 1) / NSTAGE) & 1);
            const double* vb = vbox(q);
            const double* vu = vbox(q + 1);
            const double2 U1 = lds2(vu + o11), U2 = lds2(vu + o11 + VW);
            const double2 S = lds2(vb + o11 - VW), N = lds2(vb + o11 + 2 * VW);
            const double W1 = vb[o11 - 1], W2 = vb[o11 + VW - 1], E1 = vb[o11 + 2], E2 = vb[o11 + VW + 2];
            if (RV) {
                const double NE = vb[o11 + 2 * VW + 2];  // (2X+1, 2Y+1)
                double s = 0.25 * C1.x;
                s += 0.5 * C2.x;
                s += 0.25 * N.x;
                s += 0.5 * C1.y;
                s += C2.y;
                s += 0.5 * N.y;
                s += 0.25 * E1;
                s += 0.5 * E2;
                s += 0.25 * NE;
                pv = s;
            }
            const bool mz = zgA + t <= p.gf.nz;
            double den;
            r11 = canon_residual<KM, AR>(C1.x, C1.y, W1, C2.x, S.x, U1.x, D1.x, f1.x, a1.x, p.k, den);
            r21 = canon_residual<KM, AR>(C1.y, E1, C1.x, C2.y, S.y, U1.y, D1.y, f1.y, a1.y, p.k, den);
            r12 = canon_residual<KM, AR>(C2.x, C2.y, W2, N.x, C1.x, U2.x, D2.x, f2.x, a2.x, p.k, den);
            r22 = canon_residual<KM, AR>(C2.y, E2, C2.x, N.y, C1.y, U2.y, D2.y, f2.y, a2.y, p.k, den);
            r11 = (mz && mx1 && my1) ? r11 : 0.0;
            r21 = (mz && mx2 && my1) ? r21 : 0.0;
            r12 = (mz && mx1 && my2) ? r12 : 0.0;
            r22 = (mz && mx2 && my2) ? r22 : 0.0;
            double* rs = slot(t);
            rs[(2 * ly) * Cfg::ROW_O + lx] = r11;
            rs[(2 * ly + 1) * Cfg::ROW_O + lx] = r12;
            rs[EOFF + (2 * ly) * Cfg::ROW_E + lx] = r21;
            rs[EOFF + (2 * ly + 1) * Cfg::ROW_E + lx] = r22;
            D1 = C1; D2 = C2; C1 = U1; C2 = U2;
            f1 = nf1; f2 = nf2; a1 = na1; a2 = na2;
        };

        // fac = 0.125 * w(ii) * w(jj) * w(kk), w(0) = 1, w(+-1) = 0.5, evaluated left to right like the reference
#define TERM(WI, WJ, A, B, C)                  \
    acc += (0.125 * WI * WJ * 0.5) * (A);      \
    acc += (0.125 * WI * WJ * 1.0) * (B);      \
    acc += (0.125 * WI * WJ * 0.5) * (C);
        do_plane(0);
        pvA = pv;
        const bool live = X <= p.gc.nx && Y <= p.gc.ny;
        i64 cc = gs_idx(p.gc, min(X, p.gc.nx), min(Y, p.gc.ny), zc0);
        for (int j = 0; j < nzc; j++) {
            do_plane(2 * j + 1);
            rB11 = r11; rB21 = r21; rB12 = r12; rB22 = r22;
            pvB = pv;
            do_plane(2 * j + 2);
            consumer_sync<Cfg::NCONS>();
            refill(j);
            const double* sA = slot(2 * j);
            const double* sB = slot(2 * j + 1);
            const double* sC = slot(2 * j + 2);
            const int oO = (2 * ly) * Cfg::ROW_O + lx, oE = EOFF + (2 * ly) * Cfg::ROW_E + lx;
            double acc = 0.0;
            // ii = -1: x1 = 2X-1 (odd block)
            TERM(0.5, 0.5, sA[oO], rB11, r11)
            TERM(0.5, 1.0, sA[oO + Cfg::ROW_O], rB12, r12)
            TERM(0.5, 0.5, sA[oO + 2 * Cfg::ROW_O], sB[oO + 2 * Cfg::ROW_O], sC[oO + 2 * Cfg::ROW_O])
            // ii = 0: x2 = 2X (even block)
            TERM(1.0, 0.5, sA[oE], rB21, r21)
            TERM(1.0, 1.0, sA[oE + Cfg::ROW_E], rB22, r22)
            TERM(1.0, 0.5, sA[oE + 2 * Cfg::ROW_E], sB[oE + 2 * Cfg::ROW_E], sC[oE + 2 * Cfg::ROW_E])
            // ii = +1: x3 = 2X+1 (odd block, next column)
            TERM(0.5, 0.5, sA[oO + 1], sB[oO + 1], sC[oO + 1])
            TERM(0.5, 1.0, sA[oO + Cfg::ROW_O + 1], sB[oO + Cfg::ROW_O + 1], sC[oO + Cfg::ROW_O + 1])
            TERM(0.5, 0.5, sA[oO + 2 * Cfg::ROW_O + 1], sB[oO + 2 * Cfg::ROW_O + 1], sC[oO + 2 * Cfg::ROW_O + 1])
            if (live) p.coarse[cc] = acc;
            if (RV) {
                const double rv = 0.125 * ((0.5 * pvA + pvB) + 0.5 * pv);
                if (live) {
                    p.rv1[cc] = rv;
                    p.rv2[cc] = rv;
                }
                pvA = pv;
            }
            cc += p.gc.plane;
        }
#undef TERM
    } else {
        // ---- strip thread: one fine column of the extra column / row ----
        const int s = tid - Cfg::NCOMP;
        const bool used = s < Cfg::NSTRIP_USED;
        int bx, ry;  // column offset from x = xf0-1, row offset from y = yf0-1
        if (s < Cfg::RH) {
            bx = Cfg::FX;
            ry = s;
        } else {
            bx = min(s - Cfg::RH, Cfg::FX - 1);
            ry = Cfg::FY;
        }
        const int x = xf0 - 1 + bx, y = yf0 - 1 + ry;
        const bool mxy = used && x <= p.gf.nx && y <= p.gf.ny;
        const int o = (ry + 1) * VW + bx + 2;
        const int so = (bx & 1) ? EOFF + ry * Cfg::ROW_E + (bx >> 1) : ry * Cfg::ROW_O + (bx >> 1);
        const i64 c0 = gs_idx(p.gf, min(x, p.gf.nx), min(y, p.gf.ny), zfA);

        mbar_wait(&full[0], 0);
        double D = vbox(0)[o];
        mbar_wait(&full[1 % NSTAGE], (1 / NSTAGE) & 1);
        double C = vbox(1)[o];
        double fv = mxy ? __ldg(p.f + c0) : 0.0;
        double av = (KM >= KM_NEWTON_EXP && mxy) ? __ldg(p.aux + c0) : 0.0;
        auto do_plane = [&](int t) {
            const int q = t + 1;
            double nf = 0.0, na = 0.0;
            if (t < 2 * nzc && mxy) {
                const i64 cn = c0 + (i64)(t + 1) * plane;
                nf = __ldg(p.f + cn);
                if (KM >= KM_NEWTON_EXP) na = __ldg(p.aux + cn);
            }
            mbar_wait(&full[(q + 1) % NSTAGE], ((q + 1) / NSTAGE) & 1);
            const double* vb = vbox(q);
            const double U = vbox(q + 1)[o];
            const double E = vb[o + 1], W = vb[o - 1], N = vb[o + VW], S = vb[o - VW];
            double den;
            double r = canon_residual<KM, AR>(C, E, W, N, S, U, D, fv, av, p.k, den);
            r = (mxy && zgA + t <= p.gf.nz) ? r : 0.0;
            if (used) slot(t)[so] = r;
            D = C; C = U; fv = nf; av = na;
        };
        do_plane(0);
        for (int j = 0; j < nzc; j++) {
            do_plane(2 * j + 1);
            do_plane(2 * j + 2);
            consumer_sync<Cfg::NCONS>();
        }
    }
}

}  // namespace

// ---------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------
cudaError_t kt_prolong_correct(const LaunchCtx& cx, const Geom& gf, const Geom& gc, int zb, int ze, const double* vin,
                               double* vout, const double* coarse, bool* handled)
{
    *handled = false;
    if (ze <= zb) return cudaSuccess;
    if (gf.nx < cx.tune.fast_min_n || gf.ny < 8) return cudaSuccess;  // tiny levels: the one-thread-per-point kernel
    const int npx = (gf.nx + 1) / 2, npy = (gf.ny + 1) / 2;
    ProlongFastP p{gf, gc, zb, ze, 0, vin, vout, coarse};
    const int tiles = ((npx + PBX - 1) / PBX) * ((npy + PBY - 1) / PBY);
    int occ = 8;
    if (cudaError_t e = prolong_occ(&occ)) return e;
    if (cx.tune.wave) p.zchunk = gs_pick_zchunk_waves(cx, tiles, ze - zb, occ, 2, 8, 128, true);
    else p.zchunk = gs_pick_zchunk(cx, tiles, ze - zb, 64);
    p.zchunk += p.zchunk & 1;  // even chunks: every chunk starts on the same plane parity
    dim3 grid((npx + PBX - 1) / PBX, (npy + PBY - 1) / PBY, (ze - zb + p.zchunk - 1) / p.zchunk);
    prolong_correct_zmarch<<<grid, dim3(PBX, PBY), 0, cx.stream>>>(p);
    *handled = true;
    if (cx.launches) ++*cx.launches;
    return cudaGetLastError();
}

// coarse(interior of local coarse planes [zcb, zce)) = full weighting of fine; needs fine planes 2Z-1 .. 2Z+1 of each
cudaError_t kt_restrict(const LaunchCtx& cx, const Geom& gf, const Geom& gc, int zcb, int zce, const double* fine, double* coarse,
                        double* coarse2, bool* handled)
{
    *handled = false;
    if (zce <= zcb) return cudaSuccess;
    if (gc.nx < 24 || gc.ny < 8) return cudaSuccess;  // tiny levels: the one-thread-per-point kernel
    int occ = 8;
    if (cudaError_t e = gs_dev_once(g_restrict_once, &occ,
                                    [](int* o) { return cudaOccupancyMaxActiveBlocksPerMultiprocessor(o, restrict_zmarch, RBX * RBY, 0); }))
        return e;
    RestrictFastP p{gf, gc, zcb, zce, 0, fine, coarse, coarse2};
    const int tx = (gc.nx + RBX - 1) / RBX, ty = (gc.ny + RBY - 1) / RBY;
    p.zchunk = cx.tune.wave ? gs_pick_zchunk_waves(cx, tx * ty, zce - zcb, occ, 1, 4, 64, false) : gs_pick_zchunk(cx, tx * ty, zce - zcb, 32);
    dim3 grid(tx, ty, (zce - zcb + p.zchunk - 1) / p.zchunk);
    restrict_zmarch<<<grid, dim3(RBX, RBY), 0, cx.stream>>>(p);
    *handled = true;
    if (cx.launches) ++*cx.launches;
    return cudaGetLastError();
}

namespace {

template <int KM, int AR, int CX, int CY, int NSTAGE, bool RV = false>
cudaError_t launch_rr(const LaunchCtx& cx, ResidRestrictP p, bool* handled)
{
    using Cfg = RRCfg<CX, CY, NSTAGE>;
    static DevOnce once;
    int occ = 1;
    cudaError_t pe = gs_dev_once(once, &occ, [](int* o) {
        cudaError_t e = cudaFuncSetAttribute(resid_restrict_tma<KM, AR, CX, CY, NSTAGE, RV>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::SMEM);
        if (e == cudaSuccess)
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(o, resid_restrict_tma<KM, AR, CX, CY, NSTAGE, RV>, Cfg::THREADS, Cfg::SMEM);
        return e;
    });
    if (pe != cudaSuccess) return pe;
    if (!p.v) return cudaSuccess;  // preparation call
    CUtensorMap mv;
    if (!gs_tensor_map(p.v, p.gf, Cfg::VW, Cfg::VH, &mv)) return cudaSuccess;  // not handled
    const int tx = (p.gc.nx + CX - 1) / CX, ty = (p.gc.ny + CY - 1) / CY;
    // a chunk of coarse planes re-reads three fine planes (1.5 coarse-plane times) and fills the ring
    if (cx.tune.wave) p.zchunk = gs_pick_zchunk_waves(cx, tx * ty, p.zce - p.zcb, occ, 3, 4, 64, false);
    else p.zchunk = gs_pick_zchunk(cx, tx * ty, p.zce - p.zcb, 32);
    dim3 grid(tx, ty, (p.zce - p.zcb + p.zchunk - 1) / p.zchunk);
    resid_restrict_tma<KM, AR, CX, CY, NSTAGE, RV><<<grid, Cfg::THREADS, Cfg::SMEM, cx.stream>>>(mv, p);
    *handled = true;
    return cudaGetLastError();
}

template <int KM, int AR>
cudaError_t dispatch_rr(const LaunchCtx& cx, const ResidRestrictP& p, int cfg, bool* handled)
{
    if constexpr (KM == KM_NONLINEAR) {  // FAS: R(v) in the same pass (the six-stage ring only)
        if (p.rv1 || !p.v) {
            cudaError_t e = launch_rr<KM, AR, 32, 8, 6, true>(cx, p, handled);
            if (e != cudaSuccess || p.v) return e;
        }
    }
    if (p.rv1) return cudaSuccess;  // not handled
    if (cfg == 1) return launch_rr<KM, AR, 32, 8, 6>(cx, p, handled);
    return launch_rr<KM, AR, 32, 8, 5>(cx, p, handled);
}

}  // namespace

// f_{l+1}(coarse interior planes [zcb, zce)) = R(f_l - A_l v_l); the fine level must hold v planes
// 2*Zc-2 .. 2*Zc+2 of every coarse plane Zc it is asked for
cudaError_t kt_resid_restrict(const LaunchCtx& cx, int km, int ar, const Geom& gf, const Geom& gc, int zcb, int zce,
                              const PointK& k, const double* v, const double* f, const double* aux, double* coarse, bool* handled,
                              double* rv1, double* rv2)
{
    *handled = false;
    int cfg = 0;
    if (v) {
        if (zce <= zcb) return cudaSuccess;
        if (gc.nx < 24 || gc.ny < 8 || zce - zcb < 4) return cudaSuccess;  // small levels: unfused simple kernels
        cfg = cx.tune.rr_variant;
    }
    ResidRestrictP p{gf, gc, k, zcb, zce, 0, v, f, aux, coarse, rv1, rv2};
    cudaError_t e = cudaSuccess;
    for (int c = v ? cfg : 0; c <= (v ? cfg : 1) && e == cudaSuccess; c++) {
#define CALL(KMV) AR_SWITCH(ar, e = (dispatch_rr<KMV, ARV>(cx, p, c, handled)))
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

cudaError_t kt_prepare()
{
    LaunchCtx cx{};
    Geom g{};
    PointK k{};
    bool handled;
    cudaError_t e = cudaSuccess;
    for (int km = 0; km < 4 && e == cudaSuccess; km++)
        for (int pw = 0; pw < 8 && e == cudaSuccess; pw++)
            e = kt_resid_restrict(cx, km, pw, g, g, 0, 0, k, nullptr, nullptr, nullptr, nullptr, &handled);
    int occ;
    if (e == cudaSuccess) e = prolong_occ(&occ);
    return e;
}
