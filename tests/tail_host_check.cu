// Test harness (CPU only, built by tests/test_tail_host.py): runs the coarse-tail RECORDER and its operator
// interpreter (kernels_simple.cu: ks_* with LaunchCtx::rec set, tail_do_op) over HOST arrays.  The translation unit
// under test is included as a whole so that its internals are reachable; the only thing substituted is how a finished
// program is executed (TailRec::exec): a loop over the 1024 thread indices instead of the single-CTA kernel launch.
// Operators inside one program have no cross-point dependences, so running the thread indices one after the other
// gives what the CTA gives.  The V-cycle below queues the operators in the order gs_api.cu enqueue_vcycle does for
// levels that use the one-thread-per-point kernels (src/cpu/CpuSolver.cpp:85-139).
// Nothing here is linked into the library.
#include "../gpu-solve_b200/csrc/kernels_simple.cu"

#include <math.h>

#include <utility>
#include <vector>

namespace {

int g_programs = 0, g_ops = 0;

cudaError_t exec_host(const LaunchCtx&, const TailProg& P, int km)
{
    g_programs++;
    g_ops += P.nops;
    for (int i = 0; i < P.nops; i++)
        for (int tid = 0; tid < TAIL_THREADS; tid++) switch (km) {
            case KM_LINEAR: tail_do_op<KM_LINEAR>(P, P.op[i], tid, TAIL_THREADS); break;
            case KM_NONLINEAR: tail_do_op<KM_NONLINEAR>(P, P.op[i], tid, TAIL_THREADS); break;
            case KM_NEWTON_EXP: tail_do_op<KM_NEWTON_EXP>(P, P.op[i], tid, TAIL_THREADS); break;
            default: tail_do_op<KM_NEWTON_COEF>(P, P.op[i], tid, TAIL_THREADS); break;
            }
    return cudaSuccess;
}

struct HLevel {
    Geom g;
    PointK k;
    bool pow2, vzero;
    size_t count;
    std::vector<double> bv, bt, bf, br, brest, bnv, bcoef;
    double *v, *vtmp;
};

bool pow2_of(double x)
{
    int e;
    return x > 0 && frexp(x, &e) == 0.5;
}

void to_padded(const Geom& g, const double* dense, double* a)
{
    for (int z = 0; z < g.Zg; z++)
        for (int y = 0; y < g.Y; y++)
            for (int x = 0; x < g.X; x++) a[gs_idx(g, x, y, z)] = dense[x + (size_t)g.X * (y + (size_t)g.Y * z)];
}

void to_dense(const Geom& g, const double* a, double* dense)
{
    for (int z = 0; z < g.Zg; z++)
        for (int y = 0; y < g.Y; y++)
            for (int x = 0; x < g.X; x++) dense[x + (size_t)g.X * (y + (size_t)g.Y * z)] = a[gs_idx(g, x, y, z)];
}

}  // namespace

#define CK(expr)                          \
    do {                                  \
        if ((expr) != cudaSuccess) return -1; \
    } while (0)

// V-cycles of the whole hierarchy through the recorder, from level `tail` down (levels above it are executed operator
// by operator through the same interpreter, i.e. as programs of one operator).
//   km_coef: mode 2 only, 1 = precomputed coefficient (KM_NEWTON_COEF), 0 = exp in the sweep (KM_NEWTON_EXP)
//   nv:      mode 2 only, dense newtonV of every level
//   stats:   [programs executed, operators executed]
extern "C" int tail_host_vcycle(const gs_params* p, int tail, int km_coef, int cycles, const double* v0, const double* f0,
                                const double* const* nv, double* v_out, int* stats)
{
    const uint64_t m = std::min(std::min(p->dim[0], p->dim[1]), p->dim[2]);
    const int nlev = (int)floor(log((double)m) / log(2.0)) + 1;
    std::vector<HLevel> lev(nlev);
    uint64_t n[3] = {p->dim[0], p->dim[1], p->dim[2]};
    for (int l = 0; l < nlev; l++) {
        HLevel& L = lev[l];
        Geom& g = L.g;
        g.nx = (int)n[0], g.ny = (int)n[1], g.nz = (int)n[2];
        g.X = g.nx + 2, g.Y = g.ny + 2, g.Zg = g.nz + 2;
        g.z0 = 0, g.nzl = g.Zg;
        g.pitch = ((GS_XOFF + g.X + 15) / 16) * 16;
        g.plane = g.pitch * g.Y;
        L.count = (size_t)g.plane * g.nzl;
        const double h = 1.0 / (double)(n[1] + 1);
        for (int i = 0; i < 7; i++) L.k.c[i] = p->stencil_val[i];
        L.k.hh = h * h;
        L.k.inv_hh = 1.0 / L.k.hh;
        L.k.alpha = (h * h) / p->stencil_val[0];
        L.k.preFac = p->stencil_val[0] / (h * h);
        L.k.omega = p->omega;
        L.k.gamma = p->gamma;
        L.pow2 = pow2_of(L.k.hh);
        L.vzero = false;
        for (auto* b : {&L.bv, &L.bt, &L.bf, &L.br, &L.brest, &L.bnv, &L.bcoef}) b->assign(L.count, 0.0);
        L.v = L.bv.data();
        L.vtmp = L.bt.data();
        if (nv) to_padded(g, nv[l], L.bnv.data());
        for (int d = 0; d < 3; d++) n[d] /= 2;
    }
    to_padded(lev[0].g, v0, lev[0].v);
    to_padded(lev[0].g, v0, lev[0].vtmp);
    to_padded(lev[0].g, f0, lev[0].bf.data());
    StencilOff so;
    for (int i = 0; i < 7; i++) so.dx[i] = p->stencil_off[i][0], so.dy[i] = p->stencil_off[i][1], so.dz[i] = p->stencil_off[i][2];

    const int km = p->mode == 0 ? KM_LINEAR : p->mode == 1 ? KM_NONLINEAR : km_coef ? KM_NEWTON_COEF : KM_NEWTON_EXP;
    TailRec* rec = ks_tail_new();
    rec->exec = exec_host;
    LaunchCtx cx;
    memset(&cx, 0, sizeof cx);
    g_programs = g_ops = 0;
    const int last = nlev - 1;
    const int pre = (int)p->pre, post = (int)p->post;

    auto aux = [&](HLevel& L) -> const double* { return km == KM_NEWTON_EXP ? L.bnv.data() : km == KM_NEWTON_COEF ? L.bcoef.data() : nullptr; };
    // outside the tail every operator is its own program
    auto begin1 = [&](bool in_tail) { if (!in_tail) ks_tail_begin(cx, rec); };
    auto end1 = [&](bool in_tail) { return in_tail ? cudaSuccess : ks_tail_end(cx); };
    auto smooth = [&](int l, int iters, bool in_tail) -> int {
        HLevel& L = lev[l];
        begin1(in_tail);
        if (L.vzero) {
            L.vzero = false;
            CK(ks_zero(cx, L.v, L.count));
        }
        for (int i = 0; i < iters; i++) {
            CK(ks_smooth(cx, km, L.pow2, L.g, 1, L.g.nz + 1, L.k, so, L.v, L.vtmp, L.bf.data(), aux(L)));
            std::swap(L.v, L.vtmp);
        }
        CK(end1(in_tail));
        return 0;
    };

    if (km == KM_NEWTON_COEF) {  // gs_api.cu run_vcycle: ensure_coef on every level before the cycle
        ks_tail_begin(cx, rec);
        for (int l = 0; l < nlev; l++)
            CK(ks_newton_coef(cx, lev[l].g, 0, lev[l].g.nzl, p->gamma, lev[l].bnv.data(), lev[l].bcoef.data()));
        CK(ks_tail_end(cx));
    }
    for (int c = 0; c < cycles; c++) {
        bool in_tail = false;
        for (int i = 0; i < last; i++) {
            HLevel &F = lev[i], &C = lev[i + 1];
            if (i == tail) {
                ks_tail_begin(cx, rec);
                in_tail = true;
            }
            if (smooth(i, pre, in_tail)) return -1;
            begin1(in_tail);
            CK(ks_residual(cx, km, F.pow2, F.g, 1, F.g.nz + 1, F.k, so, F.v, F.bf.data(), aux(F), F.br.data()));
            CK(ks_restrict(cx, F.g, C.g, 1, C.g.nz + 1, F.br.data(), C.bf.data(), nullptr));
            if (p->mode != 1) {
                C.vzero = true;
            } else {
                CK(ks_restrict(cx, F.g, C.g, 1, C.g.nz + 1, F.v, C.brest.data(), C.v));
                CK(ks_fas_rhs(cx, C.pow2, C.g, 1, C.g.nz + 1, C.k, so, C.brest.data(), C.bf.data(), C.br.data()));
            }
            CK(end1(in_tail));
        }
        if (smooth(last, pre + post, in_tail)) return -1;
        if ((pre + post) & 1) {
            HLevel& L = lev[last];
            begin1(in_tail);
            CK(ks_copy(cx, L.vtmp, L.v, L.count));
            std::swap(L.v, L.vtmp);
            CK(end1(in_tail));
        }
        for (int i = last; i > 0; i--) {
            HLevel &F = lev[i - 1], &C = lev[i];
            begin1(in_tail);
            if (p->mode == 1) CK(ks_axpy(cx, C.g, 0, C.g.nzl, C.v, C.brest.data(), -1));
            const bool flip = ((pre + post) & 1) != 0;
            CK(ks_prolong_correct(cx, F.g, C.g, 1, F.g.nz + 1, F.v, flip ? F.vtmp : F.v, C.v, nullptr));
            if (flip) std::swap(F.v, F.vtmp);
            CK(end1(in_tail));
            if (smooth(i - 1, post, in_tail)) return -1;
            if (i - 1 == tail) {
                CK(ks_tail_end(cx));
                in_tail = false;
            }
        }
    }
    to_dense(lev[0].g, lev[0].v, v_out);
    if (stats) {
        stats[0] = g_programs;
        stats[1] = g_ops;
    }
    ks_tail_free(rec);
    return 0;
}
