// ref_harness.cpp — TEST INFRASTRUCTURE, not product code.
//
// Puts the UNMODIFIED reference CPU path (src/cpu/*.cpp, src/Timer.cpp, compiled
// where they lie under $GPUSOLVE_REF by oracle/Makefile) behind oracle_api.h so
// the C restatement (gs_oracle.c) and the CUDA backend can be checked against
// the reference itself at full fp64 precision instead of its 6-digit stdout.
//
// The level operators are private statics (src/cpu/CpuSolver.h:10-15,
// src/cpu/NewtonSolver.h:8-10); access control is not part of the mangled name,
// so this one TU sees the class declarations with `private` spelled `public`
// (after every standard header is in) and links against the stock objects.
// No reference source is copied: only its headers are included at build time.
#include <algorithm>
#include <array>
#include <assert.h>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <string>
#include <tuple>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#define private public
#include "cpu/CpuGridData.h"
#include "cpu/CpuSolver.h"
#include "cpu/NewtonSolver.h"
#undef private

#include "oracle_api.h"

namespace {

struct Ref {
    CpuGridData* grid;
};

CpuGridData& G(void* g) { return *static_cast<Ref*>(g)->grid; }

Vector3* field_of(CpuGridData& grid, int level, int field)
{
    if (field == GS_FIELD_NEWTONF) return level == 0 ? &grid.newtonF : nullptr;
    if (level < 0 || level >= (int)grid.numLevels()) return nullptr;
    CpuGridData::LevelData& L = grid.getLevel(level);
    switch (field) {
    case GS_FIELD_V: return &L.v;
    case GS_FIELD_F: return &L.f;
    case GS_FIELD_R: return &L.r;
    case GS_FIELD_E: return &L.e;
    case GS_FIELD_RESTV: return &L.restV;
    case GS_FIELD_NEWTONV: return &L.newtonV;
    }
    return nullptr;
}

// the early-exit test exactly as spelled in src/cpu/CpuSolver.cpp:39 and
// src/cpu/NewtonSolver.cpp:37
bool reached(double res, double initial, double tol) { return res <= initial / (1.0 / tol); }

// CpuSolver::solve (src/cpu/CpuSolver.cpp:12-43) stepped through its own pieces so
// the norms are available at full precision
int solve_stepped(CpuGridData& grid, double* norms, int cap)
{
    double initial = CpuSolver::compResidual(grid, 0);
    if (norms && cap > 0) norms[0] = initial;
    int cycles = 0;
    for (std::size_t i = 0; i < grid.maxiter; i++) {
        double res = CpuSolver::vcycle(grid);
        cycles++;
        if (norms && cycles < cap) norms[cycles] = res;
        if (reached(res, initial, grid.tol)) break;
    }
    return cycles;
}

// NewtonSolver::findError (src/cpu/NewtonSolver.cpp:83-108) stepped, returns inner cycles
int find_error_stepped(CpuGridData& grid)
{
    for (std::size_t i = 1; i + 1 < grid.numLevels(); i++)
        CpuSolver::restrict(grid.getLevel(i - 1).newtonV, grid.getLevel(i).newtonV);
    const std::size_t it = grid.maxiter;
    const double tol = grid.tol;
    grid.maxiter = 10;
    grid.tol = 0.1;
    int cycles = solve_stepped(grid, nullptr, 0);
    grid.maxiter = it;
    grid.tol = tol;
    grid.getLevel(0).newtonV += grid.getLevel(0).v;
    return cycles;
}

} // namespace

extern "C" {

const char* gso_kind(void) { return "reference"; }

int gso_create(const gs_params* p, void** out)
{
    GridParams gp;
    gp.maxiter = p->maxiter;
    gp.tol = p->tol;
    gp.omega = p->omega;
    gp.gamma = p->gamma;
    gp.gridDim = {(std::size_t)p->dim[0], (std::size_t)p->dim[1], (std::size_t)p->dim[2]};
    gp.h = 1.0 / (gp.gridDim[1] + 1); // src/main.cpp:84
    gp.preSmoothing = p->pre;
    gp.postSmoothing = p->post;
    for (int i = 0; i < 7; i++) {
        gp.stencil.values[i] = p->stencil_val[i];
        gp.stencil.offsets[i] = std::make_tuple(p->stencil_off[i][0], p->stencil_off[i][1], p->stencil_off[i][2]);
    }
    gp.mode = static_cast<GridParams::Mode>(p->mode);
    gp.printProgress = p->print_progress != 0;
    Ref* r = new Ref;
    r->grid = new CpuGridData(gp);
    *out = r;
    return 0;
}

void gso_destroy(void* g)
{
    if (!g) return;
    delete static_cast<Ref*>(g)->grid;
    delete static_cast<Ref*>(g);
}

int gso_num_levels(void* g) { return (int)G(g).numLevels(); }

int gso_level_info(void* g, int level, uint64_t dim[3], double* h)
{
    if (level < 0 || level >= (int)G(g).numLevels()) return -1;
    const auto& L = G(g).getLevel(level);
    for (int i = 0; i < 3; i++) dim[i] = L.levelDim[i];
    *h = L.h;
    return 0;
}

int gso_get_field(void* g, int level, int field, double* out)
{
    Vector3* v = field_of(G(g), level, field);
    if (!v || v->flatSize() == 0) return -1;
    const std::size_t X = v->getXdim(), Y = v->getYdim(), Z = v->getZdim();
#pragma omp parallel for
    for (std::int64_t z = 0; z < (std::int64_t)Z; z++)
        for (std::size_t y = 0; y < Y; y++)
            for (std::size_t x = 0; x < X; x++) out[x + X * (y + Y * z)] = v->get(x, y, z);
    return 0;
}

int gso_set_field(void* g, int level, int field, const double* in)
{
    Vector3* v = field_of(G(g), level, field);
    if (!v) return -1;
    if (field == GS_FIELD_NEWTONF && v->flatSize() == 0) *v = G(g).getLevel(0).f; // allocate
    const std::size_t X = v->getXdim(), Y = v->getYdim(), Z = v->getZdim();
#pragma omp parallel for
    for (std::int64_t z = 0; z < (std::int64_t)Z; z++)
        for (std::size_t y = 0; y < Y; y++)
            for (std::size_t x = 0; x < X; x++) v->set(x, y, z, in[x + X * (y + Y * z)]);
    return 0;
}

double gso_residual(void* g, int level) { return CpuSolver::compResidual(G(g), level); }
void gso_smooth(void* g, int level, uint64_t iters) { CpuSolver::jacobi(G(g), level, iters); }

void gso_restrict(void* g, int level, int src_field, int dst_field)
{
    CpuSolver::restrict(*field_of(G(g), level, src_field), *field_of(G(g), level + 1, dst_field));
}

void gso_apply_stencil(void* g, int level)
{
    CpuSolver::applyStencil(G(g), level, G(g).getLevel(level).restV);
}

void gso_interpolate(void* g, int level) { CpuSolver::interpolate(G(g), level); }

void gso_axpy(void* g, int level, int dst_field, int src_field, int sign)
{
    Vector3& d = *field_of(G(g), level, dst_field);
    const Vector3& s = *field_of(G(g), level, src_field);
    if (sign >= 0) d += s; else d -= s;
}

double gso_vcycle(void* g) { return CpuSolver::vcycle(G(g)); }
int gso_solve(void* g, double* norms, int cap) { return solve_stepped(G(g), norms, cap); }

void gso_newton_begin(void* g) { G(g).newtonF = G(g).getLevel(0).f; } // NewtonSolver.cpp:12
double gso_newton_compF(void* g) { return NewtonSolver::compF(G(g)); }
int gso_newton_find_error(void* g) { return find_error_stepped(G(g)); }

// NewtonSolver::solve (src/cpu/NewtonSolver.cpp:10-44) stepped
int gso_newton_solve(void* g, double* norms, int cap, int* inner_cycles)
{
    CpuGridData& grid = G(g);
    grid.newtonF = grid.getLevel(0).f;
    double initial = NewtonSolver::compF(grid);
    if (norms && cap > 0) norms[0] = initial;
    int iters = 0, inner = 0;
    for (std::size_t i = 0; i < grid.maxiter; i++) {
        NewtonSolver::compF(grid);
        grid.getLevel(0).v.fill(0.0);
        inner += find_error_stepped(grid);
        double res = NewtonSolver::compF(grid);
        iters++;
        if (norms && iters < cap) norms[iters] = res;
        if (reached(res, initial, grid.tol)) break;
    }
    if (inner_cycles) *inner_cycles = inner;
    return iters;
}

// the reference's own entry points, untouched control flow, its own stdout
// (only in this library; used to validate the stepped loops above)
void gso_ref_native_solve(void* g)
{
    CpuGridData& grid = G(g);
    if (grid.mode == GridParams::NEWTON) NewtonSolver::solve(grid);
    else CpuSolver::solve(grid);
    std::cout.flush();
}

double gso_norm_exact(void* g, int level, int field)
{
    Vector3* v = field_of(G(g), level, field);
    if (!v) return -1.0;
    const std::size_t X = v->getXdim(), Y = v->getYdim(), Z = v->getZdim();
    long double s = 0.0L, c = 0.0L;
    for (std::size_t x = 1; x + 1 < X; x++)
        for (std::size_t y = 1; y + 1 < Y; y++)
            for (std::size_t z = 1; z + 1 < Z; z++) {
                long double a = v->get(x, y, z);
                long double t = a * a - c;
                long double u = s + t;
                c = (u - s) - t;
                s = u;
            }
    return (double)sqrtl(s);
}

void gso_set_threads(int n)
{
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

} // extern "C"
