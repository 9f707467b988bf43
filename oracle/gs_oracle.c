/*
 * gs_oracle.c — TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C restatement of the reference's CPU multigrid path (GpuSolve-cpu), the
 * CPU oracle the CUDA backend is checked against.  It is NOT a fallback: nothing
 * under gpu-solve_b200/ links, loads or calls it.
 *
 * Parity: PINNED.  tests/test_oracle.py checks this file bit for bit against the
 * reference's own translation units (oracle/_ref/libgpusolve_ref.so, built by
 * oracle/Makefile from $GPUSOLVE_REF) and against golden vectors generated from
 * them (tests/golden/, generator tests/golden/make_golden.py).  The reference
 * ships no tests or golden vectors of its own.
 *
 * Every value is produced with the same fp64 operations in the same order as the
 * reference, so fields are bit-identical; only the order of the norm's running
 * sum differs (the reference's own norm depends on OMP_NUM_THREADS).  Compile
 * without FMA contraction (-ffp-contract=off, baseline x86-64) like the
 * reference's effective flags.
 *
 * Storage here is x-fastest dense, idx = x + X*(y + Y*z); the reference's
 * Vector3 is z-fastest (src/cpu/Vector3.cpp:16).  Everything is compared by
 * logical (x,y,z).  Citations are relative to the reference root.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "oracle_api.h"

typedef struct {
    uint64_t n[3];     /* interior points, LevelData::levelDim */
    size_t X, Y, Z;    /* array dims = n + 2                   */
    size_t count;
    double h;
    double *v, *restV, *newtonV, *f, *r, *e;
} level_t;

typedef struct {
    gs_params p;
    int nlev;
    level_t* lev;
    double* newtonF;
} grid_t;

#define IDX(L, x, y, z) ((size_t)(x) + (L)->X * ((size_t)(y) + (L)->Y * (size_t)(z)))

static double* zeros(size_t n) { return (double*)calloc(n ? n : 1, sizeof(double)); }

/* src/cpu/CpuGridData.cpp:7-12 */
static double rhs_f0(double x)
{
    return (100 * x * (x - 1.0) * x * (x - 1.0) * x * (x - 1.0) * x * (x - 1.0));
}
static double rhs_f2(double x)
{
    return (100.0 * 4.0 * (x - 1.0) * (x - 1.0) * x * x * (14.0 * x * x - 14.0 * x + 3));
}

const char* gso_kind(void) { return "port"; }

/* CpuGridData::CpuGridData, src/cpu/CpuGridData.cpp:15-80 */
int gso_create(const gs_params* p, void** out)
{
    grid_t* g = (grid_t*)calloc(1, sizeof(grid_t));
    g->p = *p;
    g->p.h = 1.0 / (double)(p->dim[1] + 1); /* src/main.cpp:84 */
    uint64_t m = p->dim[0] < p->dim[1] ? p->dim[0] : p->dim[1];
    if (p->dim[2] < m) m = p->dim[2];
    /* :19  floor(log(min)/log 2) + 1 */
    g->nlev = (int)floor(log((double)m) / log(2.0)) + 1;
    g->lev = (level_t*)calloc((size_t)g->nlev, sizeof(level_t));
    for (int l = 0; l < g->nlev; l++) {
        level_t* L = &g->lev[l];
        for (int d = 0; d < 3; d++) L->n[d] = l == 0 ? p->dim[d] : g->lev[l - 1].n[d] / 2; /* :23-29 */
        L->X = L->n[0] + 2;
        L->Y = L->n[1] + 2;
        L->Z = L->n[2] + 2;
        L->count = L->X * L->Y * L->Z;
        L->v = zeros(L->count);
        L->restV = zeros(L->count);
        L->newtonV = zeros(L->count);
        L->f = zeros(L->count);
        L->r = zeros(L->count);
        L->e = (l + 1 != g->nlev) ? zeros(L->count) : NULL; /* :37-39 */
        L->h = 1.0 / (double)(L->n[1] + 1);                  /* :41 (Y dimension) */
    }
    level_t* L0 = &g->lev[0];
    const double h = g->p.h;
    if (p->mode == GS_MODE_LINEAR) {
        /* :47-58 — sampled at the ZERO-BASED interior index */
        for (int64_t k = 0; k < (int64_t)L0->n[2]; k++)
            for (int64_t j = 0; j < (int64_t)L0->n[1]; j++)
                for (int64_t i = 0; i < (int64_t)L0->n[0]; i++) {
                    double x = (int)i * h, y = (int)j * h, z = (int)k * h;
                    double val = -(rhs_f2(x) * rhs_f0(y) * rhs_f0(z) + rhs_f0(x) * rhs_f2(y) * rhs_f0(z) +
                                   rhs_f0(x) * rhs_f0(y) * rhs_f2(z));
                    L0->f[IDX(L0, i + 1, j + 1, k + 1)] = val;
                }
    } else {
        /* :62-76 — sampled at the ARRAY index, halo included */
        const double gamma = p->gamma;
        for (int64_t k = 0; k < (int64_t)L0->Z; k++)
            for (int64_t j = 0; j < (int64_t)L0->Y; j++)
                for (int64_t i = 0; i < (int64_t)L0->X; i++) {
                    double x = (int)i * h, y = (int)j * h, z = (int)k * h;
                    double val = 2.0 * ((y - y * y) * (z - z * z) + (x - x * x) * (z - z * z) + (x - x * x) * (y - y * y)) +
                                 gamma * (x - x * x) * (y - y * y) * (z - z * z) *
                                     exp((x - x * x) * (y - y * y) * (z - z * z));
                    L0->f[IDX(L0, i, j, k)] = val;
                }
    }
    *out = g;
    return 0;
}

void gso_destroy(void* gv)
{
    grid_t* g = (grid_t*)gv;
    if (!g) return;
    for (int l = 0; l < g->nlev; l++) {
        level_t* L = &g->lev[l];
        free(L->v); free(L->restV); free(L->newtonV); free(L->f); free(L->r); free(L->e);
    }
    free(g->lev);
    free(g->newtonF);
    free(g);
}

int gso_num_levels(void* gv) { return ((grid_t*)gv)->nlev; }

int gso_level_info(void* gv, int level, uint64_t dim[3], double* h)
{
    grid_t* g = (grid_t*)gv;
    if (level < 0 || level >= g->nlev) return -1;
    for (int d = 0; d < 3; d++) dim[d] = g->lev[level].n[d];
    *h = g->lev[level].h;
    return 0;
}

static double** field_slot(grid_t* g, int level, int field)
{
    if (field == GS_FIELD_NEWTONF) return level == 0 ? &g->newtonF : NULL;
    if (level < 0 || level >= g->nlev) return NULL;
    level_t* L = &g->lev[level];
    switch (field) {
    case GS_FIELD_V: return &L->v;
    case GS_FIELD_F: return &L->f;
    case GS_FIELD_R: return &L->r;
    case GS_FIELD_E: return &L->e;
    case GS_FIELD_RESTV: return &L->restV;
    case GS_FIELD_NEWTONV: return &L->newtonV;
    }
    return NULL;
}

int gso_get_field(void* gv, int level, int field, double* out)
{
    grid_t* g = (grid_t*)gv;
    double** s = field_slot(g, level, field);
    if (!s || !*s) return -1;
    memcpy(out, *s, g->lev[level].count * sizeof(double));
    return 0;
}

int gso_set_field(void* gv, int level, int field, const double* in)
{
    grid_t* g = (grid_t*)gv;
    double** s = field_slot(g, level, field);
    if (!s) return -1;
    if (!*s) {
        if (field != GS_FIELD_NEWTONF) return -1;
        *s = zeros(g->lev[0].count);
    }
    memcpy(*s, in, g->lev[level].count * sizeof(double));
    return 0;
}

/* sum_i c_i * a(p + o_i), i ascending, starting from 0.0 — the inner loop of
 * src/cpu/CpuSolver.cpp:56-60 (also :195-199, src/cpu/NewtonSolver.cpp:59-62) */
static inline double taps(const grid_t* g, const level_t* L, const double* a, size_t x, size_t y, size_t z)
{
    double s = 0.0;
    for (int i = 0; i < 7; i++) {
        double t = a[IDX(L, x + g->p.stencil_off[i][0], y + g->p.stencil_off[i][1], z + g->p.stencil_off[i][2])];
        s += g->p.stencil_val[i] * t;
    }
    return s;
}

/* plane-wise partial sums combined in plane order: independent of thread count */
static double ordered_sum(const double* part, size_t n)
{
    double s = 0.0;
    for (size_t i = 0; i < n; i++) s += part[i];
    return s;
}

/* CpuSolver::compResidual, src/cpu/CpuSolver.cpp:45-83 */
double gso_residual(void* gv, int level)
{
    grid_t* g = (grid_t*)gv;
    level_t* L = &g->lev[level];
    const int mode = g->p.mode;
    const double gamma = g->p.gamma;
    double* part = zeros(L->Z);
#pragma omp parallel for schedule(static)
    for (int64_t z = 1; z <= (int64_t)L->n[2]; z++) {
        double acc = 0.0;
        for (size_t y = 1; y <= L->n[1]; y++)
            for (size_t x = 1; x <= L->n[0]; x++) {
                const size_t c = IDX(L, x, y, z);
                double s = taps(g, L, L->v, x, y, (size_t)z);
                s /= L->h * L->h; /* :61 */
                if (mode == GS_MODE_NEWTON) { /* :63-66 */
                    double ex = exp(L->newtonV[c]);
                    s += gamma * (1 + L->newtonV[c]) * L->v[c] * ex;
                } else if (mode == GS_MODE_NONLINEAR) { /* :67-72 */
                    double ex = exp(L->v[c]);
                    double nl = gamma * L->v[c] * ex;
                    s += nl;
                }
                double r = L->f[c] - s; /* :74 */
                L->r[c] = r;
                acc += r * r;
            }
        part[z] = acc;
    }
    double tot = ordered_sum(part, L->Z);
    free(part);
    return sqrt(tot);
}

/* CpuSolver::jacobi, src/cpu/CpuSolver.cpp:141-179 */
void gso_smooth(void* gv, int level, uint64_t iters)
{
    grid_t* g = (grid_t*)gv;
    level_t* L = &g->lev[level];
    const int mode = g->p.mode;
    const double gamma = g->p.gamma, omega = g->p.omega;
    const double preFac = g->p.stencil_val[0] / (L->h * L->h); /* :144 */
    const double alpha = (L->h * L->h) / g->p.stencil_val[0];  /* :145 */
    for (uint64_t it = 0; it < iters; it++) {
        gso_residual(gv, level); /* :149 */
#pragma omp parallel for schedule(static)
        for (int64_t z = 1; z <= (int64_t)L->n[2]; z++)
            for (size_t y = 1; y <= L->n[1]; y++)
                for (size_t x = 1; x <= L->n[0]; x++) {
                    const size_t c = IDX(L, x, y, z);
                    double nv;
                    if (mode == GS_MODE_LINEAR) {
                        nv = L->v[c] + omega * (alpha * L->r[c]); /* :158 */
                    } else if (mode == GS_MODE_NONLINEAR) {
                        double ex = exp(L->v[c]); /* :161-164 */
                        double den = preFac + gamma * (1 + L->v[c]) * ex;
                        nv = L->v[c] + omega * (L->r[c] / den);
                    } else {
                        double ex = exp(L->newtonV[c]); /* :168-171 */
                        double den = preFac + gamma * (1 + L->newtonV[c]) * ex;
                        nv = L->v[c] + omega * (L->r[c] / den);
                    }
                    L->v[c] = nv;
                }
    }
}

/* CpuSolver::applyStencil on restV -> r, src/cpu/CpuSolver.cpp:182-209 */
void gso_apply_stencil(void* gv, int level)
{
    grid_t* g = (grid_t*)gv;
    level_t* L = &g->lev[level];
    const double gamma = g->p.gamma;
#pragma omp parallel for schedule(static)
    for (int64_t z = 1; z <= (int64_t)L->n[2]; z++)
        for (size_t y = 1; y <= L->n[1]; y++)
            for (size_t x = 1; x <= L->n[0]; x++) {
                const size_t c = IDX(L, x, y, z);
                double s = taps(g, L, L->restV, x, y, (size_t)z);
                s /= L->h * L->h;
                double nl = gamma * L->restV[c] * exp(L->restV[c]); /* :202 */
                s += nl;
                L->r[c] = s;
            }
}

/* CpuSolver::restrict, src/cpu/CpuSolver.cpp:211-238: 27-point full weighting,
 * running sum with the x offset outermost and the z offset innermost */
static void restrict_arrays(const level_t* F, const double* fine, const level_t* C, double* coarse)
{
#pragma omp parallel for schedule(static)
    for (int64_t z = 1; z < (int64_t)C->Z - 1; z++)
        for (size_t y = 1; y + 1 < C->Y; y++)
            for (size_t x = 1; x + 1 < C->X; x++) {
                const size_t xc = 2 * x, yc = 2 * y, zc = 2 * (size_t)z;
                double acc = 0.0;
                for (int ii = -1; ii < 2; ii++)
                    for (int jj = -1; jj < 2; jj++)
                        for (int kk = -1; kk < 2; kk++) {
                            double fac = 0.125 * ((2.0 - abs(ii)) / 2.0) * ((2.0 - abs(jj)) / 2.0) *
                                         ((2.0 - abs(kk)) / 2.0); /* :227 */
                            acc += fac * fine[IDX(F, xc + ii, yc + jj, zc + kk)];
                        }
                coarse[IDX(C, x, y, z)] = acc;
            }
}

void gso_restrict(void* gv, int level, int src_field, int dst_field)
{
    grid_t* g = (grid_t*)gv;
    double** s = field_slot(g, level, src_field);
    double** d = field_slot(g, level + 1, dst_field);
    restrict_arrays(&g->lev[level], *s, &g->lev[level + 1], *d);
}

/* CpuSolver::interpolate, src/cpu/CpuSolver.cpp:240-290: inject, then x, y, z
 * averaging passes, each reading what the previous one wrote */
void gso_interpolate(void* gv, int level)
{
    grid_t* g = (grid_t*)gv;
    level_t* F = &g->lev[level];
    level_t* C = &g->lev[level + 1];
    const double* cv = C->v;
    double* e = F->e;
    /* :247-255 */
#pragma omp parallel for schedule(static)
    for (int64_t z = 0; z < (int64_t)F->Z - 1; z += 2)
        for (size_t y = 0; y + 1 < F->Y; y += 2)
            for (size_t x = 0; x + 1 < F->X; x += 2) e[IDX(F, x, y, z)] = cv[IDX(C, x / 2, y / 2, z / 2)];
    /* :258-266 */
#pragma omp parallel for schedule(static)
    for (int64_t z = 0; z < (int64_t)F->Z; z += 2)
        for (size_t y = 0; y < F->Y; y += 2)
            for (size_t x = 0; x + 2 < F->X; x += 2)
                e[IDX(F, x + 1, y, z)] = 0.5 * e[IDX(F, x, y, z)] + 0.5 * e[IDX(F, x + 2, y, z)];
    /* :269-277 */
#pragma omp parallel for schedule(static)
    for (int64_t z = 0; z < (int64_t)F->Z; z += 2)
        for (size_t y = 0; y + 2 < F->Y; y += 2)
            for (size_t x = 0; x < F->X; x++)
                e[IDX(F, x, y + 1, z)] = 0.5 * e[IDX(F, x, y, z)] + 0.5 * e[IDX(F, x, y + 2, z)];
    /* :280-288 */
#pragma omp parallel for schedule(static)
    for (int64_t z = 0; z < (int64_t)F->Z - 2; z += 2)
        for (size_t y = 0; y < F->Y; y++)
            for (size_t x = 0; x < F->X; x++)
                e[IDX(F, x, y, z + 1)] = 0.5 * e[IDX(F, x, y, z)] + 0.5 * e[IDX(F, x, y, z + 2)];
}

/* Vector3::operator+= / -=, src/cpu/Vector3.cpp:34-54: whole array, halo included */
void gso_axpy(void* gv, int level, int dst_field, int src_field, int sign)
{
    grid_t* g = (grid_t*)gv;
    double* d = *field_slot(g, level, dst_field);
    const double* s = *field_slot(g, level, src_field);
    const size_t n = g->lev[level].count;
    if (sign >= 0)
        for (size_t i = 0; i < n; i++) d[i] += s[i];
    else
        for (size_t i = 0; i < n; i++) d[i] -= s[i];
}

/* CpuSolver::vcycle, src/cpu/CpuSolver.cpp:85-139 */
double gso_vcycle(void* gv)
{
    grid_t* g = (grid_t*)gv;
    const int last = g->nlev - 1;
    for (int i = 0; i < last; i++) {
        gso_smooth(gv, i, g->p.pre);       /* :88 */
        gso_residual(gv, i);               /* :93 */
        gso_restrict(gv, i, GS_FIELD_R, GS_FIELD_F); /* :98 */
        level_t* N = &g->lev[i + 1];
        if (g->p.mode != GS_MODE_NONLINEAR) {
            memset(N->v, 0, N->count * sizeof(double)); /* :100-101 */
        } else {
            gso_restrict(gv, i, GS_FIELD_V, GS_FIELD_RESTV); /* :106 */
            gso_restrict(gv, i, GS_FIELD_V, GS_FIELD_V);     /* :107 */
            gso_apply_stencil(gv, i + 1);                    /* :110 */
            gso_axpy(gv, i + 1, GS_FIELD_F, GS_FIELD_R, +1); /* :112 */
        }
    }
    gso_smooth(gv, last, g->p.pre + g->p.post); /* :117 */
    for (int i = last; i > 0; i--) {
        if (g->p.mode == GS_MODE_NONLINEAR) gso_axpy(gv, i, GS_FIELD_V, GS_FIELD_RESTV, -1); /* :121-125 */
        gso_interpolate(gv, i - 1);                                                          /* :128 */
        gso_axpy(gv, i - 1, GS_FIELD_V, GS_FIELD_E, +1);                                     /* :131-132 */
        gso_smooth(gv, i - 1, g->p.post);                                                    /* :134 */
    }
    return gso_residual(gv, 0); /* :138 */
}

/* the early-exit test as spelled at src/cpu/CpuSolver.cpp:39 */
static int reached(double res, double initial, double tol) { return res <= initial / (1.0 / tol); }

/* CpuSolver::solve, src/cpu/CpuSolver.cpp:12-43 */
int gso_solve(void* gv, double* norms, int cap)
{
    grid_t* g = (grid_t*)gv;
    double initial = gso_residual(gv, 0); /* :15 */
    if (norms && cap > 0) norms[0] = initial;
    int cycles = 0;
    for (uint64_t i = 0; i < g->p.maxiter; i++) {
        double res = gso_vcycle(gv);
        cycles++;
        if (norms && cycles < cap) norms[cycles] = res;
        if (reached(res, initial, g->p.tol)) break;
    }
    return cycles;
}

/* NewtonSolver::solve line 12: newtonF = level0.f */
void gso_newton_begin(void* gv)
{
    grid_t* g = (grid_t*)gv;
    if (!g->newtonF) g->newtonF = zeros(g->lev[0].count);
    memcpy(g->newtonF, g->lev[0].f, g->lev[0].count * sizeof(double));
}

/* NewtonSolver::compF, src/cpu/NewtonSolver.cpp:48-81 */
double gso_newton_compF(void* gv)
{
    grid_t* g = (grid_t*)gv;
    level_t* L = &g->lev[0];
    const double gamma = g->p.gamma;
    double* part = zeros(L->Z);
#pragma omp parallel for schedule(static)
    for (int64_t z = 1; z <= (int64_t)L->n[2]; z++) {
        double acc = 0.0;
        for (size_t y = 1; y <= L->n[1]; y++)
            for (size_t x = 1; x <= L->n[0]; x++) {
                const size_t c = IDX(L, x, y, z);
                double s = taps(g, L, L->newtonV, x, y, (size_t)z);
                s /= L->h * L->h;                  /* :64 */
                double ex = exp(L->newtonV[c]);    /* :67 */
                double nl = gamma * L->newtonV[c] * ex;
                s += nl;
                double f = g->newtonF[c] - s;      /* :71 */
                L->f[c] = f;
                acc += f * f;
            }
        part[z] = acc;
    }
    double tot = ordered_sum(part, L->Z);
    free(part);
    return sqrt(tot);
}

/* NewtonSolver::findError, src/cpu/NewtonSolver.cpp:83-108.  newtonV is NOT
 * restricted to the coarsest level (loop bound numLevels()-1, :88). */
int gso_newton_find_error(void* gv)
{
    grid_t* g = (grid_t*)gv;
    for (int i = 1; i + 1 < g->nlev; i++) gso_restrict(gv, i - 1, GS_FIELD_NEWTONV, GS_FIELD_NEWTONV);
    const uint64_t it = g->p.maxiter;
    const double tol = g->p.tol;
    g->p.maxiter = 10; /* :97 */
    g->p.tol = 0.1;    /* :98 */
    int cycles = gso_solve(gv, NULL, 0);
    g->p.maxiter = it;
    g->p.tol = tol;
    gso_axpy(gv, 0, GS_FIELD_NEWTONV, GS_FIELD_V, +1); /* :106-107 */
    return cycles;
}

/* NewtonSolver::solve, src/cpu/NewtonSolver.cpp:10-44 */
int gso_newton_solve(void* gv, double* norms, int cap, int* inner_cycles)
{
    grid_t* g = (grid_t*)gv;
    gso_newton_begin(gv);
    double initial = gso_newton_compF(gv); /* :15 */
    if (norms && cap > 0) norms[0] = initial;
    int iters = 0, inner = 0;
    for (uint64_t i = 0; i < g->p.maxiter; i++) {
        gso_newton_compF(gv);                                         /* :21 */
        memset(g->lev[0].v, 0, g->lev[0].count * sizeof(double));     /* :22 */
        inner += gso_newton_find_error(gv);                           /* :24 */
        double res = gso_newton_compF(gv);                            /* :26 */
        iters++;
        if (norms && iters < cap) norms[iters] = res;
        if (reached(res, initial, g->p.tol)) break;                   /* :37 */
    }
    if (inner_cycles) *inner_cycles = inner;
    return iters;
}

double gso_norm_exact(void* gv, int level, int field)
{
    grid_t* g = (grid_t*)gv;
    double** sl = field_slot(g, level, field);
    if (!sl || !*sl) return -1.0;
    const level_t* L = &g->lev[level];
    const double* a = *sl;
    long double s = 0.0L, c = 0.0L;
    for (size_t z = 1; z + 1 < L->Z; z++)
        for (size_t y = 1; y + 1 < L->Y; y++)
            for (size_t x = 1; x + 1 < L->X; x++) {
                long double q = a[IDX(L, x, y, z)];
                long double t = q * q - c;
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
