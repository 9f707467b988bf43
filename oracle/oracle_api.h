/*
 * oracle_api.h — TEST INFRASTRUCTURE, not product code.
 *
 * One C API, two implementations that must agree bit for bit:
 *   oracle/gs_oracle.c      -> oracle/libgs_oracle.so        plain-C restatement of the
 *                              reference CPU path (each function cites file:line)
 *   oracle/ref_harness.cpp  -> oracle/_ref/libgpusolve_ref.so  the UNMODIFIED reference
 *                              translation units compiled where they lie under
 *                              $GPUSOLVE_REF (default /root/reference) behind this API
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load either library.  The CUDA product path never does.
 *
 * Host buffers are dense (nx+2)(ny+2)(nz+2) doubles, x fastest:
 * idx = x + X*(y + Y*z) — i.e. addressed by logical (x,y,z), independent of the
 * reference's own z-fastest Vector3 storage (src/cpu/Vector3.cpp:16).
 */
#ifndef GS_ORACLE_API_H
#define GS_ORACLE_API_H
#include "../include/gpusolve_cuda.h"   /* gs_params, GS_FIELD_*, GS_MODE_* only */

#ifdef __cplusplus
extern "C" {
#endif

const char* gso_kind(void);   /* "port" (C restatement) or "reference" (real TUs) */

int    gso_create(const gs_params* p, void** out);          /* CpuGridData ctor  */
void   gso_destroy(void* g);
int    gso_num_levels(void* g);
int    gso_level_info(void* g, int level, uint64_t dim[3], double* h);
int    gso_get_field(void* g, int level, int field, double* out);
int    gso_set_field(void* g, int level, int field, const double* in);

double gso_residual(void* g, int level);                     /* compResidual      */
void   gso_smooth(void* g, int level, uint64_t iters);       /* jacobi            */
void   gso_restrict(void* g, int level, int src_field, int dst_field); /* restrict level -> level+1 */
void   gso_apply_stencil(void* g, int level);                /* r = A(restV), mode 1 */
void   gso_interpolate(void* g, int level);                  /* e_level = P v_{level+1} */
void   gso_axpy(void* g, int level, int dst_field, int src_field, int sign); /* dst (+|-)= src */
double gso_vcycle(void* g);                                  /* CpuSolver::vcycle */
/* CpuSolver::solve loop; norms[0] = initial, norms[1..] per cycle; returns cycles run */
int    gso_solve(void* g, double* norms, int cap);

void   gso_newton_begin(void* g);                            /* newtonF = level0.f */
double gso_newton_compF(void* g);
int    gso_newton_find_error(void* g);                       /* returns inner V-cycles run */
/* NewtonSolver::solve loop; norms[0] = initial ||F||, norms[1..] per iteration */
int    gso_newton_solve(void* g, double* norms, int cap, int* inner_cycles);

/* sqrt of a compensated long-double sum of squares over the interior of a field:
 * the norm yardstick (the reference's printed norm depends on OMP thread count) */
double gso_norm_exact(void* g, int level, int field);
void   gso_set_threads(int n);                               /* omp_set_num_threads */

#ifdef __cplusplus
}
#endif
#endif
