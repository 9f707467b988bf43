/*
 * gpusolve_cuda.h — C ABI of the B200 (sm_100a) CUDA backend for GpuSolve's
 * geometric-multigrid V-cycle hot path.
 *
 * The reference (Bricktricker/gpu-solve) has no plugin/FFI interface: a backend
 * is "construct a grid object from GridParams, then call solve()" and is picked
 * at compile time in src/main.cpp:5-13,88-111.  This header is the C-ABI a
 * `src/cuda` backend exposes to that driver; every entry point names the
 * reference interface it stands in for (paths relative to the reference root).
 *
 * Conventions
 *   - every function returns GS_OK (0) or a negative GS_E* code; the message of
 *     the last failure on the calling thread is gs_last_error().
 *   - no C++ types, no exceptions cross the boundary; one host thread drives a
 *     handle; the library owns device memory, streams and (multi-GPU) comms.
 *   - a field is addressed by LOGICAL (x,y,z) with 0 <= x < dim[0]+2 etc.; the
 *     device layout (x-fastest, padded pitch) never leaks.  Host buffers are
 *     dense (dim[0]+2)*(dim[1]+2)*(dim[2]+2) doubles in the requested layout.
 */
#ifndef GPUSOLVE_CUDA_H
#define GPUSOLVE_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GS_OK            0
#define GS_EINVAL       -1   /* bad argument / bad config                      */
#define GS_ECUDA        -2   /* CUDA runtime or driver error                   */
#define GS_ENOMEM       -3   /* device or host allocation failed               */
#define GS_EIO          -4   /* config / dump file could not be read/written   */
#define GS_ECOMM        -5   /* NCCL / multi-GPU communication error           */
#define GS_EUNSUPPORTED -6   /* valid request the backend does not implement   */

/* GridParams::Mode, src/gridParams.h:29-33 */
#define GS_MODE_LINEAR    0
#define GS_MODE_NONLINEAR 1
#define GS_MODE_NEWTON    2

/* LevelData members, src/cpu/CpuGridData.h:9-18, plus CpuGridData::newtonF (:36) */
#define GS_FIELD_V        0
#define GS_FIELD_F        1
#define GS_FIELD_R        2
#define GS_FIELD_E        3   /* never materialised by the fused kernels: get only via gs_prolong */
#define GS_FIELD_RESTV    4
#define GS_FIELD_NEWTONV  5
#define GS_FIELD_NEWTONF  6   /* level 0 only */

/* host buffer layouts for gs_get_field / gs_set_field */
#define GS_LAYOUT_X_FASTEST 0 /* idx = x + X*(y + Y*z)   (src/sycl/SyclBuffer.h:16-19) */
#define GS_LAYOUT_Z_FASTEST 1 /* idx = z + Z*(y + Y*x)   (src/cpu/Vector3.cpp:16,24)   */

/* POD mirror of GridParams + Stencil, src/gridParams.h:7-47
 * (std::array / std::tuple are not C-ABI safe).  `h` is filled by
 * gs_parse_config as 1/(dim[1]+1) exactly like src/main.cpp:84; gs_create
 * recomputes the per-level h the way src/cpu/CpuGridData.cpp:41 does. */
typedef struct gs_params {
    uint64_t maxiter;           /* config line 1                               */
    double   tol;               /* line 2  (used: src/cpu/CpuSolver.cpp:39)    */
    double   omega;             /* line 9, relaxation coefficient              */
    double   gamma;             /* line 10, non-linear weight                  */
    double   h;                 /* 1/(dim[1]+1)                                */
    uint64_t dim[3];            /* lines 3-5: INTERIOR points per axis         */
    uint64_t pre, post;         /* lines 7-8: smoothing sweeps                 */
    double   stencil_val[7];    /* line 11; [0] is taken as the centre tap     */
    int32_t  stencil_off[7][3]; /* lines 12-14 transposed: [tap][x,y,z]        */
    int32_t  mode;              /* line 6: GS_MODE_*                           */
    int32_t  print_progress;    /* GridParams::printProgress                   */
} gs_params;

typedef struct gs_handle gs_handle;

/* Result of gs_solve: what the reference only prints (src/cpu/CpuSolver.cpp:17,28;
 * src/cpu/NewtonSolver.cpp:16,27). */
typedef struct gs_solve_info {
    double   initial_residual;  /* "Inital residual" / "Inital newton residual" */
    double   final_residual;    /* norm after the last cycle / Newton iteration */
    uint64_t iterations;        /* V-cycles (modes 0/1) or Newton iterations    */
    uint64_t inner_cycles;      /* mode 2: total inner V-cycles run             */
    int32_t  converged;         /* 1 if the early-exit test fired               */
    int32_t  reserved;
} gs_solve_info;

/* Per-operator device time accumulated since gs_reset_timing (CUDA events on
 * the library's stream; replaces the ms-resolution src/Timer.cpp:7-27). */
#define GS_OP_SMOOTH   0
#define GS_OP_RESIDUAL 1
#define GS_OP_RESTRICT 2   /* residual+restriction (and FAS coarse RHS)        */
#define GS_OP_PROLONG  3   /* prolongation+correction                          */
#define GS_OP_OTHER    4   /* fills, axpys, Newton coefficient precompute      */
#define GS_OP_COUNT    5

/* ---- errors / info ------------------------------------------------------- */
const char* gs_last_error(void);
int  gs_version(void);
/* number of CUDA devices visible, or a negative error */
int  gs_device_count(void);

/* ---- config file: src/main.cpp:33-85 (14 whitespace separated lines,
 * README.md:19-33).  Adds validation the reference lacks (mode, offsets in
 * {-1,0,1}, dims >= 1). */
int  gs_parse_config(const char* path, gs_params* out);

/* ---- lifetime: replaces `CpuGridData grid(params)` (src/cpu/CpuGridData.cpp:15-80)
 * resp. `SyclGridData + initBuffers` (src/sycl/SyclGridData.cpp:14-147):
 * level hierarchy, zero-initialised arrays, analytic right-hand side. */
int  gs_create(const gs_params* p, int device, gs_handle** out);
/* Multi-GPU (new; the reference is single-device): one process per GPU, this
 * rank owns a z-slab of every fine level.  `nccl_id` is the 128-byte
 * ncclUniqueId produced by gs_comm_unique_id on rank 0 and broadcast by the
 * host (torch.distributed / MPI / a file).  world == 1 is gs_create. */
int  gs_comm_unique_id(void* buf128);
int  gs_create_slab(const gs_params* p, int device, int rank, int world,
                    const void* nccl_id, gs_handle** out);
int  gs_destroy(gs_handle* h);

/* slab decomposition used by gs_create_slab, pure host arithmetic (no GPU):
 * array planes [*z_begin, *z_end) of `level` owned by `rank`; levels at or
 * below *replicated_from are held whole by every rank. */
int  gs_partition(const gs_params* p, int world, int rank, int level,
                  uint64_t* z_begin, uint64_t* z_end, int* replicated_from);

/* ---- introspection: CpuGridData::numLevels / getLevel (src/cpu/CpuGridData.h:22-33) */
int  gs_num_levels(const gs_handle* h);
int  gs_get_level_info(const gs_handle* h, int level, uint64_t dim[3], double* hval);

/* ---- field access: Vector3::get/set (src/cpu/Vector3.cpp:14-27) for whole
 * arrays.  Under gs_create_slab every rank returns/accepts the whole logical
 * array (gathered / scattered internally). */
int  gs_get_field(gs_handle* h, int level, int field, double* host, int layout);
int  gs_set_field(gs_handle* h, int level, int field, const double* host, int layout);
/* Vector3::dump text format (src/cpu/Vector3.cpp:56-78), readable by plotter.py:10-26 */
int  gs_dump_field(gs_handle* h, int level, int field, const char* path);

/* ---- level operators (private statics of CpuSolver, src/cpu/CpuSolver.h:10-15) */
/* CpuSolver::jacobi(grid, level, iters)            src/cpu/CpuSolver.cpp:141-179 */
int  gs_smooth(gs_handle* h, int level, uint64_t iters);
/* CpuSolver::compResidual(grid, level): r stored, returns sqrt(sum r^2)  :45-83 */
int  gs_residual(gs_handle* h, int level, double* norm);
/* gs_residual for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same inputs and outputs as
 * gs_residual (r = f - A(v) stored on the level, *norm = sqrt(sum r^2)), but the call path allocates nothing: the scratch
 * of the norm reduction is a static __device__ array inside the library (512 KB, held by every GPU that loads the library
 * until the process exits) instead of memory from cudaMalloc.  Limits: the level may use at most 65536 reduction slots,
 * ceil(nx/64) * ceil(ny/4) * nz <= 65536 (e.g. up to 255^3; larger levels return GS_EINVAL with a message), and must not
 * be split into slabs (GS_EUNSUPPORTED).  All calls on a device share the one array, which keeps the last call's contents
 * (the entry point clears what must start at zero): calls that overlap on the GPU — two handles on two streams of one
 * device — race and must be serialised by the caller. */
int  gs_residual_static(gs_handle* h, int level, double* norm);
/* CpuSolver::restrict(level.src_field -> (level+1).dst_field)          :211-238 */
int  gs_restrict(gs_handle* h, int level, int src_field, int dst_field);
/* fused compResidual(level) ; restrict(r -> f_{level+1})            :93-98      */
int  gs_restrict_residual(gs_handle* h, int level);
/* FAS coarse right-hand side, mode 1 only: restV,v_{l+1} = R(v_l);
 * f_{l+1} += A_{l+1}(restV_{l+1})                       :106-112, :182-209      */
int  gs_fas_coarse_rhs(gs_handle* h, int level);
/* CpuSolver::interpolate(grid, level) materialised into host memory (dense,
 * layout as given): e_level = P v_{level+1}                             :240-290 */
int  gs_prolong(gs_handle* h, int level, double* host_e, int layout);
/* fused (mode 1: v_{l+1} -= restV_{l+1};) interpolate(level); v_level += e :121-132 */
int  gs_prolong_correct(gs_handle* h, int level);
/* CpuSolver::vcycle(grid): one V(pre,post) cycle, returns final ||r_0||  :85-139 */
int  gs_vcycle(gs_handle* h, double* norm);

/* ---- Newton (mode 2), src/cpu/NewtonSolver.cpp */
/* NewtonSolver::compF: f_0 = newtonF - N(newtonV), returns ||f_0||       :48-81  */
int  gs_newton_compF(gs_handle* h, double* norm);
/* one iteration of the loop in NewtonSolver::solve (:21-26): compF, v_0 = 0,
 * findError (:83-108), compF; *norm = the second compF.  The first call also
 * does the `newtonF = f_0` snapshot of :12 if gs_solve has not. */
int  gs_newton_step(gs_handle* h, double* norm, uint64_t* inner_cycles);

/* ---- whole solve: CpuSolver::solve (:12-43) for modes 0/1, NewtonSolver::solve
 * (src/cpu/NewtonSolver.cpp:10-44) for mode 2.  Prints the reference's stdout
 * lines when params.print_progress != 0.  `info` may be NULL. */
int  gs_solve(gs_handle* h, gs_solve_info* info);
/* residual history of the last gs_solve: [0] initial, [1..] per cycle */
int  gs_get_history(const gs_handle* h, double* norms, int cap, int* count);

/* ---- timing / tuning */
int  gs_reset_timing(gs_handle* h);
int  gs_get_timing(gs_handle* h, double ms[GS_OP_COUNT], uint64_t launches[GS_OP_COUNT]);
/* the same, for one level (the level an operator is called ON: restriction from l to l+1 counts on l) */
int  gs_get_level_timing(gs_handle* h, int level, double ms[GS_OP_COUNT], uint64_t launches[GS_OP_COUNT]);
/* Jacobi sweeps applied to `level` by its GS_OP_SMOOTH launches since gs_reset_timing (a two-sweeps-per-pass
 * launch of kernels_fused.cu counts 2; sweeps produced together with a residual norm count under GS_OP_RESIDUAL) */
int  gs_get_level_sweeps(gs_handle* h, int level, uint64_t* sweeps);
/* total kernels launched by this handle since creation */
uint64_t gs_launch_count(const gs_handle* h);
/* options a caller of the backend may want (all also settable as GPUSOLVE_<NAME> in the environment):
 *          "kernels" = 0 general one-thread-per-point kernels, 1 tuned sm_100a kernels (default);
 *          "graph"   = 0/1 capture the V-cycle in a CUDA graph (default 1);
 *          "timing"  = 0/1 per-operator CUDA events (default 0; disables the graph);
 *          "timers"  = 0/1 partial-timer tail in the reference Timer's format after each cycle of a solve.
 * Kernel tuning knobs and A/B switches go through the same call but are not part of the backend contract:
 * they are listed in gpusolve_cuda_tuning.h. */
int  gs_set_option(gs_handle* h, const char* name, int64_t value);
/* the CUDA stream all work of this handle is issued on (a cudaStream_t) */
void* gs_stream(gs_handle* h);
/* block the host until all queued work of the handle has finished */
int  gs_synchronize(gs_handle* h);

#ifdef __cplusplus
}
#endif
#endif /* GPUSOLVE_CUDA_H */
