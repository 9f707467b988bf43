/*
 * gpusolve_cuda_tuning.h — introspection and tuning hooks of the CUDA backend used by its own
 * tests, bench.py and tools/.  NOT part of the backend contract a GpuSolve driver binds
 * (that is gpusolve_cuda.h, which mirrors the reference's construct-then-solve interface);
 * nothing here has a counterpart in the reference.
 */
#ifndef GPUSOLVE_CUDA_TUNING_H
#define GPUSOLVE_CUDA_TUNING_H

#include "gpusolve_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

/* z-chunk length the z-marching kernels use for `planes` planes of a level cut into `tiles` xy tiles, pure host
 * arithmetic (no GPU): the chunk count that minimises (waves of sm_count * ctas_per_sm resident CTAs) x (chunk +
 * overhead planes), so grids fill whole waves.  Introspection for tests and tuning; returns the chunk (>= 1). */
int  gs_plan_zchunk(int tiles, int planes, int sm_count, int ctas_per_sm, int overhead,
                    int min_chunk, int max_chunk, int even);

/* slabs: time `reps` back-to-back refreshes of `depth` (1 or 2) ghost planes of the level's v (CUDA events on the handle's
 * stream, after one warm-up); collective, the fields do not change.  *ms_per_exchange = time / reps. */
int  gs_time_halo_exchange(gs_handle* h, int level, int depth, int reps, double* ms_per_exchange);

/* gs_set_option names beyond the four documented in gpusolve_cuda.h (every value gives the same bits; defaults in
 * gpu-solve_b200/csrc/gs_api.cu create_impl):
 *   "smooth_variant", "zchunk", "fast_min_n"   single-sweep kernels (kernels_fast.cu)
 *   "fuse2" = 0 one Jacobi sweep per pass, 1 two sweeps per pass (kernels_fused.cu), "fuse2_min_n", "zchunk2"
 *   "fuse_norm" = 0/1 produce the next pre-smoothing sweeps together with the V-cycle's final residual norm
 *   "unit_taps", "rcp_div", "wave", "vzero" = 0/1 arithmetic / scheduling shortcuts (A/B switches, default 1)
 *   "fuse_pc" = 0/1/2 prolongation + correction folded into the first post-smoothing pass (default 1: linear mode, levels of >= "pc_min_n" = 384 points per axis; 2: every mode and level)
 *   "device_stop" = 0/1 inner Newton solves as one conditional CUDA graph with the stopping test on the device
 *   "fuse_rv" = 0/1 FAS cycles: R(v) -> restV, v formed by the fused residual+restriction pass (default 1; equal within rounding)
 *   "pdl" = 0/1 small-level kernels launched with programmatic stream serialization (default 1)
 *   "s2_ty" = 0/10/12 tile height of the two-sweep kernel (0: per mode)
 *   "p2p" = 0/1 slabs: ghost planes pushed through peer-mapped memory instead of NCCL send/recv */

#ifdef __cplusplus
}
#endif
#endif /* GPUSOLVE_CUDA_TUNING_H */
