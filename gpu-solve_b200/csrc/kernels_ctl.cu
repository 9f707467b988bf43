// kernels_ctl.cu — device-side stopping test of the solve loops.
//
// CpuSolver::solve (src/cpu/CpuSolver.cpp:12-43) and the inner solve of NewtonSolver::findError
// (src/cpu/NewtonSolver.cpp:94-100: at most 10 cycles, tol 0.1) read the residual norm on the host after every
// V-cycle and stop once `res <= initialResidual / (1.0 / tol)`.  Here the loop is a CUDA conditional graph node
// (WHILE): its body is the captured V-cycle followed by stop_check, a one-thread kernel that evaluates the same
// IEEE expression on the device (fp64 sqrt and divide are correctly rounded, the library is built with
// -fmad=false) and sets the loop condition with cudaGraphSetConditional.  The host sees one graph launch and one
// copy of the residual history per solve; the number of cycles run equals the reference's by construction.
#include "gs_device.cuh"

namespace {

__global__ void stop_init_kernel(StopCtl* c, const double* sumsq, cudaGraphConditionalHandle cond, unsigned maxiter)
{
    const double r = sqrt(*sumsq);  // "Inital residual", CpuSolver.cpp:16
    c->initial = r;
    c->count = 0;
    c->hist[0] = r;
    cudaGraphSetConditional(cond, maxiter > 0 ? 1u : 0u);
}

__global__ void stop_check_kernel(StopCtl* c, const double* sumsq, cudaGraphConditionalHandle cond, unsigned maxiter, double tol)
{
    const double res = sqrt(*sumsq);  // CpuSolver::vcycle returns compResidual's sqrt, :138 / :82
    const unsigned n = c->count + 1;
    c->count = n;
    c->hist[n] = res;
    const bool reached = res <= c->initial / (1.0 / tol);  // :39, spelled the same way
    cudaGraphSetConditional(cond, (!reached && n < maxiter) ? 1u : 0u);
}

}  // namespace

cudaError_t kc_stop_init(const LaunchCtx& cx, StopCtl* c, unsigned long long cond, unsigned maxiter)
{
    stop_init_kernel<<<1, 1, 0, cx.stream>>>(c, cx.norm.result, (cudaGraphConditionalHandle)cond, maxiter);
    if (cx.launches) ++*cx.launches;
    return cudaGetLastError();
}

cudaError_t kc_stop_check(const LaunchCtx& cx, StopCtl* c, unsigned long long cond, unsigned maxiter, double tol)
{
    stop_check_kernel<<<1, 1, 0, cx.stream>>>(c, cx.norm.result, (cudaGraphConditionalHandle)cond, maxiter, tol);
    if (cx.launches) ++*cx.launches;
    return cudaGetLastError();
}
