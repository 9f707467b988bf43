// kernels_comm.cu — ghost-plane exchange of the z-slab decomposition through PEER-MAPPED memory (new: the reference is
// single-device, src/sycl/ContextHandles.h:38-44).  One process per GPU; every rank's device arena is mapped into every
// other rank's address space with CUDA IPC (gs_comm.cpp), so a halo refresh is a copy kernel whose stores travel over
// NVLink 5 / NVSwitch straight into the neighbour's ghost planes — no NCCL collective, no staging, capturable in the
// V-cycle's CUDA graph like any other kernel.
//
// Protocol (per ordered pair of ranks; all ranks issue the same sequence of exchanges, SPMD):
//   every rank keeps, for each peer, a control word triple in its own arena: `epoch` (exchanges it has started with that
//   peer), `ack_in` and `sig_in` (written remotely by the peer).  Exchange number e = epoch + 1:
//     p2p_push   1. tell the peer "all my work queued before this exchange is complete" (peer.ack_in = e): whatever it
//                   read from its ghost planes in my name is no longer needed by me... precisely: the peer may now
//                   overwrite MY ghost planes, because every kernel of mine that read them precedes this one in stream order;
//                2. wait for the peer's same message (ack_in >= e): now I may overwrite ITS ghost planes;
//                3. copy my boundary planes into its ghost planes (16-byte stores to the peer mapping);
//                4. fence, and the last CTA to finish publishes sig_in = e at the peer and epoch = e here.
//                5. that same CTA then spins until the peer's data of exchange e has landed here (sig_in >= e): when the
//                   kernel ends, the consumer queued behind it may read the ghost planes.  One launch per exchange.
//   Both sides send their ack before waiting for the other's, so the handshake cannot deadlock; a spin that exceeds
//   ~4 s sets an error word instead of hanging the GPU (the host checks it at the next synchronisation).
// The same kernel carries the gather of the first replicated coarse level (every rank pushes the coarse planes it
// computed to all peers) and the residual-norm reduction (one double per rank into every peer's slot array, summed in
// rank order by the last CTA of the push: deterministic and bit-identical on all ranks).
#include "gs_comm.h"
#include "gs_device.cuh"

namespace {

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

constexpr long long SPIN_LIMIT = 1ll << 33;  // clock64 ticks (~4.4 s at 1.9 GHz)

// returns false on timeout
__device__ __forceinline__ bool spin_until(const unsigned long long* flag, unsigned long long want, unsigned long long* err)
{
    const long long t0 = clock64();
    while (ld_acquire_sys(flag) < want) {
        if (clock64() - t0 > SPIN_LIMIT) {
            *err = 1ull;
            return false;
        }
        __nanosleep(64);
    }
    return true;
}

__global__ void __launch_bounds__(256) p2p_push(P2PPush q)
{
    __shared__ unsigned long long e_sh[GS_MAX_PEERS];
    __shared__ int last;
    const int tid = threadIdx.x;
    if (tid < q.npeer) {
        const P2PPeerOp& o = q.p[tid];
        const unsigned long long e = q.ctl[o.rank].epoch + 1;
        e_sh[tid] = e;
        if (blockIdx.x == 0) st_release_sys(&o.remote->ack_in, e);
        spin_until(&q.ctl[o.rank].ack_in, e, q.err);
    }
    __syncthreads();
    const long long gtid = (long long)blockIdx.x * blockDim.x + tid, gsz = (long long)gridDim.x * blockDim.x;
    for (int s = 0; s < q.npeer; s++) {
        const P2PPeerOp& o = q.p[s];
        for (int k = 0; k < 2; k++) {
            const double* __restrict__ src = o.src[k];
            double* __restrict__ dst = o.dst[k];
            const long long n = o.count[k];
            if (n <= 0) continue;
            if (((n | (long long)(((size_t)src | (size_t)dst) >> 3)) & 1) == 0) {  // even count, 16-byte aligned ends
                // eight independent 16-byte loads in flight per thread before the eight remote stores: the copy is bound by
                // load latency x bytes in flight (~10 MB over the grid), and the NVLink stores are fire-and-forget
                const double2* s2 = reinterpret_cast<const double2*>(src);
                double2* d2 = reinterpret_cast<double2*>(dst);
                const long long n2 = n / 2;
                long long i = gtid;
                for (; i + 7 * gsz < n2; i += 8 * gsz) {
                    double2 v[8];
#pragma unroll
                    for (int u = 0; u < 8; u++) v[u] = __ldcs(s2 + i + u * gsz);
#pragma unroll
                    for (int u = 0; u < 8; u++) d2[i + u * gsz] = v[u];
                }
                for (; i < n2; i += gsz) d2[i] = __ldcs(s2 + i);
            } else {
                for (long long i = gtid; i < n; i += gsz) dst[i] = src[i];
            }
        }
    }
    __threadfence_system();
    __syncthreads();
    if (tid == 0) {
        const unsigned t = atomicAdd(q.ticket, 1u);
        last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!last) return;
    __threadfence_system();
    if (tid < q.npeer) {
        const P2PPeerOp& o = q.p[tid];
        st_release_sys(&o.remote->sig_in, e_sh[tid]);
        q.ctl[o.rank].epoch = e_sh[tid];
        // push + arrival in one launch: the last CTA stays until the peers' data of this exchange has landed here
        spin_until(&q.ctl[o.rank].sig_in, e_sh[tid], q.err);
    }
    if (q.norm_result) {
        // every rank adds the partial sums in rank order: deterministic and bit-identical on all ranks
        __syncthreads();
        if (tid == 0) {
            double s = 0.0;
            for (int r = 0; r < q.world; r++) s += (r == q.me) ? *q.norm_result : __ldcg(q.norm_slots + r);
            *q.norm_result = s;
        }
    }
    if (tid == 0) *q.ticket = 0u;
}

}  // namespace

cudaError_t kc_p2p_push(const P2PPush& q, long long total_doubles, int sm_count, cudaStream_t s)
{
    // enough CTAs to fill the NVLink pipes (two per SM at most); small exchanges take few CTAs so the handshake stays cheap
    long long want = (total_doubles / 2 + 256 * 8 - 1) / (256 * 8);
    int grid = (int)(want < 1 ? 1 : (want > 2 * sm_count ? 2 * sm_count : want));
    p2p_push<<<grid, 256, 0, s>>>(q);
    return cudaGetLastError();
}
