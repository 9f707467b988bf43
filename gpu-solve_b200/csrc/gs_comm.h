// gs_comm.h — multi-GPU plumbing of the slab decomposition (new: the reference is
// single-device).  One process per GPU; NCCL over NVLink 5 / NVSwitch carries the
// one- or two-plane halo exchanges, the norm all-reduce and the gather of the first
// replicated coarse level.  NCCL is bound at run time (dlopen "libnccl.so.2") so a
// single-GPU build has no dependency on it and a torch process shares torch's copy.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

struct GsComm;

// writes a 128-byte ncclUniqueId; returns nullptr or an error message
const char* gs_comm_make_id(void* buf128);
// bounds: nlev x (world+1) slab boundaries (global planes), row-major
GsComm* gs_comm_create(int rank, int world, const void* id128, int nlev, const int* bounds, int replicated_from);
void gs_comm_destroy(GsComm* c);
const char* gs_comm_error(GsComm* c);  // "" when healthy

// refresh `depth` ghost planes above and below the owned LOCAL range [lo, hi) of a slab array
int gs_comm_halo(GsComm* c, double* a, long long plane, int lo, int hi, int depth, int zlo, int zhi, int Zg, cudaStream_t s);
int gs_comm_allreduce_sum(GsComm* c, double* dev, int n, cudaStream_t s);
// after restricting partitioned `fine_level` onto a replicated coarse array: every rank
// broadcasts the coarse planes it computed so that all ranks hold the whole array
int gs_comm_gather_planes(GsComm* c, double* coarse, long long plane, int fine_level, cudaStream_t s);
// whole logical array of a partitioned level into dense host memory on every rank
int gs_comm_gather_to_host(GsComm* c, const double* slab, long long plane, long long pitch, int z0, int level, double* dense,
                           size_t X, size_t Y, size_t Z, cudaStream_t s);

// ---- peer-mapped transport (kernels_comm.cu): ghost planes, coarse gather and the norm reduction without NCCL ----
constexpr int GS_MAX_PEERS = 15;  // world <= 16

// control words of one (this rank, peer) pair; they live in this rank's arena, the peer writes ack_in / sig_in remotely
struct P2PCtl {
    unsigned long long sig_in;  // exchanges of the peer whose data has landed here
    unsigned long long ack_in;  // exchanges the peer has begun (everything it queued before them is complete)
    unsigned long long epoch;   // exchanges this rank has started with the peer
    unsigned long long pad;
};
// the control block at the start of every rank's arena
struct P2PBlock {
    P2PCtl ctl[GS_MAX_PEERS + 1];      // indexed by peer rank
    double norm_slots[GS_MAX_PEERS + 1];  // partial sums of squares, written by rank r into slot r of every peer
    unsigned long long err;             // set by a spin that timed out
    unsigned ticket, pad;
};
struct P2PPeerOp {
    int rank;              // the peer
    P2PCtl* remote;        // the peer's control words for THIS rank (peer mapping)
    const double* src[2];  // up to two segments: local source ...
    double* dst[2];        // ... and destination in the peer's arena
    long long count[2];    // doubles
};
struct P2PPush {
    int npeer;
    P2PCtl* ctl;           // this rank's control words, indexed by peer rank
    unsigned long long* err;
    unsigned* ticket;
    // norm reduction: after the arrival, *norm_result = sum over ranks in rank order (own partial = *norm_result)
    const double* norm_slots;
    double* norm_result;
    int world, me;
    P2PPeerOp p[GS_MAX_PEERS];
};
cudaError_t kc_p2p_push(const P2PPush& q, long long total_doubles, int sm_count, cudaStream_t s);

// Map every rank's arena into this process (CUDA IPC; the handles and the per-rank array offsets travel by ncclAllGather).
// `offs[k]` = byte offset inside the arena of registered array k (the same k names the same logical array on every rank);
// the arena must start with a zeroed P2PBlock.  Returns 0 and enables the peer path, or -1 (peer access unavailable: the
// NCCL path stays in use; gs_comm_error tells why).
int gs_comm_enable_p2p(GsComm* c, void* arena, size_t arena_bytes, const long long* offs, int noff, int sm_count, cudaStream_t s);
bool gs_comm_p2p(const GsComm* c);
// the peer-path twins of gs_comm_halo / gs_comm_gather_planes / gs_comm_allreduce_sum.  `k` = registered index of the array
// `a` points into; z0_dn / z0_up = global plane of LOCAL plane 0 of the neighbour below / above on this level.
// Every call is ONE kernel: handshake, push of the boundary planes into the peers' ghost planes, and the wait for the peers'
// data of the same exchange.
int gs_comm_halo_p2p(GsComm* c, int k, double* a, long long plane, int lo, int hi, int depth, int zlo, int zhi, int Zg, int z0_me,
                     int z0_dn, int z0_up, cudaStream_t s);
int gs_comm_gather_planes_p2p(GsComm* c, int k, double* coarse, long long plane, int fine_level, cudaStream_t s);
int gs_comm_norm_p2p(GsComm* c, double* dev_result, cudaStream_t s);
// nonzero once a device-side wait has timed out (checked after synchronising)
int gs_comm_p2p_failed(GsComm* c);
// collective: every rank unmaps its peers' arenas (call before freeing the own arena)
void gs_comm_disable_p2p(GsComm* c, cudaStream_t s);
