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
