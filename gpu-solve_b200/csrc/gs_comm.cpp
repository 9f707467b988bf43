// gs_comm.cpp — see gs_comm.h
#include "gs_comm.h"

#include <dlfcn.h>
#include <nccl.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "gs_internal.h"

namespace {

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string err;
};

NcclApi& api()
{
    static NcclApi a;
    if (a.lib || !a.err.empty()) return a;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        a.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (a.lib) break;
    }
    if (!a.lib) {
        a.err = std::string("cannot load NCCL: ") + dlerror();
        return a;
    }
#define BIND(field, sym)                                                     \
    do {                                                                     \
        *(void**)(&a.field) = dlsym(a.lib, sym);                             \
        if (!a.field) a.err = std::string("NCCL symbol missing: ") + sym;    \
    } while (0)
    BIND(GetUniqueId, "ncclGetUniqueId");
    BIND(CommInitRank, "ncclCommInitRank");
    BIND(CommDestroy, "ncclCommDestroy");
    BIND(Send, "ncclSend");
    BIND(Recv, "ncclRecv");
    BIND(GroupStart, "ncclGroupStart");
    BIND(GroupEnd, "ncclGroupEnd");
    BIND(AllReduce, "ncclAllReduce");
    BIND(Broadcast, "ncclBroadcast");
    BIND(GetErrorString, "ncclGetErrorString");
#undef BIND
    return a;
}

}  // namespace

struct GsComm {
    int rank, world, nlev, replicated_from;
    std::vector<int> bounds;  // nlev x (world+1)
    ncclComm_t comm = nullptr;
    std::string err;
    double* staging = nullptr;
    size_t staging_count = 0;
    int b(int level, int r) const { return bounds[(size_t)level * (world + 1) + r]; }
};

#define NC(c, expr)                                                                     \
    do {                                                                                \
        ncclResult_t r__ = (expr);                                                      \
        if (r__ != ncclSuccess) {                                                       \
            (c)->err = std::string(#expr) + ": " + api().GetErrorString(r__);           \
            return -1;                                                                  \
        }                                                                               \
    } while (0)

const char* gs_comm_make_id(void* buf128)
{
    NcclApi& a = api();
    if (!a.err.empty()) return a.err.c_str();
    ncclUniqueId id;
    ncclResult_t r = a.GetUniqueId(&id);
    if (r != ncclSuccess) {
        a.err = std::string("ncclGetUniqueId: ") + a.GetErrorString(r);
        return a.err.c_str();
    }
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    memcpy(buf128, &id, 128);
    return nullptr;
}

GsComm* gs_comm_create(int rank, int world, const void* id128, int nlev, const int* bounds, int replicated_from)
{
    GsComm* c = new GsComm();
    c->rank = rank;
    c->world = world;
    c->nlev = nlev;
    c->replicated_from = replicated_from;
    c->bounds.assign(bounds, bounds + (size_t)nlev * (world + 1));
    NcclApi& a = api();
    if (!a.err.empty()) {
        c->err = a.err;
        return c;
    }
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    ncclResult_t r = a.CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) c->err = std::string("ncclCommInitRank: ") + a.GetErrorString(r);
    return c;
}

void gs_comm_destroy(GsComm* c)
{
    if (!c) return;
    if (c->comm) api().CommDestroy(c->comm);
    if (c->staging) cudaFree(c->staging);
    delete c;
}

const char* gs_comm_error(GsComm* c) { return c ? c->err.c_str() : "no communicator"; }

int gs_comm_halo(GsComm* c, double* a, long long plane, int lo, int hi, int depth, int zlo, int zhi, int Zg, cudaStream_t s)
{
    NcclApi& n = api();
    const size_t cnt = (size_t)depth * (size_t)plane;
    const bool up = c->rank + 1 < c->world && zhi < Zg;  // a neighbour above
    const bool dn = c->rank > 0 && zlo > 0;              // a neighbour below
    if (hi - lo < depth) {
        c->err = "slab thinner than the halo depth";
        return -1;
    }
    NC(c, n.GroupStart());
    if (up) {
        NC(c, n.Send(a + (size_t)(hi - depth) * plane, cnt, ncclDouble, c->rank + 1, c->comm, s));
        NC(c, n.Recv(a + (size_t)hi * plane, cnt, ncclDouble, c->rank + 1, c->comm, s));
    }
    if (dn) {
        NC(c, n.Send(a + (size_t)lo * plane, cnt, ncclDouble, c->rank - 1, c->comm, s));
        NC(c, n.Recv(a + (size_t)(lo - depth) * plane, cnt, ncclDouble, c->rank - 1, c->comm, s));
    }
    NC(c, n.GroupEnd());
    return 0;
}

int gs_comm_allreduce_sum(GsComm* c, double* dev, int cnt, cudaStream_t s)
{
    NC(c, api().AllReduce(dev, dev, (size_t)cnt, ncclDouble, ncclSum, c->comm, s));
    return 0;
}

int gs_comm_gather_planes(GsComm* c, double* coarse, long long plane, int fine_level, cudaStream_t s)
{
    NcclApi& n = api();
    const int Zc = c->b(fine_level + 1, c->world);
    NC(c, n.GroupStart());
    for (int r = 0; r < c->world; r++) {
        // coarse planes whose centre fine plane 2*Zc lies in rank r's fine slab, interior only
        int zb = (c->b(fine_level, r) + 1) / 2, ze = (c->b(fine_level, r + 1) + 1) / 2;
        if (zb < 1) zb = 1;
        if (ze > Zc - 1) ze = Zc - 1;
        if (ze <= zb) continue;
        double* p = coarse + (size_t)zb * plane;
        NC(c, n.Broadcast(p, p, (size_t)(ze - zb) * (size_t)plane, ncclDouble, r, c->comm, s));
    }
    NC(c, n.GroupEnd());
    return 0;
}

int gs_comm_gather_to_host(GsComm* c, const double* slab, long long plane, long long pitch, int z0, int level, double* dense,
                           size_t X, size_t Y, size_t Z, cudaStream_t s)
{
    NcclApi& n = api();
    const size_t need = (size_t)plane * Z;
    if (c->staging_count < need) {
        if (c->staging) cudaFree(c->staging);
        c->staging = nullptr;
        c->staging_count = 0;
        if (cudaMalloc((void**)&c->staging, need * sizeof(double)) != cudaSuccess) {
            c->err = "staging allocation failed";
            return -1;
        }
        c->staging_count = need;
    }
    const int lo = c->b(level, c->rank), hi = c->b(level, c->rank + 1);
    if (hi > lo &&
        cudaMemcpyAsync(c->staging + (size_t)lo * plane, slab + (size_t)(lo - z0) * plane, (size_t)(hi - lo) * plane * sizeof(double),
                        cudaMemcpyDeviceToDevice, s) != cudaSuccess) {
        c->err = "staging copy failed";
        return -1;
    }
    NC(c, n.GroupStart());
    for (int r = 0; r < c->world; r++) {
        const int a = c->b(level, r), b = c->b(level, r + 1);
        if (b <= a) continue;
        double* p = c->staging + (size_t)a * plane;
        NC(c, n.Broadcast(p, p, (size_t)(b - a) * (size_t)plane, ncclDouble, r, c->comm, s));
    }
    NC(c, n.GroupEnd());
    if (cudaStreamSynchronize(s) != cudaSuccess ||
        cudaMemcpy2D(dense, X * sizeof(double), c->staging + GS_XOFF, pitch * sizeof(double), X * sizeof(double), Y * Z,
                     cudaMemcpyDeviceToHost) != cudaSuccess) {
        c->err = "gather copy to host failed";
        return -1;
    }
    return 0;
}
