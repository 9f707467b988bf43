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
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
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
    BIND(AllGather, "ncclAllGather");
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
    // peer-mapped transport
    bool p2p = false;
    int sm_count = 148;
    P2PBlock* block = nullptr;               // this rank's control block (start of its arena)
    std::vector<char*> peer_base;            // [world] peer arenas mapped here (own entry: the local arena)
    std::vector<std::vector<long long>> peer_offs;  // [world][noff]
    unsigned long long* h_err = nullptr;     // pinned, mapped: a device-side wait that timed out writes 1 here
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

// ---------------------------------------------------------------------------------
// peer-mapped transport (CUDA IPC + kernels_comm.cu)
// ---------------------------------------------------------------------------------
namespace {

// true on every rank iff `ok` is true on every rank
bool all_ranks_ok(GsComm* c, bool ok, cudaStream_t s)
{
    int* d = nullptr;
    int v = ok ? 1 : 0;
    if (cudaMalloc((void**)&d, sizeof(int)) != cudaSuccess) return false;
    bool good = cudaMemcpyAsync(d, &v, sizeof(int), cudaMemcpyHostToDevice, s) == cudaSuccess &&
                api().AllReduce(d, d, 1, ncclInt, ncclMin, c->comm, s) == ncclSuccess &&
                cudaMemcpyAsync(&v, d, sizeof(int), cudaMemcpyDeviceToHost, s) == cudaSuccess && cudaStreamSynchronize(s) == cudaSuccess;
    cudaFree(d);
    return good && v == 1;
}

}  // namespace

bool gs_comm_p2p(const GsComm* c) { return c && c->p2p; }

int gs_comm_enable_p2p(GsComm* c, void* arena, size_t arena_bytes, const long long* offs, int noff, int sm_count, cudaStream_t s)
{
    NcclApi& n = api();
    if (c->world > GS_MAX_PEERS + 1) {
        c->err = "peer transport supports at most 16 ranks";
        return -1;
    }
    // payload per rank: IPC handle (64 B), arena bytes, array count, the array offsets
    const size_t per = sizeof(cudaIpcMemHandle_t) + 16 + (size_t)noff * 8;
    std::vector<char> mine(per, 0), all(per * c->world, 0);
    cudaIpcMemHandle_t mh;
    bool ok = cudaIpcGetMemHandle(&mh, arena) == cudaSuccess;
    if (!ok) cudaGetLastError();
    memcpy(mine.data(), &mh, sizeof mh);
    const long long ab = (long long)arena_bytes, no = noff;
    memcpy(mine.data() + sizeof mh, &ab, 8);
    memcpy(mine.data() + sizeof mh + 8, &no, 8);
    memcpy(mine.data() + sizeof mh + 16, offs, (size_t)noff * 8);
    char *dsend = nullptr, *drecv = nullptr;
    if (cudaMalloc((void**)&dsend, per) != cudaSuccess || cudaMalloc((void**)&drecv, per * c->world) != cudaSuccess) {
        if (dsend) cudaFree(dsend);
        c->err = "peer transport: staging allocation failed";
        return -1;
    }
    bool xch = cudaMemcpyAsync(dsend, mine.data(), per, cudaMemcpyHostToDevice, s) == cudaSuccess &&
               n.AllGather(dsend, drecv, per, ncclChar, c->comm, s) == ncclSuccess &&
               cudaMemcpyAsync(all.data(), drecv, per * c->world, cudaMemcpyDeviceToHost, s) == cudaSuccess &&
               cudaStreamSynchronize(s) == cudaSuccess;
    cudaFree(dsend);
    cudaFree(drecv);
    if (!xch) {
        c->err = "peer transport: handle exchange failed";
        return -1;  // a failed collective is not recoverable rank by rank; the caller treats it as a setup error
    }
    c->peer_base.assign(c->world, nullptr);
    c->peer_offs.assign(c->world, std::vector<long long>());
    std::string why;
    for (int r = 0; r < c->world && ok; r++) {
        const char* rec = all.data() + (size_t)r * per;
        long long rno;
        memcpy(&rno, rec + sizeof mh + 8, 8);
        if (rno != noff) {
            ok = false;
            why = "ranks disagree about the number of arrays";
            break;
        }
        c->peer_offs[r].resize(noff);
        memcpy(c->peer_offs[r].data(), rec + sizeof mh + 16, (size_t)noff * 8);
        if (r == c->rank) {
            c->peer_base[r] = (char*)arena;
            continue;
        }
        cudaIpcMemHandle_t ph;
        memcpy(&ph, rec, sizeof ph);
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, ph, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            ok = false;
            why = std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e);
            break;
        }
        c->peer_base[r] = (char*)p;
    }
    if (ok && !c->h_err) {
        ok = cudaHostAlloc((void**)&c->h_err, sizeof(unsigned long long), cudaHostAllocMapped) == cudaSuccess;
        if (ok) *c->h_err = 0;
        else why = "pinned error word";
    }
    // every rank takes the same path: peer transport only if it works everywhere
    if (!all_ranks_ok(c, ok, s)) {
        for (int r = 0; r < c->world; r++)
            if (r != c->rank && r < (int)c->peer_base.size() && c->peer_base[r]) cudaIpcCloseMemHandle(c->peer_base[r]);
        c->peer_base.clear();
        c->err = "peer transport unavailable" + (why.empty() ? std::string(" on another rank") : ": " + why);
        return -1;
    }
    c->block = (P2PBlock*)arena;
    c->sm_count = sm_count;
    c->p2p = true;
    return 0;
}

void gs_comm_disable_p2p(GsComm* c, cudaStream_t s)
{
    if (!c || !c->p2p) return;
    cudaStreamSynchronize(s);
    all_ranks_ok(c, true, s);  // nobody is still writing into a peer
    for (int r = 0; r < c->world; r++)
        if (r != c->rank && c->peer_base[r]) cudaIpcCloseMemHandle(c->peer_base[r]);
    all_ranks_ok(c, true, s);  // everybody has unmapped: the arenas may be freed
    c->p2p = false;
    if (c->h_err) cudaFreeHost(c->h_err);
    c->h_err = nullptr;
}

int gs_comm_p2p_failed(GsComm* c) { return c && c->h_err && *c->h_err != 0; }

namespace {

double* peer_array(GsComm* c, int r, int k) { return (double*)(c->peer_base[r] + c->peer_offs[r][k]); }
P2PCtl* peer_ctl_for_me(GsComm* c, int r) { return &((P2PBlock*)c->peer_base[r])->ctl[c->rank]; }

void push_header(GsComm* c, P2PPush* q)
{
    memset(q, 0, sizeof *q);
    q->ctl = c->block->ctl;
    q->err = c->h_err;
    q->ticket = &c->block->ticket;
}

}  // namespace

int gs_comm_halo_p2p(GsComm* c, int k, double* a, long long plane, int lo, int hi, int depth, int zlo, int zhi, int Zg, int z0_me,
                     int z0_dn, int z0_up, cudaStream_t s)
{
    const bool up = c->rank + 1 < c->world && zhi < Zg;
    const bool dn = c->rank > 0 && zlo > 0;
    if (hi - lo < depth) {
        c->err = "slab thinner than the halo depth";
        return -1;
    }
    if (!up && !dn) return 0;
    P2PPush q;
    push_header(c, &q);
    const long long cnt = (long long)depth * plane;
    if (up) {
        P2PPeerOp& o = q.p[q.npeer++];
        o.rank = c->rank + 1;
        o.remote = peer_ctl_for_me(c, o.rank);
        o.src[0] = a + (size_t)(hi - depth) * plane;
        o.dst[0] = peer_array(c, o.rank, k) + (size_t)(z0_me + hi - depth - z0_up) * plane;
        o.count[0] = cnt;
    }
    if (dn) {
        P2PPeerOp& o = q.p[q.npeer++];
        o.rank = c->rank - 1;
        o.remote = peer_ctl_for_me(c, o.rank);
        o.src[0] = a + (size_t)lo * plane;
        o.dst[0] = peer_array(c, o.rank, k) + (size_t)(z0_me + lo - z0_dn) * plane;
        o.count[0] = cnt;
    }
    if (kc_p2p_push(q, cnt * q.npeer, c->sm_count, s) != cudaSuccess) {
        c->err = "peer push launch failed";
        return -1;
    }
    return 0;
}

int gs_comm_gather_planes_p2p(GsComm* c, int k, double* coarse, long long plane, int fine_level, cudaStream_t s)
{
    const int Zc = c->b(fine_level + 1, c->world);
    // the coarse planes whose centre fine plane lies in this rank's fine slab, interior only (as gs_comm_gather_planes)
    int zb = (c->b(fine_level, c->rank) + 1) / 2, ze = (c->b(fine_level, c->rank + 1) + 1) / 2;
    if (zb < 1) zb = 1;
    if (ze > Zc - 1) ze = Zc - 1;
    P2PPush q;
    push_header(c, &q);
    const long long cnt = ze > zb ? (long long)(ze - zb) * plane : 0;
    for (int r = 0; r < c->world; r++) {
        if (r == c->rank) continue;
        P2PPeerOp& o = q.p[q.npeer++];
        o.rank = r;
        o.remote = peer_ctl_for_me(c, r);
        o.src[0] = coarse + (size_t)zb * plane;
        o.dst[0] = peer_array(c, r, k) + (size_t)zb * plane;  // replicated level: every rank holds it from plane 0
        o.count[0] = cnt;
    }
    if (kc_p2p_push(q, cnt * q.npeer, c->sm_count, s) != cudaSuccess) {
        c->err = "peer gather launch failed";
        return -1;
    }
    return 0;
}

int gs_comm_norm_p2p(GsComm* c, double* dev_result, cudaStream_t s)
{
    P2PPush q;
    push_header(c, &q);
    q.norm_slots = c->block->norm_slots;
    q.norm_result = dev_result;
    q.world = c->world;
    q.me = c->rank;
    for (int r = 0; r < c->world; r++) {
        if (r == c->rank) continue;
        P2PPeerOp& o = q.p[q.npeer++];
        o.rank = r;
        o.remote = peer_ctl_for_me(c, r);
        o.src[0] = dev_result;
        o.dst[0] = &((P2PBlock*)c->peer_base[r])->norm_slots[c->rank];
        o.count[0] = 1;
    }
    if (kc_p2p_push(q, 1, c->sm_count, s) != cudaSuccess) {
        c->err = "peer norm launch failed";
        return -1;
    }
    return 0;
}
