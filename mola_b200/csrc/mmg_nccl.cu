// mmg_nccl.cu — replicate a built map from one rank to all ranks with NCCL
// (ncclBroadcast of the flat, pointer-free map buffers over NVLink/NVSwitch).
// NCCL is resolved at run time with dlopen (the process usually has torch's
// libnccl.so.2 loaded already), so the library has no link-time NCCL dependency.
#include <dlfcn.h>
#include <nccl.h>

#include <cstring>

#include "mmg_host.h"
using namespace mmg;

struct mmg_comm
{
    ncclComm_t comm   = nullptr;
    int        rank   = 0, nranks = 1, device = 0;
};

namespace
{
struct NcclApi
{
    void* h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*);
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    const char* (*GetErrorString)(ncclResult_t);
};
NcclApi* api()
{
    static NcclApi a;
    static bool    tried = false;
    if (!tried)
    {
        tried = true;
        for (const char* name : {"libnccl.so.2", "libnccl.so"})
            if ((a.h = dlopen(name, RTLD_NOW | RTLD_GLOBAL))) break;
        if (a.h)
        {
#define SYM(f) *(void**)(&a.f) = dlsym(a.h, "nccl" #f)
            SYM(GetUniqueId), SYM(CommInitRank), SYM(CommDestroy), SYM(Broadcast), SYM(GroupStart), SYM(GroupEnd), SYM(GetErrorString);
#undef SYM
            if (!a.GetUniqueId || !a.CommInitRank || !a.Broadcast || !a.GroupStart || !a.GroupEnd) a.h = nullptr;
        }
    }
    return a.h ? &a : nullptr;
}
#define MMG_NCCL(expr)                                                                           \
    do {                                                                                         \
        ncclResult_t r__ = (expr);                                                               \
        if (r__ != ncclSuccess)                                                                  \
        {                                                                                        \
            set_error("%s failed: %s", #expr, A->GetErrorString ? A->GetErrorString(r__) : "?"); \
            return MMG_ERR_NCCL;                                                                 \
        }                                                                                        \
    } while (0)

struct BcastHeader
{
    int        kind;
    mmg_params prm;
    u64        cap, rows;
    u32        nvox_hi, free_n;
    u64        trees_points, trees_blocks;
};
}  // namespace

namespace mmg
{
int trees_bcast(mmg_map* m, int root, int rank, void* nccl_comm, cudaStream_t st, u64 npoints, u64 nblocks,
                int (*bcast)(void*, void*, size_t, int, cudaStream_t));
}

static int do_bcast(void* comm, void* buf, size_t bytes, int root, cudaStream_t st)
{
    NcclApi* A = api();
    if (!bytes) return MMG_OK;
    MMG_NCCL(A->Broadcast(buf, buf, bytes, ncclUint8, root, (ncclComm_t)comm, st));
    return MMG_OK;
}

extern "C" {
int mmg_comm_unique_id(uint8_t out_id[128])
{
    NcclApi* A = api();
    if (!A)
    {
        set_error("libnccl.so.2 not found");
        return MMG_ERR_NCCL;
    }
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    ncclUniqueId id;
    MMG_NCCL(A->GetUniqueId(&id));
    memcpy(out_id, &id, 128);
    return MMG_OK;
}
int mmg_comm_init(const uint8_t id_bytes[128], int rank, int nranks, int device, mmg_comm** out)
{
    NcclApi* A = api();
    if (!A)
    {
        set_error("libnccl.so.2 not found");
        return MMG_ERR_NCCL;
    }
    if (!out || !id_bytes || rank < 0 || rank >= nranks) return MMG_ERR_INVALID_ARG;
    MMG_CUDA(cudaSetDevice(device));
    ncclUniqueId id;
    memcpy(&id, id_bytes, 128);
    mmg_comm* c = new mmg_comm();
    c->rank = rank, c->nranks = nranks, c->device = device;
    ncclResult_t r = A->CommInitRank(&c->comm, nranks, id, rank);
    if (r != ncclSuccess)
    {
        set_error("ncclCommInitRank failed: %s", A->GetErrorString ? A->GetErrorString(r) : "?");
        delete c;
        return MMG_ERR_NCCL;
    }
    *out = c;
    return MMG_OK;
}
int mmg_comm_destroy(mmg_comm* c)
{
    if (!c) return MMG_OK;
    NcclApi* A = api();
    if (A && c->comm && A->CommDestroy) A->CommDestroy(c->comm);
    delete c;
    return MMG_OK;
}

int mmg_bcast(mmg_map* m, mmg_comm* c, int root, void* stream)
{
    NcclApi* A = api();
    if (!A || !m || !c) return MMG_ERR_INVALID_ARG;
    MMG_CUDA(cudaSetDevice(m->device));
    if (int rc0 = settle(m)) return rc0;
    cudaStream_t st = (cudaStream_t)stream;
    // 1) header: sizes and parameters
    BcastHeader* dh = (BcastHeader*)m->scratch_get(S_OUT, sizeof(BcastHeader) + 64);
    if (!dh) return MMG_ERR_OUT_OF_MEM;
    BcastHeader h;
    memset(&h, 0, sizeof(h));
    if (c->rank == root)
    {
        h.kind = m->kind, h.prm = m->prm, h.cap = m->cap, h.rows = m->h_nvox_hi, h.nvox_hi = m->h_nvox_hi, h.free_n = m->h_free_n;
        if (m->kind == MMG_SPARSE_TREES)
        {
            uint64_t a = 0, b = 0;
            trees_num_points(m, &a), trees_num_blocks(m, &b);
            h.trees_points = a, h.trees_blocks = b;
        }
        MMG_CUDA(cudaMemcpyAsync(dh, &h, sizeof(h), cudaMemcpyHostToDevice, st));
    }
    int rc = do_bcast(c->comm, dh, sizeof(h), root, st);
    if (rc) return rc;
    MMG_CUDA(cudaMemcpyAsync(&h, dh, sizeof(h), cudaMemcpyDeviceToHost, st));
    MMG_CUDA(cudaStreamSynchronize(st));
    if (h.kind != m->kind)
    {
        set_error("mmg_bcast: map kinds differ across ranks");
        return MMG_ERR_INVALID_ARG;
    }
    if (m->kind == MMG_SPARSE_TREES)
    {
        if (c->rank != root)
        {
            m->prm = h.prm;
            m->inv = 1.0f / h.prm.voxel_size;
            trees_set_grid(m);
        }
        return trees_bcast(m, root, c->rank, c->comm, st, h.trees_points, h.trees_blocks, do_bcast);
    }
    // 2) receivers size their buffers like the root's
    if (c->rank != root)
    {
        m->prm = h.prm;
        m->inv = 1.0f / h.prm.voxel_size;
        if (m->cap != h.cap && (rc = alloc_table(m, h.cap, st))) return rc;
        if ((rc = ensure_rows(m, h.rows, st))) return rc;
        // rows the receiver had filled beyond the root's extent would otherwise be handed out again with stale counts
        if (m->h_nvox_hi > h.rows)
        {
            const u64 a = h.rows, k = m->h_nvox_hi - h.rows;
            MMG_CUDA(cudaMemsetAsync(m->dev.cnt + a, 0, k * sizeof(u32), st));
            MMG_CUDA(cudaMemsetAsync(m->dev.vkey + a, 0xFF, k * sizeof(u64), st));
            MMG_CUDA(cudaMemsetAsync(m->dev.fill + a, 0, k * sizeof(u32), st));
            if (m->kind == MMG_SPARSE_VOXEL) MMG_CUDA(cudaMemsetAsync(m->dev.mean + a, 0, k * sizeof(float4), st));
        }
        m->h_nvox_hi    = h.nvox_hi;
        m->h_free_n     = h.free_n;
        m->points_dirty = true;
    }
    // 3) the flat buffers, one group
    const u64 cp = (u64)m->row_points();
    MMG_NCCL(A->GroupStart());
    rc = do_bcast(c->comm, m->dev.table, h.cap * sizeof(Entry), root, st);
    if (!rc) rc = do_bcast(c->comm, m->dev.pts, h.rows * cp * sizeof(float4), root, st);
    if (!rc) rc = do_bcast(c->comm, m->dev.cnt, h.rows * sizeof(u32), root, st);
    if (!rc) rc = do_bcast(c->comm, m->dev.vkey, h.rows * sizeof(u64), root, st);
    if (!rc && h.free_n) rc = do_bcast(c->comm, m->dev.free_list, (u64)h.free_n * sizeof(u32), root, st);
    if (!rc && m->kind == MMG_SPARSE_VOXEL) rc = do_bcast(c->comm, m->dev.mean, h.rows * sizeof(float4), root, st);
    if (!rc) rc = do_bcast(c->comm, m->dev.ctr, sizeof(Counters), root, st);
    MMG_NCCL(A->GroupEnd());
    if (rc) return rc;
    MMG_CUDA(cudaStreamSynchronize(st));
    return MMG_OK;
}
}
