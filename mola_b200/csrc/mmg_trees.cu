// mmg_trees.cu — host side of the SparseTreesPointCloud flavour (kernels: mmg_trees.cuh).
// Reference behaviour: mola_metric_maps/src/SparseTreesPointCloud.cpp:298-307 (insertObservation +
// eraseGridsFartherThan :961-1000), :916-959 (internal_insertPointCloud3D), :613-713 (queries),
// :715-741 (boundingBox).
#include "mmg_hits.cuh"
#include "mmg_trees.cuh"
#include "mmg_trees_radius.cuh"
#include "mmg_vm.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

using namespace mmg;

namespace mmg
{
struct TreesState
{
    TreesDev       dev{};
    u64            bcap = 0, brows = 0, ccap = 0, crows = 0;
    u32            brow_hi = 0, bfree_n = 0, crow_hi = 0, cfree_n = 0, ncells = 0;
    u64            npoints = 0;
    TreesCounters* h_ctr   = nullptr;  // pinned mirror
    uint4*         cwork = nullptr;  // one work word set per cell-table slot (k_trees_assign / k_trees_rows), sized like ctab
    u64            cwork_cap = 0;
    cudaStream_t   pending_stream = nullptr;
    VmArray        vm[8];            // bkey, bcnt, l1, l0, bfree, cpts, cnext, cfree when they live in mapped virtual ranges
    bool           use_vm = false;
    size_t         vm_limit = 0;     // address space reserved per array  // stream of the insert whose counters are still on their way (m->ctr_pending)
};
}  // namespace mmg

namespace
{
// MMG_TREES_TRACE=1: one stderr line per growth event (what grew, to what size, host time)
struct GrowTrace
{
    const char* what;
    u64         n;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    GrowTrace(const char* w, u64 n_) : what(w), n(n_) {}
    ~GrowTrace()
    {
        static const bool on = getenv("MMG_TREES_TRACE") != nullptr;
        if (on) fprintf(stderr, "[trees] grow %s -> %llu: %.3f ms\n", what, (unsigned long long)n, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    }
};
// Grows a device array to newn elements, new elements filled with fill_byte. Preferred: physical memory mapped behind
// a virtual range reserved once (mmg_vm.h) - the old contents stay where they are, nothing is copied or freed, the stream is
// not synchronised and kernels in flight keep valid pointers. Without the driver API (or MMG_NO_VM=1): allocate, copy, free.
template <typename T>
int grow(mmg_map* m, int slot, T** p, u64 oldn, u64 newn, int fill_byte, cudaStream_t st)
{
    TreesState* s = m->trees;
    VmArray&    a = s->vm[slot];
    if (s->use_vm && !a.base && !*p)
    {   // first allocation of this array: reserve address space for anything the device could hold
        // (bytes per row relative to the largest array of its group: l0 = 4 KB per block row, cpts = 512 B per chain row)
        static const size_t div[8] = {512, 1024, 64, 1, 1024, 1, 128, 128};
        if (vm_reserve(a, s->vm_limit / div[slot] + (size_t(64) << 20), m->device) != 0) a = VmArray();
    }
    if (a.base)
    {
        if (vm_ensure(a, newn * sizeof(T)) != 0)
        {
            set_error("mapping %llu B of device memory failed", (unsigned long long)(newn * sizeof(T)));
            return MMG_ERR_OUT_OF_MEM;
        }
        *p = reinterpret_cast<T*>(a.base);
        MMG_CUDA(cudaMemsetAsync(*p + oldn, fill_byte, (newn - oldn) * sizeof(T), st));
        return MMG_OK;
    }
    T* q = nullptr;
    if (cudaMalloc(&q, newn * sizeof(T)) != cudaSuccess)
    {
        set_error("cudaMalloc(%llu B) failed", (unsigned long long)(newn * sizeof(T)));
        cudaGetLastError();
        return MMG_ERR_OUT_OF_MEM;
    }
    if (oldn) MMG_CUDA(cudaMemcpyAsync(q, *p, oldn * sizeof(T), cudaMemcpyDeviceToDevice, st));
    MMG_CUDA(cudaMemsetAsync(q + oldn, fill_byte, (newn - oldn) * sizeof(T), st));
    MMG_CUDA(cudaStreamSynchronize(st));
    if (*p) cudaFree(*p);
    *p = q;
    return MMG_OK;
}
template <typename T>
void release(mmg_map* m, int slot, T** p)
{
    VmArray& a = m->trees->vm[slot];
    if (a.base) vm_release(a);
    else if (*p) cudaFree(*p);
    *p = nullptr;
}
int new_table(Entry** tab, u64 cap)
{
    Entry* t = nullptr;
    if (cudaMalloc(&t, cap * sizeof(Entry)) != cudaSuccess)
    {
        set_error("cudaMalloc table (%llu entries) failed", (unsigned long long)cap);
        cudaGetLastError();
        return MMG_ERR_OUT_OF_MEM;
    }
    MMG_CUDA(cudaMemset(t, 0xFF, cap * sizeof(Entry)));
    *tab = t;
    return MMG_OK;
}
u64 pow2_for(u64 entries, u64 minimum)
{
    u64 cap = minimum;
    while (cap < 2 * entries) cap <<= 1;
    return cap;
}
int read_ctr(mmg_map* m, cudaStream_t st)
{
    TreesState* s = m->trees;
    MMG_CUDA(cudaMemcpyAsync(s->h_ctr, s->dev.ctr, sizeof(TreesCounters), cudaMemcpyDeviceToHost, st));
    MMG_CUDA(cudaStreamSynchronize(st));
    s->brow_hi = s->h_ctr->brow_hi, s->bfree_n = s->h_ctr->bfree_n, s->crow_hi = s->h_ctr->crow_hi, s->cfree_n = s->h_ctr->cfree_n;
    s->ncells = s->h_ctr->ncells, s->npoints = s->h_ctr->npoints;
    return MMG_OK;
}
int ensure_block_rows(mmg_map* m, u64 rows, cudaStream_t st)
{
    TreesState* s = m->trees;
    if (rows <= s->brows) return MMG_OK;
    const u64 nc = std::max<u64>(rows + rows / 2, 256);
    int       rc;
    GrowTrace tr("block rows", nc);
    if ((rc = grow(m, 0, &s->dev.bkey, s->brows, nc, 0xFF, st))) return rc;
    if ((rc = grow(m, 1, &s->dev.bcnt, s->brows, nc, 0, st))) return rc;
    if ((rc = grow(m, 2, &s->dev.l1, s->brows * 16, nc * 16, 0, st))) return rc;
    if ((rc = grow(m, 3, &s->dev.l0, s->brows * 1024, nc * 1024, 0, st))) return rc;
    if ((rc = grow(m, 4, &s->dev.bfree, s->brows, nc, 0, st))) return rc;
    s->brows = nc;
    return MMG_OK;
}
int ensure_chain_rows(mmg_map* m, u64 rows, cudaStream_t st)
{
    TreesState* s = m->trees;
    if (rows <= s->crows) return MMG_OK;
    const u64 nc = std::max<u64>(2 * rows, 65536);  // 32 MB of rows at least, doubling
    int       rc;
    GrowTrace tr("chain rows", nc);
    if ((rc = grow(m, 5, &s->dev.cpts, s->crows * TR_ROW, nc * TR_ROW, 0, st))) return rc;
    if ((rc = grow(m, 6, &s->dev.cnext, s->crows, nc, 0xFF, st))) return rc;
    if ((rc = grow(m, 7, &s->dev.cfree, s->crows, nc, 0, st))) return rc;
    s->crows = nc;
    return MMG_OK;
}
// block table for `entries` keys at load <= 0.5; rebuilt from the live rows when it grows
int ensure_btab(mmg_map* m, u64 entries, cudaStream_t st)
{
    TreesState* s   = m->trees;
    const u64   cap = pow2_for(entries, 1024);
    if (cap <= s->bcap) return MMG_OK;
    GrowTrace tr("block table", cap);
    MMG_CUDA(cudaStreamSynchronize(st));
    Entry* t  = nullptr;
    int    rc = new_table(&t, cap);
    if (rc) return rc;
    if (s->dev.btab) cudaFree(s->dev.btab);
    s->dev.btab = t, s->dev.bmask = cap - 1, s->bcap = cap;
    if (s->brow_hi) MMG_LAUNCH(k_trees_rebuild_btab, (s->brow_hi + 255) / 256, 256, 0, st, s->dev, s->brow_hi);
    return MMG_OK;
}
// cell table: grow (or rebuild in place after an eviction) by rehashing the old table
int rehash_ctab(mmg_map* m, u64 cap, cudaStream_t st)
{
    TreesState* s = m->trees;
    GrowTrace   tr("cell table", cap);
    MMG_CUDA(cudaStreamSynchronize(st));
    Entry*    old    = s->dev.ctab;
    const u64 oldcap = s->ccap;
    Entry*    t      = nullptr;
    int       rc     = new_table(&t, cap);
    if (rc) return rc;
    s->dev.ctab = t, s->dev.cmask = cap - 1, s->ccap = cap;
    if (old)
    {
        MMG_LAUNCH(k_trees_rehash_ctab, (unsigned)((oldcap + 255) / 256), 256, 0, st, s->dev, (const Entry*)old, oldcap);
        MMG_CUDA(cudaStreamSynchronize(st));
        cudaFree(old);
    }
    return MMG_OK;
}
int ensure_ctab(mmg_map* m, u64 entries, cudaStream_t st)
{
    // a rehash costs two allocations, two host round trips and a pass over the old table: start at 2^18 slots (4 MB) and
    // quadruple, so that a map build of millions of cells rehashes a handful of times
    const u64 cap = pow2_for(entries, 1ull << 18);
    return cap <= m->trees->ccap ? MMG_OK : rehash_ctab(m, m->trees->ccap ? 2 * cap : cap, st);
}
void set_grid(mmg_map* m)
{  // setGridProperties, SparseTreesPointCloud.cpp:176-189
    TreesState* s = m->trees;
    s->dev.g      = m->prm.voxel_size;
    s->dev.ginv   = 1.0f / s->dev.g;
    s->dev.ghalf  = 0.5f * s->dev.g;
    s->dev.bbox   = m->dev.ctr->bbox;
}
}  // namespace

int trees_create(mmg_map* m)
{
    TreesState* s = new TreesState();
    m->trees      = s;
    MMG_CUDA(cudaMalloc(&s->dev.ctr, sizeof(TreesCounters)));
    MMG_CUDA(cudaMemset(s->dev.ctr, 0, sizeof(TreesCounters)));
    MMG_CUDA(cudaMallocHost(&s->h_ctr, sizeof(TreesCounters)));
    s->use_vm = m->use_vm, s->vm_limit = m->vm_limit;  // decided once per map in mmg_create
    set_grid(m);
    int rc = ensure_btab(m, 256, 0);
    if (!rc) rc = ensure_ctab(m, 1024, 0);
    return rc;
}
int trees_destroy(mmg_map* m)
{
    TreesState* s = m->trees;
    if (!s) return MMG_OK;
    cudaDeviceSynchronize();  // unmapping does not wait for kernels the way cudaFree does
    cudaFree(s->dev.btab), cudaFree(s->dev.ctab), cudaFree(s->dev.ctr);
    release(m, 0, &s->dev.bkey), release(m, 1, &s->dev.bcnt), release(m, 2, &s->dev.l1), release(m, 3, &s->dev.l0), release(m, 4, &s->dev.bfree);
    release(m, 5, &s->dev.cpts), release(m, 6, &s->dev.cnext), release(m, 7, &s->dev.cfree);
    if (s->cwork) cudaFree(s->cwork);
    if (s->h_ctr) cudaFreeHost(s->h_ctr);
    delete s;
    m->trees = nullptr;
    return MMG_OK;
}
int trees_clear(mmg_map* m)
{
    TreesState* s = m->trees;
    MMG_CUDA(cudaMemset(s->dev.btab, 0xFF, s->bcap * sizeof(Entry)));
    MMG_CUDA(cudaMemset(s->dev.ctab, 0xFF, s->ccap * sizeof(Entry)));
    if (s->brows)
    {
        MMG_CUDA(cudaMemset(s->dev.bkey, 0xFF, s->brows * sizeof(u64)));
        MMG_CUDA(cudaMemset(s->dev.bcnt, 0, s->brows * sizeof(u32)));
        MMG_CUDA(cudaMemset(s->dev.l1, 0, s->brows * 16 * sizeof(u32)));
        MMG_CUDA(cudaMemset(s->dev.l0, 0, s->brows * 1024 * sizeof(u32)));
    }
    MMG_CUDA(cudaMemset(s->dev.ctr, 0, sizeof(TreesCounters)));
    s->brow_hi = s->bfree_n = s->crow_hi = s->cfree_n = s->ncells = 0;
    s->npoints = 0;
    return MMG_OK;
}
int trees_set_grid(mmg_map* m)
{
    set_grid(m);
    return MMG_OK;
}
int trees_clone(const mmg_map* src, mmg_map* dst)
{
    const TreesState* a = src->trees;
    TreesState*       b = dst->trees;
    int               rc;
    if (a->bcap != b->bcap)
    {
        cudaFree(b->dev.btab);
        b->dev.btab = nullptr;
        if ((rc = new_table(&b->dev.btab, a->bcap))) return rc;
        b->bcap = a->bcap, b->dev.bmask = a->bcap - 1;
    }
    if (a->ccap != b->ccap)
    {
        cudaFree(b->dev.ctab);
        b->dev.ctab = nullptr;
        if ((rc = new_table(&b->dev.ctab, a->ccap))) return rc;
        b->ccap = a->ccap, b->dev.cmask = a->ccap - 1;
    }
    if ((rc = ensure_block_rows(dst, a->brows, 0)) || (rc = ensure_chain_rows(dst, a->crows, 0))) return rc;
    const cudaMemcpyKind k = cudaMemcpyDeviceToDevice;
    cudaMemcpy(b->dev.btab, a->dev.btab, a->bcap * sizeof(Entry), k);
    cudaMemcpy(b->dev.ctab, a->dev.ctab, a->ccap * sizeof(Entry), k);
    if (a->brows)
    {
        cudaMemcpy(b->dev.bkey, a->dev.bkey, a->brows * sizeof(u64), k);
        cudaMemcpy(b->dev.bcnt, a->dev.bcnt, a->brows * sizeof(u32), k);
        cudaMemcpy(b->dev.l1, a->dev.l1, a->brows * 16 * sizeof(u32), k);
        cudaMemcpy(b->dev.l0, a->dev.l0, a->brows * 1024 * sizeof(u32), k);
        cudaMemcpy(b->dev.bfree, a->dev.bfree, a->brows * sizeof(u32), k);
    }
    if (a->crows)
    {
        cudaMemcpy(b->dev.cpts, a->dev.cpts, a->crows * TR_ROW * sizeof(float4), k);
        cudaMemcpy(b->dev.cnext, a->dev.cnext, a->crows * sizeof(u32), k);
        cudaMemcpy(b->dev.cfree, a->dev.cfree, a->crows * sizeof(u32), k);
    }
    cudaMemcpy(b->dev.ctr, a->dev.ctr, sizeof(TreesCounters), k);
    cudaMemcpy(dst->dev.ctr, src->dev.ctr, sizeof(Counters), k);
    b->brow_hi = a->brow_hi, b->bfree_n = a->bfree_n, b->crow_hi = a->crow_hi, b->cfree_n = a->cfree_n, b->ncells = a->ncells, b->npoints = a->npoints;
    MMG_CUDA(cudaGetLastError());
    return MMG_OK;
}

// ------------------------------------------------------------------ insert
// nn_single_search on the device: the thread-per-query fast path, then the warp-per-query branch and bound for the
// queries it could not certify (no host round trip: the second kernel reads the list length on the device).
static int launch_trees_nn1(mmg_map* m, const float* qx, const float* qy, const float* qz, size_t nq, u32 stride, const PoseD& T, int use_pose,
                            float notfound_d2, uint8_t* df, float* dxyz, float* dd2, u64* did, cudaStream_t st)
{
    static const bool slow_only = getenv("MMG_TREES_NN_SLOW") != nullptr;  // A/B switch for the profiles
    const unsigned    nb        = persistent_blocks(m, nq, 8, 8);
    if (slow_only)
    {
        MMG_LAUNCH(k_trees_nn1, nb, 256, 0, st, m->trees->dev, qx, qy, qz, (u32)nq, stride, T, use_pose, notfound_d2, df, dxyz, dd2, did, (const u32*)nullptr,
                   (const u32*)nullptr);
        return MMG_OK;
    }
    u32* todo = (u32*)m->scratch_get(S_TREES_TODO, (nq + 1) * sizeof(u32));
    if (!todo) return MMG_ERR_OUT_OF_MEM;
    MMG_CUDA(cudaMemsetAsync(todo + nq, 0, sizeof(u32), st));
    MMG_LAUNCH(k_trees_nn1_fast, (unsigned)((nq + 255) / 256), 256, 0, st, m->trees->dev, qx, qy, qz, (u32)nq, stride, T, use_pose, notfound_d2, df, dxyz, dd2,
               did, todo, todo + nq);
    MMG_LAUNCH(k_trees_nn1, nb, 256, 0, st, m->trees->dev, qx, qy, qz, (u32)nq, stride, T, use_pose, notfound_d2, df, dxyz, dd2, did, (const u32*)todo,
               (const u32*)(todo + nq));
    return MMG_OK;
}

int trees_insert(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, int has_robot_pose, int mem,
                 cudaStream_t st)
{
    TreesState*       s = m->trees;
    const mmg_params& P = m->prm;
    int               rc;
    // eraseGridsFartherThan(:961-1000); the option is held in a float member (:893)
    if (m->ctr_pending.load(std::memory_order_acquire) && ((rc = settle(m)))) return rc;
    const bool no_clearance = (has_robot_pose & MMG_INSERT_NO_CLEARANCE) != 0;
    has_robot_pose &= MMG_INSERT_ROBOT_POSE;
    if (has_robot_pose && pose12 && P.remove_voxels_farther_than > 0 && s->brow_hi > s->bfree_n)
    {
        const float dist = (float)P.remove_voxels_farther_than;
        const float t[3] = {(float)pose12[9], (float)pose12[10], (float)pose12[11]};
        const int   c[3] = {(int)((t[0] + s->dev.ghalf) * s->dev.ginv), (int)((t[1] + s->dev.ghalf) * s->dev.ginv), (int)((t[2] + s->dev.ghalf) * s->dev.ginv)};
        const int   d    = (int)std::ceil(dist * s->dev.ginv);
        const u32   before = s->bfree_n;
        MMG_LAUNCH(k_trees_evict, (unsigned)(((u64)s->brow_hi * 32 + 255) / 256), 256, 0, st, s->dev, s->brow_hi, c[0], c[1], c[2], d);
        if ((rc = read_ctr(m, st))) return rc;
        if (s->bfree_n != before)
        {
            MMG_CUDA(cudaMemsetAsync(s->dev.btab, 0xFF, s->bcap * sizeof(Entry), st));
            MMG_LAUNCH(k_trees_rebuild_btab, (s->brow_hi + 255) / 256, 256, 0, st, s->dev, s->brow_hi);
            if ((rc = rehash_ctab(m, s->ccap, st))) return rc;  // drops the cells of the erased blocks
            if ((rc = read_ctr(m, st))) return rc;
        }
    }
    if (n == 0) return MMG_OK;
    if (n >= (1ull << 31))
    {
        set_error("insert batch too large (%zu points; limit 2^31-1)", n);
        return MMG_ERR_INVALID_ARG;
    }
    const float *dx, *dy, *dz;
    if ((rc = stage_in(m, xs, ys, zs, n, mem, st, &dx, &dy, &dz))) return rc;
    const u32      N  = (u32)n;
    const unsigned gN = (N + 255) / 256;
    float4*        gp = (float4*)m->scratch_get(S_GP, n * sizeof(float4));
    float4*        acc = (float4*)m->scratch_get(S_TREES0, n * sizeof(float4));
    // u32 scratch: d2 | flag | pos (+ scan scratch) | bid | cslot | seg
    const size_t w1   = 6 * n + scan_scratch_elems(n) + 16;
    u32*         wrk  = (u32*)m->scratch_get(S_TREES1, w1 * sizeof(u32));
    u32*         sbuf = (u32*)m->scratch_get(S_SORT, sort_scratch_words(n) * sizeof(u32));
    if (!gp || !acc || !wrk || !sbuf) return MMG_ERR_OUT_OF_MEM;
    float* d2   = (float*)wrk;
    u32 *  flag = wrk + n, *pos = wrk + 2 * n, *bid = wrk + 3 * n + 1, *cslot = wrk + 4 * n + 1, *seg = wrk + 5 * n + 1, *bsum = wrk + 6 * n + 2;
    const SortScratch ss = sort_scratch_at(sbuf, n);

    // phase 1: compose + clearance test against the pre-batch map, compaction of the kept points, block find-or-insert.
    // One host round trip per scan: the counters read here size everything phase 2 needs.
    const PoseD T   = make_pose(pose12);
    const float thr = no_clearance ? -1.0f : P.minimum_points_clearance * P.minimum_points_clearance;
    // the public insertPoint() appends without looking at the map (MMG_INSERT_NO_CLEARANCE): every d2 = 0 passes "d2 > -1"
    if (no_clearance)
    {
        MMG_LAUNCH(k_trees_compose, gN, 256, 0, st, dx, dy, dz, N, T, gp);
        MMG_CUDA(cudaMemsetAsync(d2, 0, n * sizeof(float), st));
    }
    else
    {
        u32* todo = (u32*)m->scratch_get(S_TREES_TODO, (n + 1) * sizeof(u32));
        if (!todo) return MMG_ERR_OUT_OF_MEM;
        MMG_LAUNCH(k_trees_clear, (unsigned)(((u64)N * TRC_LANES + 255) / 256), 256, 0, st, s->dev, dx, dy, dz, N, T, thr, gp, d2, todo);
        // the points the 3x3x3 cell region could not decide (none at centimetre clearances): exact search
        const float* gq = (const float*)gp;
        MMG_LAUNCH(k_trees_nn1, persistent_blocks(m, n, 8, 8), 256, 0, st, s->dev, gq, gq + 1, gq + 2, N, 4u, T, 0, -1.0f, (uint8_t*)nullptr, (float*)nullptr,
                   d2, (u64*)nullptr, (const u32*)todo, (const u32*)&s->dev.ctr->ntodo);
    }
    if (no_clearance) MMG_CUDA(cudaMemsetAsync(&s->dev.ctr->oob, 0, sizeof(u32), st));  // (k_trees_clear does it otherwise)
    MMG_LAUNCH(k_trees_flags, gN, 256, 0, st, s->dev, gp, d2, N, thr, flag);
    exclusive_scan<u32, u32>(flag, pos, n, &s->dev.ctr->nacc, bsum, st);
    if ((rc = ensure_btab(m, (u64)(s->brow_hi - s->bfree_n) + n, st))) return rc;  // every kept point could open a block
    MMG_LAUNCH(k_trees_locate, gN, 256, 0, st, s->dev, gp, flag, pos, N, acc, bid);
    if ((rc = read_ctr(m, st))) return rc;  // synchronises: kept points, block rows in use, out-of-range count
    const u32 nacc = s->h_ctr->nacc, oob = s->h_ctr->oob;
    if (nacc)
    {
        // phase 2 (no further round trip): block-local indices, cells, chain rows. The rows a batch can take are bounded
        // by one per touched cell plus one per 32 points, so the capacity is reserved before the kernels run.
        if ((rc = ensure_block_rows(m, s->brow_hi, st))) return rc;
        if ((rc = ensure_ctab(m, (u64)s->ncells + nacc, st))) return rc;
        if ((rc = ensure_chain_rows(m, (u64)s->crow_hi + nacc + nacc / TR_ROW + 2, st))) return rc;
        if (s->cwork_cap != s->ccap)
        {
            GrowTrace tr("cell work words", s->ccap);
            // per-slot work words of the cell table (batch ordinals, then the owner's row bookkeeping); all zero between batches
            MMG_CUDA(cudaStreamSynchronize(st));
            if (s->cwork) cudaFree(s->cwork);
            s->cwork = nullptr, s->cwork_cap = 0;
            if (cudaMalloc(&s->cwork, s->ccap * sizeof(uint4)) != cudaSuccess)
            {
                set_error("cudaMalloc(%llu B) failed", (unsigned long long)(s->ccap * sizeof(uint4)));
                cudaGetLastError();
                return MMG_ERR_OUT_OF_MEM;
            }
            MMG_CUDA(cudaMemsetAsync(s->cwork, 0, s->ccap * sizeof(uint4), st));
            s->cwork_cap = s->ccap;
        }
        u32 *          sbid, *sidx;
        const unsigned gA = (nacc + 255) / 256;
        radix_sort_index(bid, nacc, s->brow_hi, ss, &sbid, &sidx, st);
        u32* ord = seg;
        MMG_CUDA(launch_pdl(k_trees_assign, gA, 256, st, s->dev, (const u32*)sbid, (const u32*)sidx, nacc, acc, s->cwork, cslot, ord));
        MMG_CUDA(launch_pdl(k_trees_rows, gA, 256, st, s->dev, (const u32*)sbid, (const u32*)sidx, (const u32*)cslot, (const u32*)ord, nacc,
                            (const float4*)acc, s->cwork));
        MMG_CUDA(launch_pdl(k_trees_write, gA, 256, st, s->dev, (const u32*)sidx, (const u32*)cslot, (const u32*)ord, nacc, (const float4*)acc,
                            (const uint4*)s->cwork));
        // the counters the append moves (chain rows, cells, points) reach the host mirrors lazily: settle()
        MMG_CUDA(cudaMemcpyAsync(s->h_ctr, s->dev.ctr, sizeof(TreesCounters), cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaEventRecord(m->ev_ctr, st));
        s->pending_stream = st;
        m->ctr_pending.store(true, std::memory_order_release);
    }
    MMG_CUDA(cudaGetLastError());
    if (oob)
    {
        set_error("%u of %zu points have a block index outside +-2^20 and were skipped", oob, n);
        return MMG_ERR_OUT_OF_RANGE;
    }
    return MMG_OK;
}

// The host mirrors after an insert that returned without waiting for its last kernels (called by mmg::settle()).
int trees_settle(mmg_map* m)
{
    TreesState* s = m->trees;
    MMG_CUDA(cudaEventSynchronize(m->ev_ctr));
    s->brow_hi = s->h_ctr->brow_hi, s->bfree_n = s->h_ctr->bfree_n, s->crow_hi = s->h_ctr->crow_hi, s->cfree_n = s->h_ctr->cfree_n;
    s->ncells = s->h_ctr->ncells, s->npoints = s->h_ctr->npoints;
    return MMG_OK;
}

// ----------------------------------------------------------------- queries
int trees_nn_single(const mmg_map* cm, const float* xs, const float* ys, const float* zs, size_t n, size_t stride, const double* pose12,
                    uint8_t* ofound, float* oxyz, float* od2, uint64_t* oid, int mem, cudaStream_t st)
{
    mmg_map* m = const_cast<mmg_map*>(cm);
    if (n == 0) return MMG_OK;
    const size_t nq = (n + stride - 1) / stride;
    const float *dx, *dy, *dz;
    int          rc = stage_in(m, xs, ys, zs, n, mem, st, &dx, &dy, &dz);
    if (rc) return rc;
    uint8_t* df = ofound;
    float *  dxyz = oxyz, *dd2 = od2;
    u64*     did = (u64*)oid;
    if (mem == MMG_MEM_HOST)
    {
        char* ob = (char*)m->scratch_get(S_OUT, nq * 32);
        if (!ob) return MMG_ERR_OUT_OF_MEM;
        did = (u64*)ob, dxyz = (float*)(ob + nq * 8), dd2 = (float*)(ob + nq * 20), df = (uint8_t*)(ob + nq * 24);
    }
    const PoseD    T  = make_pose(pose12);
    if ((rc = launch_trees_nn1(m, dx, dy, dz, nq, (u32)stride, T, pose12 ? 1 : 0, 0.0f, df, dxyz, dd2, did, st))) return rc;
    MMG_CUDA(cudaGetLastError());
    if (mem == MMG_MEM_HOST)
    {
        MMG_CUDA(cudaMemcpyAsync(oid, did, nq * 8, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(oxyz, dxyz, nq * 12, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(od2, dd2, nq * 4, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(ofound, df, nq, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaStreamSynchronize(st));
    }
    return MMG_OK;
}

int trees_nn_multi(const mmg_map* cm, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, size_t k, uint32_t* ocount,
                   float* oxyz, float* od2, uint64_t* oid, int mem, cudaStream_t st)
{
    mmg_map* m = const_cast<mmg_map*>(cm);
    if (n == 0) return MMG_OK;
    if (k > 32)
    {
        set_error("SparseTrees nn_multiple_search: k <= 32 supported on the device (got %zu)", k);
        return MMG_ERR_UNSUPPORTED;
    }
    const float *dx, *dy, *dz;
    int          rc = stage_in(m, xs, ys, zs, n, mem, st, &dx, &dy, &dz);
    if (rc) return rc;
    u32*   dc = ocount;
    float *dxyz = oxyz, *dd2 = od2;
    u64*   did = (u64*)oid;
    if (mem == MMG_MEM_HOST)
    {
        char* ob = (char*)m->scratch_get(S_OUT, n * k * 24 + n * 4 + 64);
        if (!ob) return MMG_ERR_OUT_OF_MEM;
        did = (u64*)ob, dxyz = (float*)(ob + n * k * 8), dd2 = (float*)(ob + n * k * 20), dc = (u32*)(ob + n * k * 24);
    }
    const PoseD    T  = make_pose(pose12);
    const unsigned nb = persistent_blocks(m, n, 8, 8);
    MMG_LAUNCH(k_trees_knn, nb, 256, 0, st, m->trees->dev, dx, dy, dz, (u32)n, T, pose12 ? 1 : 0, (u32)k, dc, dxyz, dd2, did);
    MMG_CUDA(cudaGetLastError());
    if (mem == MMG_MEM_HOST)
    {
        MMG_CUDA(cudaMemcpyAsync(oid, did, n * k * 8, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(oxyz, dxyz, n * k * 12, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(od2, dd2, n * k * 4, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(ocount, dc, n * 4, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaStreamSynchronize(st));
    }
    return MMG_OK;
}

int trees_radius(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, float r2, size_t maxp_, uint64_t* ooff,
                 uint64_t* ototal, int mem, cudaStream_t st)
{
    // see mmg_trees_radius.cuh: count -> (per chunk of queries) hits with unique {d2 : rank} keys -> one thread block per
    // query sorts its keys in shared memory, applies the per-block maxPoints cut and the std::map collapse, and writes the
    // survivors at the query's reserved place; a last pass closes the gaps when anything was cut or collapsed.
    TreesState*  s = m->trees;
    const float *dx, *dy, *dz;
    int          rc = stage_in(m, xs, ys, zs, n, mem, st, &dx, &dy, &dz);
    if (rc) return rc;
    const float r    = r2 > 0 ? std::sqrt(r2) : 0.f;
    const u32   maxp = (u32)std::min<size_t>(maxp_, 0xFFFFFFFFu);
    // blocks per axis that [q - r, q + r] can touch, +1 for the double-width index 0
    const u64 per_axis = (u64)std::floor(2.0 * r / s->dev.g) + 2;
    const u64 nbq      = per_axis * per_axis * per_axis;
    if (maxp != 0 && nbq > TRR_NBQ_MAX)
    {
        set_error("nn_radius_search with maxPoints: the radius spans up to %llu blocks per query (limit %u)", (unsigned long long)nbq, TRR_NBQ_MAX);
        return MMG_ERR_INVALID_ARG;
    }
    rad_result().total = 0;
    const size_t ns = scan_scratch_elems(n);
    u32*         cnt    = (u32*)m->scratch_get(S_RCNT, (2 * n + 2) * sizeof(u32) + sizeof(TreesRadiusErr) + 64);
    u64*         off    = (u64*)m->scratch_get(S_ROFF, (n + 2 + ns) * sizeof(u64));
    u64*         boff   = (u64*)m->scratch_get(S_TREES3, (2 * (n + 2 + ns)) * sizeof(u64));
    if (!cnt || !off || !boff) return MMG_ERR_OUT_OF_MEM;
    u32*            bound = cnt + n + 1;
    u32*            next  = (u32*)(bound + n + 1);
    TreesRadiusErr* derr  = (TreesRadiusErr*)(next + 4);
    u64*            noff  = boff + (n + 2 + ns);
    MMG_CUDA(cudaMemsetAsync(off, 0, (n + 1) * sizeof(u64), st));
    u64 H = 0, B = 0;
    if (n)
    {
        const PoseD    T  = make_pose(pose12);
        const unsigned nb = persistent_blocks(m, n, 8, 8);
        MMG_CUDA(cudaMemsetAsync(next, 0, 4 * sizeof(u32) + sizeof(TreesRadiusErr), st));
        const u64 stride = nbq * (u64)maxp;
        if (maxp >= 1 && maxp <= 32 && stride <= TRK_CAP && (u64)n * stride * 24 <= ((u64)2 << 30) && g_tune[T_RADIUS_TOPK])
        {   // small maxPoints: per-block bounded k-nearest search, no hit list (mmg_trees_radius.cuh, k_trees_radius_topk)
            char* hb = (char*)m->scratch_get(S_RHITS, (u64)n * stride * 24);
            if (!hb) return MMG_ERR_OUT_OF_MEM;
            u64*   rid    = (u64*)hb;
            float* rxyz   = (float*)(hb + (u64)n * stride * 8);
            float* rd2    = (float*)(hb + (u64)n * stride * 20);
            u32*   outcnt = bound;
            MMG_LAUNCH(k_trees_radius_topk, (unsigned)std::min<size_t>((n + TRK_WARPS - 1) / TRK_WARPS, (size_t)m->num_sms * 16), TRK_WARPS * 32, 0, st, s->dev,
                       dx, dy, dz, (u32)n, T, pose12 ? 1 : 0, r2, r, maxp, (u32)stride, rxyz, rd2, rid, outcnt, derr);
            MMG_LAUNCH(k_trees_radius_stride_offsets, (unsigned)((n + 256) / 256), 256, 0, st, (u32)n, (u32)stride, boff);
            exclusive_scan<u32, u64>(outcnt, noff, n, noff + n, noff + n + 1, st);
            u64            tot = 0;
            TreesRadiusErr herr;
            MMG_CUDA(cudaMemcpyAsync(&tot, noff + n, sizeof(u64), cudaMemcpyDeviceToHost, st));
            MMG_CUDA(cudaMemcpyAsync(&herr, derr, sizeof(herr), cudaMemcpyDeviceToHost, st));
            MMG_CUDA(cudaStreamSynchronize(st));
            if (herr.rank_overflow)
            {
                set_error("nn_radius_search: the blocks around a query hold more than 2^32 points");
                return MMG_ERR_INVALID_ARG;
            }
            char* pb = (char*)m->scratch_get(S_RHITS2, std::max<u64>(tot, 1) * 24);
            if (!pb) return MMG_ERR_OUT_OF_MEM;
            u64*   pid  = (u64*)pb;
            float* pxyz = (float*)(pb + tot * 8);
            float* pd2  = (float*)(pb + tot * 20);
            MMG_LAUNCH(k_trees_radius_pack, nb, 256, 0, st, (u32)n, (const u32*)outcnt, (const u64*)boff, (const u64*)noff, (const float*)rxyz,
                       (const float*)rd2, (const u64*)rid, pxyz, pd2, pid);
            rad_result().total = tot, rad_result().id = pid, rad_result().xyz = pxyz, rad_result().d2 = pd2;
            MMG_CUDA(cudaGetLastError());
            MMG_CUDA(cudaMemcpyAsync(ooff, noff, (n + 1) * sizeof(u64), mem == MMG_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, st));
            MMG_CUDA(cudaStreamSynchronize(st));
            if (ototal) *ototal = tot;
            return MMG_OK;
        }
        MMG_LAUNCH((k_trees_radius_hits<0>), nb, 256, 0, st, s->dev, dx, dy, dz, 0u, (u32)n, T, pose12 ? 1 : 0, r2, r, cnt, (const u64*)nullptr, 0ull,
                   (u64*)nullptr, (float*)nullptr, (u32*)nullptr, derr);
        // the place reserved for a query: all its hits, or at most maxPoints per block of its index box
        const u64 capq = maxp ? std::min<u64>(nbq * (u64)maxp, 0xFFFFFFFFull) : 0;
        MMG_LAUNCH(k_trees_radius_bound, (unsigned)((n + 255) / 256), 256, 0, st, cnt, (u32)n, (u32)capq, bound);
        exclusive_scan<u32, u64>(cnt, off, n, off + n, off + n + 1, st);
        exclusive_scan<u32, u64>(bound, boff, n, boff + n, boff + n + 1, st);
        std::vector<u64> hoff(n + 1);
        TreesRadiusErr   herr;
        MMG_CUDA(cudaMemcpyAsync(hoff.data(), off, (n + 1) * sizeof(u64), cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(&B, boff + n, sizeof(u64), cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(&herr, derr, sizeof(herr), cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaStreamSynchronize(st));
        H = hoff[n];
        if (herr.rank_overflow)
        {
            set_error("nn_radius_search: the blocks around a query hold more than 2^32 points");
            return MMG_ERR_INVALID_ARG;
        }
        if (H)
        {
            char* hb = (char*)m->scratch_get(S_RHITS, B * 24);
            if (!hb) return MMG_ERR_OUT_OF_MEM;
            u64*   rid  = (u64*)hb;
            float* rxyz = (float*)(hb + B * 8);
            float* rd2  = (float*)(hb + B * 20);
            // chunks of queries: the hit scratch (24 bytes per hit) stays bounded however many hits the batch has
            const u64 limit = (u64)std::max(1, g_tune[T_RADIUS_CHUNK_HITS]);
            std::vector<std::pair<size_t, size_t>> chunks;
            u64                                    hmax = 0;
            for (size_t a = 0; a < n;)
            {
                size_t e = a + 1;
                while (e < n && hoff[e + 1] - hoff[a] <= limit) e++;
                chunks.emplace_back(a, e);
                hmax = std::max(hmax, hoff[e] - hoff[a]);
                a    = e;
            }
            // per hit: key 8 + point 12 + index 4 from the hit pass, and two (key, hit) sort buffers of 12 bytes each
            char* tb = (char*)m->scratch_get(S_QID, hmax * 48 + 64);
            if (!tb) return MMG_ERR_OUT_OF_MEM;
            u64*   hkey = (u64*)tb;
            u64*   bufk = hkey + hmax;                       // 2 * hmax keys
            float* hxyz = (float*)(tb + hmax * 24);
            u32*   hidx = (u32*)(tb + hmax * 36);
            u32*   bufp = (u32*)(tb + hmax * 40);            // 2 * hmax hit numbers
            const unsigned sblocks = (unsigned)std::min<size_t>((n + TRR_WARPS - 1) / TRR_WARPS, (size_t)m->num_sms * 6);
            const u32      nbq32   = (u32)std::max<u64>(1, std::min<u64>(nbq, TRR_NBQ_MAX));
            u32*           gtab    = (u32*)m->scratch_get(S_HSORT, (size_t)sblocks * TRR_WARPS * 2 * nbq32 * sizeof(u32) + 64);
            if (!gtab) return MMG_ERR_OUT_OF_MEM;
            u32* outcnt = bound;  // the bounds are in boff by now
            for (auto& ch : chunks)
            {
                const size_t a = ch.first, k = ch.second - ch.first;
                if (hoff[ch.second] == hoff[a])
                {
                    MMG_CUDA(cudaMemsetAsync(outcnt + a, 0, k * sizeof(u32), st));
                    continue;
                }
                MMG_LAUNCH((k_trees_radius_hits<1>), persistent_blocks(m, k, 8, 8), 256, 0, st, s->dev, dx, dy, dz, (u32)a, (u32)k, T, pose12 ? 1 : 0, r2, r,
                           (u32*)nullptr, (const u64*)off, hoff[a], hkey, hxyz, hidx, derr);
                MMG_CUDA(cudaMemsetAsync(next, 0, 3 * sizeof(u32), st));
                // short lists: a warp per query over L2-resident scratch; long ones: a block per query in shared memory
                const u32   n_small = (u32)std::max(0, g_tune[T_RADIUS_WARP_MAX]);
                const float qscale  = (r2 > 1e-30f && r2 < 1e30f && g_tune[T_RADIUS_Q16]) ? 65536.0f / r2 : 0.f;
                const u32 n_large = herr.maxcnt > n_small ? std::min<u32>(TRB_CAP, (herr.maxcnt + 1023u) & ~1023u) : n_small;
                MMG_LAUNCH(k_trees_radius_sort, (unsigned)std::min<size_t>((k + TRR_WARPS - 1) / TRR_WARPS, sblocks), TRR_THREADS, 0, st, s->dev, dx, dy, dz,
                           (u32)a, (u32)k, T, pose12 ? 1 : 0, r, maxp, (const u32*)(cnt + a), (const u64*)off, hoff[a], (const u64*)hkey,
                           (const float*)hxyz, (const u32*)hidx, bufk, bufp, hmax, gtab, nbq32, (const u64*)boff, rxyz, rd2, rid, outcnt + a, next, n_small,
                           n_large, qscale, g_tune[T_RADIUS_Q16] >= 2 ? 1 : 0);
                if (n_large > n_small)
                {
                    static bool attr_set = false;
                    if (!attr_set)
                    {
                        MMG_CUDA(cudaFuncSetAttribute(k_trees_radius_sort_block, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(TRB_CAP * 12)));
                        attr_set = true;
                    }
                    // two size classes, so that the many medium lists do not pay the shared memory of the few long ones:
                    // up to 6144 hits three blocks share an SM, above that one block has it alone
                    const u32 mid = std::min<u32>(n_large, (u32)std::max(0, g_tune[T_RADIUS_BLOCK_MID]));
                    const u32 lo[2] = {n_small, std::max(mid, n_small)}, hi[2] = {std::max(mid, n_small), n_large};
                    for (int c = 0; c < 2; c++)
                    {
                        if (hi[c] <= lo[c]) continue;
                        const size_t   smem = (size_t)hi[c] * 12;
                        const unsigned bpsm = (unsigned)std::max<size_t>(1, std::min<size_t>(4, (220 * 1024) / (smem + 26 * 1024)));
                        MMG_LAUNCH(k_trees_radius_sort_block, (unsigned)std::min<size_t>(k, (size_t)m->num_sms * bpsm), TRB_THREADS, smem, st, s->dev, dx, dy,
                                   dz, (u32)a, (u32)k, T, pose12 ? 1 : 0, r, maxp, (const u32*)(cnt + a), (const u64*)off, hoff[a], (const u64*)hkey,
                                   (const float*)hxyz, (const u32*)hidx, (const u64*)boff, rxyz, rd2, rid, outcnt + a, next + 1 + c, lo[c], hi[c], qscale, g_tune[T_RADIUS_Q16] >= 2 ? 1 : 0);
                    }
                }
            }
            exclusive_scan<u32, u64>(outcnt, noff, n, noff + n, noff + n + 1, st);
            u64 tot = 0;
            MMG_CUDA(cudaMemcpyAsync(&tot, noff + n, sizeof(u64), cudaMemcpyDeviceToHost, st));
            MMG_CUDA(cudaStreamSynchronize(st));
            if (tot == B)
            {   // nothing was cut or collapsed: the reserved places are the final ones
                rad_result().total = tot, rad_result().id = rid, rad_result().xyz = rxyz, rad_result().d2 = rd2;
                off = boff;
            }
            else
            {
                char* pb = (char*)m->scratch_get(S_RHITS2, std::max<u64>(tot, 1) * 24);
                if (!pb) return MMG_ERR_OUT_OF_MEM;
                u64*   pid  = (u64*)pb;
                float* pxyz = (float*)(pb + tot * 8);
                float* pd2  = (float*)(pb + tot * 20);
                MMG_LAUNCH(k_trees_radius_pack, persistent_blocks(m, n, 8, 8), 256, 0, st, (u32)n, (const u32*)outcnt, (const u64*)boff, (const u64*)noff,
                           (const float*)rxyz, (const float*)rd2, (const u64*)rid, pxyz, pd2, pid);
                rad_result().total = tot, rad_result().id = pid, rad_result().xyz = pxyz, rad_result().d2 = pd2;
                off = noff;
            }
            H = tot;
        }
    }
    MMG_CUDA(cudaGetLastError());
    MMG_CUDA(cudaMemcpyAsync(ooff, off, (n + 1) * sizeof(u64), mem == MMG_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, st));
    MMG_CUDA(cudaStreamSynchronize(st));
    if (ototal) *ototal = H;
    return MMG_OK;
}

// ------------------------------------------------------------------- misc
int trees_is_empty(const mmg_map* m, int* out)
{
    *out = (m->trees->brow_hi - m->trees->bfree_n) == 0;
    return MMG_OK;
}
int trees_num_blocks(const mmg_map* m, uint64_t* out)
{
    *out = m->trees->brow_hi - m->trees->bfree_n;
    return MMG_OK;
}
int trees_num_points(const mmg_map* m, uint64_t* out)
{
    *out = m->trees->npoints;
    return MMG_OK;
}
// live flags + ranks of the block rows; returns the largest block (the export row stride)
static int block_ranks(mmg_map* m, u32** rank_out, u32* maxcnt, cudaStream_t st)
{
    TreesState* s  = m->trees;
    const u32   nr = s->brow_hi;
    u32*        lr = (u32*)m->scratch_get(S_SEG, (2 * (size_t)nr + 2 + scan_scratch_elems(nr)) * sizeof(u32));
    if (!lr) return MMG_ERR_OUT_OF_MEM;
    u32* rank = lr + nr + 1;
    MMG_CUDA(cudaMemsetAsync(&s->dev.ctr->maxcnt, 0, sizeof(u32), st));
    MMG_LAUNCH(k_trees_live, (nr + 255) / 256, 256, 0, st, s->dev, nr, lr);
    exclusive_scan<u32, u32>(lr, rank, nr, (u32*)nullptr, rank + nr + 1, st);
    int rc = read_ctr(m, st);
    if (rc) return rc;
    *rank_out = rank;
    *maxcnt   = s->h_ctr->maxcnt;
    return MMG_OK;
}
int trees_row_capacity(const mmg_map* cm, uint32_t* out)
{
    mmg_map* m = const_cast<mmg_map*>(cm);
    *out       = 0;
    if (m->trees->brow_hi == m->trees->bfree_n) return MMG_OK;
    u32* rank;
    return block_ranks(m, &rank, out, 0);
}
int trees_export(const mmg_map* cm, int32_t* keys, uint32_t* counts, float* points, int mem, cudaStream_t st)
{
    mmg_map*    m = const_cast<mmg_map*>(cm);
    TreesState* s = m->trees;
    const u64   V = s->brow_hi - s->bfree_n;
    if (!V) return MMG_OK;
    u32 *rank, cap = 0;
    int  rc = block_ranks(m, &rank, &cap, st);
    if (rc) return rc;
    int32_t* dk = keys;
    u32*     dc = counts;
    float*   dp = points;
    if (mem == MMG_MEM_HOST)
    {
        char* ob = (char*)m->scratch_get(S_OUT, V * 16 + (points ? V * cap * 12 : 0) + 64);
        if (!ob) return MMG_ERR_OUT_OF_MEM;
        dk = (int32_t*)ob, dc = (u32*)(ob + V * 12), dp = points ? (float*)(ob + V * 16) : nullptr;
    }
    MMG_LAUNCH(k_trees_export_blocks, (s->brow_hi + 255) / 256, 256, 0, st, s->dev, s->brow_hi, rank, dk, dc);
    if (dp)
    {
        MMG_CUDA(cudaMemsetAsync(dp, 0, V * cap * 12, st));
        MMG_LAUNCH(k_trees_export_points, (unsigned)((s->ccap * 32 + 255) / 256), 256, 0, st, s->dev, s->ccap, rank, cap, dp, (u32*)nullptr);
    }
    MMG_CUDA(cudaGetLastError());
    if (mem == MMG_MEM_HOST)
    {
        if (keys) MMG_CUDA(cudaMemcpyAsync(keys, dk, V * 12, cudaMemcpyDeviceToHost, st));
        if (counts) MMG_CUDA(cudaMemcpyAsync(counts, dc, V * 4, cudaMemcpyDeviceToHost, st));
        if (points) MMG_CUDA(cudaMemcpyAsync(points, dp, V * cap * 12, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaStreamSynchronize(st));
    }
    return MMG_OK;
}
int trees_bbox_from_points(mmg_map* m)
{  // boundingBox() without a cached value: over all points (SparseTreesPointCloud.cpp:715-741)
    TreesState* s = m->trees;
    MMG_LAUNCH(k_trees_export_points, (unsigned)((s->ccap * 32 + 255) / 256), 256, 0, 0, s->dev, s->ccap, (const u32*)nullptr, 0u, (float*)nullptr,
               m->dev.ctr->bbox);
    MMG_CUDA(cudaGetLastError());
    return MMG_OK;
}
int trees_table_stats(mmg_map* m, mmg_table_stats* out)
{
    TreesState* s    = m->trees;
    out->capacity    = s->ccap;
    out->live_voxels = s->brow_hi - s->bfree_n;
    out->points      = s->npoints;
    out->load_factor = s->ccap ? (double)s->ncells / (double)s->ccap : 0;
    out->hbm_bytes   = s->bcap * sizeof(Entry) + s->ccap * sizeof(Entry) + s->brows * (8 + 4 + 64 + 4096 + 4) + s->crows * (TR_ROW * 16 + 8);
    return MMG_OK;
}

namespace mmg
{
// replicate from `root`: sizes first (host header already exchanged by the caller), then the flat buffers
int trees_bcast(mmg_map* m, int root, int rank, void* comm, cudaStream_t st, u64 npoints, u64 nblocks,
                int (*bcast)(void*, void*, size_t, int, cudaStream_t))
{
    TreesState* s = m->trees;
    struct Sizes
    {
        u64 bcap, brows, ccap, crows;
    } h{s->bcap, s->brow_hi, s->ccap, s->crow_hi};
    Sizes* dh = (Sizes*)m->scratch_get(S_OUT, sizeof(Sizes) + 64);
    if (!dh) return MMG_ERR_OUT_OF_MEM;
    if (rank == root) MMG_CUDA(cudaMemcpyAsync(dh, &h, sizeof(h), cudaMemcpyHostToDevice, st));
    int rc = bcast(comm, dh, sizeof(h), root, st);
    if (rc) return rc;
    MMG_CUDA(cudaMemcpyAsync(&h, dh, sizeof(h), cudaMemcpyDeviceToHost, st));
    MMG_CUDA(cudaStreamSynchronize(st));
    if (rank != root)
    {
        if (s->bcap != h.bcap)
        {
            cudaFree(s->dev.btab);
            s->dev.btab = nullptr;
            if ((rc = new_table(&s->dev.btab, h.bcap))) return rc;
            s->bcap = h.bcap, s->dev.bmask = h.bcap - 1;
        }
        if (s->ccap != h.ccap)
        {
            cudaFree(s->dev.ctab);
            s->dev.ctab = nullptr;
            if ((rc = new_table(&s->dev.ctab, h.ccap))) return rc;
            s->ccap = h.ccap, s->dev.cmask = h.ccap - 1;
        }
        if ((rc = ensure_block_rows(m, h.brows, st)) || (rc = ensure_chain_rows(m, h.crows, st))) return rc;
    }
    const TreesDev& d = s->dev;
    if ((rc = bcast(comm, d.btab, h.bcap * sizeof(Entry), root, st))) return rc;
    if ((rc = bcast(comm, d.ctab, h.ccap * sizeof(Entry), root, st))) return rc;
    if ((rc = bcast(comm, d.bkey, h.brows * sizeof(u64), root, st))) return rc;
    if ((rc = bcast(comm, d.bcnt, h.brows * sizeof(u32), root, st))) return rc;
    if ((rc = bcast(comm, d.l1, h.brows * 16 * sizeof(u32), root, st))) return rc;
    if ((rc = bcast(comm, d.l0, h.brows * 1024 * sizeof(u32), root, st))) return rc;
    if ((rc = bcast(comm, d.bfree, h.brows * sizeof(u32), root, st))) return rc;
    if ((rc = bcast(comm, d.cpts, h.crows * TR_ROW * sizeof(float4), root, st))) return rc;
    if ((rc = bcast(comm, d.cnext, h.crows * sizeof(u32), root, st))) return rc;
    if ((rc = bcast(comm, d.cfree, h.crows * sizeof(u32), root, st))) return rc;
    if ((rc = bcast(comm, d.ctr, sizeof(TreesCounters), root, st))) return rc;
    if ((rc = bcast(comm, m->dev.ctr, sizeof(Counters), root, st))) return rc;
    (void)npoints, (void)nblocks;
    return read_ctr(m, st);
}
}  // namespace mmg
