// mmg_capi.cu — the C ABI of include/mola_b200/mmg.h: handle management and the host
// side of every batched operation (staging, table/row growth, kernel sequencing).
// All compute happens in the sm_100a kernels of mmg_voxel.cuh / mmg_trees.cuh; there
// is deliberately no CPU path: without a CUDA device mmg_create fails.
#include "../../include/mola_b200/mmg.h"
#include "mmg_host.h"
#include "mmg_query.cuh"
#include "mmg_insert.cuh"

#include <algorithm>
#include <cstddef>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace mmg
{
std::atomic<unsigned long long> g_launches{0};
int                       g_tune[T_COUNT] = {0, 16, 4096, 1, 1 << 26, 0x7FFFFFFF, 6, 1024, 6144, 2, 1};
static thread_local char  g_err[512] = "";
void set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace mmg
using namespace mmg;

// ------------------------------------------------------------------ helpers
PoseD mmg::make_pose(const double* pose12)
{
    PoseD              T;
    static const double I[12] = {1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0};
    memcpy(T.m, pose12 ? pose12 : I, sizeof(T.m));
    return T;
}

namespace mmg
{
static thread_local bool          t_in_insert = false;  // set by the insert entry points: they settle lazily themselves
static thread_local ScratchScope* t_scope = nullptr;
// per-thread hit buffers of the radius searches (S_RHITS, S_RHITS2): grow-only, released when the thread ends
struct TlsRad
{
    void*     buf[2] = {nullptr, nullptr};
    size_t    cap[2] = {0, 0};
    RadResult r;
    ~TlsRad()
    {
        for (void* b : buf)
            if (b) cudaFree(b);
        cudaGetLastError();
    }
    void* get(int i, size_t bytes)
    {
        if (cap[i] < bytes)
        {
            if (buf[i]) cudaFree(buf[i]);  // synchronises the device: the previous hits are no longer in use
            buf[i] = nullptr, cap[i] = 0;
            const size_t want = bytes + bytes / 4 + 256;
            if (cudaMalloc(&buf[i], want) != cudaSuccess)
            {
                set_error("cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(cudaGetLastError()));
                return nullptr;
            }
            cap[i] = want;
        }
        return buf[i];
    }
};
static thread_local TlsRad t_rad;
RadResult&                 rad_result() { return t_rad.r; }

ScratchScope::ScratchScope(cudaStream_t st) : st_(st), prev_(t_scope) { t_scope = this; }
ScratchScope::~ScratchScope()
{
    for (void* p : all_) cudaFreeAsync(p, st_);
    t_scope = prev_;
}
void* ScratchScope::get(int slot, size_t bytes)
{
    if (slot == S_RHITS || slot == S_RHITS2) return t_rad.get(slot == S_RHITS ? 0 : 1, bytes);
    if (slot >= (int)slots_.size()) slots_.resize(slot + 1, {nullptr, 0});
    auto& s = slots_[slot];
    if (s.second >= bytes && s.first) return s.first;
    void* p = nullptr;  // an outgrown block stays alive until the scope ends: kernels already enqueued may still use it
    if (cudaMallocAsync(&p, bytes + 256, st_) != cudaSuccess)
    {
        set_error("cudaMallocAsync(%zu) failed: %s", bytes + 256, cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    all_.push_back(p);
    s = {p, bytes + 256};
    return p;
}
}  // namespace mmg

void* mmg_map::scratch_get(int slot, size_t bytes)
{
    if (mmg::t_scope) return mmg::t_scope->get(slot, bytes);
    if (slot >= (int)scratch.size()) scratch.resize(slot + 1, {nullptr, 0});
    auto& s = scratch[slot];
    if (s.second < bytes)
    {
        if (s.first)
        {
            cudaDeviceSynchronize();
            cudaFree(s.first);
        }
        const size_t want = bytes + bytes / 4 + 256;
        if (cudaMalloc(&s.first, want) != cudaSuccess)
        {
            s = {nullptr, 0};
            set_error("cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(cudaGetLastError()));
            return nullptr;
        }
        s.second = want;
    }
    return s.first;
}

int mmg_map::ensure_pipeline()
{
    if (s_in) return MMG_OK;
    MMG_CUDA(cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking));
    MMG_CUDA(cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking));
    MMG_CUDA(cudaEventCreateWithFlags(&ev_start, cudaEventDisableTiming));
    for (int i = 0; i < 8; i++)
    {
        MMG_CUDA(cudaEventCreateWithFlags(&ev_in[i], cudaEventDisableTiming));
        MMG_CUDA(cudaEventCreateWithFlags(&ev_k[i], cudaEventDisableTiming));
    }
    return MMG_OK;
}

static int init_counters(mmg_map* m, bool keep_bbox)
{
    Counters c;
    memset(&c, 0, sizeof(c));
    if (keep_bbox)
    {
        Counters old;
        MMG_CUDA(cudaMemcpy(&old, m->dev.ctr, sizeof(old), cudaMemcpyDeviceToHost));
        memcpy(c.bbox, old.bbox, sizeof(c.bbox));
    }
    else
    {
        c.bbox[0] = c.bbox[1] = c.bbox[2] = 0xFFFFFFFFu;
        c.bbox[3] = c.bbox[4] = c.bbox[5] = 0u;
    }
    MMG_CUDA(cudaMemcpy(m->dev.ctr, &c, sizeof(c), cudaMemcpyHostToDevice));
    m->h_nvox_hi = 0;
    m->h_free_n  = 0;
    return MMG_OK;
}

// Table and row store are plain cudaMalloc blocks. (Tried in round 2: the stream-ordered pool, cudaMallocAsync /
// cudaFreeAsync. Growing the pool by a few hundred MB cost 10-14 ms per growth on the B200 box against 1-2 ms for
// cudaMalloc, tools/build_trace.py; growths are rare, so the stream synchronisation cudaMalloc / cudaFree imply is cheap.)
int mmg::alloc_table(mmg_map* m, u64 cap, cudaStream_t st)
{
    Entry* t = nullptr;
    if (cap > (1ull << 32))
    {   // hash_key() yields 32 bits
        set_error("voxel table of %llu entries exceeds the 2^32-entry limit", (unsigned long long)cap);
        return MMG_ERR_OUT_OF_MEM;
    }
    if (cudaMalloc(&t, cap * sizeof(Entry)) != cudaSuccess)
    {
        set_error("cudaMalloc table (%llu entries) failed", (unsigned long long)cap);
        cudaGetLastError();
        return MMG_ERR_OUT_OF_MEM;
    }
    MMG_CUDA(cudaMemsetAsync(t, 0xFF, cap * sizeof(Entry), st));
    if (m->dev.table)
    {
        MMG_CUDA(cudaStreamSynchronize(st));
        cudaFree(m->dev.table);
    }
    m->dev.table = t;
    m->dev.mask  = cap - 1;
    m->cap       = cap;
    return MMG_OK;
}

// Makes room for `entries` keys at load factor <= 0.5 (the reference's robin_map also keeps max load 0.5); rebuilds the
// table from the live rows when it grows (exact counters needed: the callers settle() first).
static int ensure_table(mmg_map* m, u64 entries, cudaStream_t st)
{
    u64 cap = 1024;
    while (cap < 2 * entries) cap <<= 1;
    if (cap <= m->cap) return MMG_OK;
    int rc = alloc_table(m, cap, st);
    if (rc) return rc;
    if (m->h_nvox_hi) MMG_LAUNCH(k_rebuild_table, (m->h_nvox_hi + 255) / 256, 256, 0, st, m->dev, m->h_nvox_hi);
    return MMG_OK;
}

// Grows a row array to newn elements, new elements filled with fill_byte. Preferred: physical memory mapped behind a
// virtual range reserved once (mmg_vm.h): nothing is copied or freed, the stream is not synchronised, kernels in flight keep
// valid pointers. Without the driver API (or MMG_NO_VM=1): allocate, copy, free.
template <typename T>
static int grow_array(mmg_map* m, int slot, T** p, u64 oldn, u64 newn, int fill_byte, cudaStream_t st)
{
    VmArray& a = m->vm[slot];
    // (a first allocation of a gigabyte or more - the worst-case headroom of a multi-million-point batch into a new map -
    // stays with cudaMalloc: mapping page by page costs up to 3 x more there, and a store born that large rarely grows again)
    // Only the point rows (slot 0: 512 B per row, 98 % of the bytes) live in a mapped range: the small arrays copy a few MB
    // when they grow, which is cheaper than a reservation + mapping each (1.5 ms apiece on a new map).
    if (m->use_vm && slot == 0 && !a.base && !*p && newn * sizeof(T) < (size_t(1) << 30))
    {   // address space for anything the device could hold (bytes per row relative to the point rows)
        static const size_t div[6] = {1, 128, 64, 128, 128, 32};
        if (vm_reserve(a, m->vm_limit / div[slot] + (size_t(64) << 20), m->device) != 0) a = VmArray();
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
    if (*p)
    {
        MMG_CUDA(cudaStreamSynchronize(st));
        cudaFree(*p);
    }
    *p = q;
    return MMG_OK;
}
template <typename T>
static void release_array(mmg_map* m, int slot, T** p)
{
    if (m->vm[slot].base) vm_release(m->vm[slot]);
    else if (*p) cudaFree(*p);
    *p = nullptr;
}

int mmg::ensure_rows(mmg_map* m, u64 rows, cudaStream_t st)
{
    if (rows <= m->rowcap) return MMG_OK;
    // geometric growth from a floor of 256k rows (136 MB for a Hashed map, 0.08 % of the HBM): a KITTI-scale map (389k
    // voxels) grows its row store once or twice in its life, and the copies of all growths together move less than twice
    // the final size
    const u64 newcap = std::max<u64>(2 * rows, (u64)1 << 18);
    const u64 cap_pts = (u64)m->row_points();
    int       rc;
    if ((rc = grow_array(m, 0, &m->dev.pts, m->rowcap * cap_pts, newcap * cap_pts, 0, st))) return rc;
    if ((rc = grow_array(m, 1, &m->dev.cnt, m->rowcap, newcap, 0, st))) return rc;
    if ((rc = grow_array(m, 2, &m->dev.vkey, m->rowcap, newcap, 0xFF, st))) return rc;
    if ((rc = grow_array(m, 3, &m->dev.free_list, m->rowcap, newcap, 0, st))) return rc;
    if ((rc = grow_array(m, 4, &m->dev.fill, m->rowcap, newcap, 0, st))) return rc;
    if (m->kind == MMG_SPARSE_VOXEL && (rc = grow_array(m, 5, &m->dev.mean, m->rowcap, newcap, 0, st))) return rc;
    m->rowcap = newcap;
    return MMG_OK;
}

int mmg::settle(const mmg_map* cm)
{
    mmg_map* m = const_cast<mmg_map*>(cm);
    if (!m->ctr_pending.load(std::memory_order_acquire)) return MMG_OK;
    std::lock_guard<std::mutex> lk(m->ctr_mtx);
    if (!m->ctr_pending.load(std::memory_order_relaxed)) return MMG_OK;
    if (m->trees)
    {   // SparseTrees keeps its own counters (mmg_trees.cu)
        if (int rc = trees_settle(m)) return rc;
        m->ctr_pending.store(false, std::memory_order_release);
        return MMG_OK;
    }
    MMG_CUDA(cudaEventSynchronize(m->ev_ctr));
    const Counters c  = *m->h_ctr;  // written by the last block of the last k_insert_commit
    m->h_nvox_hi      = c.nvox_hi;
    m->h_free_n       = c.free_n;
    m->deferred_oob += c.oob;
    m->pending_points = 0;
    m->ctr_pending.store(false, std::memory_order_release);
    return MMG_OK;
}

int mmg::read_counters(mmg_map* m, Counters* c, cudaStream_t st)
{
    MMG_CUDA(cudaMemcpyAsync(m->h_ctr, m->dev.ctr, sizeof(Counters), cudaMemcpyDeviceToHost, st));
    MMG_CUDA(cudaStreamSynchronize(st));
    *c = *m->h_ctr;
    return MMG_OK;
}

// Stages caller SoA inputs on the device when they live in host memory.
int mmg::stage_in(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, int mem, cudaStream_t st, const float** dx,
                    const float** dy, const float** dz)
{
    if (mem == MMG_MEM_DEVICE)
    {
        *dx = xs, *dy = ys, *dz = zs;
        return MMG_OK;
    }
    float* b = (float*)m->scratch_get(S_IN, 3 * n * sizeof(float));
    if (!b) return MMG_ERR_OUT_OF_MEM;
    MMG_CUDA(cudaMemcpyAsync(b, xs, n * sizeof(float), cudaMemcpyHostToDevice, st));
    MMG_CUDA(cudaMemcpyAsync(b + n, ys, n * sizeof(float), cudaMemcpyHostToDevice, st));
    MMG_CUDA(cudaMemcpyAsync(b + 2 * n, zs, n * sizeof(float), cudaMemcpyHostToDevice, st));
    *dx = b, *dy = b + n, *dz = b + 2 * n;
    return MMG_OK;
}

unsigned mmg::persistent_blocks(const mmg_map* m, size_t work_warps, int warps_per_block, int blocks_per_sm)
{
    const size_t need = (work_warps + warps_per_block - 1) / warps_per_block;
    return (unsigned)std::max<size_t>(1, std::min<size_t>(need, (size_t)m->num_sms * blocks_per_sm));
}

// ------------------------------------------------------------------ insert
static int evict_far(mmg_map* m, const double* pose12, cudaStream_t st)
{   // far-voxel eviction, HashedVoxelPointCloud.cpp:305-326 (Hashed only)
    const mmg_params& P            = m->prm;
    const int         dist_in_grid = (int)std::ceil(P.remove_voxels_farther_than * (double)m->inv);
    const float       t[3]         = {(float)pose12[9], (float)pose12[10], (float)pose12[11]};
    const int         ox = (int)(t[0] * m->inv), oy = (int)(t[1] * m->inv), oz = (int)(t[2] * m->inv);
    MMG_LAUNCH(k_evict, (m->h_nvox_hi + 255) / 256, 256, 0, st, m->dev, m->h_nvox_hi, ox, oy, oz, dist_in_grid);
    Counters c;
    int      rc = read_counters(m, &c, st);
    if (rc) return rc;
    if (c.free_n != m->h_free_n)
    {  // something was evicted: rebuild the (tombstone-free) table from the live rows
        m->h_free_n = c.free_n;
        MMG_CUDA(cudaMemsetAsync(m->dev.table, 0xFF, m->cap * sizeof(Entry), st));
        MMG_LAUNCH(k_rebuild_table, (m->h_nvox_hi + 255) / 256, 256, 0, st, m->dev, m->h_nvox_hi);
    }
    return MMG_OK;
}

// Cap-only decimation (min_distance_between_points == 0): the sort-free path of mmg_insert.cuh.
// The batch is cut into chunks so that the worst case "every point opens a new voxel" bounds the
// table and row growth per chunk; one 64-byte counter read-back per chunk is the only sync.
template <class Tr>
static int insert_ranked(mmg_map* m, const float* dx, const float* dy, const float* dz, size_t n, const ScanSet& S0, cudaStream_t st, u32* oob_out,
                         bool async)
{
    const mmg_params& P        = m->prm;
    const u32         hard     = (u32)Tr::CAP;
    const bool        uncapped = (P.max_points_per_voxel == 0 || P.max_points_per_voxel > hard);
    const u32         eff      = uncapped ? hard : P.max_points_per_voxel;
    // oob and nseg are zero here: k_insert_commit's last block clears them after mirroring, every other writer cleans up
    const u32 seg_base = 0;
    u32       oob      = 0;
    for (size_t a = 0; a < n;)
    {
        int rc;
        // rows / voxels in use, exact or (asynchronous batches in flight) bounded from above
        u64 hi_ub   = (u64)m->h_nvox_hi + m->pending_points;
        u64 live_ub = hi_ub - m->h_free_n;
        // chunk: at least 1M points, up to 8M for big batches (worst case "every point opens a voxel": 4.3 GB of row
        // headroom, 2 % of a B200's HBM), more when the map is already large or was reserved (mmg_reserve) for more;
        // a short tail is taken along instead of paying another launch pair + counter read-back for it
        const u64 room_rows = m->rowcap > hi_ub ? m->rowcap - hi_ub : 0;
        const u64 room_tab  = m->cap / 2 > live_ub ? m->cap / 2 - live_ub : 0;
        u64       want      = std::max<u64>({(u64)1 << 20, 4 * live_ub, std::min<u64>(n - a, (u64)8 << 20), std::min(room_rows, room_tab)});
        if (n - a > want && n - a - want < want / 4) want = n - a;
        size_t k = std::min<size_t>({n - a, (size_t)RANK_MAX_CHUNK, (size_t)want});
        if (m->ctr_pending.load(std::memory_order_acquire) && (live_ub + k > m->cap / 2 || hi_ub + k > m->rowcap))
        {   // the bound no longer fits: get the exact counters before anything grows
            if ((rc = settle(m))) return rc;
            hi_ub = m->h_nvox_hi, live_ub = hi_ub - m->h_free_n;
        }
        if (hi_ub + k > m->rowcap)
        {   // the row store must grow for the worst case "every point opens a voxel": keep that within half of the free memory
            size_t free_b = 0, total_b = 0;
            if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess)
            {
                const u64 row_bytes = (u64)m->row_points() * sizeof(float4) + 40;
                const u64 fit       = free_b / 2 / (2 * row_bytes);  // ensure_rows doubles, and the old array lives until the copy is done
                if (fit > hi_ub) k = std::min<size_t>(k, std::max<u64>(fit - hi_ub, (u64)1 << 16));
            }
            cudaGetLastError();
        }
        if ((rc = ensure_table(m, live_ub + k, st))) return rc;
        if ((rc = ensure_rows(m, hi_ub + k, st))) return rc;
        u32* touched = (u32*)m->scratch_get(S_SEG, k * sizeof(u32));
        if (!touched) return MMG_ERR_OUT_OF_MEM;
        ScanSet S = S0;
        const unsigned rb = (unsigned)((k + RANK_BLOCK - 1) / RANK_BLOCK);
        if (S.poses)
            MMG_LAUNCH((k_insert_rank<Tr, true>), rb, RANK_BLOCK, 0, st, dx + a, dy + a, dz + a, (u32)k, (u32)a, S, m->inv, m->dev, eff, touched, seg_base,
                       uncapped ? 1 : 0);
        else
            MMG_LAUNCH((k_insert_rank<Tr, false>), rb, RANK_BLOCK, 0, st, dx + a, dy + a, dz + a, (u32)k, (u32)a, S, m->inv, m->dev, eff, touched, seg_base,
                       uncapped ? 1 : 0);
        const unsigned fb = persistent_blocks(m, k, 8, 8);
        launch_pdl(k_insert_commit<Tr>, fb, 256, st, m->dev, (const u32*)touched, dx + a, dy + a, dz + a, (u32)a, S, m->inv, uncapped ? 0 : 1,
                   m->kind == MMG_SPARSE_VOXEL ? 1 : 0, seg_base, m->h_ctr);
        MMG_CUDA(cudaGetLastError());
        // the commit kernel's last block mirrors the counters into the pinned block: an event marks the moment, no copy call
        MMG_CUDA(cudaEventRecord(m->ev_ctr, st));
        m->pending_points += k;
        m->ctr_pending.store(true, std::memory_order_release);
        a += k;
        if (a < n || !async)
        {   // the next chunk continues the touched-voxel list: it needs this chunk's counters
            if ((rc = settle(m))) return rc;
        }
    }
    oob             = m->deferred_oob;  // everything settled so far, earlier asynchronous batches included
    m->deferred_oob = 0;
    *oob_out        = oob;
    return MMG_OK;
}

template <class Tr>
static int insert_voxel(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, int has_robot_pose, int mem,
                        cudaStream_t st)
{
    const mmg_params& P = m->prm;
    if (m->kind == MMG_HASHED_VOXEL && has_robot_pose && pose12 && P.remove_voxels_farther_than > 0)
    {
        int rc = settle(m);  // the eviction sweeps the rows: exact counters first
        if (!rc && m->h_nvox_hi > 0) rc = evict_far(m, pose12, st);
        if (rc) return rc;
    }
    if (n == 0) return MMG_OK;
    if (n >= (1ull << 31))
    {
        set_error("insert batch too large (%zu points; limit 2^31-1)", n);
        return MMG_ERR_INVALID_ARG;
    }
    const float *dx, *dy, *dz;
    int          rc = stage_in(m, xs, ys, zs, n, mem, st, &dx, &dy, &dz);
    if (rc) return rc;
    const bool  min_dist  = (m->kind == MMG_HASHED_VOXEL) && P.min_distance_between_points > 0.f;
    const float thr       = min_dist ? P.min_distance_between_points * P.min_distance_between_points : 0.f;
    const bool  use_thr   = min_dist && thr > 0.f;
    static const bool force_sorted = getenv("MMG_INSERT_SORTED") != nullptr;  // A/B switch for the profiles
    if (!use_thr && !force_sorted)
    {
        ScanSet S{nullptr, nullptr, 1, 0.f, make_pose(pose12)};
        u32     oob = 0;
        if ((rc = insert_ranked<Tr>(m, dx, dy, dz, n, S, st, &oob, mem == MMG_MEM_DEVICE))) return rc;
        m->points_dirty = true;
        if (oob)
        {
            set_error("%u points of this batch (or of earlier asynchronous batches) have a voxel index outside +-2^20 and were skipped", oob);
            return MMG_ERR_OUT_OF_RANGE;
        }
        return MMG_OK;
    }
    if ((rc = settle(m))) return rc;
    const u64 live = m->h_nvox_hi - m->h_free_n;
    if ((rc = ensure_table(m, live + n, st))) return rc;

    const u32 nb    = (u32)sort_nblocks(n);
    float4*   gp    = (float4*)m->scratch_get(S_GP, n * sizeof(float4));
    u32*      vid   = (u32*)m->scratch_get(S_VID, n * sizeof(u32));
    u32*      sbuf  = (u32*)m->scratch_get(S_SORT, (3 * n + (size_t)RS_RADIX * nb + scan_scratch_elems((size_t)RS_RADIX * nb)) * sizeof(u32));
    u32*      segs  = (u32*)m->scratch_get(S_SEG, n * sizeof(u32));
    if (!gp || !vid || !sbuf || !segs) return MMG_ERR_OUT_OF_MEM;
    SortScratch ss{sbuf, sbuf + n, sbuf + 2 * n, sbuf + 3 * n, sbuf + 3 * n + (size_t)RS_RADIX * nb};

    const u32   hard      = (u32)Tr::CAP;
    const bool  uncapped  = (P.max_points_per_voxel == 0 || P.max_points_per_voxel > hard);
    const int   bbox_all  = (!use_thr && uncapped) ? 1 : 0;
    const int   bbox_acc  = (!use_thr && !uncapped) ? 1 : 0;

    MMG_CUDA(cudaMemsetAsync(&m->dev.ctr->nseg, 0, sizeof(u32), st));
    MMG_CUDA(cudaMemsetAsync(&m->dev.ctr->oob, 0, sizeof(u32), st));
    const PoseD T = make_pose(pose12);
    MMG_LAUNCH((k_locate<Tr>), (unsigned)((n + 255) / 256), 256, 0, st, dx, dy, dz, (u32)n, T, m->inv, m->dev, gp, vid, bbox_all);
    Counters c;
    if ((rc = read_counters(m, &c, st))) return rc;
    m->h_nvox_hi = c.nvox_hi;
    m->h_free_n  = c.free_n;
    if ((rc = ensure_rows(m, c.nvox_hi, st))) return rc;

    u32 *svid, *sidx;
    radix_sort_index(vid, n, c.nvox_hi, ss, &svid, &sidx, st);
    MMG_LAUNCH(k_segment_heads, (unsigned)((n + 255) / 256), 256, 0, st, svid, (u32)n, c.nvox_hi, segs, m->dev.ctr);
    const unsigned fb = persistent_blocks(m, n, 8, 8);
    MMG_LAUNCH((k_fold<Tr>), fb, 256, 0, st, m->dev, svid, sidx, segs, (u32)n, gp, m->inv, P.max_points_per_voxel, use_thr ? thr : 0.f, bbox_acc,
               m->kind == MMG_SPARSE_VOXEL ? 1 : 0);
    MMG_CUDA(cudaGetLastError());
    MMG_CUDA(cudaMemsetAsync(&m->dev.ctr->oob, 0, 2 * sizeof(u32), st));  // oob, nseg: the cap-only path expects them clear
    static_assert(offsetof(Counters, nseg) == offsetof(Counters, oob) + sizeof(u32), "oob and nseg are cleared by one memset");
    // this path keeps no lazy counters: calls that read the map on another stream (mmg_bbox, exports) must find it complete
    MMG_CUDA(cudaStreamSynchronize(st));
    m->points_dirty = true;
    if (c.oob)
    {
        set_error("%u of %zu points have a voxel index outside +-2^20 and were skipped", c.oob, n);
        return MMG_ERR_OUT_OF_RANGE;
    }
    return MMG_OK;
}

// ------------------------------------------------------------------ queries
struct OutBufs
{
    uint8_t* found;
    float *  xyz, *d2;
    u64*     id;
    u32*     count;
};

template <class Tr>
static int launch_nn_single(const mmg_map* m, const float* dx, const float* dy, const float* dz, size_t nq, size_t stride, const PoseD& T, int use_pose,
                             const QueryOut& o, cudaStream_t st, float bd_init = INFINITY, const MatchOut* match = nullptr, bool single_kernel = false)
{
    // `match` (pipeline only): write compacted pairings through k_nn_finish_match instead of dense results
    // A/B switch for the profiles: MMG_NN_KERNEL = tasks (default) | pruned | grouped | simple, or mmg_set_tuning("nn_kernel", 0..3)
    static const int env_variant = [] {
        const char* e = getenv("MMG_NN_KERNEL");
        return !e ? -1 : (!strcmp(e, "pruned") ? 1 : (!strcmp(e, "grouped") ? 2 : (!strcmp(e, "simple") ? 3 : 0)));
    }();
    const int kv = single_kernel ? 1 : (env_variant >= 0 ? env_variant : g_tune[T_NN_KERNEL]);
    if (kv == 0 && nq >= (size_t)g_tune[T_NN_MIN_PIPE])
    {   // three-kernel pipeline over a global task queue (mmg_query.cuh); its scratch comes from the stream-ordered pool, so
        // concurrent callers on one map never share state
        const size_t c0 = std::min<size_t>(nq, NN_CHUNK);
        const size_t off_key = 16 * c0, off_q = off_key + 8 * c0, off_cnt = off_q + 104 * c0, off_status = off_cnt + 32;
        const size_t tiles0 = (c0 + 255) / 256, wbytes = off_status + (match ? tiles0 * 8 : 0);
        char*        wb = nullptr;
        if (cudaMallocAsync((void**)&wb, wbytes, st) != cudaSuccess)
        {
            cudaGetLastError();
            set_error("cudaMallocAsync(%zu) failed", wbytes);
            return MMG_ERR_OUT_OF_MEM;
        }
        NNWork W;
        W.qpt = (float4*)wb, W.qkey = (u64*)(wb + off_key), W.count = (u32*)(wb + off_cnt);
        W.queue[0] = (u32*)(wb + off_q), W.queue[1] = W.queue[0] + 6 * c0, W.queue[2] = W.queue[1] + 12 * c0;
        for (size_t a = 0; a < nq; a += NN_CHUNK)
        {
            const u32 k = (u32)std::min<size_t>(NN_CHUNK, nq - a);
            cudaMemsetAsync(W.count, 0, match ? 32 + ((k + 255) / 256) * 8 : 16, st);  // queue tails, ticket (+ the tile status words)
            QueryOut oc{o.found + a, o.xyz + 3 * a, o.d2 + a, o.id + a};
            MMG_LAUNCH((k_nn_own<Tr>), (k + NN_OWN_BLOCK - 1) / NN_OWN_BLOCK, NN_OWN_BLOCK, 0, st, m->dev, dx + a * stride, dy + a * stride, dz + a * stride, k,
                       (u32)stride, T, use_pose, m->inv, m->prm.voxel_size, bd_init, W);
            const unsigned tb = (unsigned)std::min<size_t>(((size_t)k * 26 + 255) / 256, (size_t)m->num_sms * g_tune[T_NN_TASK_BLOCKS]);
            // the two later stages are programmatic dependent launches: scheduled while the stage before drains
            if (g_tune[T_NN_TASK_MINB] == 5) launch_pdl(k_nn_tasks<Tr, 5>, tb, 256, st, m->dev, m->inv, m->prm.voxel_size, W);
            else launch_pdl(k_nn_tasks<Tr, 6>, tb, 256, st, m->dev, m->inv, m->prm.voxel_size, W);
            if (match)
            {
                MatchOut mo = *match;
                mo.status = (u64*)(wb + off_status), mo.ticket = W.count + 3;  // count[3] is zeroed with the queue tails
                launch_pdl(k_nn_finish_match<Tr>, (k + 255) / 256, 256, st, m->dev, k, (u32)a, m->inv, W, mo);
            }
            else
                launch_pdl(k_nn_finish<Tr>, (k + 255) / 256, 256, st, m->dev, k, m->inv, W, oc);
        }
        cudaFreeAsync(wb, st);
        return MMG_OK;
    }
    const int variant = kv - 1;
    if (variant == 2 && bd_init == INFINITY)
    {
        const unsigned nb = persistent_blocks(m, nq, 8, 8);
        MMG_LAUNCH((k_nn_single<Tr>), nb, 256, 0, st, m->dev, dx, dy, dz, (u32)nq, (u32)stride, T, use_pose, m->inv, o);
    }
    else if (variant == 1 && bd_init == INFINITY)
    {
        const unsigned nb = persistent_blocks(m, (nq + 31) / 32, 8, 8);
        MMG_LAUNCH((k_nn_single_grouped<Tr>), nb, 256, 0, st, m->dev, dx, dy, dz, (u32)nq, (u32)stride, T, use_pose, m->inv, m->prm.voxel_size, o);
    }
    else
    {
        static const unsigned bs = [] {  // block size, tunable for the profiles (MMG_NN_BLOCK = 64 | 128 | 256)
            const char* e = getenv("MMG_NN_BLOCK");
            const int   v = e ? atoi(e) : 0;
            return (unsigned)((v == 64 || v == 128 || v == 256) ? v : 128);
        }();
        const unsigned nb = (unsigned)std::min<size_t>((nq + bs - 1) / bs, (size_t)m->num_sms * 64 * (256 / bs));
        // tile order: a multiplier near grid / golden ratio, coprime to the grid (MMG_NN_PERMUTE=0: launch order)
        static const bool permute = [] { const char* e = getenv("MMG_NN_PERMUTE"); return !e || atoi(e) != 0; }();
        u32 perm = 1;
        if (permute && nb > 8)
        {
            perm = (u32)(0.6180339887 * nb) | 1u;
            auto gcd = [](u32 a, u32 b) { while (b) { const u32 t = a % b; a = b; b = t; } return a; };
            while (gcd(perm, nb) != 1) perm += 2;
        }
        // tuning knobs for the profiles: MMG_NN_VARIANT = <min blocks of 256 per SM: 5 (default) | 6 | 8><p for packed f32x2 math>
        static const int tv = [] {
            const char* e = getenv("MMG_NN_VARIANT");
            if (!e) return 1;  // measured best on C2: 48 registers (5 blocks of 256 per SM), packed x/y math
            const int mb = atoi(e), pk = strchr(e, 'p') != nullptr;
            return (mb == 8 ? 2 : mb == 6 ? 1 : 0) * 2 + pk;
        }();
        switch (tv)
        {
            case 1: MMG_LAUNCH((k_nn_single_pruned<Tr, 5, true>), nb, bs, 0, st, m->dev, dx, dy, dz, (u32)nq, (u32)stride, T, use_pose, m->inv, m->prm.voxel_size, o, bd_init, perm); break;
            case 2: MMG_LAUNCH((k_nn_single_pruned<Tr, 6, false>), nb, bs, 0, st, m->dev, dx, dy, dz, (u32)nq, (u32)stride, T, use_pose, m->inv, m->prm.voxel_size, o, bd_init, perm); break;
            case 3: MMG_LAUNCH((k_nn_single_pruned<Tr, 6, true>), nb, bs, 0, st, m->dev, dx, dy, dz, (u32)nq, (u32)stride, T, use_pose, m->inv, m->prm.voxel_size, o, bd_init, perm); break;
            case 4: MMG_LAUNCH((k_nn_single_pruned<Tr, 8, false>), nb, bs, 0, st, m->dev, dx, dy, dz, (u32)nq, (u32)stride, T, use_pose, m->inv, m->prm.voxel_size, o, bd_init, perm); break;
            case 5: MMG_LAUNCH((k_nn_single_pruned<Tr, 8, true>), nb, bs, 0, st, m->dev, dx, dy, dz, (u32)nq, (u32)stride, T, use_pose, m->inv, m->prm.voxel_size, o, bd_init, perm); break;
            default: MMG_LAUNCH((k_nn_single_pruned<Tr, 5, false>), nb, bs, 0, st, m->dev, dx, dy, dz, (u32)nq, (u32)stride, T, use_pose, m->inv, m->prm.voxel_size, o, bd_init, perm); break;
        }
    }
    return MMG_OK;
}

template <class Tr>
static int nn_single_voxel(const mmg_map* cm, const float* xs, const float* ys, const float* zs, size_t n, size_t stride, const double* pose12,
                           uint8_t* ofound, float* oxyz, float* od2, uint64_t* oid, int mem, cudaStream_t st, float bd_init = INFINITY)
{
    mmg_map* m = const_cast<mmg_map*>(cm);
    if (n == 0) return MMG_OK;
    const size_t nq = (n + stride - 1) / stride;
    const PoseD  T  = make_pose(pose12);
    if (mem == MMG_MEM_DEVICE)
    {
        const int rc = launch_nn_single<Tr>(m, xs, ys, zs, nq, stride, T, pose12 ? 1 : 0, QueryOut{ofound, oxyz, od2, (u64*)oid}, st, bd_init);
        MMG_CUDA(cudaGetLastError());
        return rc;
    }
    // Pinned (page-locked, mapped) host buffers: ONE launch reads the queries and writes the results straight over PCIe
    // (coalesced loads / whole-line stores, both directions busy at once, no staging copies, no per-chunk copy calls).
    static const bool no_zero_copy = getenv("MMG_E2E_PIPELINE") != nullptr;  // A/B switch for the profiles
    if (stride == 1 && !no_zero_copy)
    {
        const void* hp[7] = {xs, ys, zs, ofound, oxyz, od2, oid};
        void*       dp[7];
        bool        all = true;
        for (int i = 0; i < 7 && all; i++)
        {
            cudaPointerAttributes at;
            all = cudaPointerGetAttributes(&at, hp[i]) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer != nullptr;
            dp[i] = all ? at.devicePointer : nullptr;
        }
        cudaGetLastError();
        if (all)
        {
            // ONE kernel (k_nn_single_pruned), not the three-stage pipeline: a kernel that reads its queries and writes its
            // results keeps both directions of the link busy at once (0.63 ms per 1M queries, link ceiling 0.476 ms,
            // tools/pcie_probe.py), while the pipeline reads in its first stage and writes in its last (0.84 ms; cutting the
            // batch into chunks that alternate between two streams did not make the stages of different chunks overlap:
            // 0.80 - 0.97 ms for chunks of 64k - 512k queries, profiles/r2f_*; queries copied in by DMA chunk by chunk on a
            // second stream with each chunk's kernel writing straight to the host buffers: 0.76 / 0.82 / 0.99 ms for 2 / 4 / 8
            // chunks - every kernel boundary waits for its posted writes to drain; queries read in place and the results
            // copied out by DMA chunk by chunk: 0.74 / 0.76 / 0.82 ms - the kernels' read requests queue behind the result
            // copy on the upstream lanes; copy engines both ways, 4 chunks on three streams: 0.71 ms)
            const int rc = launch_nn_single<Tr>(m, (const float*)dp[0], (const float*)dp[1], (const float*)dp[2], nq, 1, T, pose12 ? 1 : 0,
                                                QueryOut{(uint8_t*)dp[3], (float*)dp[4], (float*)dp[5], (u64*)dp[6]}, st, bd_init, nullptr, true);
            if (rc) return rc;
            MMG_CUDA(cudaGetLastError());
            MMG_CUDA(cudaStreamSynchronize(st));
            return MMG_OK;
        }
    }
    // Pageable host buffers: the batch is cut into up to 8 chunks and the three stages run on three streams,
    // so the H2D copy of chunk c+1, the kernel of chunk c and the D2H copy of chunk c-1 overlap
    // (PCIe is full duplex; the kernel is far shorter than either copy).
    if (stride != 1)
    {
        set_error("strided queries take device buffers");
        return MMG_ERR_INVALID_ARG;
    }
    std::lock_guard<std::mutex> pipe_lock(m->pipe_mtx);
    int rc = m->ensure_pipeline();
    if (rc) return rc;
    float* b  = (float*)m->scratch_get(S_IN, 3 * n * sizeof(float));
    char*  ob = (char*)m->scratch_get(S_OUT, nq * 32);
    if (!b || !ob) return MMG_ERR_OUT_OF_MEM;
    u64*     did  = (u64*)ob;
    float*   dxyz = (float*)(ob + nq * 8);
    float*   dd2  = (float*)(ob + nq * 20);
    uint8_t* dfn  = (uint8_t*)(ob + nq * 24);
    size_t   ch   = std::max<size_t>(65536, (nq + 3) / 4);
    ch            = (ch + 255) / 256 * 256;
    if ((nq + ch - 1) / ch > 8) ch = ((nq + 7) / 8 + 255) / 256 * 256;
    MMG_CUDA(cudaEventRecord(m->ev_start, st));
    MMG_CUDA(cudaStreamWaitEvent(m->s_in, m->ev_start, 0));
    MMG_CUDA(cudaStreamWaitEvent(m->s_out, m->ev_start, 0));
    int c = 0;
    for (size_t a = 0; a < nq; a += ch, c++)
    {
        const size_t k = std::min(ch, nq - a);
        MMG_CUDA(cudaMemcpyAsync(b + a, xs + a, k * sizeof(float), cudaMemcpyHostToDevice, m->s_in));
        MMG_CUDA(cudaMemcpyAsync(b + n + a, ys + a, k * sizeof(float), cudaMemcpyHostToDevice, m->s_in));
        MMG_CUDA(cudaMemcpyAsync(b + 2 * n + a, zs + a, k * sizeof(float), cudaMemcpyHostToDevice, m->s_in));
        MMG_CUDA(cudaEventRecord(m->ev_in[c], m->s_in));
        MMG_CUDA(cudaStreamWaitEvent(st, m->ev_in[c], 0));
        if ((rc = launch_nn_single<Tr>(m, b + a, b + n + a, b + 2 * n + a, k, 1, T, pose12 ? 1 : 0, QueryOut{dfn + a, dxyz + 3 * a, dd2 + a, did + a}, st))) return rc;
        MMG_CUDA(cudaEventRecord(m->ev_k[c], st));
        MMG_CUDA(cudaStreamWaitEvent(m->s_out, m->ev_k[c], 0));
        MMG_CUDA(cudaMemcpyAsync(oxyz + 3 * a, dxyz + 3 * a, k * 12, cudaMemcpyDeviceToHost, m->s_out));
        MMG_CUDA(cudaMemcpyAsync(oid + a, did + a, k * 8, cudaMemcpyDeviceToHost, m->s_out));
        MMG_CUDA(cudaMemcpyAsync(od2 + a, dd2 + a, k * 4, cudaMemcpyDeviceToHost, m->s_out));
        MMG_CUDA(cudaMemcpyAsync(ofound + a, dfn + a, k, cudaMemcpyDeviceToHost, m->s_out));
    }
    MMG_CUDA(cudaGetLastError());
    MMG_CUDA(cudaStreamSynchronize(m->s_out));
    MMG_CUDA(cudaStreamSynchronize(st));
    return MMG_OK;
}

template <class Tr, int K, bool REF>
static int nn_multi_voxel(const mmg_map* cm, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, size_t k,
                          uint32_t* ocount, float* oxyz, float* od2, uint64_t* oid, int mem, cudaStream_t st)
{
    mmg_map* m = const_cast<mmg_map*>(cm);
    if (n == 0) return MMG_OK;
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
    const PoseD T = make_pose(pose12);
    MMG_LAUNCH((k_nn_multi<Tr, K, REF>), (unsigned)((n + 127) / 128), 128, 0, st, m->dev, dx, dy, dz, (u32)n, T, pose12 ? 1 : 0, m->inv, m->prm.voxel_size, (u32)k, dc, dxyz,
               dd2, did);
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

template <class Tr>
static int radius_hashed(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, float r2, size_t maxp,
                         uint64_t* ooff, uint64_t* ototal, int mem, cudaStream_t st)
{
    const float *dx, *dy, *dz;
    int          rc = stage_in(m, xs, ys, zs, n, mem, st, &dx, &dy, &dz);
    if (rc) return rc;
    u32* counts = (u32*)m->scratch_get(S_RCNT, (n + 1) * sizeof(u32));
    u64* offs   = (u64*)m->scratch_get(S_ROFF, (n + 1 + scan_scratch_elems(n)) * sizeof(u64));
    if (!counts || !offs) return MMG_ERR_OUT_OF_MEM;
    const PoseD T = make_pose(pose12);
    const float r = r2 > 0 ? std::sqrt(r2) : 0.f;
    const u32   mp = (u32)std::min<size_t>(maxp, 0xFFFFFFFFu);
    const unsigned g = (unsigned)((n + 127) / 128);
    MMG_LAUNCH((k_radius_hashed<Tr, 0>), g, 128, 0, st, m->dev, dx, dy, dz, (u32)n, T, pose12 ? 1 : 0, m->inv, m->prm.voxel_size, r2, r, mp, counts, (const u64*)nullptr,
               (float*)nullptr, (float*)nullptr, (u64*)nullptr);
    exclusive_scan<u32, u64>(counts, offs, n, offs + n, offs + n + 1, st);
    u64 total = 0;
    MMG_CUDA(cudaMemcpyAsync(&total, offs + n, sizeof(u64), cudaMemcpyDeviceToHost, st));
    MMG_CUDA(cudaStreamSynchronize(st));
    rad_result().total = total;
    char* hb     = (char*)m->scratch_get(S_RHITS, std::max<u64>(total, 1) * 24);
    if (!hb) return MMG_ERR_OUT_OF_MEM;
    rad_result().id = (u64*)hb, rad_result().xyz = (float*)(hb + total * 8), rad_result().d2 = (float*)(hb + total * 20);
    if (total)
        MMG_LAUNCH((k_radius_hashed<Tr, 1>), g, 128, 0, st, m->dev, dx, dy, dz, (u32)n, T, pose12 ? 1 : 0, m->inv, m->prm.voxel_size, r2, r, mp, counts, offs, rad_result().xyz,
                   rad_result().d2, rad_result().id);
    MMG_CUDA(cudaGetLastError());
    if (mem == MMG_MEM_HOST)
        MMG_CUDA(cudaMemcpyAsync(ooff, offs, (n + 1) * sizeof(u64), cudaMemcpyDeviceToHost, st));
    else
        MMG_CUDA(cudaMemcpyAsync(ooff, offs, (n + 1) * sizeof(u64), cudaMemcpyDeviceToDevice, st));
    MMG_CUDA(cudaStreamSynchronize(st));
    if (ototal) *ototal = total;
    return MMG_OK;
}

// ------------------------------------------------------------------ C ABI
// every entry point first collects the counters of asynchronous inserts still in flight; the cap-only insert path itself
// only does so when its upper bounds no longer fit (insert_ranked)
static inline int settle_unless_insert(const mmg_map* m) { return m->ctr_pending.load(std::memory_order_acquire) && !mmg::t_in_insert ? settle(m) : MMG_OK; }
#define MMG_CHECK_MAP(m)                          \
    if (!(m))                                     \
    {                                             \
        set_error("null map handle");             \
        return MMG_ERR_INVALID_ARG;               \
    }                                             \
    MMG_CUDA(cudaSetDevice((m)->device));         \
    if (int rc__ = settle_unless_insert(m)) return rc__

// ------------------------------------------------------------------ matcher
static __global__ void k_match_threshold(uint8_t* __restrict__ found, const float* __restrict__ d2, u32 n, float max_d2)
{
    const u32 q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q < n && found[q] && !(d2[q] < max_d2)) found[q] = 0;
}
static __global__ void k_match_scatter(const uint8_t* __restrict__ found, const u32* __restrict__ offs, u32 n, const float* __restrict__ xyz,
                                       const float* __restrict__ d2, const u64* __restrict__ id, u32* __restrict__ oidx, float* __restrict__ oxyz,
                                       float* __restrict__ od2, u64* __restrict__ oid)
{
    const u32 q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n || !found[q]) return;
    const size_t o = offs[q];
    oidx[o]        = q;
    oxyz[3 * o] = xyz[3 * (size_t)q], oxyz[3 * o + 1] = xyz[3 * (size_t)q + 1], oxyz[3 * o + 2] = xyz[3 * (size_t)q + 2];
    od2[o] = d2[q];
    oid[o] = id[q];
}

extern "C" int mmg_match_batch(const mmg_map* cm, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12,
                               float max_dist_sqr, uint64_t* out_count, uint32_t* out_local_idx, float* out_xyz, float* out_d2, uint64_t* out_id,
                               int mem, void* stream)
{
    MMG_CHECK_MAP(cm);
    ScratchScope scratch_scope((cudaStream_t)stream);
    mmg_map* m = const_cast<mmg_map*>(cm);
    if (!out_count || (n && (!xs || !ys || !zs || !out_local_idx || !out_xyz || !out_d2 || !out_id)) || n >= (1ull << 31)) return MMG_ERR_INVALID_ARG;
    *out_count = 0;
    if (n == 0) return MMG_OK;
    if (m->kind == MMG_SPARSE_VOXEL)
    {  // the matcher calls nn_single_search, which throws for this map type (SparseVoxelPointCloud.cpp:695-719)
        set_error("n-search not available in this map type");
        return MMG_ERR_UNSUPPORTED;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const float *dx, *dy, *dz;
    int          rc = stage_in(m, xs, ys, zs, n, mem, st, &dx, &dy, &dz);
    if (rc) return rc;
    // "d2 < max" as the initial search radius: the largest float below max (nothing when max <= 0)
    const float bd_thr = std::isinf(max_dist_sqr) && max_dist_sqr > 0 ? INFINITY : std::nextafterf(max_dist_sqr, -INFINITY);
    static const int env_kernel = [] { const char* e = getenv("MMG_NN_KERNEL"); return e && strcmp(e, "tasks") ? 1 : 0; }();
    if (m->kind == MMG_HASHED_VOXEL && n >= (size_t)g_tune[T_NN_MIN_PIPE] && g_tune[T_NN_KERNEL] == 0 && !env_kernel)
    {   // fused: own voxel -> neighbour tasks -> compacted pairings, no dense intermediate (k_nn_finish_match)
        char* cb = nullptr;
        if (mem == MMG_MEM_HOST && !(cb = (char*)m->scratch_get(S_MATCH, n * 28 + 64))) return MMG_ERR_OUT_OF_MEM;
        u32* tot = (u32*)m->scratch_get(S_LIK, 64);
        if (!tot) return MMG_ERR_OUT_OF_MEM;
        MMG_CUDA(cudaMemsetAsync(tot, 0, sizeof(u32), st));
        MatchOut mo{};
        if (mem == MMG_MEM_HOST) mo.id = (u64*)cb, mo.xyz = (float*)(cb + n * 8), mo.d2 = (float*)(cb + n * 20), mo.idx = (u32*)(cb + n * 24);
        else mo.id = (u64*)out_id, mo.xyz = out_xyz, mo.d2 = out_d2, mo.idx = out_local_idx;
        mo.total = tot;
        rc = launch_nn_single<HashedTraits>(m, dx, dy, dz, n, 1, make_pose(pose12), pose12 ? 1 : 0, QueryOut{}, st, bd_thr, &mo);
        if (rc) return rc;
        MMG_CUDA(cudaGetLastError());
        u32 total = 0;
        MMG_CUDA(cudaMemcpyAsync(&total, tot, sizeof(u32), cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaStreamSynchronize(st));
        *out_count = total;
        if (mem == MMG_MEM_HOST && total)
        {
            MMG_CUDA(cudaMemcpyAsync(out_id, mo.id, (size_t)total * 8, cudaMemcpyDeviceToHost, st));
            MMG_CUDA(cudaMemcpyAsync(out_xyz, mo.xyz, (size_t)total * 12, cudaMemcpyDeviceToHost, st));
            MMG_CUDA(cudaMemcpyAsync(out_d2, mo.d2, (size_t)total * 4, cudaMemcpyDeviceToHost, st));
            MMG_CUDA(cudaMemcpyAsync(out_local_idx, mo.idx, (size_t)total * 4, cudaMemcpyDeviceToHost, st));
            MMG_CUDA(cudaStreamSynchronize(st));
        }
        return MMG_OK;
    }
    // dense nn_single results + scan scratch + (host callers) the compacted outputs
    const size_t dense = n * 32, scanb = (n + 2 + scan_scratch_elems(n)) * sizeof(u32), comp = mem == MMG_MEM_HOST ? n * 28 : 0;
    char*        ob    = (char*)m->scratch_get(S_MATCH, dense + scanb + comp + 256);
    if (!ob) return MMG_ERR_OUT_OF_MEM;
    u64*     id   = (u64*)ob;
    float*   xyz  = (float*)(ob + n * 8);
    float*   d2   = (float*)(ob + n * 20);
    uint8_t* fnd  = (uint8_t*)(ob + n * 24);
    u32*     offs = (u32*)(ob + ((dense + 63) / 64) * 64);
    char*    cb   = (char*)offs + ((scanb + 63) / 64) * 64;
    if (m->kind == MMG_HASHED_VOXEL)
    {
        rc = nn_single_voxel<HashedTraits>(m, dx, dy, dz, n, 1, pose12, fnd, xyz, d2, (uint64_t*)id, MMG_MEM_DEVICE, st, bd_thr);
    }
    else
    {
        rc = trees_nn_single(m, dx, dy, dz, n, 1, pose12, fnd, xyz, d2, (uint64_t*)id, MMG_MEM_DEVICE, st);
        if (!rc) MMG_LAUNCH(k_match_threshold, (unsigned)((n + 255) / 256), 256, 0, st, fnd, d2, (u32)n, max_dist_sqr);
    }
    if (rc) return rc;
    exclusive_scan<uint8_t, u32>(fnd, offs, n, offs + n, offs + n + 1, st);
    u32*   oi = out_local_idx;
    float *ox = out_xyz, *od = out_d2;
    u64*   oid = (u64*)out_id;
    if (mem == MMG_MEM_HOST) oid = (u64*)cb, ox = (float*)(cb + n * 8), od = (float*)(cb + n * 20), oi = (u32*)(cb + n * 24);
    MMG_LAUNCH(k_match_scatter, (unsigned)((n + 255) / 256), 256, 0, st, fnd, offs, (u32)n, xyz, d2, id, oi, ox, od, oid);
    MMG_CUDA(cudaGetLastError());
    u32 total = 0;
    MMG_CUDA(cudaMemcpyAsync(&total, offs + n, sizeof(u32), cudaMemcpyDeviceToHost, st));
    MMG_CUDA(cudaStreamSynchronize(st));
    *out_count = total;
    if (mem == MMG_MEM_HOST && total)
    {
        MMG_CUDA(cudaMemcpyAsync(out_id, oid, (size_t)total * 8, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(out_xyz, ox, (size_t)total * 12, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(out_d2, od, (size_t)total * 4, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(out_local_idx, oi, (size_t)total * 4, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaStreamSynchronize(st));
    }
    return MMG_OK;
}

extern "C" {

int mmg_default_params(int kind, mmg_params* p)
{
    if (!p || kind < 0 || kind > 2) return MMG_ERR_INVALID_ARG;
    memset(p, 0, sizeof(*p));
    // HashedVoxelPointCloud.h:119 (1.0), SparseVoxelPointCloud.h:158 (0.20), SparseTreesPointCloud.h:113 (10.0)
    p->voxel_size                  = kind == MMG_HASHED_VOXEL ? 1.0f : (kind == MMG_SPARSE_VOXEL ? 0.20f : 10.0f);
    p->max_points_per_voxel        = 0;
    p->remove_voxels_farther_than  = 0.0;
    p->min_distance_between_points = 0.f;
    p->minimum_points_clearance    = 0.20f;  // SparseTreesPointCloud.h:337
    p->sigma_dist                  = 0.5;
    p->max_corr_distance           = 1.0;
    p->decimation                  = 10;
    p->match_mean                  = 0;
    return MMG_OK;
}

int mmg_create(int kind, const mmg_params* p, int device, mmg_map** out)
{
    if (!p || !out || kind < 0 || kind > 2 || !(p->voxel_size > 0.f))
    {
        set_error("mmg_create: invalid argument");
        return MMG_ERR_INVALID_ARG;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    {
        cudaGetLastError();
        set_error("no CUDA device: this library has no CPU path");
        return MMG_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= ndev)
    {
        set_error("device %d out of range (%d devices)", device, ndev);
        return MMG_ERR_INVALID_ARG;
    }
    MMG_CUDA(cudaSetDevice(device));
    static const bool tuned = [] {  // MMG_TUNE="knob=value,knob=value": the mmg_set_tuning() switches from the environment (profiling runs)
        const char* e = getenv("MMG_TUNE");
        if (e)
        {
            std::string str(e);
            size_t      a = 0;
            while (a < str.size())
            {
                size_t     b  = str.find(',', a);
                const auto kv = str.substr(a, b == std::string::npos ? std::string::npos : b - a);
                const auto eq = kv.find('=');
                if (eq != std::string::npos) mmg_set_tuning(kv.substr(0, eq).c_str(), atoi(kv.c_str() + eq + 1));
                if (b == std::string::npos) break;
                a = b + 1;
            }
        }
        return true;
    }();
    (void)tuned;
    mmg_map* m = new mmg_map();
    m->kind    = kind;
    m->device  = device;
    m->prm     = *p;
    m->inv     = 1.0f / p->voxel_size;
    MMG_CUDA(cudaDeviceGetAttribute(&m->num_sms, cudaDevAttrMultiProcessorCount, device));
    MMG_CUDA(cudaMalloc(&m->dev.ctr, sizeof(Counters)));
    MMG_CUDA(cudaMallocHost(&m->h_ctr, sizeof(Counters)));
    MMG_CUDA(cudaEventCreateWithFlags(&m->ev_ctr, cudaEventDisableTiming));
    {   // the query pipelines take their scratch from the stream-ordered pool: keep freed blocks cached instead of
        // returning them to the driver at every synchronisation
        cudaMemPool_t pool;
        uint64_t      keep = UINT64_MAX;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        cudaGetLastError();
    }
    {
        static const bool no_vm = getenv("MMG_NO_VM") != nullptr;
        size_t            fr = 0, tot = 0;
        m->use_vm   = !no_vm && vm_available() && cudaMemGetInfo(&fr, &tot) == cudaSuccess;
        m->vm_limit = tot;
    }
    int rc = init_counters(m, false);
    if (!rc && kind != MMG_SPARSE_TREES) rc = alloc_table(m, 1u << 16, 0);
    cudaStreamSynchronize(0);  // the table is usable from any stream once mmg_create returns
    if (!rc && kind == MMG_SPARSE_TREES) rc = trees_create(m);
    if (rc)
    {
        mmg_destroy(m);
        return rc;
    }
    *out = m;
    return MMG_OK;
}

int mmg_destroy(mmg_map* m)
{
    if (!m) return MMG_OK;
    cudaSetDevice(m->device);
    cudaDeviceSynchronize();
    if (m->kind == MMG_SPARSE_TREES) trees_destroy(m);
    cudaFree(m->dev.table), cudaFree(m->dev.ctr);
    release_array(m, 0, &m->dev.pts), release_array(m, 1, &m->dev.cnt), release_array(m, 2, &m->dev.vkey), release_array(m, 3, &m->dev.free_list);
    release_array(m, 4, &m->dev.fill), release_array(m, 5, &m->dev.mean);
    if (m->h_ctr) cudaFreeHost(m->h_ctr);
    if (m->ev_ctr) cudaEventDestroy(m->ev_ctr);
    if (m->h_stage) cudaFreeHost(m->h_stage);
    if (m->s_in)
    {
        cudaStreamDestroy(m->s_in), cudaStreamDestroy(m->s_out), cudaEventDestroy(m->ev_start);
        for (int i = 0; i < 8; i++) cudaEventDestroy(m->ev_in[i]), cudaEventDestroy(m->ev_k[i]);
    }
    for (auto& s : m->scratch)
        if (s.first) cudaFree(s.first);
    delete m;
    return MMG_OK;
}

int mmg_clear(mmg_map* m)
{
    MMG_CHECK_MAP(m);
    MMG_CUDA(cudaDeviceSynchronize());
    if (m->kind == MMG_SPARSE_TREES) return trees_clear(m);
    // internal_clear() drops the voxels but leaves the cached bounding box (HashedVoxelPointCloud.cpp:287)
    MMG_CUDA(cudaMemset(m->dev.table, 0xFF, m->cap * sizeof(Entry)));
    if (m->rowcap)
    {
        MMG_CUDA(cudaMemset(m->dev.cnt, 0, m->rowcap * sizeof(u32)));
        MMG_CUDA(cudaMemset(m->dev.vkey, 0xFF, m->rowcap * sizeof(u64)));
    }
    return init_counters(m, true);
}

int mmg_set_params(mmg_map* m, const mmg_params* p)
{
    MMG_CHECK_MAP(m);
    if (!p || !(p->voxel_size > 0.f)) return MMG_ERR_INVALID_ARG;
    const bool resize = p->voxel_size != m->prm.voxel_size;
    m->prm            = *p;
    if (resize)
    {  // setVoxelProperties(): recompute derived fields and clear, HashedVoxelPointCloud.cpp:189-200
        m->inv = 1.0f / p->voxel_size;
        if (m->kind == MMG_SPARSE_TREES) trees_set_grid(m);
        return mmg_clear(m);
    }
    return MMG_OK;
}
int mmg_get_params(const mmg_map* m, mmg_params* p)
{
    if (!m || !p) return MMG_ERR_INVALID_ARG;
    *p = m->prm;
    return MMG_OK;
}

int mmg_reserve(mmg_map* m, uint64_t voxels)
{
    MMG_CHECK_MAP(m);
    if (m->kind == MMG_SPARSE_TREES) return MMG_OK;  // blocks grow on demand
    int rc = ensure_table(m, voxels, 0);
    if (!rc) rc = ensure_rows(m, voxels, 0);
    return rc;
}

struct InInsert
{   // marks the calling thread as inside an insert entry point (MMG_CHECK_MAP then leaves pending counters alone)
    bool prev;
    InInsert() : prev(mmg::t_in_insert) { mmg::t_in_insert = true; }
    ~InInsert() { mmg::t_in_insert = prev; }
};

int mmg_insert_batch(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, int has_robot_pose, int mem,
                     void* stream)
{
    InInsert in_insert;
    MMG_CHECK_MAP(m);
    if (n && (!xs || !ys || !zs)) return MMG_ERR_INVALID_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    switch (m->kind)
    {
        case MMG_HASHED_VOXEL: return insert_voxel<HashedTraits>(m, xs, ys, zs, n, pose12, has_robot_pose & MMG_INSERT_ROBOT_POSE, mem, st);
        case MMG_SPARSE_VOXEL: return insert_voxel<SparseVoxelTraits>(m, xs, ys, zs, n, pose12, has_robot_pose & MMG_INSERT_ROBOT_POSE, mem, st);
        default: return trees_insert(m, xs, ys, zs, n, pose12, has_robot_pose, mem, st);
    }
}

int mmg_insert_scans(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n_scans, const uint64_t* scan_offsets,
                     const double* poses12, int has_robot_pose, int mem, void* stream)
{
    InInsert in_insert;
    MMG_CHECK_MAP(m);
    if (n_scans == 0) return MMG_OK;
    if (!scan_offsets || !poses12) return MMG_ERR_INVALID_ARG;
    const size_t n = scan_offsets[n_scans] - scan_offsets[0];
    if (n && (!xs || !ys || !zs)) return MMG_ERR_INVALID_ARG;
    for (size_t s = 0; s < n_scans; s++)
        if (scan_offsets[s + 1] < scan_offsets[s])
        {
            set_error("scan_offsets must ascend");
            return MMG_ERR_INVALID_ARG;
        }
    cudaStream_t      st = (cudaStream_t)stream;
    const mmg_params& P  = m->prm;
    // One fused pass is only equivalent to scan-by-scan insertion when nothing happens between the scans
    // (no eviction) and the per-voxel rule is the cap-only one; otherwise insert scan by scan.
    const bool evicts = (has_robot_pose & MMG_INSERT_ROBOT_POSE) && P.remove_voxels_farther_than > 0 && m->kind != MMG_SPARSE_VOXEL;
    const bool fused  = m->kind != MMG_SPARSE_TREES && !evicts && !(m->kind == MMG_HASHED_VOXEL && P.min_distance_between_points > 0.f) &&
                       n < (1ull << 31) && scan_offsets[0] == 0;
    if (!fused)
    {
        int worst = MMG_OK;
        for (size_t s = 0; s < n_scans; s++)
        {
            const size_t a = scan_offsets[s], k = scan_offsets[s + 1] - a;
            const int    rc = mmg_insert_batch(m, xs + a, ys + a, zs + a, k, poses12 + 12 * s, has_robot_pose, mem, stream);
            if (rc == MMG_ERR_OUT_OF_RANGE) worst = rc;
            else if (rc) return rc;
        }
        return worst;
    }
    if (n == 0) return MMG_OK;
    const float *dx, *dy, *dz;
    int          rc = stage_in(m, xs, ys, zs, n, mem, st, &dx, &dy, &dz);
    if (rc) return rc;
    char* sb = (char*)m->scratch_get(S_SCANS, n_scans * sizeof(PoseD) + (n_scans + 1) * sizeof(u32));
    if (!sb) return MMG_ERR_OUT_OF_MEM;
    // poses + 32-bit offsets go through a pinned staging block: one async copy, no sync (insert_ranked() below ends with
    // a counter read-back on the same stream, so the block is free again when this call returns)
    const size_t sbytes = n_scans * sizeof(PoseD) + (n_scans + 1) * sizeof(u32);
    if (m->h_stage_bytes < sbytes)
    {
        if (m->h_stage) cudaFreeHost(m->h_stage);
        m->h_stage = nullptr, m->h_stage_bytes = 0;
        MMG_CUDA(cudaMallocHost(&m->h_stage, sbytes * 2));
        m->h_stage_bytes = sbytes * 2;
    }
    memcpy(m->h_stage, poses12, n_scans * sizeof(PoseD));
    u32* off = (u32*)((char*)m->h_stage + n_scans * sizeof(PoseD));
    for (size_t s = 0; s <= n_scans; s++) off[s] = (u32)scan_offsets[s];
    MMG_CUDA(cudaMemcpyAsync(sb, m->h_stage, sbytes, cudaMemcpyHostToDevice, st));
    ScanSet S{(const PoseD*)sb, (const u32*)(sb + n_scans * sizeof(PoseD)), (u32)n_scans, (float)((double)n_scans / (double)n), make_pose(nullptr)};
    u32     oob = 0;
    // the pinned staging block of poses / offsets is reused by the next call: an asynchronous return would need it to stay untouched
    if (m->kind == MMG_HASHED_VOXEL) rc = insert_ranked<HashedTraits>(m, dx, dy, dz, n, S, st, &oob, false);
    else rc = insert_ranked<SparseVoxelTraits>(m, dx, dy, dz, n, S, st, &oob, false);
    if (rc) return rc;
    m->points_dirty = true;
    if (oob)
    {
        set_error("%u of %zu points have a voxel index outside +-2^20 and were skipped", oob, n);
        return MMG_ERR_OUT_OF_RANGE;
    }
    return MMG_OK;
}

int mmg_nn_single_batch(const mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, uint8_t* out_found,
                        float* out_xyz, float* out_d2, uint64_t* out_id, int mem, void* stream)
{
    MMG_CHECK_MAP(m);
    ScratchScope scratch_scope((cudaStream_t)stream);
    if (n && (!xs || !ys || !zs || !out_found || !out_xyz || !out_d2 || !out_id)) return MMG_ERR_INVALID_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    switch (m->kind)
    {
        case MMG_HASHED_VOXEL: return nn_single_voxel<HashedTraits>(m, xs, ys, zs, n, 1, pose12, out_found, out_xyz, out_d2, out_id, mem, st);
        case MMG_SPARSE_VOXEL:
            // extension: the reference throws here (SparseVoxelPointCloud.cpp:695-719);
            // the host class keeps that behaviour unless the k-NN extension is enabled.
            return nn_single_voxel<SparseVoxelTraits>(m, xs, ys, zs, n, 1, pose12, out_found, out_xyz, out_d2, out_id, mem, st);
        default: return trees_nn_single(m, xs, ys, zs, n, 1, pose12, out_found, out_xyz, out_d2, out_id, mem, st);
    }
}

int mmg_nn_multi_batch(const mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, size_t k,
                       uint32_t* out_count, float* out_xyz, float* out_d2, uint64_t* out_id, int mem, void* stream)
{
    MMG_CHECK_MAP(m);
    ScratchScope scratch_scope((cudaStream_t)stream);
    if (k == 0)
    {  // ASSERT_(N > 0), HashedVoxelPointCloud.cpp:702
        set_error("nn_multiple_search: N must be > 0");
        return MMG_ERR_INVALID_ARG;
    }
    if (n && (!xs || !ys || !zs || !out_count || !out_xyz || !out_d2 || !out_id)) return MMG_ERR_INVALID_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    switch (m->kind)
    {
        case MMG_HASHED_VOXEL: return nn_multi_voxel<HashedTraits, 3, true>(m, xs, ys, zs, n, pose12, k, out_count, out_xyz, out_d2, out_id, mem, st);
        case MMG_SPARSE_VOXEL:
            if (k > 8)
            {
                set_error("SparseVoxel k-NN extension supports k <= 8");
                return MMG_ERR_UNSUPPORTED;
            }
            return nn_multi_voxel<SparseVoxelTraits, 8, false>(m, xs, ys, zs, n, pose12, k, out_count, out_xyz, out_d2, out_id, mem, st);
        default: return trees_nn_multi(m, xs, ys, zs, n, pose12, k, out_count, out_xyz, out_d2, out_id, mem, st);
    }
}

int mmg_nn_radius_batch(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, float radius_sqr,
                        size_t max_points, uint64_t* out_offsets, uint64_t* out_total, int mem, void* stream)
{
    MMG_CHECK_MAP(m);
    ScratchScope scratch_scope((cudaStream_t)stream);
    if (!out_offsets || (n && (!xs || !ys || !zs))) return MMG_ERR_INVALID_ARG;
    if (n >= (1ull << 31)) return MMG_ERR_INVALID_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    rad_result()    = RadResult{};
    rad_result().owner = m;
    switch (m->kind)
    {
        case MMG_HASHED_VOXEL: return radius_hashed<HashedTraits>(m, xs, ys, zs, n, pose12, radius_sqr, max_points, out_offsets, out_total, mem, st);
        case MMG_SPARSE_VOXEL: return sv_radius(m, xs, ys, zs, n, pose12, radius_sqr, max_points, out_offsets, out_total, mem, st);
        default: return trees_radius(m, xs, ys, zs, n, pose12, radius_sqr, max_points, out_offsets, out_total, mem, st);
    }
}

int mmg_nn_radius_fetch(mmg_map* m, float* out_xyz, float* out_d2, uint64_t* out_id, int mem, void* stream)
{
    MMG_CHECK_MAP(m);
    cudaStream_t st = (cudaStream_t)stream;
    // the hits of the CALLING THREAD's last mmg_nn_radius_batch on this map
    if (rad_result().owner != m)
    {
        set_error("mmg_nn_radius_fetch: this thread has no pending radius result for this map");
        return MMG_ERR_INVALID_ARG;
    }
    const u64    t  = rad_result().total;
    if (!t) return MMG_OK;
    const cudaMemcpyKind k = mem == MMG_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
    MMG_CUDA(cudaMemcpyAsync(out_xyz, rad_result().xyz, t * 12, k, st));
    MMG_CUDA(cudaMemcpyAsync(out_d2, rad_result().d2, t * 4, k, st));
    MMG_CUDA(cudaMemcpyAsync(out_id, rad_result().id, t * 8, k, st));
    if (mem == MMG_MEM_HOST) MMG_CUDA(cudaStreamSynchronize(st));
    return MMG_OK;
}

int mmg_likelihood(const mmg_map* cm, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, double* out_loglik, int mem,
                   void* stream)
{
    MMG_CHECK_MAP(cm);
    ScratchScope scratch_scope((cudaStream_t)stream);
    mmg_map* m = const_cast<mmg_map*>(cm);
    if (!out_loglik) return MMG_ERR_INVALID_ARG;
    int empty = 0;
    mmg_is_empty(m, &empty);
    if (empty || n == 0)
    {  // isEmpty() -> 0, HashedVoxelPointCloud.cpp:445; no points -> 0, :460
        *out_loglik = 0.0;
        return MMG_OK;
    }
    if (!(m->prm.sigma_dist > 0) || m->prm.decimation == 0)
    {  // ASSERT_GT_(sigma_dist, 0) :515
        set_error("likelihood: sigma_dist must be > 0 and decimation >= 1");
        return MMG_ERR_INVALID_ARG;
    }
    if (m->kind == MMG_SPARSE_VOXEL)
    {  // calls nn_single_search, which throws for this map type (SparseVoxelPointCloud.cpp:695-719)
        set_error("n-search not available in this map type");
        return MMG_ERR_UNSUPPORTED;
    }
    cudaStream_t st     = (cudaStream_t)stream;
    const size_t stride = m->prm.decimation;
    const size_t nq     = (n + stride - 1) / stride;
    // results stay on the device: stage the input ourselves, then call the device path
    const float *dx, *dy, *dz;
    int          rc = stage_in(m, xs, ys, zs, n, mem, st, &dx, &dy, &dz);
    if (rc) return rc;
    char* ob = (char*)m->scratch_get(S_LIK, nq * 32 + 64 + 257 * 8);
    if (!ob) return MMG_ERR_OUT_OF_MEM;
    uint64_t* id  = (uint64_t*)ob;
    float*   xyz = (float*)(ob + nq * 8);
    float*   d2  = (float*)(ob + nq * 20);
    uint8_t* fnd = (uint8_t*)(ob + nq * 24);
    double*  sum = (double*)(ob + ((nq * 25 + 63) / 8) * 8);
    if (m->kind == MMG_HASHED_VOXEL)
        rc = nn_single_voxel<HashedTraits>(m, dx, dy, dz, n, stride, pose12, fnd, xyz, d2, id, MMG_MEM_DEVICE, st);
    else
        rc = trees_nn_single(m, dx, dy, dz, n, stride, pose12, fnd, xyz, d2, id, MMG_MEM_DEVICE, st);
    if (rc) return rc;
    const float max_sqr = (float)(m->prm.max_corr_distance * m->prm.max_corr_distance);
    {   // fixed-shape two-level sum (deterministic): up to 256 blocks of 1024 write partials, one block adds them in order
        const unsigned lb = (unsigned)std::min<size_t>(256, (nq + 4095) / 4096);
        MMG_LAUNCH(k_likelihood_partial, lb, 1024, 0, st, fnd, d2, (u32)nq, max_sqr, sum + 1);
        MMG_LAUNCH(k_likelihood_final, 1, 256, 0, st, sum + 1, lb, sum);
    }
    double h = 0;
    MMG_CUDA(cudaMemcpyAsync(&h, sum, sizeof(double), cudaMemcpyDeviceToHost, st));
    MMG_CUDA(cudaStreamSynchronize(st));
    *out_loglik = -(h / (double)nq) / m->prm.sigma_dist;
    return MMG_OK;
}

int mmg_is_empty(const mmg_map* m, int* out_empty)
{
    if (!m || !out_empty) return MMG_ERR_INVALID_ARG;
    if (m->kind == MMG_SPARSE_TREES) return trees_is_empty(m, out_empty);
    if (int rc = settle(m)) return rc;
    *out_empty = (m->h_nvox_hi - m->h_free_n) == 0;
    return MMG_OK;
}

int mmg_bbox(mmg_map* m, float out[6])
{
    MMG_CHECK_MAP(m);
    if (!out) return MMG_ERR_INVALID_ARG;
    Counters c;
    int      rc = read_counters(m, &c, 0);
    if (rc) return rc;
    if (c.bbox[0] == 0xFFFFFFFFu && c.bbox[3] == 0u)
    {  // not cached yet: boundingBox() computes it from the voxels / blocks / points and caches it
        int empty = 0;
        mmg_is_empty(m, &empty);
        if (empty)
        {
            const u32 z = float_to_ordered(0.0f);
            for (int i = 0; i < 6; i++) c.bbox[i] = z;
            MMG_CUDA(cudaMemcpy(m->dev.ctr->bbox, c.bbox, sizeof(c.bbox), cudaMemcpyHostToDevice));
        }
        else
        {
            if (m->kind == MMG_HASHED_VOXEL)
                MMG_LAUNCH((k_bbox_from_voxels<HashedTraits>), (m->h_nvox_hi + 255) / 256, 256, 0, 0, m->dev, m->h_nvox_hi, m->prm.voxel_size,
                           m->dev.ctr->bbox);
            else if (m->kind == MMG_SPARSE_VOXEL)
                MMG_LAUNCH((k_bbox_from_voxels<SparseVoxelTraits>), (m->h_nvox_hi + 255) / 256, 256, 0, 0, m->dev, m->h_nvox_hi, m->prm.voxel_size,
                           m->dev.ctr->bbox);
            else if ((rc = trees_bbox_from_points(m)))
                return rc;
            if ((rc = read_counters(m, &c, 0))) return rc;
        }
    }
    for (int i = 0; i < 6; i++) out[i] = ordered_to_float(c.bbox[i]);
    return MMG_OK;
}

int mmg_invalidate_bbox(mmg_map* m)
{
    MMG_CHECK_MAP(m);
    u32 b[6] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u};
    MMG_CUDA(cudaMemcpy(m->dev.ctr->bbox, b, sizeof(b), cudaMemcpyHostToDevice));
    return MMG_OK;
}

int mmg_num_voxels(const mmg_map* m, uint64_t* out_v)
{
    if (!m || !out_v) return MMG_ERR_INVALID_ARG;
    if (m->kind == MMG_SPARSE_TREES) return trees_num_blocks(m, out_v);
    if (int rc = settle(m)) return rc;
    *out_v = m->h_nvox_hi - m->h_free_n;
    return MMG_OK;
}

int mmg_row_capacity(const mmg_map* m, uint32_t* out_cap)
{
    if (!m || !out_cap) return MMG_ERR_INVALID_ARG;
    if (m->kind == MMG_SPARSE_TREES) return trees_row_capacity(m, out_cap);
    *out_cap = (uint32_t)m->row_points();
    return MMG_OK;
}

static int export_voxel(const mmg_map* cm, int32_t* keys, uint32_t* counts, float* points, float* means, float* cov, int mem, cudaStream_t st)
{
    mmg_map*  m  = const_cast<mmg_map*>(cm);
    const u32 nr = m->h_nvox_hi;
    const u64 V  = m->h_nvox_hi - m->h_free_n;
    if (!V) return MMG_OK;
    const u32 cap = (u32)m->row_points();
    u32*      lr  = (u32*)m->scratch_get(S_SEG, (2 * (size_t)nr + 2 + scan_scratch_elems(nr)) * sizeof(u32));
    if (!lr) return MMG_ERR_OUT_OF_MEM;
    u32* rank = lr + nr + 1;
    MMG_LAUNCH(k_live_flags, (nr + 255) / 256, 256, 0, st, m->dev, nr, lr);
    exclusive_scan<u32, u32>(lr, rank, nr, (u32*)nullptr, rank + nr + 1, st);
    int32_t* dk = keys;
    u32*     dc = counts;
    float *  dp = points, *dm = means, *dv = cov;
    char*    ob = nullptr;
    if (mem == MMG_MEM_HOST)
    {
        const size_t bytes = V * 12 + V * 4 + (points ? V * cap * 12 : 0) + (means ? V * 12 : 0) + (cov ? V * 24 : 0) + 64;
        ob                 = (char*)m->scratch_get(S_OUT, bytes);
        if (!ob) return MMG_ERR_OUT_OF_MEM;
        char* q = ob;
        if (keys) dk = (int32_t*)q, q += V * 12;
        if (counts) dc = (u32*)q, q += V * 4;
        if (points) dp = (float*)q, q += V * cap * 12;
        if (means) dm = (float*)q, q += V * 12;
        if (cov) dv = (float*)q, q += V * 24;
    }
    const unsigned g = (unsigned)(((size_t)nr * 32 + 255) / 256);
    if (m->kind == MMG_HASHED_VOXEL)
        MMG_LAUNCH((k_export<HashedTraits>), g, 256, 0, st, m->dev, nr, rank, dk, dc, dp, (float*)nullptr, (float*)nullptr);
    else
        MMG_LAUNCH((k_export<SparseVoxelTraits>), g, 256, 0, st, m->dev, nr, rank, dk, dc, dp, dm, dv);
    MMG_CUDA(cudaGetLastError());
    if (mem == MMG_MEM_HOST)
    {
        if (keys) MMG_CUDA(cudaMemcpyAsync(keys, dk, V * 12, cudaMemcpyDeviceToHost, st));
        if (counts) MMG_CUDA(cudaMemcpyAsync(counts, dc, V * 4, cudaMemcpyDeviceToHost, st));
        if (points) MMG_CUDA(cudaMemcpyAsync(points, dp, V * cap * 12, cudaMemcpyDeviceToHost, st));
        if (means) MMG_CUDA(cudaMemcpyAsync(means, dm, V * 12, cudaMemcpyDeviceToHost, st));
        if (cov) MMG_CUDA(cudaMemcpyAsync(cov, dv, V * 24, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaStreamSynchronize(st));
    }
    return MMG_OK;
}

static __global__ void k_sum_counts(const u32* __restrict__ cnt, u32 nrows, unsigned long long* __restrict__ out)
{   // rows of evicted voxels hold 0
    unsigned long long s = 0;
    for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < nrows; i += gridDim.x * blockDim.x) s += cnt[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(MMG_FULL, s, o);
    if ((threadIdx.x & 31) == 0 && s) atomicAdd(out, s);
}

int mmg_num_points(const mmg_map* cm, uint64_t* out_n)
{
    MMG_CHECK_MAP(cm);
    ScratchScope scratch_scope(nullptr);
    mmg_map*     m = const_cast<mmg_map*>(cm);
    if (!out_n) return MMG_ERR_INVALID_ARG;
    if (m->kind == MMG_SPARSE_TREES) return trees_num_points(m, out_n);
    *out_n = 0;
    if (!m->h_nvox_hi) return MMG_OK;
    // one reduction over the per-row counts on the device, 8 bytes back (round 1 exported every count and summed on the host)
    unsigned long long* d = (unsigned long long*)m->scratch_get(S_LIK, 64);
    if (!d) return MMG_ERR_OUT_OF_MEM;
    MMG_CUDA(cudaMemsetAsync(d, 0, 8, 0));
    MMG_LAUNCH(k_sum_counts, std::min<unsigned>((m->h_nvox_hi + 255) / 256, (unsigned)m->num_sms * 8), 256, 0, 0, m->dev.cnt, m->h_nvox_hi, d);
    unsigned long long h = 0;
    MMG_CUDA(cudaMemcpyAsync(&h, d, 8, cudaMemcpyDeviceToHost, 0));
    MMG_CUDA(cudaStreamSynchronize(0));
    *out_n = h;
    return MMG_OK;
}

int mmg_export(const mmg_map* m, int32_t* out_keys, uint32_t* out_counts, float* out_points, int mem, void* stream)
{
    MMG_CHECK_MAP(m);
    ScratchScope scratch_scope((cudaStream_t)stream);
    if (m->kind == MMG_SPARSE_TREES) return trees_export(m, out_keys, out_counts, out_points, mem, (cudaStream_t)stream);
    return export_voxel(m, out_keys, out_counts, out_points, nullptr, nullptr, mem, (cudaStream_t)stream);
}

int mmg_voxel_stats(const mmg_map* m, float* out_mean, float* out_cov, int mem, void* stream)
{
    MMG_CHECK_MAP(m);
    ScratchScope scratch_scope((cudaStream_t)stream);
    if (m->kind != MMG_SPARSE_VOXEL)
    {
        set_error("voxel mean/covariance exist for SparseVoxelPointCloud only");
        return MMG_ERR_UNSUPPORTED;
    }
    return export_voxel(m, nullptr, nullptr, nullptr, out_mean, out_cov, mem, (cudaStream_t)stream);
}

__global__ void k_probe_stats(VoxelMapDev m, u32 nrows, unsigned long long* sum, u32* mx)
{
    const u32 vid = blockIdx.x * blockDim.x + threadIdx.x;
    if (vid >= nrows) return;
    const u64 key = m.vkey[vid];
    if (key == KEY_EMPTY) return;
    u64 h = hash_key(key) & m.mask;
    u32 len = 1;
    while (m.table[h].key != key)
    {
        h = (h + 1) & m.mask;
        len++;
    }
    atomicAdd(sum, (unsigned long long)len);
    atomicMax(mx, len);
}

int mmg_get_table_stats(const mmg_map* cm, mmg_table_stats* out)
{
    MMG_CHECK_MAP(cm);
    ScratchScope scratch_scope(nullptr);
    mmg_map* m = const_cast<mmg_map*>(cm);
    if (!out) return MMG_ERR_INVALID_ARG;
    memset(out, 0, sizeof(*out));
    if (m->kind == MMG_SPARSE_TREES) return trees_table_stats(m, out);
    out->capacity    = m->cap;
    out->live_voxels = m->h_nvox_hi - m->h_free_n;
    out->load_factor = m->cap ? (double)out->live_voxels / (double)m->cap : 0;
    out->hbm_bytes   = m->cap * sizeof(Entry) + m->rowcap * ((u64)m->row_points() * 16 + 4 + 8 + 4 + (m->kind == MMG_SPARSE_VOXEL ? 16 : 0));
    mmg_num_points(m, &out->points);
    if (m->h_nvox_hi)
    {
        char* b = (char*)m->scratch_get(S_OUT, 64);
        if (!b) return MMG_ERR_OUT_OF_MEM;
        MMG_CUDA(cudaMemset(b, 0, 16));
        MMG_LAUNCH(k_probe_stats, (m->h_nvox_hi + 255) / 256, 256, 0, 0, m->dev, m->h_nvox_hi, (unsigned long long*)b, (u32*)(b + 8));
        unsigned long long h[2];
        MMG_CUDA(cudaMemcpy(h, b, 16, cudaMemcpyDeviceToHost));
        out->mean_probe = out->live_voxels ? (double)h[0] / (double)out->live_voxels : 0;
        out->max_probe  = (double)(u32)h[1];
    }
    return MMG_OK;
}

int mmg_clone(const mmg_map* src, mmg_map** out)
{
    MMG_CHECK_MAP(src);
    if (!out) return MMG_ERR_INVALID_ARG;
    MMG_CUDA(cudaDeviceSynchronize());
    mmg_map* m  = nullptr;
    int      rc = mmg_create(src->kind, &src->prm, src->device, &m);
    if (rc) return rc;
    if (src->kind == MMG_SPARSE_TREES)
    {
        rc = trees_clone(src, m);
    }
    else
    {
        if ((rc = alloc_table(m, src->cap, 0)) == MMG_OK && (rc = ensure_rows(m, src->rowcap ? src->rowcap : 1, 0)) == MMG_OK)
        {
            // ensure_rows may round up; copy the source extent only
            const u64 rc_rows = src->rowcap, cp = (u64)src->row_points();
            cudaMemcpy(m->dev.table, src->dev.table, src->cap * sizeof(Entry), cudaMemcpyDeviceToDevice);
            if (rc_rows)
            {
                cudaMemcpy(m->dev.pts, src->dev.pts, rc_rows * cp * sizeof(float4), cudaMemcpyDeviceToDevice);
                cudaMemcpy(m->dev.cnt, src->dev.cnt, rc_rows * sizeof(u32), cudaMemcpyDeviceToDevice);
                cudaMemcpy(m->dev.vkey, src->dev.vkey, rc_rows * sizeof(u64), cudaMemcpyDeviceToDevice);
                cudaMemcpy(m->dev.free_list, src->dev.free_list, rc_rows * sizeof(u32), cudaMemcpyDeviceToDevice);
                if (src->kind == MMG_SPARSE_VOXEL) cudaMemcpy(m->dev.mean, src->dev.mean, rc_rows * sizeof(float4), cudaMemcpyDeviceToDevice);
            }
            cudaMemcpy(m->dev.ctr, src->dev.ctr, sizeof(Counters), cudaMemcpyDeviceToDevice);
            m->h_nvox_hi = src->h_nvox_hi;
            m->h_free_n  = src->h_free_n;
            if (cudaGetLastError() != cudaSuccess) rc = MMG_ERR_CUDA;
        }
    }
    if (rc)
    {
        mmg_destroy(m);
        return rc;
    }
    *out = m;
    return MMG_OK;
}

__global__ void k_stencil_points(VoxelMapDev m, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 n, float inv,
                                 int round_mode, unsigned long long* total)
{
    const u32 q = blockIdx.x * blockDim.x + threadIdx.x;
    u32       s = 0;
    if (q < n)
    {
        int  c[3];
        bool ok = round_mode ? (SparseVoxelTraits::to_index(qx[q], inv, c[0]) & SparseVoxelTraits::to_index(qy[q], inv, c[1]) & SparseVoxelTraits::to_index(qz[q], inv, c[2]))
                             : (HashedTraits::to_index(qx[q], inv, c[0]) & HashedTraits::to_index(qy[q], inv, c[1]) & HashedTraits::to_index(qz[q], inv, c[2]));
        if (ok)
            for (int x = c[0] - 1; x <= c[0] + 1; x++)
                for (int y = c[1] - 1; y <= c[1] + 1; y++)
                    for (int z = c[2] - 1; z <= c[2] + 1; z++)
                    {
                        u32 vid, cnt;
                        if (index_in_range(x, y, z) && probe_ro(m.table, m.mask, pack_key(x, y, z), vid, cnt)) s += cnt;
                    }
    }
    s = __reduce_add_sync(MMG_FULL, s);
    if ((threadIdx.x & 31) == 0 && s) atomicAdd(total, (unsigned long long)s);
}

int mmg_mean_stencil_points(const mmg_map* cm, const float* xs, const float* ys, const float* zs, size_t n, int mem, void* stream, double* out)
{
    MMG_CHECK_MAP(cm);
    ScratchScope scratch_scope((cudaStream_t)stream);
    mmg_map* m = const_cast<mmg_map*>(cm);
    if (!out || m->kind == MMG_SPARSE_TREES) return MMG_ERR_INVALID_ARG;
    *out = 0;
    if (!n) return MMG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const float *dx, *dy, *dz;
    int          rc = stage_in(m, xs, ys, zs, n, mem, st, &dx, &dy, &dz);
    if (rc) return rc;
    unsigned long long* d = (unsigned long long*)m->scratch_get(S_LIK, 64);
    if (!d) return MMG_ERR_OUT_OF_MEM;
    MMG_CUDA(cudaMemsetAsync(d, 0, 8, st));
    MMG_LAUNCH(k_stencil_points, (unsigned)((n + 255) / 256), 256, 0, st, m->dev, dx, dy, dz, (u32)n, m->inv, m->kind == MMG_SPARSE_VOXEL ? 1 : 0, d);
    unsigned long long h = 0;
    MMG_CUDA(cudaMemcpyAsync(&h, d, 8, cudaMemcpyDeviceToHost, st));
    MMG_CUDA(cudaStreamSynchronize(st));
    *out = (double)h / (double)n;
    return MMG_OK;
}

int mmg_set_tuning(const char* key, int value)
{
    static const char* names[T_COUNT] = {"nn_kernel", "nn_task_blocks", "nn_min_pipe", "nn_permute", "radius_chunk_hits", "sv_split_hits", "nn_task_minb", "radius_warp_max", "radius_block_mid", "radius_q16", "radius_topk"};
    for (int i = 0; key && i < T_COUNT; i++)
        if (!strcmp(key, names[i]))
        {
            g_tune[i] = value;
            return MMG_OK;
        }
    set_error("mmg_set_tuning: unknown knob");
    return MMG_ERR_INVALID_ARG;
}

uint64_t    mmg_kernel_launch_count(void) { return g_launches; }
const char* mmg_last_error(void) { return g_err; }
const char* mmg_version(void) { return "mola_b200 maps 0.1 (sm_100a)"; }
}
