// mmg_voxel.cuh — kernels of the two voxel-hashed maps (HashedVoxelPointCloud and
// SparseVoxelPointCloud flavours, selected by a traits type; see mmg_device.cuh).
//
// HBM layout (all flat, pointer-free, so a map replicates with plain broadcasts):
//   table : Entry[cap]           open addressing, linear probing, 16 B entries {key, vid, cnt}
//   pts   : float4[rowcap*CAP]   row vid holds the voxel's points in slot (= insertion) order
//   cnt   : u32[rowcap]          points stored in row vid
//   vkey  : u64[rowcap]          packed voxel index of row vid (KEY_EMPTY = free row)
//   mean  : float4[rowcap]       running mean (SparseVoxel only)
#pragma once
#include "mmg_device.cuh"
#include "mmg_sort.cuh"

namespace mmg
{
struct Counters
{
    u32 nvox_hi;   // rows ever allocated (high-water mark)
    u32 free_n;    // entries in the free-row stack
    u32 oob;       // points whose voxel index left the packed-key range
    u32 nseg;      // segments of the current insert batch
    u32 bbox[6];   // ordered-uint min xyz / max xyz of points that passed the insertion checks
    u32 nlive_out; // export cursor
    u32 done;      // blocks of k_insert_commit that have finished (the last one mirrors the counters to the host)
    u32 pad[4];
};

struct VoxelMapDev
{
    Entry*  table;
    u64     mask;  // cap - 1
    float4* pts;
    u32*    cnt;
    u64*    vkey;
    float4* mean;
    u32*    free_list;
    u32*    fill;  // per row: free slots handed out by the running insert batch (zero outside k_insert_rank/_commit)
    Counters* ctr;
};

// ------------------------------------------------------------ table upkeep
__device__ __forceinline__ u32 alloc_row(const VoxelMapDev& m)
{
    const int f = (int)atomicSub(&m.ctr->free_n, 1u);
    if (f > 0) return m.free_list[f - 1];
    atomicAdd(&m.ctr->free_n, 1u);
    return atomicAdd(&m.ctr->nvox_hi, 1u);
}

// Find-or-insert with 64-bit atomicCAS on the key; the winner publishes the row id
// afterwards, losers that meet the key before that spin on the (volatile) vid.
__device__ __forceinline__ u32 find_or_insert(const VoxelMapDev& m, u64 key)
{
    u64 h = hash_key(key) & m.mask;
    for (;;)
    {
        Entry* e = m.table + h;
        u64    k = *reinterpret_cast<volatile u64*>(&e->key);
        if (k == KEY_EMPTY)
        {
            k = atomicCAS(reinterpret_cast<u64*>(&e->key), KEY_EMPTY, key);
            if (k == KEY_EMPTY)
            {
                const u32 vid = alloc_row(m);
                e->cnt        = 0;
                __threadfence();
                atomicExch(&e->vid, vid);
                return vid;
            }
        }
        if (k == key)
        {
            u32 v;
            while ((v = *reinterpret_cast<volatile u32*>(&e->vid)) == VID_INVALID) {}
            return v;
        }
        h = (h + 1) & m.mask;
    }
}

__device__ __forceinline__ void warp_bbox_atomics(Counters* ctr, bool ok, float x, float y, float z)
{
    u32 mn[3] = {ok ? float_to_ordered(x) : 0xFFFFFFFFu, ok ? float_to_ordered(y) : 0xFFFFFFFFu, ok ? float_to_ordered(z) : 0xFFFFFFFFu};
    u32 mx[3] = {ok ? float_to_ordered(x) : 0u, ok ? float_to_ordered(y) : 0u, ok ? float_to_ordered(z) : 0u};
#pragma unroll
    for (int a = 0; a < 3; a++)
    {
        mn[a] = __reduce_min_sync(MMG_FULL, mn[a]);
        mx[a] = __reduce_max_sync(MMG_FULL, mx[a]);
    }
    if (lane_id() == 0)
    {
#pragma unroll
        for (int a = 0; a < 3; a++)
        {
            if (mn[a] != 0xFFFFFFFFu) atomicMin(&ctr->bbox[a], mn[a]);
            if (mx[a] != 0u) atomicMax(&ctr->bbox[3 + a], mx[a]);
        }
    }
}

// K1: compose + voxel index + find-or-insert. One thread per point; equal keys in a
// warp are aggregated with __match_any_sync so one lane probes for all of them.
template <class Tr>
__global__ void __launch_bounds__(256)
    k_locate(const float* __restrict__ xs, const float* __restrict__ ys, const float* __restrict__ zs, u32 n, PoseD pose, float inv,
             VoxelMapDev m, float4* __restrict__ gp, u32* __restrict__ vid_out, int bbox_all)
{
    const u32  i     = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < n;
    float3     g     = make_float3(0.f, 0.f, 0.f);
    int        cx = 0, cy = 0, cz = 0;
    bool       ok = false;
    if (valid)
    {
        g  = compose_point(pose, xs[i], ys[i], zs[i]);
        ok = Tr::to_index(g.x, inv, cx) & Tr::to_index(g.y, inv, cy) & Tr::to_index(g.z, inv, cz);
        gp[i] = make_float4(g.x, g.y, g.z, 0.f);
        if (!ok) atomicAdd(&m.ctr->oob, 1u);
    }
    const u64 key    = ok ? pack_key(cx, cy, cz) : KEY_EMPTY;
    const u32 peers  = __match_any_sync(MMG_FULL, key);
    const int leader = __ffs(peers) - 1;
    u32       vid    = VID_INVALID;
    if (ok && lane_id() == leader) vid = find_or_insert(m, key);
    vid = __shfl_sync(MMG_FULL, vid, leader);
    if (valid) vid_out[i] = ok ? vid : VID_INVALID;
    if (bbox_all) warp_bbox_atomics(m.ctr, ok, g.x, g.y, g.z);
}

// Segment heads of the sorted (vid, index) list -> seg_start[] (any order).
static __global__ void k_segment_heads(const u32* __restrict__ svid, u32 n, u32 invalid_from, u32* __restrict__ seg_start, Counters* ctr)
{
    const u32 j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const u32 v = svid[j];
    if (v >= invalid_from) return;
    if (j == 0 || svid[j - 1] != v) seg_start[atomicAdd(&ctr->nseg, 1u)] = j;
}

template <class Tr>
__device__ __forceinline__ void publish_count(const VoxelMapDev& m, u32 vid, u64 key, u32 newcnt)
{
    m.cnt[vid]  = newcnt;
    m.vkey[vid] = key;
    u64 h       = hash_key(key) & m.mask;
    for (;;)
    {
        if (m.table[h].key == key)
        {
            m.table[h].cnt = newcnt;
            return;
        }
        h = (h + 1) & m.mask;
    }
}

// K3: one warp per touched voxel folds that voxel's batch points in input order:
//   accept while n < eff_cap (max_points_per_voxel / hard cap), optional
//   min-distance test against the points already in the voxel (sequential),
//   SparseVoxel running mean. Restates insertPoint(): HashedVoxelPointCloud.cpp:588-626,
//   SparseVoxelPointCloud.cpp:193-204,638-655.
template <class Tr>
__global__ void __launch_bounds__(256)
    k_fold(VoxelMapDev m, const u32* __restrict__ svid, const u32* __restrict__ sidx, const u32* __restrict__ seg_start, u32 n,
           const float4* __restrict__ gp, float inv, u32 max_ppv /*0 = none*/, float min_dist_sqr /*<=0 = off*/, int bbox_accepted,
           int with_mean)
{
    constexpr int CAP    = Tr::CAP;
    const int     lane   = lane_id();
    const u32     nwarps = (gridDim.x * blockDim.x) >> 5;
    const u32     nseg   = m.ctr->nseg;
    for (u32 s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < nseg; s += nwarps)
    {
        const u32 j0  = seg_start[s];
        const u32 vid = svid[j0];
        u32       cnt = m.cnt[vid];
        const u32 n0  = cnt;
        // the voxel key, recomputed from the first point of the segment
        const float4 p0 = gp[sidx[j0]];
        int          cx, cy, cz;
        Tr::to_index(p0.x, inv, cx);
        Tr::to_index(p0.y, inv, cy);
        Tr::to_index(p0.z, inv, cz);
        const u64 key = pack_key(cx, cy, cz);
        u32 bmn[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, bmx[3] = {0u, 0u, 0u};
        float4 mean = make_float4(0.f, 0.f, 0.f, 0.f);
        if (with_mean && n0 > 0) mean = m.mean[vid];

        if (!(min_dist_sqr > 0.f))
        {
            // closed form: the first (eff_cap - n0) points of the segment are stored
            const u32  eff  = (max_ppv != 0 && max_ppv < (u32)CAP) ? max_ppv : (u32)CAP;
            const u32  room = n0 < eff ? eff - n0 : 0u;
            const u32  j    = j0 + lane;
            const bool take = (u32)lane < room && j < n && svid[j] == vid;
            float4     p    = make_float4(0.f, 0.f, 0.f, 0.f);
            if (take)
            {
                p = gp[sidx[j]];
                p.w = 0.f;
                m.pts[(size_t)vid * CAP + n0 + lane] = p;
            }
            const u32 macc = __popc(__ballot_sync(MMG_FULL, take));
            if (bbox_accepted) warp_bbox_atomics(m.ctr, take, p.x, p.y, p.z);
            if (with_mean)
                for (u32 t = 0; t < macc; t++)
                {
                    const float px = __shfl_sync(MMG_FULL, p.x, t), py = __shfl_sync(MMG_FULL, p.y, t), pz = __shfl_sync(MMG_FULL, p.z, t);
                    const float nf = (float)(n0 + t), k = __fdiv_rn(1.0f, (float)(n0 + t + 1));
                    mean.x = __fmul_rn(__fadd_rn(__fmul_rn(nf, mean.x), px), k);
                    mean.y = __fmul_rn(__fadd_rn(__fmul_rn(nf, mean.y), py), k);
                    mean.z = __fmul_rn(__fadd_rn(__fmul_rn(nf, mean.z), pz), k);
                }
            cnt = n0 + macc;
        }
        else
        {
            // sequential: lane l holds stored point l; every candidate is tested against all of them
            float4 mine = make_float4(0.f, 0.f, 0.f, 0.f);
            if ((u32)lane < n0) mine = m.pts[(size_t)vid * CAP + lane];
            bool done = false;
            for (u32 j = j0; !done; j += 32)
            {
                const u32  jj  = j + lane;
                const bool has = jj < n && svid[jj] == vid;
                const u32  nh  = __popc(__ballot_sync(MMG_FULL, has));
                float4     c   = make_float4(0.f, 0.f, 0.f, 0.f);
                if (has) c = gp[sidx[jj]];
                for (u32 t = 0; t < nh; t++)
                {
                    if (max_ppv != 0 && cnt >= max_ppv)
                    {
                        done = true;
                        break;
                    }
                    const float px = __shfl_sync(MMG_FULL, c.x, t), py = __shfl_sync(MMG_FULL, c.y, t), pz = __shfl_sync(MMG_FULL, c.z, t);
                    if (cnt > 0)
                    {
                        const bool close = (u32)lane < cnt && sqr_dist(mine.x, mine.y, mine.z, px, py, pz) < min_dist_sqr;
                        if (__any_sync(MMG_FULL, close)) continue;
                    }
                    // passed both tests: counts for the bounding box even if the 32-slot row is full
                    bmn[0] = min(bmn[0], float_to_ordered(px)), bmn[1] = min(bmn[1], float_to_ordered(py)), bmn[2] = min(bmn[2], float_to_ordered(pz));
                    bmx[0] = max(bmx[0], float_to_ordered(px)), bmx[1] = max(bmx[1], float_to_ordered(py)), bmx[2] = max(bmx[2], float_to_ordered(pz));
                    if (cnt < (u32)CAP)
                    {
                        if ((u32)lane == cnt)
                        {
                            mine = make_float4(px, py, pz, 0.f);
                            m.pts[(size_t)vid * CAP + cnt] = mine;
                        }
                        cnt++;
                    }
                }
                if (nh < 32) done = true;
            }
            if (lane == 0 && bmn[0] != 0xFFFFFFFFu)
            {
#pragma unroll
                for (int a = 0; a < 3; a++)
                {
                    atomicMin(&m.ctr->bbox[a], bmn[a]);
                    atomicMax(&m.ctr->bbox[3 + a], bmx[a]);
                }
            }
        }
        if (lane == 0)
        {
            publish_count<Tr>(m, vid, key, cnt);
            if (with_mean) m.mean[vid] = mean;
        }
    }
}

// Table rebuild from the live rows (after growth or eviction).
static __global__ void k_rebuild_table(VoxelMapDev m, u32 nrows)
{
    const u32 vid = blockIdx.x * blockDim.x + threadIdx.x;
    if (vid >= nrows) return;
    const u64 key = m.vkey[vid];
    if (key == KEY_EMPTY) return;
    u64 h = hash_key(key) & m.mask;
    for (;;)
    {
        Entry* e = m.table + h;
        if (atomicCAS(reinterpret_cast<u64*>(&e->key), KEY_EMPTY, key) == KEY_EMPTY)
        {
            e->vid = vid;
            e->cnt = m.cnt[vid];
            return;
        }
        h = (h + 1) & m.mask;
    }
}

// Far-voxel eviction (Chebyshev distance in voxel indices; HashedVoxelPointCloud.cpp:305-326).
static __global__ void k_evict(VoxelMapDev m, u32 nrows, int ox, int oy, int oz, int dist_in_grid)
{
    const u32 vid = blockIdx.x * blockDim.x + threadIdx.x;
    if (vid >= nrows) return;
    const u64 key = m.vkey[vid];
    if (key == KEY_EMPTY) return;
    int cx, cy, cz;
    unpack_key(key, cx, cy, cz);
    const int d = max(max(abs(cx - ox), abs(cy - oy)), abs(cz - oz));
    if (d > dist_in_grid)
    {
        m.vkey[vid] = KEY_EMPTY;
        m.cnt[vid]  = 0;
        m.free_list[atomicAdd(&m.ctr->free_n, 1u)] = vid;
    }
}

// Export of the live rows in row-id order: live[vid] flags -> exclusive scan -> rank[vid].
static __global__ void k_live_flags(VoxelMapDev m, u32 nrows, u32* __restrict__ live)
{
    const u32 vid = blockIdx.x * blockDim.x + threadIdx.x;
    if (vid < nrows) live[vid] = (m.vkey[vid] != KEY_EMPTY) ? 1u : 0u;
}
// One warp per row. Optional outputs: points (V*CAP*3), means (V*3), cov (V*6; extension N4:
// 1/(n-1) sum (p-mean)(p-mean)^T accumulated in slot order in fp32, zero when n < 2).
template <class Tr>
__global__ void k_export(VoxelMapDev m, u32 nrows, const u32* __restrict__ rank, int32_t* __restrict__ keys, u32* __restrict__ counts,
                         float* __restrict__ points, float* __restrict__ means, float* __restrict__ cov)
{
    constexpr int CAP  = Tr::CAP;
    const int     lane = lane_id();
    const u32     vid  = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (vid >= nrows) return;
    const u64 key = m.vkey[vid];
    if (key == KEY_EMPTY) return;
    const size_t o = rank[vid];
    const u32    c = m.cnt[vid];
    float4       p = make_float4(0.f, 0.f, 0.f, 0.f);
    if (lane < CAP && (u32)lane < c) p = m.pts[(size_t)vid * CAP + lane];
    if (lane == 0)
    {
        if (keys)
        {
            int cx, cy, cz;
            unpack_key(key, cx, cy, cz);
            keys[3 * o] = cx, keys[3 * o + 1] = cy, keys[3 * o + 2] = cz;
        }
        if (counts) counts[o] = c;
    }
    if (points && lane < CAP)
    {
        float* row = points + (o * CAP + lane) * 3;
        row[0] = p.x, row[1] = p.y, row[2] = p.z;
    }
    if (means || cov)
    {
        const float4 mu = m.mean[vid];
        if (means && lane == 0) means[3 * o] = mu.x, means[3 * o + 1] = mu.y, means[3 * o + 2] = mu.z;
        if (cov)
        {
            float a[6] = {0, 0, 0, 0, 0, 0};
            for (u32 s = 0; s < c; s++)
            {
                const float dx = __fsub_rn(__shfl_sync(MMG_FULL, p.x, s), mu.x), dy = __fsub_rn(__shfl_sync(MMG_FULL, p.y, s), mu.y),
                            dz = __fsub_rn(__shfl_sync(MMG_FULL, p.z, s), mu.z);
                a[0] = __fadd_rn(a[0], __fmul_rn(dx, dx)), a[1] = __fadd_rn(a[1], __fmul_rn(dx, dy)), a[2] = __fadd_rn(a[2], __fmul_rn(dx, dz));
                a[3] = __fadd_rn(a[3], __fmul_rn(dy, dy)), a[4] = __fadd_rn(a[4], __fmul_rn(dy, dz)), a[5] = __fadd_rn(a[5], __fmul_rn(dz, dz));
            }
            const float k = c > 1 ? __fdiv_rn(1.0f, (float)(c - 1)) : 0.0f;
            if (lane < 6) cov[6 * o + lane] = __fmul_rn(a[lane], k);
        }
    }
}

// Bounding box from voxel corners (Hashed boundingBox() :872-902) or from the 32^3
// outer blocks (SparseVoxel boundingBox(), SparseVoxelPointCloud.cpp:800-832).
template <class Tr>
__global__ void k_bbox_from_voxels(VoxelMapDev m, u32 nrows, float vsize, u32* __restrict__ out6)
{
    const u32 vid = blockIdx.x * blockDim.x + threadIdx.x;
    bool      ok  = false;
    float     lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    if (vid < nrows)
    {
        const u64 key = m.vkey[vid];
        if (key != KEY_EMPTY)
        {
            ok = true;
            int c[3];
            unpack_key(key, c[0], c[1], c[2]);
#pragma unroll
            for (int a = 0; a < 3; a++)
            {
                if (Tr::CAP == 32)
                {
                    lo[a] = __fmul_rn((float)c[a], vsize);
                    hi[a] = __fadd_rn(lo[a], __fmul_rn(1.0f, vsize));
                }
                else
                {
                    const float half = __fmul_rn(0.5f, vsize);
                    const float gmh  = __fsub_rn(__fmul_rn(1.0f, __fmul_rn(vsize, 32.0f)), __fmul_rn(1.0f, half));
                    const float ctr  = __fmul_rn((float)(c[a] & ~31), vsize);
                    lo[a]            = __fsub_rn(ctr, __fmul_rn(1.0f, half));
                    hi[a]            = __fadd_rn(ctr, gmh);
                }
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 3; a++)
    {
        // both corners feed both min and max, exactly like two updateWithPoint calls
        const u32 l = ok ? float_to_ordered(fminf(lo[a], hi[a])) : 0xFFFFFFFFu;
        const u32 h = ok ? float_to_ordered(fmaxf(lo[a], hi[a])) : 0u;
        const u32 rl = __reduce_min_sync(MMG_FULL, l), rh = __reduce_max_sync(MMG_FULL, h);
        if (lane_id() == 0)
        {
            if (rl != 0xFFFFFFFFu) atomicMin(&out6[a], rl);
            if (rh != 0u) atomicMax(&out6[3 + a], rh);
        }
    }
}

// ----------------------------------------------------------------- queries
struct QueryOut
{
    uint8_t* found;
    float*   xyz;
    float*   d2;
    u64*     id;
};

// nn_single_search: one warp per query. Lanes 0..26 probe the 27 stencil cells
// (lane = (dx+1)*9 + (dy+1)*3 + (dz+1), i.e. the reference's cx,cy,cz loop order,
// HashedVoxelPointCloud.cpp:758-761); then the warp walks the non-empty cells in that
// order, lane l loading slot l of the row with one LDG.128, and keeps per lane the
// first strict minimum. The final shuffle reduction minimises (d2, cell*32+slot), which
// is exactly "first candidate in scan order wins ties" (strict '<' at :728).
template <class Tr>
__global__ void __launch_bounds__(256)
    k_nn_single(VoxelMapDev m, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 n, u32 stride,
                PoseD pose, int use_pose, float inv, QueryOut out)
{
    constexpr int CAP    = Tr::CAP;
    const int     lane   = lane_id();
    const u32     nwarps = (gridDim.x * blockDim.x) >> 5;
    const int     ddx = lane / 9 - 1, ddy = (lane / 3) % 3 - 1, ddz = lane % 3 - 1;
    for (u32 q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; q < n; q += nwarps)
    {
        const size_t qi = (size_t)q * stride;
        float3       p  = make_float3(qx[qi], qy[qi], qz[qi]);
        if (use_pose) p = compose_point(pose, p.x, p.y, p.z);
        int  cx, cy, cz;
        bool ok = Tr::to_index(p.x, inv, cx) & Tr::to_index(p.y, inv, cy) & Tr::to_index(p.z, inv, cz);
        u32  vid = 0, cnt = 0;
        if (ok && lane < 27)
        {
            const int x = cx + ddx, y = cy + ddy, z = cz + ddz;
            if (index_in_range(x, y, z))
                if (!probe_ro(m.table, m.mask, pack_key(x, y, z), vid, cnt)) cnt = 0;
        }
        u32    cells = __ballot_sync(MMG_FULL, cnt > 0);
        float  bd    = __int_as_float(0x7f800000);
        u32    br    = 0xFFFFFFFFu;
        float4 bp    = make_float4(0.f, 0.f, 0.f, 0.f);
        while (cells)
        {
            const int b = __ffs(cells) - 1;
            cells &= cells - 1;
            const u32 v = __shfl_sync(MMG_FULL, vid, b), c = __shfl_sync(MMG_FULL, cnt, b);
            if ((u32)lane < c)
            {
                const float4 t = __ldg(&m.pts[(size_t)v * CAP + lane]);
                const float  d = sqr_dist(t.x, t.y, t.z, p.x, p.y, p.z);
                if (d < bd || br == 0xFFFFFFFFu)
                {
                    bd = d;
                    br = (u32)b * 32u + (u32)lane;
                    bp = t;
                }
            }
        }
        float rd = bd;
        u32   rr = br;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
        {
            const float od = __shfl_xor_sync(MMG_FULL, rd, o);
            const u32   orr = __shfl_xor_sync(MMG_FULL, rr, o);
            if (orr != 0xFFFFFFFFu && (rr == 0xFFFFFFFFu || od < rd || (od == rd && orr < rr)))
            {
                rd = od;
                rr = orr;
            }
        }
        if (lane == 0)
        {
            out.found[q] = (rr != 0xFFFFFFFFu) ? 1 : 0;
            if (rr == 0xFFFFFFFFu)
            {
                out.xyz[3 * (size_t)q] = 0.f, out.xyz[3 * (size_t)q + 1] = 0.f, out.xyz[3 * (size_t)q + 2] = 0.f;
                out.d2[q] = 0.f;
                out.id[q] = 0;
            }
        }
        if (rr != 0xFFFFFFFFu && br == rr)
        {
            const int b = (int)(rr >> 5), slot = (int)(rr & 31);
            out.xyz[3 * (size_t)q] = bp.x, out.xyz[3 * (size_t)q + 1] = bp.y, out.xyz[3 * (size_t)q + 2] = bp.z;
            out.d2[q] = rd;
            out.id[q] = plain_id<Tr::SLOT_BITS>(cx + b / 9 - 1, cy + (b / 3) % 3 - 1, cz + b % 3 - 1, slot);
        }
    }
}

// Candidate visitor shared by the per-thread kernels: calls f(d2, point, cell index
// triple, slot) for every stored point of every cell in [a, b], cx outermost.
template <class Tr, class F>
__device__ __forceinline__ void visit_cells(const VoxelMapDev& m, int ax, int ay, int az, int bx, int by, int bz, F&& f)
{
    constexpr int CAP = Tr::CAP;
    for (int x = ax; x <= bx; x++)
        for (int y = ay; y <= by; y++)
            for (int z = az; z <= bz; z++)
            {
                if (!index_in_range(x, y, z)) continue;
                u32 vid, cnt;
                if (!probe_ro(m.table, m.mask, pack_key(x, y, z), vid, cnt)) continue;
                for (u32 s = 0; s < cnt; s++) f(x, y, z, (int)s, vid, __ldg(&m.pts[(size_t)vid * CAP + s]));
            }
}

// Conservative squared distance from coordinate q to the slab of cell index c along one axis (the cell's
// coordinate span widened by a few ulps and the square scaled down, like voxel_lower_bound()).
template <class Tr>
__device__ __forceinline__ float axis_gap(int c, float vsize, float q)
{
    float lo, hi;
    Tr::cell_span(c, lo, hi);
    const float eps = (fabsf((float)c) + 2.0f) * vsize * 1e-6f;
    const float d   = fmaxf(0.f, fmaxf(lo * vsize - eps - q, q - (hi * vsize + eps)));
    return d * d * 0.9999f;
}

// visit_cells in the same (reference) order, but a cell is skipped, before its hash probe, when its box lies
// farther from p than bound(): exact as long as every candidate with d2 > bound() is a no-op for f at that moment.
template <class Tr, class B, class F>
__device__ __forceinline__ void visit_cells_bounded(const VoxelMapDev& m, int ax, int ay, int az, int bx, int by, int bz, const float3& p, float vsize,
                                                    B&& bound, F&& f)
{
    constexpr int CAP = Tr::CAP;
    for (int x = ax; x <= bx; x++)
    {
        const float gx = axis_gap<Tr>(x, vsize, p.x);
        if (gx > bound()) continue;
        for (int y = ay; y <= by; y++)
        {
            const float gxy = gx + axis_gap<Tr>(y, vsize, p.y);
            if (gxy > bound()) continue;
            for (int z = az; z <= bz; z++)
            {
                if (gxy + axis_gap<Tr>(z, vsize, p.z) > bound()) continue;
                if (!index_in_range(x, y, z)) continue;
                u32 vid, cnt;
                if (!probe_ro(m.table, m.mask, pack_key(x, y, z), vid, cnt)) continue;
                for (u32 s = 0; s < cnt; s++) f(x, y, z, (int)s, vid, __ldg(&m.pts[(size_t)vid * CAP + s]));
            }
        }
    }
}

struct Hit3
{
    float x, y, z, d2;
    u64   id;
};
// The reference's K-slot sorted insert with its forward-copy shift,
// HashedVoxelPointCloud.cpp:720-741. EXACT=false gives a true stable top-K instead.
template <int K, bool REF_SHIFT>
struct Slots
{
    Hit3 m[K];
    int  found = 0;
    __device__ __forceinline__ void offer(float d2, float x, float y, float z, u64 id)
    {
        int i = 0;
        while (i < found && !(d2 < m[i].d2)) i++;
        if (i >= K) return;
        const int last = min(found + 1, K);
        if (REF_SHIFT)
            for (int j = i + 1; j < last; j++) m[j] = m[j - 1];
        else
            for (int j = last - 1; j > i; j--) m[j] = m[j - 1];
        m[i] = {x, y, z, d2, id};
        if (found < K) found++;
    }
};

// nn_multiple_search (Hashed: K=3 with the reference's shift; SparseVoxel extension: K=8 true top-k).
template <class Tr, int K, bool REF_SHIFT>
__global__ void __launch_bounds__(128)
    k_nn_multi(VoxelMapDev m, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 n, PoseD pose,
               int use_pose, float inv, float vsize, u32 kreq, u32* __restrict__ ocount, float* __restrict__ oxyz, float* __restrict__ od2,
               u64* __restrict__ oid)
{
    const u32 q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    float3 p = make_float3(qx[q], qy[q], qz[q]);
    if (use_pose) p = compose_point(pose, p.x, p.y, p.z);
    int cx, cy, cz;
    const bool ok = Tr::to_index(p.x, inv, cx) & Tr::to_index(p.y, inv, cy) & Tr::to_index(p.z, inv, cz);
    Slots<K, REF_SHIFT> s;
    // Reference visiting order (the K=3 shift defect makes the third slot order dependent). A candidate that is
    // not strictly below the current last slot leaves the slots untouched (offer() walks past all K), and slot values never
    // grow, so once K slots are filled any cell whose box is farther than slot K-1 can be skipped without changing anything.
    if (ok)
        visit_cells_bounded<Tr>(
            m, cx - 1, cy - 1, cz - 1, cx + 1, cy + 1, cz + 1, p, vsize,
            [&]() { return s.found == K ? s.m[K - 1].d2 : __int_as_float(0x7f800000); },
            [&](int x, int y, int z, int slot, u32, const float4& t) {
                s.offer(sqr_dist(t.x, t.y, t.z, p.x, p.y, p.z), t.x, t.y, t.z, plain_id<Tr::SLOT_BITS>(x, y, z, slot));
            });
    const u32 nout = min(kreq, (u32)s.found);
    ocount[q]      = nout;
    for (u32 i = 0; i < nout; i++)
    {
        const size_t o = (size_t)q * kreq + i;
        oxyz[3 * o] = s.m[i].x, oxyz[3 * o + 1] = s.m[i].y, oxyz[3 * o + 2] = s.m[i].z;
        od2[o] = s.m[i].d2, oid[o] = s.m[i].id;
    }
}

// nn_radius_search, Hashed flavour (HashedVoxelPointCloud.cpp:773-870). Pass 0 counts,
// pass 1 fills at offsets[q]. maxPoints==0: all hits in scan order; else the 3-slot buffer.
template <class Tr, int PASS>
__global__ void __launch_bounds__(128)
    k_radius_hashed(VoxelMapDev m, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 n, PoseD pose,
                    int use_pose, float inv, float vsize, float r2, float r, u32 maxp, u32* __restrict__ counts, const u64* __restrict__ offsets,
                    float* __restrict__ oxyz, float* __restrict__ od2, u64* __restrict__ oid)
{
    const u32 q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    float3 p = make_float3(qx[q], qy[q], qz[q]);
    if (use_pose) p = compose_point(pose, p.x, p.y, p.z);
    int  a[3], b[3];
    bool ok = Tr::to_index(__fsub_rn(p.x, r), inv, a[0]) & Tr::to_index(__fsub_rn(p.y, r), inv, a[1]) & Tr::to_index(__fsub_rn(p.z, r), inv, a[2]) &
              Tr::to_index(__fadd_rn(p.x, r), inv, b[0]) & Tr::to_index(__fadd_rn(p.y, r), inv, b[1]) & Tr::to_index(__fadd_rn(p.z, r), inv, b[2]);
    u32  c = 0;
    u64  o = PASS ? offsets[q] : 0;
    Slots<3, true> s;
    // cells of the reference's index box that lie wholly outside the sphere hold only d > r2 candidates, which the
    // reference drops (:832): skipping them changes nothing
    if (ok && r2 > 0.f)
        visit_cells_bounded<Tr>(m, a[0], a[1], a[2], b[0], b[1], b[2], p, vsize, [&]() { return r2; }, [&](int x, int y, int z, int slot, u32, const float4& t) {
            const float d = sqr_dist(t.x, t.y, t.z, p.x, p.y, p.z);
            if (d > r2) return;
            if (maxp != 0) { s.offer(d, t.x, t.y, t.z, plain_id<Tr::SLOT_BITS>(x, y, z, slot)); return; }
            if (PASS)
            {
                oxyz[3 * (o + c)] = t.x, oxyz[3 * (o + c) + 1] = t.y, oxyz[3 * (o + c) + 2] = t.z;
                od2[o + c] = d, oid[o + c] = plain_id<Tr::SLOT_BITS>(x, y, z, slot);
            }
            c++;
        });
    if (maxp != 0)
    {
        c = min(maxp, (u32)s.found);
        if (PASS)
            for (u32 i = 0; i < c; i++)
            {
                oxyz[3 * (o + i)] = s.m[i].x, oxyz[3 * (o + i) + 1] = s.m[i].y, oxyz[3 * (o + i) + 2] = s.m[i].z;
                od2[o + i] = s.m[i].d2, oid[o + i] = s.m[i].id;
            }
    }
    if (!PASS) counts[q] = c;
}

// Likelihood tail: sum over the decimated queries of min(d2, max_sqr) in fp64. Fixed summation shape (block b takes the
// elements b*1024+t, stride gridDim*1024; warp shuffles, then the warps, then the blocks in order): the same inputs give
// the same bits on every run.
__device__ __forceinline__ double block_sum_1024(double s, double* sm)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(MMG_FULL, s, o);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
    __syncthreads();
    double r = 0;
    if (threadIdx.x < 32)
    {
        r = threadIdx.x < (blockDim.x >> 5) ? sm[threadIdx.x] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) r += __shfl_down_sync(MMG_FULL, r, o);
    }
    return r;
}
static __global__ void __launch_bounds__(1024) k_likelihood_partial(const uint8_t* __restrict__ found, const float* __restrict__ d2, u32 n, float max_sqr,
                                                                    double* __restrict__ partial)
{
    __shared__ double sm[32];
    double            s = 0;
    for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        if (found[i]) s += (double)fminf(d2[i], max_sqr);
    s = block_sum_1024(s, sm);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
}
static __global__ void __launch_bounds__(256) k_likelihood_final(const double* __restrict__ partial, u32 nb, double* __restrict__ out)
{
    __shared__ double sm[32];
    double            s = threadIdx.x < nb ? partial[threadIdx.x] : 0.0;
    s                   = block_sum_1024(s, sm);
    if (threadIdx.x == 0) *out = s;
}

}  // namespace mmg
