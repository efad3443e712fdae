// mmg_insert.cuh — sort-free batched insertion for the cap-only decimation mode
// (min_distance_between_points == 0, the reference default).
//
// Reference semantics (insertPoint called once per point in input order,
// HashedVoxelPointCloud.cpp:588-626, SparseVoxelPointCloud.cpp:638-655): a voxel holding n0
// points keeps the FIRST room = eff_cap - n0 points of the batch that fall into it, in input
// order. So per voxel the job is "the room smallest input indices, sorted" — a bounded top-k,
// not a sort of the whole batch.
//
// k_insert_rank (one thread per point): compose, voxel index, warp-aggregated find-or-insert,
// then every point that can still make it offers its index to the voxel's FREE row slots with a
// cascade of atomicMax: slot s ends up holding the s-th smallest index offered (each slot keeps
// the maximum of ~index it ever sees and the displaced value moves on to the next slot; the
// multiset {slots} U {carried} is conserved, so the result is independent of the thread
// schedule). The free slots' float4.w lanes serve as the scratch (0 = empty; stored points keep
// w == 0), so the ranking needs no memory of its own. Three exact prunings keep it cheap:
// full voxels drop their points right after the probe; a point with w earlier same-voxel peers
// in its warp starts at slot w (>= its final rank); a point later than the voxel's last free
// slot's current holder is dropped (slots only ever move to earlier indices); and the prefix of
// slots that already hold earlier points is skipped with plain loads before the first atomic.
//
// k_insert_commit (one warp per touched voxel): turns the ranked indices into stored points
// (compose is recomputed, it is deterministic), running mean for SparseVoxel, bounding box,
// count publication. Both kernels take several scans with one pose each (global index = input
// order), which equals inserting the scans one after the other as long as no eviction runs
// between them.
#pragma once
#include "mmg_voxel.cuh"

namespace mmg
{
struct ScanSet
{
    const PoseD* poses;    // nscans poses (device), or nullptr: `single` applies to every point
    const u32*   offsets;  // nscans + 1 ascending point offsets (device)
    u32          nscans;
    PoseD        single;
};

__device__ __forceinline__ const PoseD& scan_pose(const ScanSet& S, u32 gi)
{
    if (!S.poses) return S.single;
    u32 lo = 0, hi = S.nscans;  // last scan with offsets[s] <= gi
    while (hi - lo > 1)
    {
        const u32 mid = (lo + hi) >> 1;
        if (__ldg(S.offsets + mid) <= gi) lo = mid; else hi = mid;
    }
    return S.poses[lo];
}

// scan of point gi, given the scan of an earlier point of the same block (scans are long, so this is 0 or 1 step)
__device__ __forceinline__ u32 scan_from(const ScanSet& S, u32 s0, u32 gi)
{
    while (s0 + 1 < S.nscans && __ldg(S.offsets + s0 + 1) <= gi) s0++;
    return s0;
}
__device__ __forceinline__ u32 scan_index(const ScanSet& S, u32 gi)
{
    u32 lo = 0, hi = S.nscans;
    while (hi - lo > 1)
    {
        const u32 mid = (lo + hi) >> 1;
        if (__ldg(S.offsets + mid) <= gi) lo = mid; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ u32* slot_w(float4* p) { return reinterpret_cast<u32*>(p) + 3; }

// find-or-insert that also returns the voxel's published point count. One 16-byte L2 load (ld.cg) per probe returns
// key, row id and count together. A voxel created by THIS batch has count 0 until k_insert_commit runs, so a reader that
// has to wait for the creator's row id (or still sees the 0xFFFFFFFF of the never-used entry) knows the count is 0.
__device__ __forceinline__ u32 find_or_insert_cnt(const VoxelMapDev& m, u64 key, u32& cnt)
{
    u64 h = hash_key(key) & m.mask;
    for (;;)
    {
        Entry*      e = m.table + h;
        const uint4 v = __ldcg(reinterpret_cast<const uint4*>(e));
        u64         k = (u64)v.x | ((u64)v.y << 32);
        if (k == key && v.z != VID_INVALID)
        {
            cnt = v.w == 0xFFFFFFFFu ? 0u : v.w;
            return v.z;
        }
        if (k == KEY_EMPTY)
        {
            k = atomicCAS(reinterpret_cast<u64*>(&e->key), KEY_EMPTY, key);
            if (k == KEY_EMPTY)
            {
                const u32 vid = alloc_row(m);
                e->cnt        = 0;
                __threadfence();
                atomicExch(&e->vid, vid);
                cnt = 0;
                return vid;
            }
        }
        if (k == key)
        {   // being created by another thread of this batch
            u32 r;
            while ((r = *reinterpret_cast<volatile u32*>(&e->vid)) == VID_INVALID) {}
            cnt = 0;
            return r;
        }
        h = (h + 1) & m.mask;
    }
}

// block-wide bounding-box update: warp REDUX, shared memory, six atomics per block
__device__ __forceinline__ void block_bbox_atomics(Counters* ctr, bool ok, float x, float y, float z, u32 (*sm)[6])
{
    u32 v[6] = {ok ? float_to_ordered(x) : 0xFFFFFFFFu, ok ? float_to_ordered(y) : 0xFFFFFFFFu, ok ? float_to_ordered(z) : 0xFFFFFFFFu,
                ok ? float_to_ordered(x) : 0u,          ok ? float_to_ordered(y) : 0u,          ok ? float_to_ordered(z) : 0u};
#pragma unroll
    for (int a = 0; a < 3; a++)
    {
        v[a]     = __reduce_min_sync(MMG_FULL, v[a]);
        v[3 + a] = __reduce_max_sync(MMG_FULL, v[3 + a]);
    }
    const int w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    if (lane_id() == 0)
    {
#pragma unroll
        for (int a = 0; a < 6; a++) sm[w][a] = v[a];
    }
    __syncthreads();
    if (threadIdx.x < 6)
    {
        const int a = threadIdx.x;
        u32       r = sm[0][a];
        for (int i = 1; i < nw; i++) r = a < 3 ? min(r, sm[i][a]) : max(r, sm[i][a]);
        if (a < 3) { if (r != 0xFFFFFFFFu) atomicMin(&ctr->bbox[a], r); }
        else       { if (r != 0u) atomicMax(&ctr->bbox[a], r); }
    }
}

template <class Tr>
__global__ void __launch_bounds__(256)
    k_insert_rank(const float* __restrict__ xs, const float* __restrict__ ys, const float* __restrict__ zs, u32 n, u32 base, ScanSet S, float inv,
                  VoxelMapDev m, u32 eff_cap, u32* __restrict__ touched, int bbox_all)
{
    constexpr int  CAP = Tr::CAP;
    __shared__ u32 sm_bbox[8][6];
    __shared__ u32 sm_scan;
    const u32      i     = blockIdx.x * blockDim.x + threadIdx.x;
    const bool     valid = i < n;
    float3         g     = make_float3(0.f, 0.f, 0.f);
    int            cx = 0, cy = 0, cz = 0;
    bool           ok = false;
    if (S.poses)
    {   // one binary search per block instead of one per point
        if (threadIdx.x == 0) sm_scan = scan_index(S, base + blockIdx.x * blockDim.x);
        __syncthreads();
    }
    if (valid)
    {
        const float lx = xs[i], ly = ys[i], lz = zs[i];
        g = compose_point(S.poses ? S.poses[scan_from(S, sm_scan, base + i)] : S.single, lx, ly, lz);
        ok = Tr::to_index(g.x, inv, cx) & Tr::to_index(g.y, inv, cy) & Tr::to_index(g.z, inv, cz);
        if (!ok) atomicAdd(&m.ctr->oob, 1u);
    }
    // before the probes: every warp reaches the barrier inside at once (after them it would wait for the slowest prober)
    if (bbox_all) block_bbox_atomics(m.ctr, ok, g.x, g.y, g.z, sm_bbox);
    const int lane   = lane_id();
    const u64 key    = ok ? pack_key(cx, cy, cz) : (KEY_EMPTY - 1 - (u64)lane);  // invalid lanes never match each other
    const u32 peers  = __match_any_sync(MMG_FULL, key);
    const int leader = __ffs(peers) - 1;
    u32       vid = VID_INVALID, n0 = 0;
    if (ok && lane == leader) vid = find_or_insert_cnt(m, key, n0);
    vid = __shfl_sync(MMG_FULL, vid, leader);
    n0  = __shfl_sync(MMG_FULL, n0, leader);
    if (ok && n0 < eff_cap)
    {
        const u32 room  = eff_cap - n0;
        u32       s     = __popc(peers & ((1u << lane) - 1u));  // earlier same-voxel points of this warp
        float4*   row   = m.pts + (size_t)vid * CAP + n0;
        u32       x     = ~i;                                   // larger = earlier; never 0 (i < 2^31)
        if (s < room)
        {
            // Plain loads first (issued together): the last free slot (a point later than its holder can never get in: slots
            // only move to earlier indices) and the prefix of slots already held by earlier points (a slot value only ever
            // grows, so this point can never settle there, and that many earlier points of the voxel exist: its final rank is
            // at least that). Points arriving in input order then need ONE atomic instead of walking the filled prefix.
            u32 wl = *reinterpret_cast<volatile u32*>(slot_w(row + room - 1));
            for (;;)
            {
                u32 w[4];
#pragma unroll
                for (int k = 0; k < 4; k++) w[k] = (s + k < room) ? *reinterpret_cast<volatile u32*>(slot_w(row + s + k)) : 0u;
                if (wl > x)
                {
                    s = room;
                    break;
                }
                int adv = 0;
#pragma unroll
                for (int k = 0; k < 4; k++)
                    if (adv == k && w[k] > x) adv = k + 1;
                s += adv;
                if (adv < 4 || s >= room) break;
            }
            for (; s < room; s++)
            {
                const u32 old = atomicMax(slot_w(row + s), x);
                if (old == 0u)
                {   // settled in an empty slot; the one thread that fills slot 0 first lists the voxel
                    if (s == 0) touched[atomicAdd(&m.ctr->nseg, 1u)] = vid;
                    break;
                }
                if (old < x) x = old;  // displaced a later point: carry it on
            }
        }
    }
}

template <class Tr>
__global__ void __launch_bounds__(256)
    k_insert_commit(VoxelMapDev m, const u32* __restrict__ touched, const float* __restrict__ xs, const float* __restrict__ ys,
                    const float* __restrict__ zs, u32 base, ScanSet S, float inv, int bbox_accepted, int with_mean)
{
    constexpr int CAP    = Tr::CAP;
    const int     lane   = lane_id();
    const u32     nwarps = (gridDim.x * blockDim.x) >> 5;
    const u32     nseg   = m.ctr->nseg;
    for (u32 t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < nseg; t += nwarps)
    {
        const u32 vid  = touched[t];
        const u32 n0   = m.cnt[vid];
        float4*   slot = m.pts + (size_t)vid * CAP + n0 + lane;
        u32       w    = 0;
        if (n0 + (u32)lane < (u32)CAP) w = *slot_w(slot);
        const bool take = w != 0u;
        float4     p    = make_float4(0.f, 0.f, 0.f, 0.f);
        if (take)
        {
            const u32    i = ~w;
            const float3 g = compose_point(scan_pose(S, base + i), xs[i], ys[i], zs[i]);
            p              = make_float4(g.x, g.y, g.z, 0.f);
            *slot          = p;  // also clears the ranking scratch in .w
        }
        const u32 macc = __popc(__ballot_sync(MMG_FULL, take));
        if (bbox_accepted) warp_bbox_atomics(m.ctr, take, p.x, p.y, p.z);
        if (with_mean)
        {   // VoxelData::insertPoint running mean, SparseVoxelPointCloud.cpp:193-204 (order fixed)
            float4 mean = make_float4(0.f, 0.f, 0.f, 0.f);
            if (n0 > 0) mean = m.mean[vid];
            for (u32 a = 0; a < macc; a++)
            {
                const float px = __shfl_sync(MMG_FULL, p.x, a), py = __shfl_sync(MMG_FULL, p.y, a), pz = __shfl_sync(MMG_FULL, p.z, a);
                const float nf = (float)(n0 + a), k = __fdiv_rn(1.0f, (float)(n0 + a + 1));
                mean.x = __fmul_rn(__fadd_rn(__fmul_rn(nf, mean.x), px), k);
                mean.y = __fmul_rn(__fadd_rn(__fmul_rn(nf, mean.y), py), k);
                mean.z = __fmul_rn(__fadd_rn(__fmul_rn(nf, mean.z), pz), k);
            }
            if (lane == 0) m.mean[vid] = mean;
        }
        if (lane == 0)
        {   // lane 0 always holds a point: the voxel was listed by the thread that filled this slot
            int cx, cy, cz;
            Tr::to_index(p.x, inv, cx);
            Tr::to_index(p.y, inv, cy);
            Tr::to_index(p.z, inv, cz);
            publish_count<Tr>(m, vid, pack_key(cx, cy, cz), n0 + macc);
        }
    }
}
}  // namespace mmg
