// mmg_insert.cuh — sort-free batched insertion for the cap-only decimation mode
// (min_distance_between_points == 0, the reference default).
//
// Reference semantics (insertPoint called once per point in input order,
// HashedVoxelPointCloud.cpp:588-626, SparseVoxelPointCloud.cpp:638-655): a voxel holding n0
// points keeps the FIRST room = eff_cap - n0 points of the batch that fall into it, in input
// order. So per voxel the job is "the room smallest input indices, sorted" — a bounded top-k,
// not a sort of the whole batch.
//
// k_insert_rank (one thread per point, the lanes of a warp = 32 consecutive points): compose,
// voxel index, warp-aggregated find-or-insert, then the voxel's FREE row slots are used as an
// UNSORTED bounded set of the earliest input indices seen so far. The float4.w lanes of the free
// slots are the scratch (code = ~index, larger = earlier, 0 = empty; stored points keep w == 0),
// so the ranking needs no memory of its own. For one voxel at a time the whole warp snapshots the
// slot codes (lane L = free slot L, one coalesced pass over the row), the voxel's points of this
// warp (ascending lane = ascending index) each claim the currently smallest code of the
// snapshot if their own is larger (REDUX min + ballot, all in registers), and every claimer then
// issues ONE compare-and-swap {seen code -> own code}. A replaced code simply vanishes: at that
// moment cap codes larger than it have been offered, so it cannot be among the voxel's first cap
// points; the same holds for a rejected point. Both facts are about the multiset of offered
// indices, not about the (possibly stale) snapshot, hence the result is independent of the thread
// schedule. A failed CAS (another warp changed that slot) retries on a fresh snapshot. Empty
// slots are the smallest code and ties go to the lowest slot, so the filled slots always form a
// prefix. Two exact prunings keep the common case to the probe alone: full voxels drop their
// points right after it, and once every free slot is taken the smallest code of the set (it only
// ever grows) is published next to the count in the hash entry ({26-bit code, 6-bit count}), so
// later points are rejected by the 16-byte probe load itself.
// (The previous scheme kept the slots sorted with a cascade of atomicMax; points of concurrently
// running warps arrive out of order, and every displaced code walked the rest of the row with
// one dependent L2 atomic per slot — 31 % of the kernel's stall samples, profiles/r1i_*.)
//
// k_insert_commit (one warp per touched voxel): sorts the <= cap codes of the row (rank by
// counting), turns them into stored points in input order (compose is recomputed, it is
// deterministic), running mean for SparseVoxel, bounding box, count publication (which also
// clears the entry's code). Both kernels take several scans with one pose each (global index =
// input order), which equals inserting the scans one after the other as long as no eviction runs
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
    float        scans_per_point;  // nscans / total points: first guess of a point's scan (scans usually have one size)
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

// scan of point gi: equal-size guess, verified; binary search otherwise
__device__ __forceinline__ u32 scan_guess(const ScanSet& S, u32 gi)
{
    const u32 s = min(S.nscans - 1u, (u32)((float)gi * S.scans_per_point));
    if (__ldg(S.offsets + s) <= gi && gi < __ldg(S.offsets + s + 1)) return s;
    return scan_index(S, gi);
}

// Bounding box of all valid points, barrier-free: the current box is read once (a stale copy is only ever smaller),
// and only warps holding a point outside it reduce and issue atomics — none in steady state.
struct BoxNow
{
    uint4 a;  // min x, y, z, max x   (ordered-uint form)
    uint2 b;  // max y, z
};
__device__ __forceinline__ BoxNow load_box(const Counters* ctr)
{
    BoxNow B;
    // plain (L1-cached) loads: one L2 request per SM and L1 lifetime instead of one per warp on the counters' hot line
    B.a = *reinterpret_cast<const uint4*>(&ctr->bbox[0]);
    B.b = *reinterpret_cast<const uint2*>(&ctr->bbox[4]);
    return B;
}
__device__ __forceinline__ void warp_bbox_extend(Counters* ctr, const BoxNow& B, bool ok, float x, float y, float z)
{
    const u32  ox = float_to_ordered(x), oy = float_to_ordered(y), oz = float_to_ordered(z);
    const bool ext = ok && (ox < B.a.x || oy < B.a.y || oz < B.a.z || ox > B.a.w || oy > B.b.x || oz > B.b.y);
    if (!__any_sync(MMG_FULL, ext)) return;
    const u32 mn[3] = {__reduce_min_sync(MMG_FULL, ok ? ox : 0xFFFFFFFFu), __reduce_min_sync(MMG_FULL, ok ? oy : 0xFFFFFFFFu),
                       __reduce_min_sync(MMG_FULL, ok ? oz : 0xFFFFFFFFu)};
    const u32 mx[3] = {__reduce_max_sync(MMG_FULL, ok ? ox : 0u), __reduce_max_sync(MMG_FULL, ok ? oy : 0u), __reduce_max_sync(MMG_FULL, ok ? oz : 0u)};
    if (lane_id() == 0)
    {
        if (mn[0] < B.a.x) atomicMin(&ctr->bbox[0], mn[0]);
        if (mn[1] < B.a.y) atomicMin(&ctr->bbox[1], mn[1]);
        if (mn[2] < B.a.z) atomicMin(&ctr->bbox[2], mn[2]);
        if (mx[0] > B.a.w) atomicMax(&ctr->bbox[3], mx[0]);
        if (mx[1] > B.b.x) atomicMax(&ctr->bbox[4], mx[1]);
        if (mx[2] > B.b.y) atomicMax(&ctr->bbox[5], mx[2]);
    }
}

__device__ __forceinline__ u32* slot_w(float4* p) { return reinterpret_cast<u32*>(p) + 3; }

// Entry::cnt while an insert batch runs: bits 0-5 = the voxel's published point count n0 (<= 32), bits 6-31 = the
// low 26 bits of the smallest slot code once all free slots are taken (0 = not yet). k_insert_commit rewrites the
// word with the new plain count, so nothing outside the two insert kernels ever sees the code.
constexpr u32 RANK_CNT_BITS  = 6;
constexpr u32 RANK_CODE_MASK = (1u << 26) - 1;
constexpr u32 RANK_MAX_CHUNK = 1u << 25;  // chunk-local indices stay below 2^26 - 1, so a 26-bit code is never 0

// Probe for the insert kernel. One 16-byte L2 load (ld.cg) per probe returns key, row id and count word together.
// Outcomes: PROBE_FOUND (row id and count word valid), PROBE_CLAIMED (this thread won the empty entry: it still has to
// get a row and publish it), PROBE_PENDING (another thread of this batch won it and has not published the row yet).
// Nothing waits in here: rows are allocated and published after the loop with one counter update per warp, and only
// then do the pending lanes wait (a creator that had to wait for its own warp's pending lanes first could be waited on
// in a circle). A voxel created by THIS batch has count 0 until k_insert_commit runs.
enum { PROBE_NONE = 0, PROBE_FOUND, PROBE_CLAIMED, PROBE_PENDING };
__device__ __forceinline__ int probe_or_claim(const VoxelMapDev& m, u64 key, u32& vid, u32& cntword, u32& ent)
{
    u64 h = hash_key(key) & m.mask;
    for (;;)
    {
        Entry*      e = m.table + h;
        const uint4 v = __ldcg(reinterpret_cast<const uint4*>(e));
        u64         k = (u64)v.x | ((u64)v.y << 32);
        ent           = (u32)h;
        if (k == key && v.z != VID_INVALID)
        {
            cntword = v.w == 0xFFFFFFFFu ? 0u : v.w;
            vid     = v.z;
            return PROBE_FOUND;
        }
        if (k == KEY_EMPTY)
        {
            k = atomicCAS(reinterpret_cast<u64*>(&e->key), KEY_EMPTY, key);
            if (k == KEY_EMPTY) return PROBE_CLAIMED;
        }
        if (k == key) return PROBE_PENDING;
        h = (h + 1) & m.mask;
    }
}

constexpr int RANK_BLOCK = 128;

template <class Tr, bool MULTI>
__global__ void __launch_bounds__(RANK_BLOCK)
    k_insert_rank(const float* __restrict__ xs, const float* __restrict__ ys, const float* __restrict__ zs, u32 n, u32 base, const ScanSet S, float inv,
                  VoxelMapDev m, u32 eff_cap, u32* __restrict__ touched, u32 seg_base, int bbox_all)
{
    constexpr int CAP   = Tr::CAP;
    const u32     i     = blockIdx.x * blockDim.x + threadIdx.x;
    const int     lane  = lane_id();
    const bool    valid = i < n;
    pdl_trigger();  // k_insert_commit (a programmatic dependent launch) may be scheduled while this grid drains
    float         lx = 0.f, ly = 0.f, lz = 0.f;
    if (valid) lx = __ldg(xs + i), ly = __ldg(ys + i), lz = __ldg(zs + i);  // in flight while the scan lookup runs
    BoxNow box;
    if (bbox_all) box = load_box(m.ctr);
    float3 g  = make_float3(0.f, 0.f, 0.f);
    int    cx = 0, cy = 0, cz = 0;
    bool   ok = false;
    if (valid)
    {
        if constexpr (MULTI)
        {   // one lookup per warp (uniform), then at most a step or two per lane
            const u32 sw = scan_guess(S, base + i - (u32)lane);
            g            = compose_point(S.poses[scan_from(S, sw, base + i)], lx, ly, lz);
        }
        else
            g = compose_point(S.single, lx, ly, lz);
        ok = Tr::to_index(g.x, inv, cx) & Tr::to_index(g.y, inv, cy) & Tr::to_index(g.z, inv, cz);
        if (!ok) atomicAdd(&m.ctr->oob, 1u);
    }
    if (bbox_all) warp_bbox_extend(m.ctr, box, ok, g.x, g.y, g.z);
    const u64 key    = ok ? pack_key(cx, cy, cz) : (KEY_EMPTY - 1 - (u64)lane);  // invalid lanes never match each other
    const u32 peers  = __match_any_sync(MMG_FULL, key);
    const int leader = __ffs(peers) - 1;
    u32       vid = VID_INVALID, cw = 0, ent = 0;
    int       st  = PROBE_NONE;
    if (ok && lane == leader) st = probe_or_claim(m, key, vid, cw, ent);
    {   // new voxels of this warp: one update of the row counter for all of them (the free-row stack only holds
        // something after evictions; its slow path then takes rows one by one), then publish; the counters' cache line
        // sees one atomic per warp instead of three per voxel
        const u32 claimed = __ballot_sync(MMG_FULL, st == PROBE_CLAIMED);
        if (claimed)
        {
            const bool stack_empty = *reinterpret_cast<volatile u32*>(&m.ctr->free_n) == 0u;
            u32        row0        = 0;
            if (stack_empty && lane == __ffs(claimed) - 1) row0 = atomicAdd(&m.ctr->nvox_hi, (u32)__popc(claimed));
            row0 = __shfl_sync(MMG_FULL, row0, __ffs(claimed) - 1);
            if (st == PROBE_CLAIMED)
            {
                vid      = stack_empty ? row0 + (u32)__popc(claimed & ((1u << lane) - 1u)) : alloc_row(m);
                Entry* e = m.table + ent;
                e->cnt   = 0;
                __threadfence();
                atomicExch(&e->vid, vid);
            }
        }
        if (st == PROBE_PENDING)
            while ((vid = *reinterpret_cast<volatile u32*>(&m.table[ent].vid)) == VID_INVALID) {}
    }
    vid = __shfl_sync(MMG_FULL, vid, leader);
    cw  = __shfl_sync(MMG_FULL, cw, leader);
    ent = __shfl_sync(MMG_FULL, ent, leader);
    const u32 n0    = cw & ((1u << RANK_CNT_BITS) - 1u);
    const u32 field = cw >> RANK_CNT_BITS;  // 0: free slots still being handed out; 1: all handed out; else the set's minimum code
    const u32 x     = ~i;                   // slot code: larger = earlier; never 0 (i < 2^26)
    const u32 room  = eff_cap > n0 ? eff_cap - n0 : 0u;
    // still in the race: voxel not full, fewer than `room` earlier points of the voxel in this very warp, and not later
    // than the voxel's published bound (every slot then holds an earlier point, and slot codes only grow)
    bool alive = ok && (u32)__popc(peers & ((1u << lane) - 1u)) < room && (field <= 1u || (x & RANK_CODE_MASK) >= field);
    // Hand-out phase: the voxel's points of this warp take consecutive free slots with ONE atomicAdd on the row's
    // reservation counter (all voxels of the warp in parallel) and store their codes; no two points ever meet in a slot.
    {
        const u32 want  = __ballot_sync(MMG_FULL, alive && field == 0u) & peers;  // this voxel's candidates in the warp
        const int first = __ffs(want) - 1;
        u32       pos   = 0;
        if (want && lane == first) pos = atomicAdd(m.fill + vid, (u32)__popc(want));
        pos = __shfl_sync(MMG_FULL, pos, first < 0 ? lane : first);
        bool lists = false;
        if (want >> lane & 1u)
        {
            const u32 s = pos + (u32)__popc(want & ((1u << lane) - 1u));
            if (s < room)
            {
                *reinterpret_cast<volatile u32*>(slot_w(m.pts + (size_t)vid * CAP + n0 + s)) = x;
                alive = false;
                lists = s == 0;  // the point that gets the first free slot lists the voxel for k_insert_commit
            }
            // the group that takes the last free slot tells later probes to skip the counter
            if (lane == first && pos < room && pos + (u32)__popc(want) >= room) atomicMax(&m.table[ent].cnt, (1u << RANK_CNT_BITS) | n0);
        }
        const u32 lm = __ballot_sync(MMG_FULL, lists);
        if (lm)
        {   // one list-cursor update per warp
            u32 at = 0;
            if (lane == __ffs(lm) - 1) at = atomicAdd(&m.ctr->nseg, (u32)__popc(lm));
            at = __shfl_sync(MMG_FULL, at, __ffs(lm) - 1);
            if (lists) touched[at + (u32)__popc(lm & ((1u << lane) - 1u)) - seg_base] = vid;
        }
    }
    // Replacement phase: only for points that found every free slot handed out (voxels hit by several warps at once:
    // the later point may have arrived first). One voxel of the warp at a time, all 32 lanes cooperating.
    u32       pend = __ballot_sync(MMG_FULL, alive);
    const u32 x0   = ~(i - (u32)lane);  // code of lane 0; lane L holds x0 - L
    while (pend)
    {
        const int gl    = __ffs(pend) - 1;
        const u32 gpeer = __shfl_sync(MMG_FULL, peers, gl);
        const u32 gvid  = __shfl_sync(MMG_FULL, vid, gl);
        const u32 gn0   = __shfl_sync(MMG_FULL, n0, gl);
        const u32 groom = eff_cap - gn0;
        float4*   row   = m.pts + (size_t)gvid * CAP + gn0;
        // snapshot: lane L <- code of free slot L (lanes past the room hold +inf and are never the minimum). Every slot
        // has an owner by now; a code still on its way (0) is awaited.
        u32 v;
        do v = (u32)lane < groom ? *reinterpret_cast<volatile u32*>(slot_w(row + lane)) : 0xFFFFFFFFu;
        while (__any_sync(MMG_FULL, v == 0u));
        const u32 todo   = gpeer & pend;
        int       myslot = -1;
        u32       myexp  = 0;
        for (u32 t = todo; t; t &= t - 1)
        {   // ascending lane = ascending input index = descending code
            const int pl = __ffs(t) - 1;
            const u32 xp = x0 - (u32)pl;
            const u32 mn = __reduce_min_sync(MMG_FULL, v);
            if (xp <= mn) break;  // `room` earlier points are already there: this one and all later ones are out
            const int sl = __ffs(__ballot_sync(MMG_FULL, v == mn)) - 1;
            if (lane == pl) myslot = sl, myexp = mn;
            if (lane == sl) v = xp;
        }
        bool fail = false;
        if (myslot >= 0) fail = atomicCAS(slot_w(row + myslot), myexp, x) != myexp;
        const u32 failed = __ballot_sync(MMG_FULL, fail);
        pend &= ~(todo & ~failed);  // placed or rejected; a failed claim retries on a fresh snapshot
        if (!failed)
        {   // v is a lower bound of every slot's code from now on: publish the set's minimum
            const u32 mn = __reduce_min_sync(MMG_FULL, v);
            if (lane == gl) atomicMax(&m.table[ent].cnt, ((mn & RANK_CODE_MASK) << RANK_CNT_BITS) | gn0);
        }
    }
}

template <class Tr>
__global__ void __launch_bounds__(256)
    k_insert_commit(VoxelMapDev m, const u32* __restrict__ touched, const float* __restrict__ xs, const float* __restrict__ ys,
                    const float* __restrict__ zs, u32 base, ScanSet S, float inv, int bbox_accepted, int with_mean, u32 seg_base,
                    Counters* __restrict__ host_mirror)
{
    constexpr int CAP    = Tr::CAP;
    const int     lane   = lane_id();
    const u32     nwarps = (gridDim.x * blockDim.x) >> 5;
    pdl_wait();  // everything k_insert_rank wrote is visible from here on
    const u32     nseg   = m.ctr->nseg - seg_base;  // the list cursor restarts at 0 with every chunk (cleared by the last block below)
    for (u32 t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < nseg; t += nwarps)
    {
        const u32 vid  = touched[t];
        const u32 n0   = m.cnt[vid];
        if (lane == 0) m.fill[vid] = 0;
        float4*   slot = m.pts + (size_t)vid * CAP + n0 + lane;
        u32       w0   = 0;
        if (n0 + (u32)lane < (u32)CAP) w0 = *slot_w(slot);
        // the filled slots are a prefix of the free ones but in no particular order: lane r takes the r-th largest code
        const u32 macc = __popc(__ballot_sync(MMG_FULL, w0 != 0u));
        u32       rank = 0, w = 0;
        for (u32 a = 0; a < macc; a++) rank += __shfl_sync(MMG_FULL, w0, a) > w0 ? 1u : 0u;
        for (u32 a = 0; a < macc; a++)
        {
            const u32 wa = __shfl_sync(MMG_FULL, w0, a), ra = __shfl_sync(MMG_FULL, rank, a);
            if (ra == (u32)lane) w = wa;
        }
        const bool take = (u32)lane < macc;
        float4     p    = make_float4(0.f, 0.f, 0.f, 0.f);
        if (take)
        {
            const u32    i = ~w;
            const float3 g = compose_point(S.poses ? S.poses[scan_guess(S, base + i)] : S.single, __ldg(xs + i), __ldg(ys + i), __ldg(zs + i));
            p              = make_float4(g.x, g.y, g.z, 0.f);
            *slot          = p;  // also clears the ranking scratch in .w
        }
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
    // the last block to finish mirrors the counters into pinned host memory: the host only has to wait for the stream
    __syncthreads();
    if (threadIdx.x == 0)
    {
        __threadfence();
        if (atomicAdd(&m.ctr->done, 1u) == gridDim.x - 1)
        {
            m.ctr->done = 0;
            __threadfence();
            volatile u32*       dst = reinterpret_cast<volatile u32*>(host_mirror);
            const volatile u32* src = reinterpret_cast<const volatile u32*>(m.ctr);
            for (int k = 0; k < (int)(sizeof(Counters) / sizeof(u32)); k++) dst[k] = src[k];
            __threadfence_system();
            // the host has this chunk's out-of-range count and list length in the mirror: restart both for the next chunk
            // (no memset launch per call; every other block has passed its last read of nseg when `done` is complete)
            m.ctr->oob  = 0;
            m.ctr->nseg = 0;
        }
    }
}
}  // namespace mmg
