// mmg_trees_radius.cuh — SparseTreesPointCloud::nn_radius_search on the device (SparseTreesPointCloud.cpp:645-713).
//
// Reference semantics: for every block of the index box coordToOuterIdx(q - r) .. (q + r), in cx, cy, cz order, the block's
// CSimplePointsMap::nn_radius_search returns its hits (d2 < r2) ascending in (d2, block-local index), cut at maxPoints when
// that is non-zero; all of them go through one std::map<float, ...> keyed by d2, so equal squared distances overwrite each
// other (the LAST one inserted survives: later block, or higher index inside a block) and the output is
// ascending in d2, not cut again.
//
// Device form (round 2; round 1 ran eight global radix-sort passes over the whole hit list, 0.35 % of HBM peak at r = 2 m):
//   k_trees_radius_hits<0>  one warp per query walks the blocks with the occupancy-grid branch and bound and counts hits
//   k_trees_radius_hits<1>  same walk, writes per hit a UNIQUE 64-bit key {d2 bits : rank} with
//                           rank = (points of the blocks visited before this one) + block-local index, i.e. the order the
//                           reference inserts into its map, plus the point and its index
//   k_trees_radius_sort     one thread block per query: bitonic sort of the query's keys in shared memory (the hits of
//                           one query, ~5000 at r = 2 m on the 50 M-point map, never leave the SM), the maxPoints cut per
//                           block as one warp's in-order walk with per-block counters, keep-last-of-equal-d2, and the
//                           survivors written once, in order, at the query's place of the result
// Queries with more hits than fit in shared memory take the same code over global-memory scratch.
#pragma once
#include "mmg_trees.cuh"

namespace mmg
{
constexpr int TRR_THREADS = 512;
constexpr u32 TRR_SMEM_CAP = 16384;  // keys of one query held in shared memory (10 bytes each)
constexpr u32 TRR_NBQ_MAX  = 1024;   // blocks per query the maxPoints walk keeps counters for

struct TreesRadiusErr
{
    u32 rank_overflow;  // the blocks of one query hold >= 2^32 points (cannot happen below 4 G points per neighbourhood)
    u32 maxcnt;         // largest per-query hit count (chooses the sort kernel's shared-memory size)
};

template <int PASS>
struct RadiusVisitor2
{
    float  q[3], r2;
    u32    c, rank0;
    u64    base;
    u64*   okey;
    float* oxyz;
    u32*   oidx;
    __device__ __forceinline__ float bound() const { return r2; }
    __device__ __forceinline__ void  row(const float4& p, bool valid)
    {
        const float d   = valid ? sqr_dist(q[0], q[1], q[2], p.x, p.y, p.z) : 0.f;
        const bool  hit = valid && d < r2;  // CSimplePointsMap::nn_radius_search keeps d2 < r2 (strict)
        const u32   m   = __ballot_sync(MMG_FULL, hit);
        if (PASS && hit)
        {
            const u64 o   = base + c + __popc(m & ((1u << lane_id()) - 1));
            const u32 idx = __float_as_uint(p.w);
            okey[o]       = ((u64)__float_as_uint(d) << 32) | (u64)(rank0 + idx);
            oxyz[3 * o] = p.x, oxyz[3 * o + 1] = p.y, oxyz[3 * o + 2] = p.z;
            oidx[o] = idx;
        }
        c += __popc(m);
    }
    __device__ __forceinline__ void cell_done() {}
};

// the index box of a query; false when a corner leaves the packed-key range (the reference would index such blocks too,
// but none can exist in the map)
__device__ __forceinline__ bool trees_radius_box(const TreesDev& t, const float3& p, float r, int (&a)[3], int (&b)[3])
{
    return trees_index(t, __fsub_rn(p.x, r), a[0]) & trees_index(t, __fsub_rn(p.y, r), a[1]) & trees_index(t, __fsub_rn(p.z, r), a[2]) &
           trees_index(t, __fadd_rn(p.x, r), b[0]) & trees_index(t, __fadd_rn(p.y, r), b[1]) & trees_index(t, __fadd_rn(p.z, r), b[2]);
}

// Queries [q0, q0 + n) of the batch. PASS 0: cnt[q] (relative index). PASS 1: hits at off[q] (relative to the chunk).
template <int PASS>
__global__ void __launch_bounds__(256)
    k_trees_radius_hits(TreesDev t, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 q0, u32 n, PoseD pose,
                        int use_pose, float r2, float r, u32* __restrict__ cnt, const u64* __restrict__ off, u64 off_base, u64* __restrict__ okey,
                        float* __restrict__ oxyz, u32* __restrict__ oidx, TreesRadiusErr* __restrict__ err)
{
    const int lane   = lane_id();
    const u32 nwarps = (gridDim.x * blockDim.x) >> 5;
    for (u32 w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < n; w += nwarps)
    {
        const u32 q = q0 + w;
        float3    p = make_float3(qx[q], qy[q], qz[q]);
        if (use_pose) p = compose_point(pose, p.x, p.y, p.z);
        int                 a[3], b[3];
        const bool          ok = trees_radius_box(t, p, r, a, b);
        RadiusVisitor2<PASS> v;
        v.q[0] = p.x, v.q[1] = p.y, v.q[2] = p.z, v.r2 = r2, v.c = 0, v.rank0 = 0;
        v.base = PASS ? off[q] - off_base : 0, v.okey = okey, v.oxyz = oxyz, v.oidx = oidx;
        if (ok && r2 > 0.f)
            for (int x = a[0]; x <= b[0]; x++)
                for (int y = a[1]; y <= b[1]; y++)
                    for (int z = a[2]; z <= b[2]; z++)
                    {
                        u32 bid, ecnt;
                        if (!index_in_range(x, y, z) || !probe_ro(t.btab, t.bmask, pack_key(x, y, z), bid, ecnt)) continue;
                        const BlockGeom g = trees_geom(t, x, y, z);
                        trees_traverse(t, bid, g, v.q, v);
                        const u32 bc = __ldg(&t.bcnt[bid]);
                        if (!PASS && v.rank0 + bc < v.rank0 && lane == 0) atomicAdd(&err->rank_overflow, 1u);
                        v.rank0 += bc;
                    }
        if (!PASS && lane == 0)
        {
            cnt[w] = v.c;
            if (v.c > TRR_SMEM_CAP / 4) atomicMax(&err->maxcnt, v.c);
        }
    }
}

static __global__ void k_trees_radius_bound(const u32* __restrict__ cnt, u32 n, u32 cap_per_query /*0 = none*/, u32* __restrict__ bound)
{
    const u32 q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q < n) bound[q] = cap_per_query ? min(cnt[q], cap_per_query) : cnt[q];
}

// in-place bitonic sort of P = 2^k (key, payload) pairs by one thread block; keys are unique except for the padding
template <class PosT>
__device__ __forceinline__ void block_bitonic(u64* __restrict__ key, PosT* __restrict__ pos, u32 P)
{
    for (u32 k = 2; k <= P; k <<= 1)
        for (u32 j = k >> 1; j > 0; j >>= 1)
        {
            for (u32 i = threadIdx.x; i < (P >> 1); i += blockDim.x)
            {
                const u32  lo = ((i & ~(j - 1)) << 1) | (i & (j - 1)), hi = lo | j;
                const bool up = (lo & k) == 0;
                const u64  a = key[lo], b = key[hi];
                if ((a > b) == up)
                {
                    key[lo] = b, key[hi] = a;
                    const PosT t = pos[lo];
                    pos[lo]      = pos[hi];
                    pos[hi]      = t;
                }
            }
            __syncthreads();
        }
}

// One thread block per query (grid-stride over the chunk's queries, heaviest-agnostic: the sort cost is per query).
// skey/spos: shared memory for up to smem_cap keys; gkey/gpos: global scratch, gcap keys per thread block, for the queries
// that do not fit.
template <class PosT, u32 FLAG_BITS>
__device__ __forceinline__ void trees_radius_finish(const TreesDev& t, u64* __restrict__ key, PosT* __restrict__ pos, u32 N, u32 maxp, u32* s_rank0,
                                                    u32* s_cnt, u32 nblk, const float* __restrict__ hxyz, const u32* __restrict__ hidx, u64 obase,
                                                    float* __restrict__ rxyz, float* __restrict__ rd2, u64* __restrict__ rid, u32* __restrict__ outcnt_q,
                                                    u32* s_scan)
{
    constexpr PosT ACC = (PosT)1 << (FLAG_BITS - 1), SURV = (PosT)1 << (FLAG_BITS - 2), IDX = SURV - 1;
    const int      lane = lane_id(), wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    if (maxp == 0)
    {   // every hit enters the map: the last of each run of equal d2 survives
        for (u32 i = threadIdx.x; i < N; i += blockDim.x)
            if (i + 1 >= N || (u32)(key[i] >> 32) != (u32)(key[i + 1] >> 32)) pos[i] |= SURV;
    }
    else
    {
        // the per-block cut: in ascending (d2, rank) order a hit is taken while its block has handed out fewer than maxPoints;
        // one warp walks the sorted list 32 at a time with per-block counters (ranks of one block are consecutive, so the
        // block of a hit is found by searching the blocks' first ranks)
        if (wid == 0)
        {
            u32 pend = 0xFFFFFFFFu, pend_d2 = 0;  // last accepted hit so far: survives unless the next accepted one has its d2
            for (u32 base = 0; base < N; base += 32)
            {
                const u32  i     = base + lane;
                const bool valid = i < N;
                const u64  kv    = valid ? key[i] : 0;
                u32        ord   = 0xFFFFFFFFu;
                if (valid)
                {
                    const u32 rk = (u32)kv;
                    u32       lo = 0, hi = nblk;  // last block whose first rank <= rk
                    while (hi - lo > 1)
                    {
                        const u32 mid = (lo + hi) >> 1;
                        if (s_rank0[mid] <= rk) lo = mid; else hi = mid;
                    }
                    ord = lo;
                }
                const u32 peers = __match_any_sync(MMG_FULL, ord);
                bool      acc   = false;
                if (valid)
                {
                    const u32 before = s_cnt[ord] + __popc(peers & ((1u << lane) - 1));
                    acc              = before < maxp;
                }
                __syncwarp();
                if (valid && lane == __ffs(peers) - 1) s_cnt[ord] += __popc(peers);
                __syncwarp();
                const u32 am = __ballot_sync(MMG_FULL, acc);
                if (am)
                {
                    const u32 d2  = (u32)(kv >> 32);
                    const int fst = __ffs(am) - 1, lst = 31 - __clz(am);
                    const u32 d2f = __shfl_sync(MMG_FULL, d2, fst);
                    if (lane == 0 && pend != 0xFFFFFFFFu && pend_d2 != d2f) pos[pend] |= SURV;
                    const u32 above = am & ~((2u << lane) - 1u);  // accepted lanes after this one
                    const int nxt   = above ? __ffs(above) - 1 : lane;
                    const u32 d2n   = __shfl_sync(MMG_FULL, d2, nxt);
                    if (acc)
                    {
                        pos[i] |= ACC;
                        if (above && d2n != d2) pos[i] |= SURV;
                    }
                    pend    = base + (u32)lst;
                    pend_d2 = __shfl_sync(MMG_FULL, d2, lst);
                }
            }
            if (lane == 0 && pend != 0xFFFFFFFFu) pos[pend] |= SURV;
        }
    }
    __syncthreads();
    // ordered compaction of the survivors to the query's place in the result
    u32 running = 0;
    for (u32 base = 0; base < N; base += blockDim.x)
    {
        const u32  i = base + threadIdx.x;
        const bool s = i < N && (pos[i] & SURV);
        const u32  m = __ballot_sync(MMG_FULL, s);
        if (lane == 0) s_scan[wid] = __popc(m);
        __syncthreads();
        u32 before = running, total = running;
        for (int k = 0; k < nw; k++)
        {
            const u32 c = s_scan[k];
            if (k < wid) before += c;
            total += c;
        }
        if (s)
        {
            const u64 o = obase + before + __popc(m & ((1u << lane) - 1));
            const u32 h = (u32)(pos[i] & IDX);
            rxyz[3 * o] = hxyz[3 * (size_t)h], rxyz[3 * o + 1] = hxyz[3 * (size_t)h + 1], rxyz[3 * o + 2] = hxyz[3 * (size_t)h + 2];
            rd2[o] = __uint_as_float((u32)(key[i] >> 32));
            rid[o] = (u64)hidx[h];
        }
        running = total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *outcnt_q = running;
}

__global__ void __launch_bounds__(TRR_THREADS)
    k_trees_radius_sort(TreesDev t, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 q0, u32 n, PoseD pose,
                        int use_pose, float r, u32 maxp, const u32* __restrict__ cnt /*chunk relative*/, const u64* __restrict__ off, u64 off_base,
                        const u64* __restrict__ key, const float* __restrict__ hxyz, const u32* __restrict__ hidx, u64* __restrict__ gkey, u32* __restrict__ gpos,
                        u32 gcap /*keys per thread block in gkey / gpos*/, u32 smem_cap,
                        const u64* __restrict__ boff, float* __restrict__ rxyz, float* __restrict__ rd2, u64* __restrict__ rid, u32* __restrict__ outcnt,
                        u32* __restrict__ next_query)
{
    extern __shared__ __align__(16) unsigned char smem[];
    u64*            skey = reinterpret_cast<u64*>(smem);
    unsigned short* spos = reinterpret_cast<unsigned short*>(smem + (size_t)smem_cap * 8);
    __shared__ u32  s_rank0[TRR_NBQ_MAX + 1], s_cnt[TRR_NBQ_MAX], s_scan[TRR_THREADS / 32], s_q, s_nblk;
    for (;;)
    {
        __syncthreads();
        if (threadIdx.x == 0) s_q = atomicAdd(next_query, 1u);
        __syncthreads();
        const u32 w = s_q;
        if (w >= n) break;
        const u32 q = q0 + w, N = cnt[w];
        if (N == 0)
        {
            if (threadIdx.x == 0) outcnt[w] = 0;
            continue;
        }
        const u64 h0 = off[q] - off_base;
        if (maxp != 0 && threadIdx.x == 0)
        {   // first rank of every block of the query's index box, in visiting order (blocks that do not exist hold no rank)
            float3 p = make_float3(qx[q], qy[q], qz[q]);
            if (use_pose) p = compose_point(pose, p.x, p.y, p.z);
            int a[3], b[3];
            trees_radius_box(t, p, r, a, b);
            u32 nb = 0, rank0 = 0;
            for (int x = a[0]; x <= b[0]; x++)
                for (int y = a[1]; y <= b[1]; y++)
                    for (int z = a[2]; z <= b[2]; z++)
                    {
                        u32 bid, ecnt;
                        if (!index_in_range(x, y, z) || !probe_ro(t.btab, t.bmask, pack_key(x, y, z), bid, ecnt)) continue;
                        if (nb < TRR_NBQ_MAX) s_rank0[nb] = rank0, s_cnt[nb] = 0;
                        rank0 += __ldg(&t.bcnt[bid]);
                        nb++;
                    }
            s_nblk = min(nb, TRR_NBQ_MAX);
        }
        u32 P = 1;
        while (P < N) P <<= 1;
        if (P <= smem_cap)
        {
            for (u32 i = threadIdx.x; i < P; i += blockDim.x)
            {
                skey[i] = i < N ? key[h0 + i] : ~0ull;
                spos[i] = (unsigned short)(i < N ? i : 0);
            }
            __syncthreads();
            block_bitonic<unsigned short>(skey, spos, P);
            trees_radius_finish<unsigned short, 16>(t, skey, spos, N, maxp, s_rank0, s_cnt, s_nblk, hxyz + 3 * h0, hidx + h0, boff[q], rxyz, rd2, rid,
                                                    outcnt + w, s_scan);
        }
        else
        {   // larger than shared memory: the same steps over this thread block's own slice of global scratch
            u64* gk = gkey + (size_t)blockIdx.x * gcap;
            u32* gp = gpos + (size_t)blockIdx.x * gcap;
            for (u32 i = threadIdx.x; i < P; i += blockDim.x)
            {
                gk[i] = i < N ? key[h0 + i] : ~0ull;
                gp[i] = i < N ? i : 0;
            }
            __syncthreads();
            block_bitonic<u32>(gk, gp, P);
            trees_radius_finish<u32, 32>(t, gk, gp, N, maxp, s_rank0, s_cnt, s_nblk, hxyz + 3 * h0, hidx + h0, boff[q], rxyz, rd2, rid, outcnt + w, s_scan);
        }
    }
}

// moves the survivors of every query from its reserved place (boff) to its final place (noff); one warp per query
static __global__ void k_trees_radius_pack(u32 n, const u32* __restrict__ outcnt, const u64* __restrict__ boff, const u64* __restrict__ noff,
                                           const float* __restrict__ sxyz, const float* __restrict__ sd2, const u64* __restrict__ sid,
                                           float* __restrict__ dxyz, float* __restrict__ dd2, u64* __restrict__ did)
{
    const int lane   = lane_id();
    const u32 nwarps = (gridDim.x * blockDim.x) >> 5;
    for (u32 q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; q < n; q += nwarps)
    {
        const u32 c = outcnt[q];
        const u64 s = boff[q], d = noff[q];
        for (u32 i = lane; i < c; i += 32)
        {
            dxyz[3 * (d + i)] = sxyz[3 * (s + i)], dxyz[3 * (d + i) + 1] = sxyz[3 * (s + i) + 1], dxyz[3 * (d + i) + 2] = sxyz[3 * (s + i) + 2];
            dd2[d + i] = sd2[s + i];
            did[d + i] = sid[s + i];
        }
    }
}
}  // namespace mmg
