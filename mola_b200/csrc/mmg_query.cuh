// mmg_query.cuh — the batched nn_single_search kernel (the headline kernel).
//
// One warp owns 32 consecutive queries. Queries of the chunk that fall into the same
// voxel (LiDAR scan order makes that the common case) are grouped with a ballot; the 27
// stencil cells are probed ONCE per group (lane = (dx+1)*9 + (dy+1)*3 + (dz+1), the
// reference's cx,cy,cz loop order, HashedVoxelPointCloud.cpp:758-761) and every row of
// points is loaded ONCE per pass (lane l <- slot l, one LDG.128) and compared against up
// to G queries of the group held in registers. Per lane the candidates arrive in
// ascending (cell, slot) order, so a strict '<' keeps the first minimum; two REDUX
// reductions (min d2 bits, then min rank among the lanes holding that d2) give exactly
// the reference's "first candidate in scan order wins ties" (strict '<' at :728).
// d2 is a sum of squares of finite values, so its bit pattern orders like the float.
#pragma once
#include "mmg_voxel.cuh"

namespace mmg
{
// One pass over the stencil for G queries of the same voxel. `cells` = non-empty stencil cells
// (bit b = lane b's cell), (bx, by, bz) = the group's voxel. The queries' own voxel goes first,
// then only cells whose box can still hold a point at least as close as the current best of
// one of the G queries (exact pruning: skipped cells hold strictly farther points only, so the
// result, including the scan-order tie-break, is the one of the exhaustive walk).
template <class Tr, int G>
__device__ __forceinline__ void nn_group_pass(const VoxelMapDev& m, u32 cells, u32 vid, u32 cnt, int lane, int bx, int by, int bz, float vsize,
                                              const float (&gx)[G], const float (&gy)[G], const float (&gz)[G], float (&bd)[G], u32 (&br)[G],
                                              u32 (&wb)[G])
{
    constexpr int CAP = Tr::CAP;
    float         lb[G];
    {
        const int x = bx + lane / 9 - 1, y = by + (lane / 3) % 3 - 1, z = bz + lane % 3 - 1;
#pragma unroll
        for (int g = 0; g < G; g++) lb[g] = voxel_lower_bound<Tr>(x, y, z, vsize, gx[g], gy[g], gz[g]);
    }
#pragma unroll
    for (int g = 0; g < G; g++) wb[g] = 0x7f800000u;
    int b = (cells >> 13) & 1u ? 13 : -1;  // the queries' own voxel
    for (;;)
    {
        if (b < 0)
        {
            bool can = false;
#pragma unroll
            for (int g = 0; g < G; g++) can |= !(lb[g] > __uint_as_float(wb[g]));
            const u32 cand = __ballot_sync(MMG_FULL, can) & cells;
            if (!cand) break;
            b = __ffs(cand) - 1;
        }
        cells &= ~(1u << b);
        const u32 v = __shfl_sync(MMG_FULL, vid, b), c = __shfl_sync(MMG_FULL, cnt, b);
        if ((u32)lane < c)
        {
            const float4 t    = __ldg(&m.pts[(size_t)v * CAP + lane]);
            const u32    rank = (u32)b * 32u + (u32)lane;
#pragma unroll
            for (int g = 0; g < G; g++)
            {
                const float d = sqr_dist(t.x, t.y, t.z, gx[g], gy[g], gz[g]);
                if (d < bd[g] || (d == bd[g] && rank < br[g]))
                {
                    bd[g] = d;
                    br[g] = rank;
                }
            }
        }
#pragma unroll
        for (int g = 0; g < G; g++) wb[g] = __reduce_min_sync(MMG_FULL, __float_as_uint(bd[g]));
        b = -1;
    }
}

template <class Tr>
__global__ void __launch_bounds__(256)
    k_nn_single_grouped(VoxelMapDev m, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 n, u32 stride,
                        PoseD pose, int use_pose, float inv, float vsize, QueryOut out)
{
    constexpr int CAP     = Tr::CAP;
    constexpr int G       = 4;
    const int     lane    = lane_id();
    const u32     nwarps  = (gridDim.x * blockDim.x) >> 5;
    const u32     nchunks = (n + 31) >> 5;
    const int     ddx = lane / 9 - 1, ddy = (lane / 3) % 3 - 1, ddz = lane % 3 - 1;
    const float   INF = __int_as_float(0x7f800000);
    for (u32 chunk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; chunk < nchunks; chunk += nwarps)
    {
        const u32  q     = chunk * 32u + (u32)lane;
        const bool valid = q < n;
        float3     p     = make_float3(0.f, 0.f, 0.f);
        int        cx = 0, cy = 0, cz = 0;
        bool       ok = false;
        if (valid)
        {
            const size_t qi = (size_t)q * stride;
            p               = make_float3(qx[qi], qy[qi], qz[qi]);
            if (use_pose) p = compose_point(pose, p.x, p.y, p.z);
            ok = Tr::to_index(p.x, inv, cx) & Tr::to_index(p.y, inv, cy) & Tr::to_index(p.z, inv, cz);
        }
        const u64 key = ok ? pack_key(cx, cy, cz) : KEY_EMPTY;
        // this lane's own result, filled in when its group is processed
        float rx = 0.f, ry = 0.f, rz = 0.f, rd = 0.f;
        u64   rid    = 0;
        u32   rfound = 0;
        u32   remaining = __ballot_sync(MMG_FULL, ok);
        while (remaining)
        {
            const int L    = __ffs(remaining) - 1;
            const u64 keyL = __shfl_sync(MMG_FULL, key, L);
            u32       group = __ballot_sync(MMG_FULL, key == keyL);
            remaining &= ~group;
            int bx, by, bz;
            unpack_key(keyL, bx, by, bz);
            u32 vid = 0, cnt = 0;
            if (lane < 27)
            {
                const int x = bx + ddx, y = by + ddy, z = bz + ddz;
                if (index_in_range(x, y, z))
                    if (!probe_ro(m.table, m.mask, pack_key(x, y, z), vid, cnt)) cnt = 0;
            }
            const u32 cells = __ballot_sync(MMG_FULL, cnt > 0);
            if (!cells) continue;
            while (group)
            {
                int   gl[G];
                float gx[G], gy[G], gz[G], bd[G];
                u32   br[G];
                int   ng = 0;
#pragma unroll
                for (int g = 0; g < G; g++)
                {
                    if (group)
                    {
                        gl[g] = __ffs(group) - 1;
                        group &= group - 1;
                        ng = g + 1;
                    }
                    else
                        gl[g] = gl[0];
                    gx[g] = __shfl_sync(MMG_FULL, p.x, gl[g]);
                    gy[g] = __shfl_sync(MMG_FULL, p.y, gl[g]);
                    gz[g] = __shfl_sync(MMG_FULL, p.z, gl[g]);
                    bd[g] = INF;
                    br[g] = 0xFFFFFFFFu;
                }
                u32 wb[G];
                if (ng == 1)
                {
                    float ax[1] = {gx[0]}, ay[1] = {gy[0]}, az[1] = {gz[0]}, ad[1] = {INF};
                    u32   ar[1] = {0xFFFFFFFFu}, aw[1];
                    nn_group_pass<Tr, 1>(m, cells, vid, cnt, lane, bx, by, bz, vsize, ax, ay, az, ad, ar, aw);
                    bd[0] = ad[0], br[0] = ar[0], wb[0] = aw[0];
                }
                else
                    nn_group_pass<Tr, G>(m, cells, vid, cnt, lane, bx, by, bz, vsize, gx, gy, gz, bd, br, wb);
#pragma unroll
                for (int g = 0; g < G; g++)
                {
                    if (g >= ng) break;
                    const u32 dmin = wb[g];
                    const u32 rr   = __reduce_min_sync(MMG_FULL, __float_as_uint(bd[g]) == dmin ? br[g] : 0xFFFFFFFFu);
                    if (rr == 0xFFFFFFFFu) continue;  // cannot happen when cells != 0; kept for safety
                    const int    b = (int)(rr >> 5), slot = (int)(rr & 31);
                    const u32    v = __shfl_sync(MMG_FULL, vid, b);
                    const float4 t = __ldg(&m.pts[(size_t)v * CAP + slot]);  // same address in every lane: one broadcast load
                    if (lane == gl[g])
                    {
                        rx = t.x, ry = t.y, rz = t.z;
                        rd     = __uint_as_float(dmin);
                        rid    = plain_id<Tr::SLOT_BITS>(bx + b / 9 - 1, by + (b / 3) % 3 - 1, bz + b % 3 - 1, slot);
                        rfound = 1;
                    }
                }
            }
        }
        if (valid)
        {
            out.found[q] = (uint8_t)rfound;
            out.d2[q]    = rd;
            out.id[q]    = rid;
            out.xyz[3 * (size_t)q] = rx, out.xyz[3 * (size_t)q + 1] = ry, out.xyz[3 * (size_t)q + 2] = rz;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// k_nn_single_pruned: one THREAD per query, exact branch-and-bound over the 27-cell stencil.
//
// The reference visits all 27 cells; the answer is the minimum of (d2, scan rank), so any visiting
// order and any pruning that only drops cells whose every point is strictly farther than the
// current best give the same result. The query's own voxel goes first (it nearly always holds the
// answer for ICP-style queries), then a neighbour is probed only if its box can still hold a
// point at least as close. The box distance separates per axis, so it costs three table look-ups
// in registers per cell. Consecutive LiDAR points fall into the same voxel, so the lanes of a warp
// mostly read the same row and the loads coalesce into broadcasts; every lane does useful distance
// evaluations (no idle slots as in the row-per-warp layout), which is what this instruction-issue
// bound kernel needs (ncu: profiles/r1b_*).
template <class Tr>
__device__ __forceinline__ void nn_axis_gaps(int c, float vsize, float q, float (&gap)[3])
{
    // squared distance from q to the slab of cell c-1, c, c+1 along one axis (conservative)
#pragma unroll
    for (int k = 0; k < 3; k++)
    {
        float lo, hi;
        Tr::cell_span(c + k - 1, lo, hi);
        const float eps = (fabsf((float)(c + k - 1)) + 2.0f) * vsize * 1e-6f;
        const float d   = fmaxf(0.f, fmaxf(lo * vsize - eps - q, q - (hi * vsize + eps)));
        gap[k]          = d * d * 0.9999f;
    }
}

// (dx*dx + dy*dy) + dz*dz with the x/y lanes as one packed FADD2 + FMUL2 (sm_100 f32x2, IEEE round-to-nearest per lane:
// bit-identical to sqr_dist); t.x/t.y arrive as an aligned register pair from the LDG.128, so no moves are needed.
__device__ __forceinline__ float sqr_dist_packed(const float4& t, unsigned long long qxy, float qz)
{
    unsigned long long a, d, mm;
    float              mx, my;
    asm("mov.b64 %0, {%1,%2};" : "=l"(a) : "f"(t.x), "f"(t.y));
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(qxy));
    asm("mul.rn.f32x2 %0, %1, %1;" : "=l"(mm) : "l"(d));
    asm("mov.b64 {%0,%1}, %2;" : "=f"(mx), "=f"(my) : "l"(mm));
    const float dz = __fsub_rn(t.z, qz);
    return __fadd_rn(__fadd_rn(mx, my), __fmul_rn(dz, dz));
}

// visits one stencil cell b (0..26) for this thread's query; merges into (bd, br, bv)
template <class Tr, bool PACKED = false>
__device__ __forceinline__ void nn_visit_cell(const VoxelMapDev& m, int b, int cx, int cy, int cz, const float3& p, float& bd, u32& br, u32& bv)
{
    constexpr int CAP = Tr::CAP;
    const int     x = cx + b / 9 - 1, y = cy + (b / 3) % 3 - 1, z = cz + b % 3 - 1;
    if (!index_in_range(x, y, z)) return;
    u32 vid, cnt;
    if (!probe_ro(m.table, m.mask, pack_key(x, y, z), vid, cnt)) return;
    const float4* row = m.pts + (size_t)vid * CAP;
    // inside a cell the slots come in ascending order, so a strict '<' keeps the first minimum
    // a row spans up to four 128-byte lines: ask for the later ones now, so the walk below does not pay one L2 round
    // trip per line
    if (cnt > 8) asm volatile("prefetch.global.L1 [%0];" ::"l"(row + 8));
    if (cnt > 16) asm volatile("prefetch.global.L1 [%0];" ::"l"(row + 16));
    if (cnt > 24) asm volatile("prefetch.global.L1 [%0];" ::"l"(row + 24));
    float cd = __int_as_float(0x7f800000);
    u32   cs = 0;
    unsigned long long qxy = 0;
    if (PACKED) asm("mov.b64 %0, {%1,%2};" : "=l"(qxy) : "f"(p.x), "f"(p.y));
#pragma unroll 4
    for (u32 s = 0; s < cnt; s++)
    {
        const float4 t = __ldg(row + s);
        const float  d = PACKED ? sqr_dist_packed(t, qxy, p.z) : sqr_dist(t.x, t.y, t.z, p.x, p.y, p.z);
        if (d < cd) cd = d, cs = s;
    }
    const u32 rank = (u32)b * 32u + cs;
    if (cnt && (cd < bd || (cd == bd && rank < br))) bd = cd, br = rank, bv = vid;
}

template <class Tr, int MINB = 1, bool PACKED = false>
__global__ void __launch_bounds__(256, MINB)
    k_nn_single_pruned(VoxelMapDev m, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 n, u32 stride,
                       PoseD pose, int use_pose, float inv, float vsize, QueryOut out, float bd_init, u32 perm_mul)
{
    // bd_init = +inf for nn_single_search; the matcher passes the largest float below its squared distance
    // threshold, which turns "d2 < threshold" into the initial branch-and-bound radius (mmg_match_batch)
    constexpr int CAP  = Tr::CAP;
    const int     lane = lane_id();
    // warp-uniform trip count: the coalesced result stores below shuffle across the whole warp
    // Blocks take their 128-query tiles in a scattered order (tile = block * perm_mul mod grid, perm_mul coprime to the
    // grid): a LiDAR batch has its heavy queries (dense near-field rings) in contiguous runs, and in launch order the
    // last run would be the kernel's tail with most SMs idle (ncu: SM busy 200k..350k of 357k cycles without this).
    const u32 tile = (u32)(((u64)blockIdx.x * perm_mul) % gridDim.x);
    for (u32 q0 = tile * blockDim.x + (threadIdx.x & ~31u); q0 < n; q0 += gridDim.x * blockDim.x)
    {
        const u32  q     = q0 + (u32)lane;
        const bool valid = q < n;
        float3     p     = make_float3(0.f, 0.f, 0.f);
        int        cx = 0, cy = 0, cz = 0;
        bool       ok = false;
        if (valid)
        {
            const size_t qi = (size_t)q * stride;
            p               = make_float3(qx[qi], qy[qi], qz[qi]);
            if (use_pose) p = compose_point(pose, p.x, p.y, p.z);
            ok = Tr::to_index(p.x, inv, cx) & Tr::to_index(p.y, inv, cy) & Tr::to_index(p.z, inv, cz);
        }
        float bd = bd_init;
        u32   br = 0xFFFFFFFFu, bv = 0;
        if (ok)
        {
            float gx[3], gy[3], gz[3];
            nn_axis_gaps<Tr>(cx, vsize, p.x, gx);
            nn_axis_gaps<Tr>(cy, vsize, p.y, gy);
            nn_axis_gaps<Tr>(cz, vsize, p.z, gz);
            // the query's own voxel first: it nearly always holds the answer and sets the bound
            nn_visit_cell<Tr, PACKED>(m, 13, cx, cy, cz, p, bd, br, bv);
            // this thread's own work list: the neighbours whose box can still matter
            u32 todo = 0;
#pragma unroll
            for (int b = 0; b < 27; b++)
                if (b != 13 && !(gx[b / 9] + gy[(b / 3) % 3] + gz[b % 3] > bd)) todo |= 1u << b;
            // face neighbours first, then edges, then corners: the closer boxes tighten the bound soonest, so fewer of the
            // later ones survive the re-check below (any order gives the same answer: the merge is on (d2, scan rank))
            constexpr u32 FACE = 0x415410u, EDGE = 0x2aa8aaau;
            while (todo)
            {
                const u32 tier = (todo & FACE) ? (todo & FACE) : ((todo & EDGE) ? (todo & EDGE) : todo);
                const int b    = __ffs(tier) - 1;
                todo &= ~(1u << b);
                const int   ix = b / 9, iy = (b / 3) % 3, iz = b % 3;
                const float lb = (ix == 0 ? gx[0] : ix == 1 ? gx[1] : gx[2]) + (iy == 0 ? gy[0] : iy == 1 ? gy[1] : gy[2]) +
                                 (iz == 0 ? gz[0] : iz == 1 ? gz[1] : gz[2]);
                if (lb > bd) continue;  // the bound has tightened since the list was made
                nn_visit_cell<Tr, PACKED>(m, b, cx, cy, cz, p, bd, br, bv);
            }
        }
        const bool found = br != 0xFFFFFFFFu;
        float4     t     = make_float4(0.f, 0.f, 0.f, 0.f);
        u64        id    = 0;
        if (found)
        {
            const int b = (int)(br >> 5), slot = (int)(br & 31);
            t  = __ldg(m.pts + (size_t)bv * CAP + slot);
            id = plain_id<Tr::SLOT_BITS>(cx + b / 9 - 1, cy + (b / 3) % 3 - 1, cz + b % 3 - 1, slot);
        }
        if (valid)
        {
            out.found[q] = found ? 1 : 0;
            out.d2[q]    = found ? bd : 0.f;
            out.id[q]    = id;
        }
        // xyz is an array of structs: transpose through shuffles so the warp writes its 96 floats as three fully
        // coalesced 128-byte stores (matters most when `out` is mapped host memory: whole PCIe write bursts)
#pragma unroll
        for (int k = 0; k < 3; k++)
        {
            const int   w = 32 * k + lane, src = w / 3, c = w - 3 * src;
            const float vx = __shfl_sync(MMG_FULL, t.x, src), vy = __shfl_sync(MMG_FULL, t.y, src), vz = __shfl_sync(MMG_FULL, t.z, src);
            if (q0 + (u32)src < n) out.xyz[3 * (size_t)q0 + w] = c == 0 ? vx : (c == 1 ? vy : vz);
        }
    }
}
}  // namespace mmg

namespace mmg
{
// ---------------------------------------------------------------------------------------------
// nn_single_search as a three-kernel pipeline over a global task queue (the default since round 2).
//
// k_nn_single_pruned (above) does the whole branch-and-bound in one thread per query. Its two measured losses (ncu,
// round 1: 17 of 32 lanes active, SMs busy 74 % of the launch) both come from the neighbour visits: a few lanes of a
// warp have many cells left while the others idle, and runs of such queries form the kernel's tail. Pooling the visits
// per warp (tried first, round 2a) balances lanes but serialises a heavy tile's visits into hundreds of dependent
// rounds. So the visits leave the warp altogether:
//   k_nn_own    one thread per query: compose, voxel index, the query's OWN voxel (sets the bound; lanes of a warp
//               mostly share the row, so the loads are broadcasts). Stores the composed point and the packed result
//               key {d2 bits : cell*32+slot} per query and appends one 4-byte task {cell : query} for every neighbour
//               cell whose box can still hold a point at least as close, to the face / edge / corner queue. One
//               atomicAdd per block and queue; inside a warp the tasks are laid out cell-major, so consecutive tasks
//               are consecutive scan points looking at the same neighbour cell: the same row more often than not.
//   k_nn_tasks  one thread per task, all queues in tier order: re-check the cell's lower bound against the query's
//               CURRENT key (tightened by every task that finished before), probe, walk the row, 64-bit atomicMin.
//               Every task is at most one 32-point row, so a plain grid-stride loop is balanced by construction and
//               every lane has work.
//   k_nn_finish one thread per query: unpack the key, fetch the winning point, write the four outputs coalesced.
// The key order (d2, cell*32+slot) is exactly the reference's "first candidate in scan order wins ties" (strict '<'
// at HashedVoxelPointCloud.cpp:728); pruning only drops cells whose every point is strictly farther than the bound,
// so the result is the exhaustive walk's, bit for bit, whatever order the tasks run in.
struct NNWork
{
    float4* qpt;       // composed query points; w = row id of the query's own voxel (bit pattern) when it has one
    u64*    qkey;      // per query {d2 bits : rank}; rank 0xFFFFFFFF = nothing found yet
    u32*    queue[3];  // face / edge / corner tasks: cell (5 bits) << 27 | query (27 bits)
    u32*    count;     // [0..2] queue tails, [3] ticket of k_nn_finish_match
};
constexpr u32 NN_FACE = 0x415410u, NN_EDGE = 0x2aa8aaau, NN_CORNER = 0x7FFFFFFu & ~(NN_FACE | NN_EDGE | (1u << 13));
constexpr u32 NN_CHUNK = 1u << 20;  // queries per pipeline pass (task ids hold 27 bits; bounds the worst-case queue size)
constexpr int NN_OWN_BLOCK = 256;

// squared distance from q to the two outer slabs of cell c's axis neighbours (c - 1 and c + 1); the query lies inside cell c,
// whose own slab contributes 0
template <class Tr>
__device__ __forceinline__ void nn_axis_gaps2(int c, float vsize, float q, float& below, float& above)
{
    float lo, hi, t;
    Tr::cell_span(c - 1, t, hi);  // upper end of the cell below
    Tr::cell_span(c + 1, lo, t);  // lower end of the cell above
    const float e0 = (fabsf((float)(c - 1)) + 2.0f) * vsize * 1e-6f, e1 = (fabsf((float)(c + 1)) + 2.0f) * vsize * 1e-6f;
    const float d0 = fmaxf(0.f, q - (hi * vsize + e0)), d1 = fmaxf(0.f, lo * vsize - e1 - q);
    below = d0 * d0 * 0.9999f, above = d1 * d1 * 0.9999f;
}

// nearest of the first `cnt` slots of a row: strict '<' in ascending slot order keeps the first minimum
__device__ __forceinline__ void row_nearest(const float4* __restrict__ row, u32 cnt, const float3& p, float& cd, u32& cs)
{
    unsigned long long qxy;
    asm("mov.b64 %0, {%1,%2};" : "=l"(qxy) : "f"(p.x), "f"(p.y));
    cd = __int_as_float(0x7f800000);
    cs = 0;
#pragma unroll 2
    for (u32 s = 0; s < cnt; s += 2)
    {
        float4 a, b;
        ld_row_pair(row + s, a, b);
        const float d0 = sqr_dist_packed(a, qxy, p.z);
        if (d0 < cd) cd = d0, cs = s;
        const float d1 = sqr_dist_packed(b, qxy, p.z);
        if (s + 1 < cnt && d1 < cd) cd = d1, cs = s + 1;
    }
}

template <class Tr>
__global__ void __launch_bounds__(NN_OWN_BLOCK)
    k_nn_own(VoxelMapDev m, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 n, u32 stride, PoseD pose,
             int use_pose, float inv, float vsize, float bd_init, NNWork w)
{
    constexpr int CAP = Tr::CAP;
    constexpr int NW  = NN_OWN_BLOCK / 32;
    pdl_trigger();  // k_nn_tasks may be scheduled as soon as the last blocks of this grid are resident; it waits before reading
    __shared__ u32 s_cnt[3][NW];
    __shared__ u32 s_base[3];
    const int  lane = lane_id(), wid = threadIdx.x >> 5;
    const u32  q     = blockIdx.x * NN_OWN_BLOCK + threadIdx.x;
    const bool valid = q < n;
    float3     p     = make_float3(0.f, 0.f, 0.f);
    int        cx = 0, cy = 0, cz = 0;
    bool       ok = false;
    if (valid)
    {
        const size_t qi = (size_t)q * stride;
        p               = make_float3(qx[qi], qy[qi], qz[qi]);
        if (use_pose) p = compose_point(pose, p.x, p.y, p.z);
        ok = Tr::to_index(p.x, inv, cx) & Tr::to_index(p.y, inv, cy) & Tr::to_index(p.z, inv, cz);
    }
    float bd = bd_init;
    u32   br = 0xFFFFFFFFu, bv = 0, t = 0;
    if (ok)
    {
        u32 vid, cnt;
        if (probe_ro(m.table, m.mask, pack_key(cx, cy, cz), vid, cnt) && cnt)
        {   // the query's own voxel: rank 13 * 32 + slot
            float cd;
            u32   cs;
            row_nearest(m.pts + (size_t)vid * CAP, cnt, p, cd, cs);
            if (cd < bd || (cd == bd && 13u * 32u + cs < br)) bd = cd, br = 13u * 32u + cs, bv = vid;
        }
        // the neighbour cells whose box can still hold a point at least as close: the box distance separates per axis, and
        // the query's own slab contributes nothing
        float g[3][3];
        nn_axis_gaps2<Tr>(cx, vsize, p.x, g[0][0], g[0][2]);
        nn_axis_gaps2<Tr>(cy, vsize, p.y, g[1][0], g[1][2]);
        nn_axis_gaps2<Tr>(cz, vsize, p.z, g[2][0], g[2][2]);
        g[0][1] = g[1][1] = g[2][1] = 0.f;
#pragma unroll
        for (int b = 0; b < 27; b++)
            if (b != 13 && !(g[0][b / 9] + g[1][(b / 3) % 3] + g[2][b % 3] > bd)) t |= 1u << b;
    }
    if (valid)
    {
        w.qpt[q]  = make_float4(p.x, p.y, p.z, __uint_as_float(bv));
        w.qkey[q] = ((u64)__float_as_uint(bd) << 32) | (u64)br;
    }
    // queue space: per-warp counts -> one atomicAdd per block and queue
    const u32 TIER[3] = {NN_FACE, NN_EDGE, NN_CORNER};
    u32       wcnt[3];
#pragma unroll
    for (int k = 0; k < 3; k++)
    {
        wcnt[k] = __reduce_add_sync(MMG_FULL, (u32)__popc(t & TIER[k]));
        if (lane == 0) s_cnt[k][wid] = wcnt[k];
    }
    __syncthreads();
    if (threadIdx.x < 3)
    {
        u32 tot = 0;
#pragma unroll
        for (int i = 0; i < NW; i++) tot += s_cnt[threadIdx.x][i];
        s_base[threadIdx.x] = tot ? atomicAdd(&w.count[threadIdx.x], tot) : 0u;
    }
    __syncthreads();
    const u32 any = __reduce_or_sync(MMG_FULL, t);
    if (!any) return;
    const u32 lt = (1u << lane) - 1u;
#pragma unroll
    for (int k = 0; k < 3; k++)
    {
        if (!wcnt[k]) continue;
        u32 o = s_base[k];
        for (int i = 0; i < wid; i++) o += s_cnt[k][i];
        u32* Q = w.queue[k];
        for (u32 r = any & TIER[k]; r; r &= r - 1)
        {   // cell-major inside the warp: consecutive tasks are consecutive scan points looking at the same neighbour cell
            const int b   = __ffs(r) - 1;
            const u32 bal = __ballot_sync(MMG_FULL, (t >> b) & 1u);
            if ((t >> b) & 1u) Q[o + __popc(bal & lt)] = ((u32)b << 27) | q;
            o += __popc(bal);
        }
    }
}

// One thread per task, the three queues in tier order (face cells, then edges, then corners: the closer cells have
// tightened a query's bound before its farther ones are looked at). MINB = resident blocks of 256 per SM the kernel is
// compiled for (register budget): the work is bound by L1 tag look-ups and by latency (task -> query record -> table probe
// -> row are four dependent loads), so more resident warps hide more of it. A lane whose task is pruned by the re-check
// or finds no voxel fetches its next task right away instead of idling while the others walk their rows.
template <class Tr, int MINB>
__global__ void __launch_bounds__(256, MINB)
    k_nn_tasks(VoxelMapDev m, float inv, float vsize, NNWork w)
{
    constexpr int CAP     = Tr::CAP;
    const u32     gtid    = blockIdx.x * blockDim.x + threadIdx.x, gstride = gridDim.x * blockDim.x;
    pdl_trigger();
    pdl_wait();  // everything k_nn_own wrote (query records, keys, queues, tails) is visible from here on
#pragma unroll
    for (int k = 0; k < 3; k++)
    {   // unrolled: the queue pointers stay in registers (a dynamic index would put the parameter block on the stack)
        const u32  cnt = w.count[k];
        const u32* Q   = w.queue[k];
        u32        i   = gtid;
        for (;;)
        {
            bool   live = false;
            u32    q = 0, c = 0, vid = 0;
            int    b = 0;
            float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
            u64    key0 = 0;
#pragma unroll 1
            while (i < cnt)
            {
                const u32 tk = __ldg(Q + i);
                i += gstride;
                q    = tk & 0x7FFFFFFu;
                b    = (int)(tk >> 27);
                p    = __ldg(w.qpt + q);
                key0 = __ldcg(w.qkey + q);  // L2: sees what the tasks that finished so far have found
                int cx, cy, cz;
                Tr::to_index(p.x, inv, cx), Tr::to_index(p.y, inv, cy), Tr::to_index(p.z, inv, cz);
                const int x = cx + b / 9 - 1, y = cy + (b / 3) % 3 - 1, z = cz + b % 3 - 1;
                if (!index_in_range(x, y, z)) continue;
                if (voxel_lower_bound<Tr>(x, y, z, vsize, p.x, p.y, p.z) > __uint_as_float((u32)(key0 >> 32))) continue;
                if (!probe_ro(m.table, m.mask, pack_key(x, y, z), vid, c) || !c) continue;
                live = true;
                break;
            }
            if (!__any_sync(MMG_FULL, live)) break;
            if (live)
            {
                float cd;
                u32   cs;
                row_nearest(m.pts + (size_t)vid * CAP, c, make_float3(p.x, p.y, p.z), cd, cs);
                const u64 key = ((u64)__float_as_uint(cd) << 32) | (u64)((u32)b * 32u + cs);
                if (key < key0) atomicMin(w.qkey + q, key);
            }
        }
    }
}

template <class Tr>
__global__ void __launch_bounds__(256)
    k_nn_finish(VoxelMapDev m, u32 n, float inv, NNWork w, QueryOut out)
{
    constexpr int CAP  = Tr::CAP;
    const int     lane = lane_id();
    pdl_wait();
    const u32     q0   = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u;
    const u32     q    = q0 + (u32)lane;
    if (q0 >= n) return;
    const bool valid = q < n;
    float4     t     = make_float4(0.f, 0.f, 0.f, 0.f);
    u64        id    = 0;
    float      d2    = 0.f;
    bool       found = false;
    if (valid)
    {
        const u64 key = w.qkey[q];
        const u32 br  = (u32)key;
        found         = br != 0xFFFFFFFFu;
        if (found)
        {
            const float4 p = w.qpt[q];
            int          cx, cy, cz;
            Tr::to_index(p.x, inv, cx), Tr::to_index(p.y, inv, cy), Tr::to_index(p.z, inv, cz);
            const int b = (int)(br >> 5), slot = (int)(br & 31);
            const int x = cx + b / 9 - 1, y = cy + (b / 3) % 3 - 1, z = cz + b % 3 - 1;
            u32       vid = __float_as_uint(p.w), c;  // own voxel: its row id came along with the query record
            if (b != 13) probe_ro(m.table, m.mask, pack_key(x, y, z), vid, c);
            t  = __ldg(m.pts + (size_t)vid * CAP + slot);
            id = plain_id<Tr::SLOT_BITS>(x, y, z, slot);
            d2 = __uint_as_float((u32)(key >> 32));
        }
        out.found[q] = found ? 1 : 0;
        out.d2[q]    = d2;
        out.id[q]    = id;
    }
    // xyz is an array of structs: transpose through shuffles so the warp writes its 96 floats as three fully coalesced
    // 128-byte stores (matters most when `out` is mapped host memory: whole PCIe write bursts)
#pragma unroll
    for (int k = 0; k < 3; k++)
    {
        const int   ww = 32 * k + lane, src = ww / 3, c = ww - 3 * src;
        const float vx = __shfl_sync(MMG_FULL, t.x, src), vy = __shfl_sync(MMG_FULL, t.y, src), vz = __shfl_sync(MMG_FULL, t.z, src);
        if (q0 + (u32)src < n) out.xyz[3 * (size_t)q0 + ww] = c == 0 ? vx : (c == 1 ? vy : vz);
    }
}

// k_nn_finish_match: the last stage of the pipeline for mmg_match_batch. The distance threshold is
// already the initial search radius of k_nn_own, so "found" means "paired"; instead of dense per-query results this stage
// writes the pairings only, compacted in local-point order, in the same launch: a single-pass scan over tiles of 256
// queries with decoupled look-back (each tile publishes its count, then sums its predecessors' counts until it meets one
// whose inclusive prefix is known). 28 bytes leave the SM per KEPT pairing and nothing else.
struct MatchOut
{
    u32*   idx;    // local point index of each pairing
    float* xyz;    // map point
    float* d2;
    u64*   id;
    u64*   status; // one word per tile, zeroed before the launch: flag (2 bits) | count
    u32*   ticket; // tile dispenser, zeroed before the launch
    u32*   total;  // running number of pairings: read by tile 0 (earlier chunks of the batch), written by the last tile
};
template <class Tr>
__global__ void __launch_bounds__(256)
    k_nn_finish_match(VoxelMapDev m, u32 n, u32 q_base, float inv, NNWork w, MatchOut o)
{
    constexpr int CAP = Tr::CAP;
    constexpr u64 FLAG_A = 1ull << 62, FLAG_P = 2ull << 62, VAL = FLAG_A - 1;
    __shared__ u32 s_tile, s_wsum[8], s_base;
    const int      lane = lane_id(), wid = threadIdx.x >> 5;
    pdl_wait();
    if (threadIdx.x == 0) s_tile = atomicAdd(o.ticket, 1u);
    __syncthreads();
    const u32  tile  = s_tile;
    const u32  q     = tile * 256u + threadIdx.x;
    const bool valid = q < n;
    u64        key   = 0;
    bool       kept  = false;
    if (valid)
    {
        key  = w.qkey[q];
        kept = (u32)key != 0xFFFFFFFFu;
    }
    const u32 bal = __ballot_sync(MMG_FULL, kept);
    if (lane == 0) s_wsum[wid] = __popc(bal);
    __syncthreads();
    u32 before = __popc(bal & ((1u << lane) - 1u)), tile_total = 0;
#pragma unroll
    for (int k = 0; k < 8; k++)
    {
        const u32 c = s_wsum[k];
        if (k < wid) before += c;
        tile_total += c;
    }
    if (wid == 0)
    {
        u32 excl = 0;
        if (tile == 0)
        {
            excl = *o.total;
            if (lane == 0) atomicExch(o.status, FLAG_P | (u64)(excl + tile_total));
        }
        else
        {
            if (lane == 0) atomicExch(o.status + tile, FLAG_A | (u64)tile_total);
            int look = (int)tile - 1;
            for (;;)
            {
                const int i = look - lane;
                u64       v = FLAG_P;  // tiles before the first count as a known prefix of 0
                if (i >= 0)
                    do {
                        v = *reinterpret_cast<volatile u64*>(o.status + i);
                    } while ((v >> 62) == 0);
                const u32 pm = __ballot_sync(MMG_FULL, (v >> 62) == 2);
                if (pm)
                {
                    const int first = __ffs(pm) - 1;  // nearest predecessor with an inclusive prefix
                    excl += __reduce_add_sync(MMG_FULL, lane <= first ? (u32)(v & VAL) : 0u);
                    break;
                }
                excl += __reduce_add_sync(MMG_FULL, (u32)(v & VAL));
                look -= 32;
            }
            if (lane == 0) atomicExch(o.status + tile, FLAG_P | (u64)(excl + tile_total));
        }
        if (lane == 0)
        {
            s_base = excl;
            if (tile == (n + 255u) / 256u - 1u) *o.total = excl + tile_total;
        }
    }
    __syncthreads();
    if (kept)
    {
        const u32    br = (u32)key;
        const float4 p  = w.qpt[q];
        int          cx, cy, cz;
        Tr::to_index(p.x, inv, cx), Tr::to_index(p.y, inv, cy), Tr::to_index(p.z, inv, cz);
        const int b = (int)(br >> 5), slot = (int)(br & 31);
        const int x = cx + b / 9 - 1, y = cy + (b / 3) % 3 - 1, z = cz + b % 3 - 1;
        u32       vid = __float_as_uint(p.w), c;
        if (b != 13) probe_ro(m.table, m.mask, pack_key(x, y, z), vid, c);
        const float4 t  = __ldg(m.pts + (size_t)vid * CAP + slot);
        const size_t at = (size_t)s_base + before;
        o.idx[at] = q_base + q;
        o.xyz[3 * at] = t.x, o.xyz[3 * at + 1] = t.y, o.xyz[3 * at + 2] = t.z;
        o.d2[at] = __uint_as_float((u32)(key >> 32));
        o.id[at] = plain_id<Tr::SLOT_BITS>(x, y, z, slot);
    }
}
}  // namespace mmg
