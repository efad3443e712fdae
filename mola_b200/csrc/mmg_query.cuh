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
template <class Tr, int G>
__device__ __forceinline__ void nn_group_pass(const VoxelMapDev& m, u32 cells, u32 vid, u32 cnt, int lane, const float (&gx)[G],
                                              const float (&gy)[G], const float (&gz)[G], float (&bd)[G], u32 (&br)[G])
{
    constexpr int CAP = Tr::CAP;
    while (cells)
    {
        const int b = __ffs(cells) - 1;
        cells &= cells - 1;
        const u32 v = __shfl_sync(MMG_FULL, vid, b), c = __shfl_sync(MMG_FULL, cnt, b);
        if ((u32)lane < c)
        {
            const float4 t    = __ldg(&m.pts[(size_t)v * CAP + lane]);
            const u32    rank = (u32)b * 32u + (u32)lane;
#pragma unroll
            for (int g = 0; g < G; g++)
            {
                const float d = sqr_dist(t.x, t.y, t.z, gx[g], gy[g], gz[g]);
                if (d < bd[g])
                {
                    bd[g] = d;
                    br[g] = rank;
                }
            }
        }
    }
}

template <class Tr>
__global__ void __launch_bounds__(256)
    k_nn_single_grouped(VoxelMapDev m, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 n, u32 stride,
                        PoseD pose, int use_pose, float inv, QueryOut out)
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
                if (ng == 1)
                {
                    float ax[1] = {gx[0]}, ay[1] = {gy[0]}, az[1] = {gz[0]}, ad[1] = {INF};
                    u32   ar[1] = {0xFFFFFFFFu};
                    nn_group_pass<Tr, 1>(m, cells, vid, cnt, lane, ax, ay, az, ad, ar);
                    bd[0] = ad[0], br[0] = ar[0];
                }
                else
                    nn_group_pass<Tr, G>(m, cells, vid, cnt, lane, gx, gy, gz, bd, br);
#pragma unroll
                for (int g = 0; g < G; g++)
                {
                    if (g >= ng) break;
                    const u32 db   = __float_as_uint(bd[g]);
                    const u32 dmin = __reduce_min_sync(MMG_FULL, db);
                    const u32 rr   = __reduce_min_sync(MMG_FULL, db == dmin ? br[g] : 0xFFFFFFFFu);
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
}  // namespace mmg
