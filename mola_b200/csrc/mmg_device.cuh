// mmg_device.cuh — device-side primitives shared by the map kernels (sm_100a).
//
// Exact-arithmetic helpers: every fp32 result that the reference computes on the
// host without FMA contraction is computed here with __fmul_rn/__fadd_rn/__fsub_rn
// so nvcc cannot fuse it, in the same association order:
//   index    = (int32)(x * voxel_size_inv)          HashedVoxelPointCloud.h:76-83
//   d2       = (dx*dx + dy*dy) + dz*dz              TPoint3Df::sqrNorm (MRPT)
//   compose  = ((R0*x + R1*y) + R2*z) + t  in fp64  CPose3D::composePoint (MRPT, N3)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mmg
{
typedef unsigned long long u64;
typedef uint32_t           u32;

constexpr u64 KEY_EMPTY    = 0xFFFFFFFFFFFFFFFFull;
constexpr u32 VID_INVALID  = 0xFFFFFFFFu;
constexpr int IDX_LIMIT    = 1 << 20;  // voxel indices must satisfy |c| < 2^20
constexpr int FULL_MASK_I  = -1;
#define MMG_FULL 0xFFFFFFFFu

// 16-byte hash-table entry: one LDG.128 returns key, row id and point count.
struct __align__(16) Entry
{
    u64 key;
    u32 vid;
    u32 cnt;
};

struct PoseD
{
    double m[12];  // row-major R then t
};

__device__ __forceinline__ float3 compose_point(const PoseD& T, float lx, float ly, float lz)
{
    const double x = (double)lx, y = (double)ly, z = (double)lz;
    const double gx = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T.m[0], x), __dmul_rn(T.m[1], y)), __dmul_rn(T.m[2], z)), T.m[9]);
    const double gy = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T.m[3], x), __dmul_rn(T.m[4], y)), __dmul_rn(T.m[5], z)), T.m[10]);
    const double gz = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T.m[6], x), __dmul_rn(T.m[7], y)), __dmul_rn(T.m[8], z)), T.m[11]);
    return make_float3(__double2float_rn(gx), __double2float_rn(gy), __double2float_rn(gz));
}

__device__ __forceinline__ float sqr_dist(float ax, float ay, float az, float bx, float by, float bz)
{
    const float dx = __fsub_rn(ax, bx), dy = __fsub_rn(ay, by), dz = __fsub_rn(az, bz);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

// Voxel-map flavours. CAP = points per voxel row, SLOT_BITS = id bits for the slot.
struct HashedTraits
{  // mola::HashedVoxelPointCloud: truncation toward zero
    static constexpr int CAP = 32, SLOT_BITS = 5;
    __device__ static __forceinline__ bool to_index(float v, float inv, int& c)
    {
        const float s = __fmul_rn(v, inv);
        if (!(fabsf(s) < (float)IDX_LIMIT)) return false;  // also rejects NaN
        c = __float2int_rz(s);
        return true;
    }
    // index range [lo, hi] (in voxel units) of the coordinates that truncate to c: index 0 spans (-1, 1)
    __device__ static __forceinline__ void cell_span(int c, float& lo, float& hi)
    {
        lo = (float)(c > 0 ? c : c - 1);
        hi = (float)(c < 0 ? c : c + 1);
    }
};
struct SparseVoxelTraits
{  // mola::SparseVoxelPointCloud: mrpt::round = lrint (ties to even), SparseVoxelPointCloud.h:108-115
    static constexpr int CAP = 16, SLOT_BITS = 4;
    __device__ static __forceinline__ bool to_index(float v, float inv, int& c)
    {
        const float s = __fmul_rn(v, inv);
        if (!(fabsf(s) < (float)(IDX_LIMIT - 1))) return false;
        c = __float2int_rn(s);
        return true;
    }
    __device__ static __forceinline__ void cell_span(int c, float& lo, float& hi)
    {
        lo = (float)c - 0.5f;
        hi = (float)c + 0.5f;
    }
};
// Conservative lower bound of the squared distance between q and any point stored in voxel
// (cx, cy, cz): the voxel's coordinate box is widened by a few ulps of its largest coordinate
// (the index is computed from fl(x * fl(1/v))) and the result is scaled down, so a voxel is
// only ever skipped when every point in it is strictly farther than the bound.
template <class Tr>
__device__ __forceinline__ float voxel_lower_bound(int cx, int cy, int cz, float vsize, float qx, float qy, float qz)
{
    const int   c[3] = {cx, cy, cz};
    const float q[3] = {qx, qy, qz};
    float       s    = 0.f;
#pragma unroll
    for (int a = 0; a < 3; a++)
    {
        float lo, hi;
        Tr::cell_span(c[a], lo, hi);
        const float eps = (fabsf((float)c[a]) + 2.0f) * vsize * 1e-6f;
        const float d   = fmaxf(0.f, fmaxf(lo * vsize - eps - q[a], q[a] - (hi * vsize + eps)));
        s += d * d;
    }
    return s * 0.9999f;
}

__device__ __host__ __forceinline__ bool index_in_range(int cx, int cy, int cz)
{
    return cx > -IDX_LIMIT && cx < IDX_LIMIT && cy > -IDX_LIMIT && cy < IDX_LIMIT && cz > -IDX_LIMIT && cz < IDX_LIMIT;
}
// 3 x 21 bits, biased; bit 63 stays clear so KEY_EMPTY can never be a key.
__device__ __host__ __forceinline__ u64 pack_key(int cx, int cy, int cz)
{
    return (u64)(u32)(cx + IDX_LIMIT) | ((u64)(u32)(cy + IDX_LIMIT) << 21) | ((u64)(u32)(cz + IDX_LIMIT) << 42);
}
__device__ __host__ __forceinline__ void unpack_key(u64 k, int& cx, int& cy, int& cz)
{
    cx = (int)(k & 0x1FFFFF) - IDX_LIMIT;
    cy = (int)((k >> 21) & 0x1FFFFF) - IDX_LIMIT;
    cz = (int)((k >> 42) & 0x1FFFFF) - IDX_LIMIT;
}
__device__ __host__ __forceinline__ u64 hash_key(u64 k)
{  // 32-bit multiply-xorshift mix of the two key halves (7 integer instructions; tables stay far below 2^32 entries).
   // The reference's 20-bit Teschner hash (index3d_t.h:83-101) is not observable in results and clusters badly beyond
   // 1M buckets, so the table uses its own.
    u32 h = (u32)k * 0x85EBCA6Bu ^ (u32)(k >> 32) * 0xC2B2AE35u;
    h ^= h >> 15;
    h *= 0x2C1B3C6Du;
    h ^= h >> 12;
    h *= 0x297A2D39u;
    h ^= h >> 15;
    return h;
}
// g2plain: slot bits + 3 x 20 index bits (HashedVoxelPointCloud.h:95-108, SparseVoxelPointCloud.h:127-140).
template <int SLOT_BITS>
__device__ __host__ __forceinline__ u64 plain_id(int cx, int cy, int cz, int slot)
{
    const u64 F = (1u << 20) - 1;
    return ((u64)slot & ((1u << SLOT_BITS) - 1)) | (((u64)(long long)cx & F) << SLOT_BITS) |
           (((u64)(long long)cy & F) << (SLOT_BITS + 20)) | (((u64)(long long)cz & F) << (SLOT_BITS + 40));
}

__device__ __forceinline__ uint4 ld_entry_nc(const Entry* e)
{
    return __ldg(reinterpret_cast<const uint4*>(e));
}
// Read-only probe (query kernels; the table is immutable while they run).
__device__ __forceinline__ bool probe_ro(const Entry* __restrict__ tab, u64 mask, u64 key, u32& vid, u32& cnt)
{
    u64 h = hash_key(key) & mask;
#pragma unroll 1
    for (;;)
    {
        const uint4 e = ld_entry_nc(tab + h);
        const u64   k = (u64)e.x | ((u64)e.y << 32);
        if (k == key)
        {
            vid = e.z;
            cnt = e.w;
            return true;
        }
        if (k == KEY_EMPTY) return false;
        h = (h + 1) & mask;
    }
}

// float <-> order-preserving uint (for atomicMin/atomicMax bounding boxes)
__device__ __host__ __forceinline__ u32 float_to_ordered(float f)
{
#ifdef __CUDA_ARCH__
    const u32 b = __float_as_uint(f);
#else
    union { float f; u32 u; } c; c.f = f; const u32 b = c.u;
#endif
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __host__ __forceinline__ float ordered_to_float(u32 o)
{
    const u32 b = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    union { float f; u32 u; } c; c.u = b; return c.f;
#endif
}

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

}  // namespace mmg

// Two consecutive row slots with one 256-bit load (sm_100 LDG.256): half the load instructions and half the L1 tag
// look-ups of the row walks, which ncu shows to be the busiest unit of these kernels (7 look-ups per warp load in
// k_nn_tasks, 0.6 per SM cycle). Rows are 512-byte aligned and s is even, so the pair never straddles a 32-byte boundary.
__device__ __forceinline__ void ld_row_pair(const float4* p, float4& a, float4& b)
{
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
        : "l"(p));
}
