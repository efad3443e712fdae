// mmg_trees.cuh — device side of mola::SparseTreesPointCloud.
//
// The reference keeps one CSimplePointsMap + lazily rebuilt nanoflann kd-tree per coarse
// block (grid_size, 10 m by default; SparseTreesPointCloud.h:126-234) and answers every
// query with an EXACT search inside the query's own block (nn_single/multiple,
// SparseTreesPointCloud.cpp:613-643) or inside the few blocks that a radius touches
// (:645-713). Ids are block-local insertion indices. A kd-tree that is rebuilt on the host
// after every insert is the wrong shape for a GPU; here each block carries a two-level
// occupancy grid instead and the search is a branch-and-bound over it:
//
//   btab  : Entry[bcap]          block index (3 x 21 bit) -> block row `bid`
//   bkey  : u64[brows]           packed block index of row bid (KEY_EMPTY = free)
//   bcnt  : u32[brows]           points in the block (= next block-local index)
//   l1    : u32[brows*16]        8^3 super-cell occupancy bits
//   l0    : u32[brows*1024]      32^3 fine-cell occupancy bits, 64 per super-cell
//   ctab  : Entry[ccap]          (bid << 15 | cell) -> {head row of the cell's chain, count}
//   cpts  : float4[crows*32]     chain rows: x, y, z, block-local index (bit pattern)
//   cnext : u32[crows]           next (older, full) row of the chain
//
// A query first looks at the 27 fine cells around it (seeding the bound), then tests the
// 512 super-cells and, inside the survivors, the 64 fine cells against a conservative
// box-distance bound; only cells that can still hold a better (or equal: ties go to the
// lower index) point are read. Results are exact, so they equal the reference's brute force
// restated in oracle/mmo_port.cpp (block_nn).
#pragma once
#include "mmg_device.cuh"
#include "mmg_sort.cuh"

namespace mmg
{
constexpr int TR_ROW   = 32;    // points per chain row
constexpr int TR_CELLS = 32768; // fine cells per block (32^3)

struct TreesCounters
{
    u32 brow_hi, bfree_n;  // block rows: high-water mark, free stack size
    u32 crow_hi, cfree_n;  // chain rows
    u32 ncells;            // live cell entries
    u32 oob;
    u32 nseg;
    u32 ntodo;             // points k_trees_clear left to the exact search (zeroed again by k_trees_flags)
    u32 maxcnt;            // scratch for reductions
    u32 nacc;              // points the current batch keeps (total of the flag scan)
    unsigned long long npoints;
};

struct TreesDev
{
    Entry*  btab;
    u64     bmask;
    u64*    bkey;
    u32*    bcnt;
    u32*    l1;
    u32*    l0;
    u32*    bfree;
    Entry*  ctab;
    u64     cmask;
    float4* cpts;
    u32*    cnext;
    u32*    cfree;
    TreesCounters* ctr;
    u32*    bbox;  // ordered-uint min xyz / max xyz, lives in the map's Counters block
    float   g, ginv, ghalf;
};

// coordToOuterIdx: (int32)((x + g/2) * (1/g)), truncation (SparseTreesPointCloud.h:70-77)
__device__ __forceinline__ bool trees_index(const TreesDev& t, float v, int& c)
{
    const float s = __fmul_rn(__fadd_rn(v, t.ghalf), t.ginv);
    if (!(fabsf(s) < (float)IDX_LIMIT)) return false;
    c = __float2int_rz(s);
    return true;
}
// Extent of block index k along one axis. Truncation makes index 0 twice as wide.
__device__ __forceinline__ void trees_axis(const TreesDev& t, int k, float& lo, float& size)
{
    if (k > 0)
        lo = ((float)k - 0.5f) * t.g, size = t.g;
    else if (k == 0)
        lo = -1.5f * t.g, size = 2.0f * t.g;
    else
        lo = ((float)k - 1.5f) * t.g, size = t.g;
}
struct BlockGeom
{
    float lo[3], cs[3], ics[3], margin[3];  // origin, fine-cell size, its inverse, safety margin
};
__device__ __forceinline__ BlockGeom trees_geom(const TreesDev& t, int kx, int ky, int kz)
{
    BlockGeom g;
    const int k[3] = {kx, ky, kz};
#pragma unroll
    for (int a = 0; a < 3; a++)
    {
        float size;
        trees_axis(t, k[a], g.lo[a], size);
        g.cs[a]     = size * (1.0f / 32.0f);
        g.ics[a]    = 32.0f / size;
        g.margin[a] = size * 1e-4f + (fabsf(g.lo[a]) + size) * 2e-6f;
    }
    return g;
}
// fine cell (0..31 per axis) of a point that belongs to the block
__device__ __forceinline__ int trees_cell_axis(const BlockGeom& g, int a, float v)
{
    const int c = (int)floorf((v - g.lo[a]) * g.ics[a]);
    return min(31, max(0, c));
}
// blocked cell id: super-cell (3 x 3 bit) * 64 + position inside it (3 x 2 bit)
__device__ __forceinline__ u32 trees_cell_id(int cx, int cy, int cz)
{
    return (u32)((((cx >> 2) * 8 + (cy >> 2)) * 8 + (cz >> 2)) * 64 + ((cx & 3) * 4 + (cy & 3)) * 4 + (cz & 3));
}
__device__ __forceinline__ void trees_cell_xyz(u32 id, int& cx, int& cy, int& cz)
{
    const u32 sc = id >> 6, in = id & 63;
    cx = (int)((sc >> 6) << 2 | (in >> 4));
    cy = (int)(((sc >> 3) & 7) << 2 | ((in >> 2) & 3));
    cz = (int)((sc & 7) << 2 | (in & 3));
}
// conservative lower bound of the squared distance from q to any point filed in the
// box of `n` cells starting at cell c0 (per axis)
__device__ __forceinline__ float trees_box_lb(const BlockGeom& g, const float q[3], int cx, int cy, int cz, int ncell)
{
    const int c[3] = {cx, cy, cz};
    float     s    = 0.f;
#pragma unroll
    for (int a = 0; a < 3; a++)
    {
        // points are filed by clamping, so the first/last cell also takes the (rounding-sized)
        // overshoot of points that sit on the block boundary: wider margin there
        const float mlo = (c[a] == 0) ? 8.f * g.margin[a] : g.margin[a];
        const float mhi = (c[a] + ncell >= 32) ? 8.f * g.margin[a] : g.margin[a];
        const float lo  = g.lo[a] + (float)c[a] * g.cs[a] - mlo;
        const float hi  = g.lo[a] + (float)(c[a] + ncell) * g.cs[a] + mhi;
        const float d  = fmaxf(0.f, fmaxf(lo - q[a], q[a] - hi));
        s += d * d;
    }
    return s * 0.9999f;
}
__device__ __forceinline__ u64 trees_cell_key(u32 bid, u32 cell) { return ((u64)bid << 15) | (u64)cell; }

// ------------------------------------------------------------ block traversal
// Calls V.row(point, valid) (all 32 lanes, lane l <- slot l of a chain row) for every chain
// row of every fine cell of block `bid` whose box may still hold a point with d2 <= V.bound().
// V.cell_done() runs after each cell (the visitor tightens its bound there). Control flow is
// warp-uniform throughout, so visitors may use warp collectives.
template <class V>
__device__ __forceinline__ void trees_visit_cell(const TreesDev& t, u32 bid, u32 cell, V& vis)
{
    const int lane = lane_id();
    u32       head, cnt;
    if (!probe_ro(t.ctab, t.cmask, trees_cell_key(bid, cell), head, cnt) || cnt == 0) return;
    u32 fill  = ((cnt - 1) & (TR_ROW - 1)) + 1;
    u32 nrows = (cnt + TR_ROW - 1) / TR_ROW;
    u32 row   = head;
    for (u32 r = 0; r < nrows; r++)
    {
        const bool valid = (u32)lane < fill;
        float4     p     = make_float4(0.f, 0.f, 0.f, 0.f);
        if (valid) p = __ldg(&t.cpts[(size_t)row * TR_ROW + lane]);
        vis.row(p, valid);
        row  = __ldg(&t.cnext[row]);
        fill = TR_ROW;
    }
    vis.cell_done();
}

template <class V>
__device__ __forceinline__ void trees_traverse(const TreesDev& t, u32 bid, const BlockGeom& g, const float q[3], V& vis)
{
    const int lane = lane_id();
    // 1) the 27 fine cells around the query seed the bound
    int c0[3];
#pragma unroll
    for (int a = 0; a < 3; a++) c0[a] = trees_cell_axis(g, a, q[a]);
    {
        bool has  = false;
        u32  cell = 0;
        if (lane < 27)
        {
            const int x = c0[0] + lane / 9 - 1, y = c0[1] + (lane / 3) % 3 - 1, z = c0[2] + lane % 3 - 1;
            if (x >= 0 && x < 32 && y >= 0 && y < 32 && z >= 0 && z < 32)
            {
                cell = trees_cell_id(x, y, z);
                has  = (__ldg(&t.l0[(size_t)bid * 1024 + (cell >> 5)]) >> (cell & 31)) & 1u;
            }
        }
        u32 m = __ballot_sync(MMG_FULL, has);
        while (m)
        {
            const int j = __ffs(m) - 1;
            m &= m - 1;
            trees_visit_cell(t, bid, __shfl_sync(MMG_FULL, cell, j), vis);
        }
    }
    // 2) branch and bound over the 8^3 super-cells and their 4^3 fine cells
    for (int w = 0; w < 16; w++)
    {
        const u32 l1w = __ldg(&t.l1[(size_t)bid * 16 + w]);
        if (!l1w) continue;
        float lbs = 0.f;
        bool  sc_ok = false;
        if ((l1w >> lane) & 1u)
        {
            const int sc = w * 32 + lane;
            lbs          = trees_box_lb(g, q, (sc >> 6) << 2, ((sc >> 3) & 7) << 2, (sc & 7) << 2, 4);
            sc_ok        = !(lbs > vis.bound());
        }
        u32 sm = __ballot_sync(MMG_FULL, sc_ok);
        while (sm)
        {
            const int s = __ffs(sm) - 1;
            sm &= sm - 1;
            if (__shfl_sync(MMG_FULL, lbs, s) > vis.bound()) continue;
            const int sc = w * 32 + s;
            const int bx = (sc >> 6) << 2, by = ((sc >> 3) & 7) << 2, bz = (sc & 7) << 2;
            float     lb[2];
            u32       cm[2];
#pragma unroll
            for (int h = 0; h < 2; h++)
            {
                const u32 bits = __ldg(&t.l0[(size_t)bid * 1024 + sc * 2 + h]);
                const int in = h * 32 + lane, x = bx + (in >> 4), y = by + ((in >> 2) & 3), z = bz + (in & 3);
                bool      ok = (bits >> lane) & 1u;
                lb[h]        = 0.f;
                if (ok && abs(x - c0[0]) <= 1 && abs(y - c0[1]) <= 1 && abs(z - c0[2]) <= 1) ok = false;  // seen in step 1
                if (ok)
                {
                    lb[h] = trees_box_lb(g, q, x, y, z, 1);
                    ok    = !(lb[h] > vis.bound());
                }
                cm[h] = __ballot_sync(MMG_FULL, ok);
            }
#pragma unroll
            for (int h = 0; h < 2; h++)
                while (cm[h])
                {
                    const int j = __ffs(cm[h]) - 1;
                    cm[h] &= cm[h] - 1;
                    if (__shfl_sync(MMG_FULL, lb[h], j) > vis.bound()) continue;
                    trees_visit_cell(t, bid, (u32)(sc * 64 + h * 32 + j), vis);
                }
        }
    }
}

// ------------------------------------------------------------------ visitors
struct NN1Visitor
{
    float q[3];
    float bd, wb;  // this lane's best d2, the warp's best d2
    u32   bi;
    float bx, by, bz;
    __device__ __forceinline__ void init(float x, float y, float z)
    {
        q[0] = x, q[1] = y, q[2] = z;
        bd = wb = __int_as_float(0x7f800000);
        bi      = 0xFFFFFFFFu;
        bx = by = bz = 0.f;
    }
    __device__ __forceinline__ float bound() const { return wb; }
    __device__ __forceinline__ void  row(const float4& p, bool valid)
    {
        if (!valid) return;
        const float d = sqr_dist(q[0], q[1], q[2], p.x, p.y, p.z);
        const u32   i = __float_as_uint(p.w);
        if (d < bd || (d == bd && i < bi)) bd = d, bi = i, bx = p.x, by = p.y, bz = p.z;
    }
    __device__ __forceinline__ void cell_done() { wb = __uint_as_float(__reduce_min_sync(MMG_FULL, __float_as_uint(bd))); }
};

// k <= 32 best in ascending (d2, index): lane i holds the i-th best
struct KNNVisitor
{
    float q[3];
    float ld, lx, ly, lz, wb;
    u32   li, k;
    __device__ __forceinline__ void init(float x, float y, float z, u32 k_)
    {
        q[0] = x, q[1] = y, q[2] = z, k = k_;
        ld = wb = __int_as_float(0x7f800000);
        li      = 0xFFFFFFFFu;
        lx = ly = lz = 0.f;
    }
    __device__ __forceinline__ float bound() const { return wb; }
    __device__ __forceinline__ void  row(const float4& p, bool valid)
    {
        const int   lane = lane_id();
        const float d    = valid ? sqr_dist(q[0], q[1], q[2], p.x, p.y, p.z) : 0.f;
        const u32   i    = __float_as_uint(p.w);
        float       kd   = __shfl_sync(MMG_FULL, ld, k - 1);
        u32         ki   = __shfl_sync(MMG_FULL, li, k - 1);
        u32         m    = __ballot_sync(MMG_FULL, valid && (d < kd || (d == kd && i < ki)));
        while (m)
        {
            const int j = __ffs(m) - 1;
            m &= m - 1;
            const float cd = __shfl_sync(MMG_FULL, d, j), cx = __shfl_sync(MMG_FULL, p.x, j), cy = __shfl_sync(MMG_FULL, p.y, j),
                        cz = __shfl_sync(MMG_FULL, p.z, j);
            const u32 ci  = __shfl_sync(MMG_FULL, i, j);
            const u32 pos = __popc(__ballot_sync(MMG_FULL, ld < cd || (ld == cd && li < ci)));
            if (pos >= k) continue;
            const float ud = __shfl_up_sync(MMG_FULL, ld, 1), ux = __shfl_up_sync(MMG_FULL, lx, 1), uy = __shfl_up_sync(MMG_FULL, ly, 1),
                        uz = __shfl_up_sync(MMG_FULL, lz, 1);
            const u32 ui   = __shfl_up_sync(MMG_FULL, li, 1);
            if ((u32)lane > pos) ld = ud, li = ui, lx = ux, ly = uy, lz = uz;
            if ((u32)lane == pos) ld = cd, li = ci, lx = cx, ly = cy, lz = cz;
        }
    }
    __device__ __forceinline__ void cell_done() { wb = __shfl_sync(MMG_FULL, ld, k - 1); }
};

// hits with d2 < r2 (nanoflann's strict test): counts them, or writes them at `base`
// ------------------------------------------------------------------ queries
__device__ __forceinline__ bool trees_find_block(const TreesDev& t, float x, float y, float z, int& kx, int& ky, int& kz, u32& bid)
{
    if (!(trees_index(t, x, kx) & trees_index(t, y, ky) & trees_index(t, z, kz))) return false;
    u32 cnt;
    return probe_ro(t.btab, t.bmask, pack_key(kx, ky, kz), bid, cnt);
}

// nn_single_search: exact NN inside the query's own block (SparseTreesPointCloud.cpp:613-624);
// ties to the lowest block-local index. One warp per query.
static __global__ void __launch_bounds__(256)
    k_trees_nn1(TreesDev t, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 n, u32 stride, PoseD pose,
                int use_pose, float notfound_d2, uint8_t* __restrict__ ofound, float* __restrict__ oxyz, float* __restrict__ od2, u64* __restrict__ oid,
                const u32* __restrict__ list /*nullable: the queries k_trees_nn1_fast left open*/, const u32* __restrict__ nlist)
{
    const int lane   = lane_id();
    const u32 nwarps = (gridDim.x * blockDim.x) >> 5;
    if (list) n = *nlist;
    for (u32 w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < n; w += nwarps)
    {
        const u32    q  = list ? list[w] : w;
        const size_t qi = (size_t)q * stride;
        float3       p  = make_float3(qx[qi], qy[qi], qz[qi]);
        if (use_pose) p = compose_point(pose, p.x, p.y, p.z);
        int        kx, ky, kz;
        u32        bid   = 0;
        const bool has   = trees_find_block(t, p.x, p.y, p.z, kx, ky, kz, bid);
        bool       found = false;
        if (has)
        {
            const BlockGeom g = trees_geom(t, kx, ky, kz);
            NN1Visitor      v;
            v.init(p.x, p.y, p.z);
            trees_traverse(t, bid, g, v.q, v);
            const u32 idx = __reduce_min_sync(MMG_FULL, v.bd == v.wb ? v.bi : 0xFFFFFFFFu);
            found         = idx != 0xFFFFFFFFu;
            if (found && v.bd == v.wb && v.bi == idx)
            {
                if (ofound) ofound[q] = 1;
                if (oxyz) oxyz[3 * (size_t)q] = v.bx, oxyz[3 * (size_t)q + 1] = v.by, oxyz[3 * (size_t)q + 2] = v.bz;
                od2[q] = v.bd;
                if (oid) oid[q] = (u64)idx;
            }
        }
        if (!found && lane == 0)
        {
            if (ofound) ofound[q] = 0;
            if (oxyz) oxyz[3 * (size_t)q] = 0.f, oxyz[3 * (size_t)q + 1] = 0.f, oxyz[3 * (size_t)q + 2] = 0.f;
            od2[q] = notfound_d2;  // 0 for API outputs, -1 as the "no neighbour" marker of the clearance test
            if (oid) oid[q] = 0;
        }
    }
}

// nn_single_search fast path: one THREAD per query looks only at the 27 fine cells around the query. The answer is final
// when it is provably the block-wide one: its d2 lies strictly below the squared distance from the query to the nearest
// face of that 3x3x3 cell region that is interior to the block (everything filed outside the region is farther, so neither
// a closer point nor an equal-distance tie with a lower index can exist there). ICP-style queries (a scan re-posed by a
// few centimetres against a 0.31 m cell grid) nearly always certify; the rest are listed in `todo` and answered by the
// warp-per-query branch and bound (k_trees_nn1 with that list). Same results either way.
static __global__ void __launch_bounds__(256)
    k_trees_nn1_fast(TreesDev t, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 n, u32 stride, PoseD pose,
                     int use_pose, float notfound_d2, uint8_t* __restrict__ ofound, float* __restrict__ oxyz, float* __restrict__ od2, u64* __restrict__ oid,
                     u32* __restrict__ todo, u32* __restrict__ ntodo)
{
    const u32 q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    const size_t qi = (size_t)q * stride;
    float3       p  = make_float3(qx[qi], qy[qi], qz[qi]);
    if (use_pose) p = compose_point(pose, p.x, p.y, p.z);
    int   kx, ky, kz;
    u32   bid = 0;
    float bd  = __int_as_float(0x7f800000);
    u32   bi  = 0xFFFFFFFFu;
    float bx = 0.f, by = 0.f, bz = 0.f;
    bool  final_ = true;  // no block: nothing to find (SparseTreesPointCloud.cpp:617-619)
    if (trees_find_block(t, p.x, p.y, p.z, kx, ky, kz, bid))
    {
        const BlockGeom g    = trees_geom(t, kx, ky, kz);
        const float     qv[3] = {p.x, p.y, p.z};
        int             c0[3];
#pragma unroll
        for (int a = 0; a < 3; a++) c0[a] = trees_cell_axis(g, a, qv[a]);
        // own cell first (sets the bound), then the 26 neighbours that can still matter
        for (int it = 0; it < 27; it++)
        {
            const int j = it == 0 ? 13 : (it <= 13 ? it - 1 : it);
            const int x = c0[0] + j / 9 - 1, y = c0[1] + (j / 3) % 3 - 1, z = c0[2] + j % 3 - 1;
            if (x < 0 || x > 31 || y < 0 || y > 31 || z < 0 || z > 31) continue;
            const u32 cell = trees_cell_id(x, y, z);
            if (!((__ldg(&t.l0[(size_t)bid * 1024 + (cell >> 5)]) >> (cell & 31)) & 1u)) continue;
            if (it && trees_box_lb(g, qv, x, y, z, 1) > bd) continue;
            u32 head, cnt;
            if (!probe_ro(t.ctab, t.cmask, trees_cell_key(bid, cell), head, cnt) || cnt == 0) continue;
            u32 fill = ((cnt - 1) & (TR_ROW - 1)) + 1, row = head;
            for (u32 r = 0, nrows = (cnt + TR_ROW - 1) / TR_ROW; r < nrows; r++)
            {
                const float4* rp = t.cpts + (size_t)row * TR_ROW;
                for (u32 s = 0; s < fill; s++)
                {
                    const float4 c = __ldg(rp + s);
                    const float  d = sqr_dist(qv[0], qv[1], qv[2], c.x, c.y, c.z);
                    const u32    i = __float_as_uint(c.w);
                    if (d < bd || (d == bd && i < bi)) bd = d, bi = i, bx = c.x, by = c.y, bz = c.z;
                }
                row  = __ldg(&t.cnext[row]);
                fill = TR_ROW;
            }
        }
        // squared distance to the complement of the region inside the block (conservative, like trees_box_lb)
        float out = __int_as_float(0x7f800000);
#pragma unroll
        for (int a = 0; a < 3; a++)
        {
            const int   lo_c = max(c0[a] - 1, 0), hi_c = min(c0[a] + 1, 31);
            const float mg   = 8.f * g.margin[a];
            if (lo_c > 0) out = fminf(out, fmaxf(0.f, qv[a] - (g.lo[a] + (float)lo_c * g.cs[a]) - mg));
            if (hi_c < 31) out = fminf(out, fmaxf(0.f, (g.lo[a] + (float)(hi_c + 1) * g.cs[a]) - qv[a] - mg));
        }
        final_ = bi != 0xFFFFFFFFu && bd < out * out * 0.9999f;
    }
    if (!final_)
    {
        todo[atomicAdd(ntodo, 1u)] = q;
        return;
    }
    const bool found = bi != 0xFFFFFFFFu;
    if (ofound) ofound[q] = found ? 1 : 0;
    if (oxyz) oxyz[3 * (size_t)q] = bx, oxyz[3 * (size_t)q + 1] = by, oxyz[3 * (size_t)q + 2] = bz;
    od2[q] = found ? bd : notfound_d2;
    if (oid) oid[q] = found ? (u64)bi : 0;
}

// nn_multiple_search (:626-643): k <= 32 nearest inside the block, ascending (d2, index)
static __global__ void __launch_bounds__(256)
    k_trees_knn(TreesDev t, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 n, PoseD pose, int use_pose,
                u32 k, u32* __restrict__ ocount, float* __restrict__ oxyz, float* __restrict__ od2, u64* __restrict__ oid)
{
    const int lane   = lane_id();
    const u32 nwarps = (gridDim.x * blockDim.x) >> 5;
    for (u32 q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; q < n; q += nwarps)
    {
        float3 p = make_float3(qx[q], qy[q], qz[q]);
        if (use_pose) p = compose_point(pose, p.x, p.y, p.z);
        int  kx, ky, kz;
        u32  bid = 0, nout = 0;
        if (trees_find_block(t, p.x, p.y, p.z, kx, ky, kz, bid))
        {
            const BlockGeom g = trees_geom(t, kx, ky, kz);
            KNNVisitor      v;
            v.init(p.x, p.y, p.z, k);
            trees_traverse(t, bid, g, v.q, v);
            const bool mine = (u32)lane < k && v.li != 0xFFFFFFFFu;
            nout            = __popc(__ballot_sync(MMG_FULL, mine));
            if (mine)
            {
                const size_t o = (size_t)q * k + lane;
                oxyz[3 * o] = v.lx, oxyz[3 * o + 1] = v.ly, oxyz[3 * o + 2] = v.lz;
                od2[o] = v.ld, oid[o] = (u64)v.li;
            }
        }
        if (lane == 0) ocount[q] = nout;
    }
}

// ------------------------------------------------------------------- insert
// internal_insertPointCloud3D (SparseTreesPointCloud.cpp:916-959): phase 1 composes every point
// and asks the PRE-BATCH map for its nearest neighbour (k_trees_nn1 on the composed points);
// a point is kept iff it has none or d2 > clearance^2. Phase 2 appends the kept points in input
// order, so the block-local index of a point is bcnt[block] + its rank among the kept points of
// the same block.
static __global__ void k_trees_compose(const float* __restrict__ xs, const float* __restrict__ ys, const float* __restrict__ zs, u32 n, PoseD pose,
                                       float4* __restrict__ gp)
{
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float3 g = compose_point(pose, xs[i], ys[i], zs[i]);
    gp[i]          = make_float4(g.x, g.y, g.z, 0.f);
}
// Clearance test of phase 1 in one pass, FOUR lanes per input point: compose, then ask whether the pre-batch map holds
// ANY point of the same block with d2 <= thr (= the reference's "nearest neighbour has d2 <= clearance^2": the nearest
// one is within thr iff some point is). Only the fine cells whose box can reach thr are opened, and the walk stops at
// the first row holding such a point. The four lanes walk the same cells and rows together and each looks at its own
// eight slots of a 32-slot row (four 256-bit loads), so a row costs ONE load round trip (the
// kernel is bound by the chain of dependent L2 round trips, not by bandwidth or issue). Writes the composed point and
// d2 = some distance <= thr (rejected by k_trees_flags) or -1 ("none": kept). When thr reaches past the 3x3x3 cell region
// (clearance of the order of the 0.31 m cells) the region cannot certify "none": those points go to `todo` for the
// exact search (k_trees_nn1).
constexpr int TRC_LANES = 4;
static __global__ void __launch_bounds__(256)
    k_trees_clear(TreesDev t, const float* __restrict__ xs, const float* __restrict__ ys, const float* __restrict__ zs, u32 n, PoseD pose, float thr,
                  float4* __restrict__ gp, float* __restrict__ d2, u32* __restrict__ todo)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) t.ctr->oob = 0;  // counted by k_trees_flags, read by the host after k_trees_locate
    const u32 q = (blockIdx.x * blockDim.x + threadIdx.x) / TRC_LANES;
    if (q >= n) return;  // whole groups leave together (blockDim is a multiple of TRC_LANES)
    const int      sub   = threadIdx.x & (TRC_LANES - 1);
    const unsigned gmask = ((1u << TRC_LANES) - 1u) << (lane_id() & ~(TRC_LANES - 1));
    const float3   p     = compose_point(pose, xs[q], ys[q], zs[q]);
    if (sub == 0) gp[q] = make_float4(p.x, p.y, p.z, 0.f);
    int   kx, ky, kz;
    u32   bid = 0;
    float hit = -1.0f;
    bool  sure = true;
    if (trees_find_block(t, p.x, p.y, p.z, kx, ky, kz, bid))
    {
        const BlockGeom g     = trees_geom(t, kx, ky, kz);
        const float     qv[3] = {p.x, p.y, p.z};
        int             c0[3];
#pragma unroll
        for (int a = 0; a < 3; a++) c0[a] = trees_cell_axis(g, a, qv[a]);
        for (int it = 0; it < 27 && hit < 0.f; it++)
        {
            const int j = it == 0 ? 13 : (it <= 13 ? it - 1 : it);
            const int x = c0[0] + j / 9 - 1, y = c0[1] + (j / 3) % 3 - 1, z = c0[2] + j % 3 - 1;
            if (x < 0 || x > 31 || y < 0 || y > 31 || z < 0 || z > 31) continue;
            if (it && trees_box_lb(g, qv, x, y, z, 1) > thr) continue;
            const u32 cell = trees_cell_id(x, y, z);
            if (!((__ldg(&t.l0[(size_t)bid * 1024 + (cell >> 5)]) >> (cell & 31)) & 1u)) continue;
            u32 head, cnt;
            if (!probe_ro(t.ctab, t.cmask, trees_cell_key(bid, cell), head, cnt) || cnt == 0) continue;
            u32 fill = ((cnt - 1) & (TR_ROW - 1)) + 1, row = head;
            for (u32 r = 0, nrows = (cnt + TR_ROW - 1) / TR_ROW; r < nrows && hit < 0.f; r++)
            {
                const float4* rp  = t.cpts + (size_t)row * TR_ROW + sub * 8;
                const u32     nxt = (r + 1 < nrows) ? __ldg(&t.cnext[row]) : 0u;  // in flight together with the row
                float4        c[8];  // this lane's eight slots (slots past `fill` are loaded but never looked at)
#pragma unroll
                for (int k = 0; k < 4; k++) ld_row_pair(rp + 2 * k, c[2 * k], c[2 * k + 1]);
                float mine = -1.0f;
#pragma unroll
                for (int k = 0; k < 8; k++)
                {
                    const float d = sqr_dist(qv[0], qv[1], qv[2], c[k].x, c[k].y, c[k].z);
                    if ((u32)(sub * 8 + k) < fill && d <= thr) mine = d;
                }
                // any lane's hit ends the walk for the group; every lane takes the hit of the lowest lane that has one
                const unsigned any = __ballot_sync(gmask, mine >= 0.f) & gmask;
                if (any) hit = __shfl_sync(gmask, mine, __ffs(any) - 1);
                row  = nxt;
                fill = TR_ROW;
            }
        }
        if (hit < 0.f)
        {
            float out = __int_as_float(0x7f800000);
#pragma unroll
            for (int a = 0; a < 3; a++)
            {
                const int   lo_c = max(c0[a] - 1, 0), hi_c = min(c0[a] + 1, 31);
                const float mg   = 8.f * g.margin[a];
                if (lo_c > 0) out = fminf(out, fmaxf(0.f, qv[a] - (g.lo[a] + (float)lo_c * g.cs[a]) - mg));
                if (hi_c < 31) out = fminf(out, fmaxf(0.f, (g.lo[a] + (float)(hi_c + 1) * g.cs[a]) - qv[a] - mg));
            }
            sure = thr < out * out * 0.9999f;
        }
    }
    if (sub != 0) return;
    if (sure) d2[q] = hit;
    else todo[atomicAdd(&t.ctr->ntodo, 1u)] = q;
}
static __global__ void k_trees_flags(TreesDev t, const float4* __restrict__ gp, const float* __restrict__ d2, u32 n, float thr, u32* __restrict__ flag)
{
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) t.ctr->ntodo = 0;  // the exact search before this kernel has consumed the list: ready for the next batch
    if (i >= n) return;
    const float4 p = gp[i];
    int          kx, ky, kz;
    const bool   ok = trees_index(t, p.x, kx) & trees_index(t, p.y, ky) & trees_index(t, p.z, kz);
    if (!ok) atomicAdd(&t.ctr->oob, 1u);
    const float d = d2[i];
    flag[i]       = (ok && (d < 0.f || d > thr)) ? 1u : 0u;
}
// Extends the bounding box (ordered-uint min xyz / max xyz) by the warp's valid points. The current box is read first
// (a stale copy is only ever smaller): only warps holding a point outside it reduce and issue atomics - none in steady state.
__device__ __forceinline__ void warp_bbox6(u32* bbox, bool ok, float x, float y, float z)
{
    const float v[3] = {x, y, z};
    u32         o[3];
    bool        ext = false;
#pragma unroll
    for (int a = 0; a < 3; a++)
    {
        o[a] = float_to_ordered(v[a]);
        ext |= ok && (o[a] < __ldcg(&bbox[a]) || o[a] > __ldcg(&bbox[3 + a]));
    }
    if (!__any_sync(MMG_FULL, ext)) return;
#pragma unroll
    for (int a = 0; a < 3; a++)
    {
        const u32 mn = __reduce_min_sync(MMG_FULL, ok ? o[a] : 0xFFFFFFFFu);
        const u32 mx = __reduce_max_sync(MMG_FULL, ok ? o[a] : 0u);
        if (lane_id() == 0)
        {
            if (mn != 0xFFFFFFFFu) atomicMin(&bbox[a], mn);
            if (mx != 0u) atomicMax(&bbox[3 + a], mx);
        }
    }
}
__device__ __forceinline__ u32 trees_alloc(u32* free_n, u32* hi, const u32* free_list)
{
    const int f = (int)atomicSub(free_n, 1u);
    if (f > 0) return free_list[f - 1];
    atomicAdd(free_n, 1u);
    return atomicAdd(hi, 1u);
}
// find-or-insert on a {key, vid, cnt} table; the winner of the CAS allocates the row id
template <class Alloc>
__device__ __forceinline__ u32 trees_find_or_insert(Entry* tab, u64 mask, u64 key, Alloc&& alloc, u64* slot_out)
{
    u64 h = hash_key(key) & mask;
    for (;;)
    {
        Entry* e = tab + h;
        u64    k = *reinterpret_cast<volatile u64*>(&e->key);
        if (k == KEY_EMPTY)
        {
            k = atomicCAS(reinterpret_cast<u64*>(&e->key), KEY_EMPTY, key);
            if (k == KEY_EMPTY)
            {
                const u32 vid = alloc();
                e->cnt        = 0;
                __threadfence();
                atomicExch(&e->vid, vid);
                if (slot_out) *slot_out = h;
                return vid;
            }
        }
        if (k == key)
        {
            u32 v;
            while ((v = *reinterpret_cast<volatile u32*>(&e->vid)) == VID_INVALID) {}
            if (slot_out) *slot_out = h;
            return v;
        }
        h = (h + 1) & mask;
    }
}
// compaction of the kept points (input order) + block find-or-insert
static __global__ void __launch_bounds__(256)
    k_trees_locate(TreesDev t, const float4* __restrict__ gp, const u32* __restrict__ flag, const u32* __restrict__ pos, u32 n, float4* __restrict__ acc,
                   u32* __restrict__ bid_out)
{
    const u32  i    = blockIdx.x * blockDim.x + threadIdx.x;
    const bool keep = i < n && flag[i];
    float4     p    = make_float4(0.f, 0.f, 0.f, 0.f);
    u64        key  = KEY_EMPTY;
    if (keep)
    {
        p = gp[i];
        int kx, ky, kz;
        trees_index(t, p.x, kx), trees_index(t, p.y, ky), trees_index(t, p.z, kz);
        key = pack_key(kx, ky, kz);
    }
    const u32 peers  = __match_any_sync(MMG_FULL, key);
    const int leader = __ffs(peers) - 1;
    u32       bid    = VID_INVALID;
    if (keep && lane_id() == leader)
        bid = trees_find_or_insert(t.btab, t.bmask, key, [&]() { return trees_alloc(&t.ctr->bfree_n, &t.ctr->brow_hi, t.bfree); }, nullptr);
    bid = __shfl_sync(MMG_FULL, bid, leader);
    if (keep)
    {
        acc[pos[i]]     = p;
        bid_out[pos[i]] = bid;
    }
    warp_bbox6(t.bbox, keep, p.x, p.y, p.z);
}
// first position of value b in the ascending list sbid[0, n)
__device__ __forceinline__ u32 trees_lower_bound(const u32* __restrict__ sbid, u32 n, u32 b)
{
    u32 lo = 0, hi = n;
    while (lo < hi)
    {
        const u32 mid = (lo + hi) >> 1;
        if (__ldg(sbid + mid) < b) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}
// Phase 2 of the insert, three kernels over the kept points in block order (e = position in the list sorted by block;
// the sort is stable, so e - (first position of the block) is the point's rank among the block's kept points):
//   k_trees_assign  block-local index, fine cell, cell find-or-insert, and the point's ordinal among the batch points
//                   of its cell (a counter per cell slot in `cwork[slot].w`, zero between batches);
//   k_trees_rows    ordinal 0 = the cell's owner: links the chain rows the cell's batch points need in front of the head
//                   (several rows come as one contiguous run), publishes the new head and count, and leaves
//                   {old count, old head, first new row} in cwork[slot]; the first point of a block publishes the block;
//   k_trees_write   every point stores itself at slot (old count + ordinal) of its cell.
// The order of a cell's points inside its rows is the order of the ordinals (arbitrary): queries order by
// (d2, block-local index) and the export by block-local index, never by the position in a row.
static __global__ void __launch_bounds__(256)
    k_trees_assign(TreesDev t, const u32* __restrict__ sbid, const u32* __restrict__ sidx, u32 n, float4* __restrict__ acc, uint4* __restrict__ cwork,
                   u32* __restrict__ cslot, u32* __restrict__ ord)
{
    pdl_trigger();
    pdl_wait();
    const u32  e     = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = e < n;
    u64        key   = KEY_EMPTY;
    if (valid)
    {
        const u32 bid = sbid[e];
        const u32 j   = sidx[e];
        float4    p   = acc[j];
        p.w           = __uint_as_float(t.bcnt[bid] + (e - trees_lower_bound(sbid, n, bid)));
        acc[j]        = p;
        int kx, ky, kz;
        trees_index(t, p.x, kx), trees_index(t, p.y, ky), trees_index(t, p.z, kz);
        const BlockGeom g = trees_geom(t, kx, ky, kz);
        key = trees_cell_key(bid, trees_cell_id(trees_cell_axis(g, 0, p.x), trees_cell_axis(g, 1, p.y), trees_cell_axis(g, 2, p.z)));
    }
    const u32 peers  = __match_any_sync(MMG_FULL, key);
    const int leader = __ffs(peers) - 1;
    u64       slot   = 0;
    u32       base   = 0;
    if (valid && lane_id() == leader)
    {
        trees_find_or_insert(t.ctab, t.cmask, key, [&]() { atomicAdd(&t.ctr->ncells, 1u); return 0xFFFFFFFEu; /* "no chain yet" */ }, &slot);
        base = atomicAdd(&cwork[slot].w, (u32)__popc(peers));
    }
    slot = __shfl_sync(MMG_FULL, slot, leader);
    base = __shfl_sync(MMG_FULL, base, leader);
    if (valid)
    {
        cslot[e] = (u32)slot;
        ord[e]   = base + __popc(peers & ((1u << lane_id()) - 1));
    }
}
static __global__ void __launch_bounds__(256)
    k_trees_rows(TreesDev t, const u32* __restrict__ sbid, const u32* __restrict__ sidx, const u32* __restrict__ cslot, const u32* __restrict__ ord, u32 n,
                 const float4* __restrict__ acc, uint4* __restrict__ cwork)
{
    pdl_trigger();
    pdl_wait();
    const u32 e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    if (ord[e] == 0)
    {
        const u32 slot = cslot[e];
        Entry*    en   = t.ctab + slot;
        const u32 cnt = en->cnt, len = cwork[slot].w;
        u32       head = en->vid;
        if (cnt == 0)
        {  // first points of this cell: publish it in the occupancy masks
            head           = VID_INVALID;
            const u32 bid  = (u32)(en->key >> 15), cell = (u32)(en->key & (TR_CELLS - 1));
            atomicOr(&t.l0[(size_t)bid * 1024 + (cell >> 5)], 1u << (cell & 31));
            atomicOr(&t.l1[(size_t)bid * 16 + (cell >> 11)], 1u << ((cell >> 6) & 31));
        }
        const u32 have = (cnt + TR_ROW - 1) / TR_ROW, need = (cnt + len + TR_ROW - 1) / TR_ROW - have;
        u32       first = 0;
        if (need)
        {
            first = need == 1 ? trees_alloc(&t.ctr->cfree_n, &t.ctr->crow_hi, t.cfree) : atomicAdd(&t.ctr->crow_hi, need);
            t.cnext[first] = head;
            for (u32 r = 1; r < need; r++) t.cnext[first + r] = first + r - 1;
            en->vid = first + need - 1;  // the newest row is the head and the only partially filled one
        }
        en->cnt     = cnt + len;
        cwork[slot] = make_uint4(cnt, head, first, 0u);
    }
    const u32 b = sbid[e];
    if (e != 0 && sbid[e - 1] == b) return;
    u32 lo = e + 1, hi = n;  // one past the block's last point
    while (lo < hi)
    {
        const u32 mid = (lo + hi) >> 1;
        if (__ldg(sbid + mid) <= b) lo = mid + 1;
        else hi = mid;
    }
    const float4 p = acc[sidx[e]];
    int          kx, ky, kz;
    trees_index(t, p.x, kx), trees_index(t, p.y, ky), trees_index(t, p.z, kz);
    t.bkey[b] = pack_key(kx, ky, kz);
    t.bcnt[b] += lo - e;
    atomicAdd(&t.ctr->npoints, (unsigned long long)(lo - e));
}
static __global__ void __launch_bounds__(256)
    k_trees_write(TreesDev t, const u32* __restrict__ sidx, const u32* __restrict__ cslot, const u32* __restrict__ ord, u32 n,
                  const float4* __restrict__ acc, const uint4* __restrict__ cwork)
{
    pdl_trigger();
    pdl_wait();
    const u32 e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    const uint4 w = cwork[cslot[e]];  // {old count, old head, first new row, 0}
    const u32   g = w.x + ord[e], row_ord = g / TR_ROW, have = (w.x + TR_ROW - 1) / TR_ROW;
    const u32   row = row_ord < have ? w.y : w.z + (row_ord - have);
    t.cpts[(size_t)row * TR_ROW + (g & (TR_ROW - 1))] = acc[sidx[e]];
}

// ---------------------------------------------------------------- upkeep
// eraseGridsFartherThan (SparseTreesPointCloud.cpp:961-1000): blocks outside the index box
// [c - d, c + d]^3 are dropped. One warp per block row (it also clears the row's masks).
static __global__ void __launch_bounds__(256) k_trees_evict(TreesDev t, u32 nrows, int cx, int cy, int cz, int d)
{
    const int lane = lane_id();
    const u32 b    = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (b >= nrows) return;
    const u64 key = t.bkey[b];
    if (key == KEY_EMPTY) return;
    int kx, ky, kz;
    unpack_key(key, kx, ky, kz);
    const bool inside = kx >= cx - d && ky >= cy - d && kz >= cz - d && kx <= cx + d && ky <= cy + d && kz <= cz + d;
    if (inside) return;
    for (int w = lane; w < 1024; w += 32) t.l0[(size_t)b * 1024 + w] = 0;
    if (lane < 16) t.l1[(size_t)b * 16 + lane] = 0;
    __syncwarp();
    if (lane == 0)
    {
        atomicAdd(&t.ctr->npoints, (unsigned long long)(-(long long)t.bcnt[b]));
        t.bkey[b] = KEY_EMPTY;
        t.bcnt[b] = 0;
        t.bfree[atomicAdd(&t.ctr->bfree_n, 1u)] = b;
    }
}
static __global__ void k_trees_rebuild_btab(TreesDev t, u32 nrows)
{
    const u32 b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nrows) return;
    const u64 key = t.bkey[b];
    if (key == KEY_EMPTY) return;
    u64 h = hash_key(key) & t.bmask;
    for (;;)
    {
        Entry* e = t.btab + h;
        if (atomicCAS(reinterpret_cast<u64*>(&e->key), KEY_EMPTY, key) == KEY_EMPTY)
        {
            e->vid = b;
            e->cnt = 0;
            return;
        }
        h = (h + 1) & t.bmask;
    }
}
// Moves the live cell entries of an old table into t.ctab (growth, or after an eviction: entries
// of dropped blocks are not carried over and their chain rows go back to the free stack).
static __global__ void k_trees_rehash_ctab(TreesDev t, const Entry* __restrict__ old, u64 oldcap)
{
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= oldcap) return;
    const Entry o = old[i];
    if (o.key == KEY_EMPTY) return;
    const u32 bid = (u32)(o.key >> 15);
    if (t.bkey[bid] == KEY_EMPTY)
    {
        u32 row = o.vid;
        for (u32 r = 0, nr = (o.cnt + TR_ROW - 1) / TR_ROW; r < nr; r++)
        {
            const u32 nx = t.cnext[row];
            t.cfree[atomicAdd(&t.ctr->cfree_n, 1u)] = row;
            row = nx;
        }
        atomicSub(&t.ctr->ncells, 1u);
        return;
    }
    u64 h = hash_key(o.key) & t.cmask;
    for (;;)
    {
        Entry* e = t.ctab + h;
        if (atomicCAS(reinterpret_cast<u64*>(&e->key), KEY_EMPTY, o.key) == KEY_EMPTY)
        {
            e->vid = o.vid;
            e->cnt = o.cnt;
            return;
        }
        h = (h + 1) & t.cmask;
    }
}

// ---------------------------------------------------------------- export
static __global__ void k_trees_live(TreesDev t, u32 nrows, u32* __restrict__ live)
{
    const u32 b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nrows) return;
    const bool l = t.bkey[b] != KEY_EMPTY;
    live[b]      = l ? 1u : 0u;
    if (l) atomicMax(&t.ctr->maxcnt, t.bcnt[b]);
}
static __global__ void k_trees_export_blocks(TreesDev t, u32 nrows, const u32* __restrict__ rank, int32_t* __restrict__ keys, u32* __restrict__ counts)
{
    const u32 b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nrows) return;
    const u64 key = t.bkey[b];
    if (key == KEY_EMPTY) return;
    const size_t o = rank[b];
    int          kx, ky, kz;
    unpack_key(key, kx, ky, kz);
    if (keys) keys[3 * o] = kx, keys[3 * o + 1] = ky, keys[3 * o + 2] = kz;
    if (counts) counts[o] = t.bcnt[b];
}
// one warp per cell-table slot; points land at (block rank, block-local index); optional bbox
static __global__ void __launch_bounds__(256)
    k_trees_export_points(TreesDev t, u64 ccap, const u32* __restrict__ rank, u32 cap, float* __restrict__ points, u32* bbox)
{
    const int lane = lane_id();
    const u64 i    = ((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (i >= ccap) return;
    const Entry e = t.ctab[i];
    if (e.key == KEY_EMPTY || e.cnt == 0) return;
    const u32 bid  = (u32)(e.key >> 15);
    u32       fill = ((e.cnt - 1) & (TR_ROW - 1)) + 1, row = e.vid;
    for (u32 r = 0, nr = (e.cnt + TR_ROW - 1) / TR_ROW; r < nr; r++)
    {
        const bool valid = (u32)lane < fill;
        float4     p     = make_float4(0.f, 0.f, 0.f, 0.f);
        if (valid) p = t.cpts[(size_t)row * TR_ROW + lane];
        if (valid && points)
        {
            float* o = points + ((size_t)rank[bid] * cap + __float_as_uint(p.w)) * 3;
            o[0] = p.x, o[1] = p.y, o[2] = p.z;
        }
        if (bbox) warp_bbox6(bbox, valid, p.x, p.y, p.z);
        row  = t.cnext[row];
        fill = TR_ROW;
    }
}
}  // namespace mmg
