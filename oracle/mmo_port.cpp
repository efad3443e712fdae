// mmo_port.cpp — CPU restatement ("port") of the hot path of mola_metric_maps.
//
// TEST INFRASTRUCTURE ONLY (see mmo.h). This file is the oracle the CUDA path
// is compared against, and the timed "port" CPU baseline when the reference
// itself (oracle/_ref/libmmo_ref.so) is not available. It is never linked into
// or called from the product library.
//
// Parity pinning: the restatement is checked against the reference's own code
// (oracle/_ref, the reference .cpp files compiled unmodified against
// shim/mrpt) by tests/test_oracle_vs_reference.py on seeded inputs, and against
// the committed golden vectors under tests/golden/. The MRPT arithmetic that is
// not in the reference tree (CPose3D::composePoint, mrpt::round,
// TPoint3Df::sqrNorm, CSimplePointsMap/nanoflann) follows the definitions of
// SURVEY.md note N3 — against upstream MRPT those sub-steps are "parity unpinned".
//
// Every function cites the reference lines (relative to
// /root/reference/mola_metric_maps/) whose behaviour it restates. The code is
// written from the behaviour, with our own containers; compile with
// -O2 -ffp-contract=off so no FMA contraction changes an fp32 result.
#include "mmo.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <map>
#include <set>
#include <thread>
#include <unordered_map>
#include <vector>

namespace
{
struct P3
{
    float x, y, z;
};
struct Key
{
    int32_t cx, cy, cz;
    bool    operator==(const Key& o) const { return cx == o.cx && cy == o.cy && cz == o.cz; }
    bool    operator<(const Key& o) const
    {  // lexicographic, include/mola_metric_maps/index3d_t.h:104-111
        if (cx != o.cx) return cx < o.cx;
        if (cy != o.cy) return cy < o.cy;
        return cz < o.cz;
    }
};
struct KeyHash
{
    size_t operator()(const Key& k) const
    {  // any good mix: the container's hash is not observable in results
        uint64_t h = (uint64_t)(uint32_t)k.cx * 0x9E3779B97F4A7C15ull;
        h ^= (uint64_t)(uint32_t)k.cy * 0xC2B2AE3D27D4EB4Full + (h >> 29);
        h ^= (uint64_t)(uint32_t)k.cz * 0x165667B19E3779F9ull + (h << 7);
        return (size_t)(h ^ (h >> 32));
    }
};

// MRPT CPose3D::composePoint: rotation in double, left-to-right sums, then
// rounded to float. Called from HashedVoxelPointCloud.cpp:369,404,421,529,1077.
inline P3 compose(const double* T, float lx, float ly, float lz)
{
    static const double I[12] = {1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0};
    if (!T) T = I;
    const double x = lx, y = ly, z = lz;
    const double gx = ((T[0] * x + T[1] * y) + T[2] * z) + T[9];
    const double gy = ((T[3] * x + T[4] * y) + T[5] * z) + T[10];
    const double gz = ((T[6] * x + T[7] * y) + T[8] * z) + T[11];
    return {(float)gx, (float)gy, (float)gz};
}
// TPoint3Df::sqrNorm of (a-b): (dx*dx + dy*dy) + dz*dz in fp32.
inline float sqr_dist(const P3& a, const P3& b)
{
    const float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z;
    return (dx * dx + dy * dy) + dz * dz;
}
struct BBox
{
    bool valid = false;
    P3   mn{0, 0, 0}, mx{0, 0, 0};
    void add(const P3& p)
    {  // cached_.boundingBox_ emplace(pt,pt) / updateWithPoint, HashedVoxelPointCloud.cpp:621-625
        if (!valid)
        {
            valid = true;
            mn = mx = p;
            return;
        }
        if (p.x < mn.x) mn.x = p.x;
        if (p.y < mn.y) mn.y = p.y;
        if (p.z < mn.z) mn.z = p.z;
        if (p.x > mx.x) mx.x = p.x;
        if (p.y > mx.y) mx.y = p.y;
        if (p.z > mx.z) mx.z = p.z;
    }
};

struct Hit
{
    P3       p;
    float    d2;
    uint64_t id;
};

// The 3-slot "sorted insert" of HashedVoxelPointCloud.cpp:720-741 (and :802-823),
// including its forward-copy shift (SURVEY note N1): inserting at position i
// overwrites every later live slot with the old value of slot i.
template <size_t K>
struct SortedSlots
{
    Hit    m[K];
    size_t found = 0;
    void   offer(float d2, const P3& p, uint64_t id)
    {
        size_t i = 0;
        while (i < found && !(d2 < m[i].d2)) i++;
        if (i >= K) return;
        const size_t last = std::min(found + 1, K);
        for (size_t j = i + 1; j < last; j++) m[j] = m[j - 1];
        m[i] = {p, d2, id};
        if (found < K) found++;
    }
};

struct MapBase
{
    mmo_params prm{};
    BBox       bbox;
    // radius results of the last mmo_nn_radius call
    std::vector<Hit> rad_hits;
    virtual ~MapBase() = default;
    virtual void     clear()                                                            = 0;
    virtual bool     empty() const                                                      = 0;
    virtual void     set_params(const mmo_params& p)                                    = 0;
    virtual void     insert(const float*, const float*, const float*, size_t, const double*, int) = 0;
    virtual int      nn_multi_one(const P3& q, size_t k, Hit* out, size_t& nout) const   = 0;
    virtual int      nn_single_one(const P3& q, Hit& out, bool& found) const             = 0;
    virtual int      radius_one(const P3& q, float r2, size_t maxp, std::vector<Hit>& out) const = 0;
    virtual void     compute_bbox_uncached(BBox& b) const                                = 0;
    virtual uint64_t num_voxels() const                                                  = 0;
    virtual uint64_t num_points() const                                                  = 0;
    virtual uint32_t row_cap() const                                                     = 0;
    virtual void     export_all(int32_t*, uint32_t*, float*) const                       = 0;
    virtual void     voxel_stats(float*, float*) const {}
    virtual double   stencil_points(const P3&) const { return 0; }
};

// ------------------------------------------------------------------ Hashed
struct Hashed : MapBase
{
    static constexpr uint32_t CAP = 32;  // SSO_MAX_POINTS_PER_VOXEL, HashedVoxelPointCloud.h:61
    struct Voxel
    {
        P3       pts[CAP];
        uint32_t n = 0;
    };
    std::unordered_map<Key, Voxel, KeyHash> vox;
    float                                   inv = 1.f, vsize = 1.f;

    void set_params(const mmo_params& p) override
    {  // setVoxelProperties clears the map, HashedVoxelPointCloud.cpp:189-200
        const bool resize = (p.voxel_size != prm.voxel_size);
        prm               = p;
        if (resize)
        {
            vsize = p.voxel_size;
            inv   = 1.0f / vsize;
            vox.clear();
        }
    }
    void clear() override { vox.clear(); }  // internal_clear :287 (bbox cache is NOT reset)
    bool empty() const override { return vox.empty(); }

    Key to_idx(const P3& p) const
    {  // coordToGlobalIdx: truncation toward zero, HashedVoxelPointCloud.h:76-83
        return {(int32_t)(p.x * inv), (int32_t)(p.y * inv), (int32_t)(p.z * inv)};
    }
    static uint64_t plain_id(const Key& g, int slot)
    {  // g2plain, HashedVoxelPointCloud.h:95-108: 5 slot bits + 3 x 20 index bits
        const uint64_t F = (1u << 20) - 1;
        return ((uint64_t)slot & 31u) | (((uint64_t)(int64_t)g.cx & F) << 5) |
               (((uint64_t)(int64_t)g.cy & F) << 25) | (((uint64_t)(int64_t)g.cz & F) << 45);
    }
    void insert_point(const P3& p)
    {  // insertPoint, HashedVoxelPointCloud.cpp:588-626
        Voxel&         v  = vox[to_idx(p)];  // created even if the point is then rejected
        const uint32_t n0 = v.n;
        if (prm.max_points_per_voxel != 0 && n0 >= prm.max_points_per_voxel) return;
        if (prm.min_distance_between_points > 0 && n0 != 0)
        {
            float best = sqr_dist(v.pts[0], p);
            for (uint32_t i = 1; i < n0; i++)
            {
                const float d = sqr_dist(v.pts[i], p);
                if (d < best) best = d;
            }
            const float thr = prm.min_distance_between_points * prm.min_distance_between_points;
            if (best < thr) return;
        }
        if (v.n < CAP) v.pts[v.n++] = p;  // VoxelData::insertPoint, .h:171-174 (silent drop at 32)
        bbox.add(p);                      // even when the 32-slot array dropped it
    }
    void insert(const float* xs, const float* ys, const float* zs, size_t n, const double* T,
                int has_pose) override
    {  // internal_insertObservation :289-331 then the per-point loop :419-422 / :1073-1079
        if (has_pose && T && prm.remove_voxels_farther_than > 0)
        {
            const int dist_in_grid = (int)std::ceil(prm.remove_voxels_farther_than * inv);
            const Key c            = to_idx({(float)T[9], (float)T[10], (float)T[11]});
            for (auto it = vox.begin(); it != vox.end();)
            {
                const int d = std::max({std::abs(it->first.cx - c.cx), std::abs(it->first.cy - c.cy),
                                        std::abs(it->first.cz - c.cz)});
                if (d > dist_in_grid)
                    it = vox.erase(it);
                else
                    ++it;
            }
        }
        for (size_t i = 0; i < n; i++) insert_point(compose(T, xs[i], ys[i], zs[i]));
    }
    template <size_t K>
    void knn(const P3& q, SortedSlots<K>& s) const
    {  // nn_multiple_search_impl<K> :691-771: 27-cell stencil, cx outermost, slot order
        const Key c = to_idx(q);
        for (int32_t cx = c.cx - 1; cx <= c.cx + 1; cx++)
            for (int32_t cy = c.cy - 1; cy <= c.cy + 1; cy++)
                for (int32_t cz = c.cz - 1; cz <= c.cz + 1; cz++)
                {
                    auto it = vox.find(Key{cx, cy, cz});
                    if (it == vox.end()) continue;
                    const Voxel& v = it->second;
                    for (uint32_t i = 0; i < v.n; i++)
                        s.offer(sqr_dist(v.pts[i], q), v.pts[i], plain_id(it->first, (int)i));
                }
    }
    int nn_single_one(const P3& q, Hit& out, bool& found) const override
    {  // nn_single_search :666-679 (keeps its three per-call vectors for the timed baseline)
        std::vector<P3>       r;
        std::vector<float>    d;
        std::vector<uint64_t> ids;
        SortedSlots<1>        s;
        knn<1>(q, s);
        for (size_t i = 0; i < std::min<size_t>(1, s.found); i++)
        {
            d.push_back(s.m[i].d2);
            r.push_back(s.m[i].p);
            ids.push_back(s.m[i].id);
        }
        found = !r.empty();
        if (found) out = {r[0], d[0], ids[0]};
        return 0;
    }
    int nn_multi_one(const P3& q, size_t k, Hit* out, size_t& nout) const override
    {  // nn_multiple_search :681-689 with HARD_MAX_MATCHES = 3 (:44)
        SortedSlots<3> s;
        knn<3>(q, s);
        nout = std::min<size_t>(k, s.found);
        for (size_t i = 0; i < nout; i++) out[i] = s.m[i];
        return 0;
    }
    int radius_one(const P3& q, float r2, size_t maxp, std::vector<Hit>& out) const override
    {  // nn_radius_search :773-870
        if (r2 <= 0) return 0;
        const float    r = std::sqrt(r2);
        const Key      a = to_idx({q.x - r, q.y - r, q.z - r});
        const Key      b = to_idx({q.x + r, q.y + r, q.z + r});
        SortedSlots<3> s;
        for (int32_t cx = a.cx; cx <= b.cx; cx++)
            for (int32_t cy = a.cy; cy <= b.cy; cy++)
                for (int32_t cz = a.cz; cz <= b.cz; cz++)
                {
                    auto it = vox.find(Key{cx, cy, cz});
                    if (it == vox.end()) continue;
                    const Voxel& v = it->second;
                    for (uint32_t i = 0; i < v.n; i++)
                    {
                        const float d = sqr_dist(v.pts[i], q);
                        if (d > r2) continue;
                        const uint64_t id = plain_id(it->first, (int)i);
                        if (maxp != 0)
                            s.offer(d, v.pts[i], id);
                        else
                            out.push_back({v.pts[i], d, id});
                    }
                }
        if (maxp != 0)
            for (size_t i = 0; i < std::min<size_t>(maxp, s.found); i++) out.push_back(s.m[i]);
        return 0;
    }
    void compute_bbox_uncached(BBox& b) const override
    {  // boundingBox() :872-902: voxel corners, not points
        for (const auto& kv : vox)
        {
            const P3 c{kv.first.cx * vsize, kv.first.cy * vsize, kv.first.cz * vsize};
            b.add(c);
            b.add({c.x + 1.0f * vsize, c.y + 1.0f * vsize, c.z + 1.0f * vsize});
        }
    }
    uint64_t num_voxels() const override { return vox.size(); }
    uint64_t num_points() const override
    {
        uint64_t n = 0;
        for (const auto& kv : vox) n += kv.second.n;
        return n;
    }
    uint32_t row_cap() const override { return CAP; }
    void     export_all(int32_t* keys, uint32_t* counts, float* pts) const override
    {
        std::vector<const std::pair<const Key, Voxel>*> v;
        v.reserve(vox.size());
        for (const auto& kv : vox) v.push_back(&kv);
        std::sort(v.begin(), v.end(), [](auto* a, auto* b) { return a->first < b->first; });
        for (size_t i = 0; i < v.size(); i++)
        {
            keys[3 * i] = v[i]->first.cx, keys[3 * i + 1] = v[i]->first.cy, keys[3 * i + 2] = v[i]->first.cz;
            counts[i] = v[i]->second.n;
            float* row = pts + i * CAP * 3;
            std::memset(row, 0, sizeof(float) * CAP * 3);
            for (uint32_t j = 0; j < v[i]->second.n; j++)
                row[3 * j] = v[i]->second.pts[j].x, row[3 * j + 1] = v[i]->second.pts[j].y,
                        row[3 * j + 2] = v[i]->second.pts[j].z;
        }
    }
    double stencil_points(const P3& q) const override
    {
        const Key c = to_idx(q);
        double    n = 0;
        for (int dx = -1; dx <= 1; dx++)
            for (int dy = -1; dy <= 1; dy++)
                for (int dz = -1; dz <= 1; dz++)
                {
                    auto it = vox.find(Key{c.cx + dx, c.cy + dy, c.cz + dz});
                    if (it != vox.end()) n += it->second.n;
                }
        return n;
    }
};

// ------------------------------------------------------------- SparseVoxel
struct SparseVoxel : MapBase
{
    static constexpr uint32_t CAP = 16;  // HARDLIMIT_MAX_POINTS_PER_VOXEL, SparseVoxelPointCloud.h:61
    struct Voxel
    {
        P3      pts[CAP];
        uint8_t n = 0;
        P3      mean{0, 0, 0};
    };
    std::unordered_map<Key, Voxel, KeyHash> vox;     // cells that have been touched by an insert
    std::set<Key>                           blocks;  // outer 32^3 blocks that exist (std::map keys)
    float                                   inv = 5.f, vsize = 0.2f;

    void set_params(const mmo_params& p) override
    {  // setVoxelProperties, SparseVoxelPointCloud.cpp:214-231
        const bool resize = (p.voxel_size != prm.voxel_size);
        prm               = p;
        if (resize)
        {
            vsize = p.voxel_size;
            inv   = 1.0f / vsize;
            clear();
        }
    }
    void clear() override
    {
        vox.clear();
        blocks.clear();
    }
    bool empty() const override { return blocks.empty(); }
    Key  to_idx(const P3& p) const
    {  // coordToGlobalIdx with mrpt::round (= lrint, ties to even; N3), SparseVoxelPointCloud.h:108-115
        return {(int32_t)std::lrint(p.x * inv), (int32_t)std::lrint(p.y * inv),
                (int32_t)std::lrint(p.z * inv)};
    }
    static Key outer(const Key& g) { return {g.cx & ~31, g.cy & ~31, g.cz & ~31}; }  // g2o .h:84-90
    static uint64_t plain_id(const Key& g, int slot)
    {  // g2plain, SparseVoxelPointCloud.h:127-140: 4 slot bits + 3 x 20 index bits
        const uint64_t F = (1u << 20) - 1;
        return ((uint64_t)slot & 15u) | (((uint64_t)(int64_t)g.cx & F) << 4) |
               (((uint64_t)(int64_t)g.cy & F) << 24) | (((uint64_t)(int64_t)g.cz & F) << 44);
    }
    void insert_point(const P3& p)
    {  // insertPoint, SparseVoxelPointCloud.cpp:638-655 + VoxelData::insertPoint :193-204
        const Key g = to_idx(p);
        blocks.insert(outer(g));  // the block is created by the lookup, accepted or not
        Voxel&         v  = vox[g];
        const uint32_t n0 = v.n;
        if (prm.max_points_per_voxel == 0 || n0 < prm.max_points_per_voxel)
        {
            if (v.n < CAP)
            {
                const float nf = (float)v.n;
                v.mean         = {nf * v.mean.x + p.x, nf * v.mean.y + p.y, nf * v.mean.z + p.z};
                v.pts[v.n++]   = p;
                const float k  = 1.0f / v.n;
                v.mean.x *= k, v.mean.y *= k, v.mean.z *= k;
            }
            bbox.add(p);
        }
    }
    void insert(const float* xs, const float* ys, const float* zs, size_t n, const double* T,
                int has_pose) override
    {  // internal_insertObservation: no eviction in this map type, SparseVoxelPointCloud.cpp:362-380
        for (size_t i = 0; i < n; i++) insert_point(compose(T, xs[i], ys[i], zs[i]));
    }
    // The reference throws here (:695-719). With knn_ext set (test/bench infrastructure for the extension of
    // SURVEY N4) the port answers instead: the k <= 8 lexicographically smallest (d2, scan order) over the
    // 27-cell stencil of round-based indices, cells visited cx, cy, cz ascending, slots ascending.
    bool knn_ext = false;
    int  nn_single_one(const P3& q, Hit& out, bool& found) const override
    {
        if (!knn_ext) return -5;
        size_t n = 0;
        nn_multi_one(q, 1, &out, n);
        found = n > 0;
        return 0;
    }
    int nn_multi_one(const P3& q, size_t k, Hit* out, size_t& nout) const override
    {
        if (!knn_ext) return -5;
        if (k > 8) return -5;
        const Key c = to_idx(q);
        nout        = 0;
        for (int x = c.cx - 1; x <= c.cx + 1; x++)
            for (int y = c.cy - 1; y <= c.cy + 1; y++)
                for (int z = c.cz - 1; z <= c.cz + 1; z++)
                {
                    auto it = vox.find(Key{x, y, z});
                    if (it == vox.end()) continue;
                    for (int s = 0; s < it->second.n; s++)
                    {
                        const P3&   p  = it->second.pts[s];
                        const float dx = p.x - q.x, dy = p.y - q.y, dz = p.z - q.z;
                        const float d2 = (dx * dx + dy * dy) + dz * dz;
                        size_t      i  = 0;
                        while (i < nout && !(d2 < out[i].d2)) i++;  // strict '<': the first seen wins ties
                        if (i >= k) continue;
                        const size_t last = std::min(nout + 1, k);
                        for (size_t j = last - 1; j > i; j--) out[j] = out[j - 1];
                        out[i] = {p, d2, plain_id(Key{x, y, z}, s)};
                        if (nout < k) nout++;
                    }
                }
        return 0;
    }
    int radius_one(const P3& q, float r2, size_t maxp, std::vector<Hit>& out) const override
    {  // nn_radius_search, SparseVoxelPointCloud.cpp:721-798
        if (r2 <= 0) return 0;
        const float r = std::sqrt(r2);
        const Key   a = to_idx({q.x - r, q.y - r, q.z - r});
        const Key   b = to_idx({q.x + r, q.y + r, q.z + r});
        // maxPoints != 0: candidates keyed by d2 in a std::map<float,...>; equal d2 overwrite (N2)
        std::map<float, std::pair<P3, uint64_t>> cand;
        auto add = [&](float d, const P3& p, uint64_t id) {
            if (maxp != 0)
                cand[d] = {p, id};
            else
                out.push_back({p, d, id});
        };
        for (int32_t cx = a.cx; cx <= b.cx; cx++)
            for (int32_t cy = a.cy; cy <= b.cy; cy++)
                for (int32_t cz = a.cz; cz <= b.cz; cz++)
                {
                    const Key k{cx, cy, cz};
                    auto      it = vox.find(k);
                    if (it == vox.end() || it->second.n == 0) continue;
                    const Voxel& v = it->second;
                    if (prm.match_mean)
                    {
                        const float d = sqr_dist(v.mean, q);
                        if (d > r2) continue;
                        add(d, v.mean, plain_id(k, 0));
                    }
                    else
                        for (uint32_t i = 0; i < v.n; i++)
                        {
                            const float d = sqr_dist(v.pts[i], q);
                            if (d > r2) continue;
                            add(d, v.pts[i], plain_id(k, (int)i));
                        }
                }
        if (maxp != 0)
        {
            size_t c = 0;
            for (const auto& kv : cand)
            {
                if (c++ >= maxp) break;
                out.push_back({kv.second.first, kv.first, kv.second.second});
            }
        }
        return 0;
    }
    void compute_bbox_uncached(BBox& b) const override
    {  // boundingBox(), SparseVoxelPointCloud.cpp:800-832: from the outer blocks
        const float half = 0.5f * vsize;
        const float gmh  = 1.0f * (vsize * 32) - 1.0f * half;
        for (const Key& o : blocks)
        {
            const P3 c{o.cx * vsize, o.cy * vsize, o.cz * vsize};
            b.add({c.x - 1.0f * half, c.y - 1.0f * half, c.z - 1.0f * half});
            b.add({c.x + gmh, c.y + gmh, c.z + gmh});
        }
    }
    std::vector<const std::pair<const Key, Voxel>*> sorted_nonempty() const
    {
        std::vector<const std::pair<const Key, Voxel>*> v;
        for (const auto& kv : vox)
            if (kv.second.n) v.push_back(&kv);
        std::sort(v.begin(), v.end(), [](auto* a, auto* b) { return a->first < b->first; });
        return v;
    }
    uint64_t num_voxels() const override
    {
        uint64_t c = 0;
        for (const auto& kv : vox) c += kv.second.n != 0;
        return c;
    }
    uint64_t num_points() const override
    {
        uint64_t n = 0;
        for (const auto& kv : vox) n += kv.second.n;
        return n;
    }
    uint32_t row_cap() const override { return CAP; }
    void     export_all(int32_t* keys, uint32_t* counts, float* pts) const override
    {
        auto v = sorted_nonempty();
        for (size_t i = 0; i < v.size(); i++)
        {
            keys[3 * i] = v[i]->first.cx, keys[3 * i + 1] = v[i]->first.cy, keys[3 * i + 2] = v[i]->first.cz;
            counts[i] = v[i]->second.n;
            float* row = pts + i * CAP * 3;
            std::memset(row, 0, sizeof(float) * CAP * 3);
            for (uint32_t j = 0; j < v[i]->second.n; j++)
                row[3 * j] = v[i]->second.pts[j].x, row[3 * j + 1] = v[i]->second.pts[j].y,
                        row[3 * j + 2] = v[i]->second.pts[j].z;
        }
    }
    void voxel_stats(float* mean, float* cov) const override
    {  // mean: the reference's running mean. cov: extension N4 (no reference behaviour):
       // 1/(n-1) sum (p-mean)(p-mean)^T in slot order, fp32, zero when n < 2.
        auto v = sorted_nonempty();
        for (size_t i = 0; i < v.size(); i++)
        {
            const Voxel& x = v[i]->second;
            mean[3 * i] = x.mean.x, mean[3 * i + 1] = x.mean.y, mean[3 * i + 2] = x.mean.z;
            float c[6] = {0, 0, 0, 0, 0, 0};
            for (uint32_t j = 0; j < x.n; j++)
            {
                const float dx = x.pts[j].x - x.mean.x, dy = x.pts[j].y - x.mean.y, dz = x.pts[j].z - x.mean.z;
                c[0] += dx * dx, c[1] += dx * dy, c[2] += dx * dz;
                c[3] += dy * dy, c[4] += dy * dz, c[5] += dz * dz;
            }
            const float k = x.n > 1 ? 1.0f / (float)(x.n - 1) : 0.0f;
            for (int t = 0; t < 6; t++) cov[6 * i + t] = c[t] * k;
        }
    }
};

// ------------------------------------------------------------- SparseTrees
struct SparseTrees : MapBase
{
    struct Block
    {
        std::vector<P3> pts;  // CSimplePointsMap contents in insertion order
    };
    std::map<Key, Block> blocks;  // std::map<outer_index3d_t, GridData>, SparseTreesPointCloud.h:143-144
    float                g = 10.f, ginv = 0.1f, ghalf = 5.f;

    void set_params(const mmo_params& p) override
    {  // setGridProperties, SparseTreesPointCloud.cpp:176-189
        const bool resize = (p.voxel_size != prm.voxel_size);
        prm               = p;
        if (resize)
        {
            g     = p.voxel_size;
            ginv  = 1.0f / g;
            ghalf = 0.5f * g;
            blocks.clear();
        }
    }
    void clear() override { blocks.clear(); }
    bool empty() const override { return blocks.empty(); }
    Key  to_idx(const P3& p) const
    {  // coordToOuterIdx, SparseTreesPointCloud.h:70-77 (truncation of (x + g/2)/g)
        return {(int32_t)((p.x + ghalf) * ginv), (int32_t)((p.y + ghalf) * ginv),
                (int32_t)((p.z + ghalf) * ginv)};
    }
    // In-block exact NN (MRPT CSimplePointsMap -> nanoflann in the reference; N3):
    // minimum d2, ties to the lowest block-local index.
    static bool block_nn(const Block& b, const P3& q, Hit& out)
    {
        if (b.pts.empty()) return false;
        float  best = sqr_dist(q, b.pts[0]);
        size_t bi   = 0;
        for (size_t i = 1; i < b.pts.size(); i++)
        {
            const float d = sqr_dist(q, b.pts[i]);
            if (d < best) best = d, bi = i;
        }
        out = {b.pts[bi], best, (uint64_t)bi};
        return true;
    }
    int nn_single_one(const P3& q, Hit& out, bool& found) const override
    {  // nn_single_search, SparseTreesPointCloud.cpp:613-624: the query's own block only
        auto it = blocks.find(to_idx(q));
        found   = (it != blocks.end()) && block_nn(it->second, q, out);
        return 0;
    }
    int nn_multi_one(const P3& q, size_t k, Hit* out, size_t& nout) const override
    {  // nn_multiple_search :626-643: k nearest inside the block, ascending (d2, index)
        nout    = 0;
        auto it = blocks.find(to_idx(q));
        if (it == blocks.end()) return 0;
        std::vector<Hit> all;
        const auto&      P = it->second.pts;
        all.reserve(P.size());
        for (size_t i = 0; i < P.size(); i++) all.push_back({P[i], sqr_dist(q, P[i]), (uint64_t)i});
        const size_t m = std::min(k, all.size());
        std::partial_sort(all.begin(), all.begin() + m, all.end(), [](const Hit& a, const Hit& b) {
            return a.d2 < b.d2 || (a.d2 == b.d2 && a.id < b.id);
        });
        for (size_t i = 0; i < m; i++) out[i] = all[i];
        nout = m;
        return 0;
    }
    int radius_one(const P3& q, float r2, size_t maxp, std::vector<Hit>& out) const override
    {  // nn_radius_search, SparseTreesPointCloud.cpp:645-713. Per block: hits with d2 < r2
       // (nanoflann's strict test; N3), nearest first, cut to maxPoints if non-zero; then merged
       // through a std::map<float,...> (equal d2 collapse, last block wins; N2), not re-truncated.
        if (r2 <= 0) return 0;
        const float r = std::sqrt(r2);
        const Key   a = to_idx({q.x - r, q.y - r, q.z - r});
        const Key   b = to_idx({q.x + r, q.y + r, q.z + r});
        std::map<float, std::pair<P3, uint64_t>> cand;
        for (int32_t cx = a.cx; cx <= b.cx; cx++)
            for (int32_t cy = a.cy; cy <= b.cy; cy++)
                for (int32_t cz = a.cz; cz <= b.cz; cz++)
                {
                    auto it = blocks.find(Key{cx, cy, cz});
                    if (it == blocks.end()) continue;
                    std::vector<Hit> sub;
                    const auto&      P = it->second.pts;
                    for (size_t i = 0; i < P.size(); i++)
                    {
                        const float d = sqr_dist(q, P[i]);
                        if (d < r2) sub.push_back({P[i], d, (uint64_t)i});
                    }
                    std::sort(sub.begin(), sub.end(), [](const Hit& x, const Hit& y) {
                        return x.d2 < y.d2 || (x.d2 == y.d2 && x.id < y.id);
                    });
                    if (maxp != 0 && sub.size() > maxp) sub.resize(maxp);
                    for (const Hit& h : sub) cand[h.d2] = {h.p, h.id};
                }
        for (const auto& kv : cand) out.push_back({kv.second.first, kv.first, kv.second.second});
        return 0;
    }
    void insert(const float* xs, const float* ys, const float* zs, size_t n, const double* T,
                int has_pose) override
    {  // internal_insertObservation :298-307 + eraseGridsFartherThan :961-1000
        if (has_pose && T && prm.remove_voxels_farther_than > 0)
        {
            // remove_submaps_farther_than is held in a float member; ceil(float*float)
            const float dist = (float)prm.remove_voxels_farther_than;
            const Key   c    = to_idx({(float)T[9], (float)T[10], (float)T[11]});
            const int   d    = (int)std::ceil(dist * ginv);
            for (auto it = blocks.begin(); it != blocks.end();)
            {
                const Key& k = it->first;
                const bool inside = k.cx >= c.cx - d && k.cy >= c.cy - d && k.cz >= c.cz - d &&
                                    k.cx <= c.cx + d && k.cy <= c.cy + d && k.cz <= c.cz + d;
                if (inside)
                    ++it;
                else
                    it = blocks.erase(it);
            }
        }
        // internal_insertPointCloud3D :916-959: clearance test against the pre-batch map,
        // then append the accepted points in input order.
        if (!n) return;
        std::vector<P3>      gp(n);
        std::vector<uint8_t> ins(n);
        const float thr = prm.minimum_points_clearance * prm.minimum_points_clearance;
        for (size_t i = 0; i < n; i++)
        {
            gp[i] = compose(T, xs[i], ys[i], zs[i]);
            Hit  h;
            bool f = false;
            nn_single_one(gp[i], h, f);
            ins[i] = (!f || h.d2 > thr) ? 1 : 0;
        }
        for (size_t i = 0; i < n; i++)
            if (ins[i])
            {  // insertPoint, SparseTreesPointCloud.h:224-234
                blocks[to_idx(gp[i])].pts.push_back(gp[i]);
                bbox.add(gp[i]);
            }
    }
    void compute_bbox_uncached(BBox& b) const override
    {  // boundingBox() :715-741: over all points
        for (const auto& kv : blocks)
            for (const P3& p : kv.second.pts) b.add(p);
    }
    uint64_t num_voxels() const override { return blocks.size(); }
    uint64_t num_points() const override
    {
        uint64_t n = 0;
        for (const auto& kv : blocks) n += kv.second.pts.size();
        return n;
    }
    uint32_t row_cap() const override
    {
        size_t m = 0;
        for (const auto& kv : blocks) m = std::max(m, kv.second.pts.size());
        return (uint32_t)m;
    }
    void export_all(int32_t* keys, uint32_t* counts, float* pts) const override
    {
        const uint32_t cap = row_cap();
        size_t         i   = 0;
        for (const auto& kv : blocks)
        {
            keys[3 * i] = kv.first.cx, keys[3 * i + 1] = kv.first.cy, keys[3 * i + 2] = kv.first.cz;
            counts[i]  = (uint32_t)kv.second.pts.size();
            float* row = pts + i * (size_t)cap * 3;
            std::memset(row, 0, sizeof(float) * cap * 3);
            for (size_t j = 0; j < kv.second.pts.size(); j++)
                row[3 * j] = kv.second.pts[j].x, row[3 * j + 1] = kv.second.pts[j].y,
                        row[3 * j + 2] = kv.second.pts[j].z;
            i++;
        }
    }
};

template <class F>
void shard(size_t n, int nthreads, F&& f)
{
    if (nthreads <= 1 || n < 1024)
    {
        f(0, n);
        return;
    }
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++)
    {
        const size_t a = n * t / nthreads, b = n * (t + 1) / nthreads;
        th.emplace_back([=, &f] { f(a, b); });
    }
    for (auto& t : th) t.join();
}
}  // namespace

extern "C" {
const char* mmo_impl_name(void) { return "port"; }
void*       mmo_create(int kind, const mmo_params* p)
{
    MapBase* m = nullptr;
    if (kind == 0) m = new Hashed();
    if (kind == 1) m = new SparseVoxel();
    if (kind == 2) m = new SparseTrees();
    if (!m) return nullptr;
    m->prm.voxel_size = -1.f;  // force the derived-field update
    m->set_params(*p);
    return m;
}
void mmo_destroy(void* h) { delete (MapBase*)h; }
void mmo_clear(void* h) { ((MapBase*)h)->clear(); }
void mmo_set_params(void* h, const mmo_params* p) { ((MapBase*)h)->set_params(*p); }
int  mmo_is_empty(void* h) { return ((MapBase*)h)->empty() ? 1 : 0; }
int  mmo_insert(void* h, const float* xs, const float* ys, const float* zs, size_t n, const double* T,
                int has_pose)
{
    ((MapBase*)h)->insert(xs, ys, zs, n, T, has_pose);
    return 0;
}
int mmo_nn_single(void* h, const float* xs, const float* ys, const float* zs, size_t n, uint8_t* found,
                  float* oxyz, float* od2, uint64_t* oid, int nthreads)
{
    const MapBase* m  = (MapBase*)h;
    int            rc = 0;
    shard(n, nthreads, [&](size_t a, size_t b) {
        for (size_t i = a; i < b; i++)
        {
            Hit  hit;
            bool f = false;
            int  r = m->nn_single_one({xs[i], ys[i], zs[i]}, hit, f);
            if (r) { rc = r; return; }
            found[i] = f;
            if (f)
            {
                oxyz[3 * i] = hit.p.x, oxyz[3 * i + 1] = hit.p.y, oxyz[3 * i + 2] = hit.p.z;
                od2[i] = hit.d2, oid[i] = hit.id;
            }
        }
    });
    return rc;
}
int mmo_enable_knn_extension(void* h, int on)
{
    auto* sv = dynamic_cast<SparseVoxel*>((MapBase*)h);
    if (!sv) return -5;
    sv->knn_ext = on != 0;
    return 0;
}
int mmo_nn_multi(void* h, const float* xs, const float* ys, const float* zs, size_t n, size_t k,
                 uint32_t* ocount, float* oxyz, float* od2, uint64_t* oid, int nthreads)
{
    const MapBase* m = (MapBase*)h;
    if (k == 0) return -1;  // ASSERT_(N > 0), HashedVoxelPointCloud.cpp:702
    int rc = 0;
    shard(n, nthreads, [&](size_t a, size_t b) {
        std::vector<Hit> buf(k);
        for (size_t i = a; i < b; i++)
        {
            size_t nout = 0;
            int    r    = m->nn_multi_one({xs[i], ys[i], zs[i]}, k, buf.data(), nout);
            if (r) { rc = r; return; }
            ocount[i] = (uint32_t)nout;
            for (size_t j = 0; j < nout; j++)
            {
                const size_t o = i * k + j;
                oxyz[3 * o] = buf[j].p.x, oxyz[3 * o + 1] = buf[j].p.y, oxyz[3 * o + 2] = buf[j].p.z;
                od2[o] = buf[j].d2, oid[o] = buf[j].id;
            }
        }
    });
    return rc;
}
int mmo_nn_radius(void* h, const float* xs, const float* ys, const float* zs, size_t n, float r2,
                  size_t maxp, uint64_t* off, int nthreads)
{
    MapBase* m = (MapBase*)h;
    std::vector<std::vector<Hit>> per(n);
    int rc = 0;
    shard(n, nthreads, [&](size_t a, size_t b) {
        for (size_t i = a; i < b; i++)
        {
            int r = m->radius_one({xs[i], ys[i], zs[i]}, r2, maxp, per[i]);
            if (r) { rc = r; return; }
        }
    });
    if (rc) return rc;
    m->rad_hits.clear();
    off[0] = 0;
    for (size_t i = 0; i < n; i++)
    {
        off[i + 1] = off[i] + per[i].size();
        m->rad_hits.insert(m->rad_hits.end(), per[i].begin(), per[i].end());
    }
    return 0;
}
void mmo_nn_radius_fetch(void* h, float* oxyz, float* od2, uint64_t* oid)
{
    const MapBase* m = (MapBase*)h;
    for (size_t i = 0; i < m->rad_hits.size(); i++)
    {
        oxyz[3 * i] = m->rad_hits[i].p.x, oxyz[3 * i + 1] = m->rad_hits[i].p.y,
                 oxyz[3 * i + 2] = m->rad_hits[i].p.z;
        od2[i] = m->rad_hits[i].d2, oid[i] = m->rad_hits[i].id;
    }
}
int mmo_likelihood(void* h, const float* xs, const float* ys, const float* zs, size_t n, const double* T,
                   double* out)
{  // internal_computeObservationLikelihoodPointCloud3D, HashedVoxelPointCloud.cpp:509-546
    const MapBase* m = (MapBase*)h;
    if (m->empty()) { *out = 0; return 0; }  // :445
    if (!(m->prm.sigma_dist > 0) || m->prm.decimation == 0) return -1;
    const float max_sqr = (float)(m->prm.max_corr_distance * m->prm.max_corr_distance);
    double      sum     = 0;
    size_t      cnt     = 0;
    for (size_t i = 0; i < n; i += m->prm.decimation, cnt++)
    {
        Hit  hit;
        bool f = false;
        int  r = m->nn_single_one(compose(T, xs[i], ys[i], zs[i]), hit, f);
        if (r) return r;
        if (!f) continue;
        float e = hit.d2;
        if (max_sqr < e) e = max_sqr;
        sum += (double)e;
    }
    if (cnt) sum /= (double)cnt;
    *out = -sum / m->prm.sigma_dist;
    return 0;
}
void mmo_bbox(void* h, float out[6])
{
    MapBase* m = (MapBase*)h;
    if (!m->bbox.valid)
    {
        m->bbox.valid = true;
        m->bbox.mn = m->bbox.mx = {0, 0, 0};
        if (!m->empty())
        {
            BBox b;
            m->compute_bbox_uncached(b);
            m->bbox = b;
        }
    }
    out[0] = m->bbox.mn.x, out[1] = m->bbox.mn.y, out[2] = m->bbox.mn.z;
    out[3] = m->bbox.mx.x, out[4] = m->bbox.mx.y, out[5] = m->bbox.mx.z;
}
uint64_t mmo_num_voxels(void* h) { return ((MapBase*)h)->num_voxels(); }
uint64_t mmo_num_points(void* h) { return ((MapBase*)h)->num_points(); }
uint32_t mmo_row_capacity(void* h) { return ((MapBase*)h)->row_cap(); }
void mmo_export(void* h, int32_t* k, uint32_t* c, float* p) { ((MapBase*)h)->export_all(k, c, p); }
void mmo_voxel_stats(void* h, float* mean, float* cov) { ((MapBase*)h)->voxel_stats(mean, cov); }
double mmo_mean_stencil_points(void* h, const float* xs, const float* ys, const float* zs, size_t n)
{
    const MapBase* m = (MapBase*)h;
    double         s = 0;
    for (size_t i = 0; i < n; i++) s += m->stencil_points({xs[i], ys[i], zs[i]});
    return n ? s / (double)n : 0;
}
}
