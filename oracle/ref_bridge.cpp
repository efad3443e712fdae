// ref_bridge.cpp — exposes the REFERENCE's own classes through oracle/mmo.h.
//
// TEST INFRASTRUCTURE ONLY. build_ref.sh compiles this file together with the reference
// sources where they lie (/root/reference/mola_metric_maps/src/{Hashed,SparseVoxel,
// SparseTrees}PointCloud.cpp, unmodified, plus its vendored tsl::robin_map) against the
// minimal MRPT stand-in in shim/mrpt, into oracle/_ref/libmmo_ref.so. No reference source
// is copied into this repository. What runs behind these entry points is therefore the
// reference's own insertPoint / nn_* code; only the MRPT pieces that are not in the
// reference tree (composePoint, round, sqrNorm, CSimplePointsMap) come from the shim.
#include <mola_metric_maps/HashedVoxelPointCloud.h>
#include <mola_metric_maps/SparseTreesPointCloud.h>
#include <mola_metric_maps/SparseVoxelPointCloud.h>
#include <mrpt/obs/CObservationPointCloud.h>

#include <thread>

#include "mmo.h"

namespace
{
using mrpt::math::TPoint3Df;
struct Hit
{
    TPoint3Df p;
    float     d2;
    uint64_t  id;
};
struct Handle
{
    int                                           kind = 0;
    mmo_params                                    prm{};
    std::shared_ptr<mola::HashedVoxelPointCloud>  hv;
    std::shared_ptr<mola::SparseVoxelPointCloud>  sv;
    std::shared_ptr<mola::SparseTreesPointCloud>  st;
    std::vector<Hit>                              rad;
    mrpt::maps::CMetricMap*                       base() const
    {
        return kind == 0 ? (mrpt::maps::CMetricMap*)hv.get() : kind == 1 ? (mrpt::maps::CMetricMap*)sv.get() : (mrpt::maps::CMetricMap*)st.get();
    }
    const mrpt::maps::NearestNeighborsCapable* nn() const
    {
        return kind == 0 ? (const mrpt::maps::NearestNeighborsCapable*)hv.get()
                         : kind == 1 ? (const mrpt::maps::NearestNeighborsCapable*)sv.get() : (const mrpt::maps::NearestNeighborsCapable*)st.get();
    }
    void apply(const mmo_params& p, bool create)
    {
        const bool resize = create || p.voxel_size != prm.voxel_size;
        prm               = p;
        if (kind == 0)
        {
            if (create) hv = mola::HashedVoxelPointCloud::Create(p.voxel_size);
            else if (resize) hv->setVoxelProperties(p.voxel_size);
            hv->insertionOptions.max_points_per_voxel        = p.max_points_per_voxel;
            hv->insertionOptions.remove_voxels_farther_than  = p.remove_voxels_farther_than;
            hv->insertionOptions.min_distance_between_points = p.min_distance_between_points;
            hv->likelihoodOptions.sigma_dist = p.sigma_dist, hv->likelihoodOptions.max_corr_distance = p.max_corr_distance;
            hv->likelihoodOptions.decimation = p.decimation;
        }
        else if (kind == 1)
        {
            if (create) sv = mola::SparseVoxelPointCloud::Create(p.voxel_size);
            else if (resize) sv->setVoxelProperties(p.voxel_size);
            sv->insertionOptions.max_points_per_voxel = p.max_points_per_voxel;
            sv->likelihoodOptions.sigma_dist = p.sigma_dist, sv->likelihoodOptions.max_corr_distance = p.max_corr_distance;
            sv->likelihoodOptions.decimation = p.decimation, sv->likelihoodOptions.match_mean = p.match_mean != 0;
        }
        else
        {
            if (create) st = mola::SparseTreesPointCloud::Create(p.voxel_size);
            else if (resize) st->setGridProperties(p.voxel_size);
            st->insertionOptions.minimum_points_clearance    = p.minimum_points_clearance;
            st->insertionOptions.remove_submaps_farther_than = (float)p.remove_voxels_farther_than;
            st->likelihoodOptions.sigma_dist = p.sigma_dist, st->likelihoodOptions.max_corr_distance = p.max_corr_distance;
            st->likelihoodOptions.decimation = p.decimation;
        }
    }
};
template <class F>
void shard(size_t n, int nthreads, F&& f)
{
    if (nthreads <= 1 || n < 1024) { f(0, n); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++)
    {
        const size_t a = n * t / nthreads, b = n * (t + 1) / nthreads;
        th.emplace_back([=, &f] { f(a, b); });
    }
    for (auto& t : th) t.join();
}
struct VoxRow
{
    int32_t                k[3];
    std::vector<TPoint3Df> pts;
    TPoint3Df              mean;
};
std::vector<VoxRow> rows_of(const Handle& H)
{
    std::vector<VoxRow> r;
    if (H.kind == 0)
        H.hv->visitAllVoxels([&](const mola::HashedVoxelPointCloud::global_index3d_t& i, const mola::HashedVoxelPointCloud::VoxelData& v) {
            VoxRow x{{i.cx, i.cy, i.cz}, {}, {}};
            const auto pts = v.points();
            for (size_t j = 0; j < pts.size(); j++) x.pts.push_back(pts[j]);
            r.push_back(std::move(x));
        });
    else if (H.kind == 1)
        H.sv->visitAllVoxels([&](const mola::SparseVoxelPointCloud::outer_index3d_t& o, const mola::SparseVoxelPointCloud::inner_plain_index_t pi,
                                 const mola::SparseVoxelPointCloud::VoxelData& v, const mola::SparseVoxelPointCloud::InnerGrid& g) {
            const auto pts = v.points(g);
            if (pts.empty()) return;
            const auto in = mola::SparseVoxelPointCloud::plain2i(pi);
            VoxRow     x{{o.cx + (int32_t)in.cx, o.cy + (int32_t)in.cy, o.cz + (int32_t)in.cz}, {}, v.mean()};
            for (size_t j = 0; j < pts.size(); j++) x.pts.push_back(pts[j]);
            r.push_back(std::move(x));
        });
    else
        H.st->visitAllGrids([&](const mola::SparseTreesPointCloud::outer_index3d_t& o, const mola::SparseTreesPointCloud::GridData& g) {
            VoxRow      x{{o.cx, o.cy, o.cz}, {}, {}};
            const auto& P = g.points();
            for (size_t j = 0; j < P.size(); j++)
            {
                float a, b, c;
                P.getPointFast(j, a, b, c);
                x.pts.push_back({a, b, c});
            }
            r.push_back(std::move(x));
        });
    std::sort(r.begin(), r.end(), [](const VoxRow& a, const VoxRow& b) {
        return a.k[0] != b.k[0] ? a.k[0] < b.k[0] : (a.k[1] != b.k[1] ? a.k[1] < b.k[1] : a.k[2] < b.k[2]);
    });
    return r;
}
uint32_t cap_of(const Handle& H, const std::vector<VoxRow>& r)
{
    if (H.kind == 0) return 32;
    if (H.kind == 1) return 16;
    size_t m = 0;
    for (auto& x : r) m = std::max(m, x.pts.size());
    return (uint32_t)m;
}
}  // namespace

extern "C" {
const char* mmo_impl_name(void) { return "reference"; }
void*       mmo_create(int kind, const mmo_params* p)
{
    if (kind < 0 || kind > 2) return nullptr;
    auto* H = new Handle();
    H->kind = kind;
    H->apply(*p, true);
    return H;
}
void mmo_destroy(void* h) { delete (Handle*)h; }
// Archives written / read by the reference classes' own serializeTo / serializeFrom (through the shim's CArchive):
// the byte streams the C++ host classes of mola_b200/host must be able to read and produce.
// Only in this library (the port has no archive code). mmo_ref_serialize returns the archive size; it fills `out`
// when cap is large enough.
size_t mmo_ref_serialize(void* h, uint8_t* out, size_t cap)
{
    auto*                         H = (Handle*)h;
    mrpt::serialization::CArchive ar;
    if (H->kind == 0) ar << *H->hv;
    else if (H->kind == 1) ar << *H->sv;
    else ar << *H->st;
    if (out && cap >= ar.buf.size()) std::memcpy(out, ar.buf.data(), ar.buf.size());
    return ar.buf.size();
}
int mmo_ref_deserialize(void* h, const uint8_t* buf, size_t n)
{
    auto* H = (Handle*)h;
    try
    {
        mrpt::serialization::CArchive ar;
        ar.buf.assign(buf, buf + n);
        if (H->kind == 0) ar >> *H->hv;
        else if (H->kind == 1) ar >> *H->sv;
        else ar >> *H->st;
        return ar.rpos == ar.buf.size() ? 0 : -2;
    }
    catch (const std::exception&)
    {
        return -1;
    }
}
// The reference object itself, as the mrpt::maps::CMetricMap* its plugin interface is. The C++ host-class test (built
// against the same stand-in headers) calls insertObservation / computeObservationLikelihood on it with every
// observation class and a non-identity sensorPose, next to the same call on the GPU class. Reference library only.
void* mmo_ref_metric_map(void* h) { return ((Handle*)h)->base(); }
void  mmo_clear(void* h) { ((Handle*)h)->base()->clear(); }
void mmo_set_params(void* h, const mmo_params* p) { ((Handle*)h)->apply(*p, false); }
int  mmo_is_empty(void* h) { return ((Handle*)h)->base()->isEmpty() ? 1 : 0; }

int mmo_insert(void* h, const float* xs, const float* ys, const float* zs, size_t n, const double* T, int has_pose)
{
    Handle& H = *(Handle*)h;
    mrpt::obs::CObservationPointCloud obs;
    auto                              pc = mrpt::maps::CSimplePointsMap::Create();
    pc->reserve(n);
    for (size_t i = 0; i < n; i++) pc->insertPointFast(xs[i], ys[i], zs[i]);
    obs.pointcloud = pc;
    if (!T)
    {
        H.base()->insertObservation(obs);
        return 0;
    }
    // pose without "robot pose" semantics: same public call with the eviction option switched off
    mmo_params saved = H.prm;
    if (!has_pose)
    {
        mmo_params q                 = H.prm;
        q.remove_voxels_farther_than = 0;
        H.apply(q, false);
    }
    H.base()->insertObservation(obs, mrpt::poses::CPose3D::FromRowMajor12(T));
    if (!has_pose) H.apply(saved, false);
    return 0;
}

int mmo_nn_single(void* h, const float* xs, const float* ys, const float* zs, size_t n, uint8_t* found, float* oxyz, float* od2, uint64_t* oid,
                  int nthreads)
{
    const Handle& H  = *(Handle*)h;
    int           rc = 0;
    shard(n, nthreads, [&](size_t a, size_t b) {
        try
        {
            for (size_t i = a; i < b; i++)
            {
                TPoint3Df  r;
                float      d  = 0;
                uint64_t   id = 0;
                const bool f  = H.nn()->nn_single_search(TPoint3Df(xs[i], ys[i], zs[i]), r, d, id);
                found[i]      = f;
                if (f) oxyz[3 * i] = r.x, oxyz[3 * i + 1] = r.y, oxyz[3 * i + 2] = r.z, od2[i] = d, oid[i] = id;
            }
        }
        catch (const std::exception&)
        {
            rc = -5;
        }
    });
    return rc;
}
int mmo_nn_multi(void* h, const float* xs, const float* ys, const float* zs, size_t n, size_t k, uint32_t* ocount, float* oxyz, float* od2,
                 uint64_t* oid, int nthreads)
{
    const Handle& H  = *(Handle*)h;
    int           rc = 0;
    shard(n, nthreads, [&](size_t a, size_t b) {
        try
        {
            std::vector<TPoint3Df> r;
            std::vector<float>     d;
            std::vector<uint64_t>  ids;
            for (size_t i = a; i < b; i++)
            {
                H.nn()->nn_multiple_search(TPoint3Df(xs[i], ys[i], zs[i]), k, r, d, ids);
                ocount[i] = (uint32_t)r.size();
                for (size_t j = 0; j < r.size(); j++)
                {
                    const size_t o = i * k + j;
                    oxyz[3 * o] = r[j].x, oxyz[3 * o + 1] = r[j].y, oxyz[3 * o + 2] = r[j].z, od2[o] = d[j], oid[o] = ids[j];
                }
            }
        }
        catch (const std::exception&)
        {
            rc = (k == 0) ? -1 : -5;
        }
    });
    return rc;
}
int mmo_nn_radius(void* h, const float* xs, const float* ys, const float* zs, size_t n, float r2, size_t maxp, uint64_t* off, int nthreads)
{
    Handle&                       H = *(Handle*)h;
    std::vector<std::vector<Hit>> per(n);
    int                           rc = 0;
    shard(n, nthreads, [&](size_t a, size_t b) {
        try
        {
            std::vector<TPoint3Df> r;
            std::vector<float>     d;
            std::vector<uint64_t>  ids;
            for (size_t i = a; i < b; i++)
            {
                H.nn()->nn_radius_search(TPoint3Df(xs[i], ys[i], zs[i]), r2, r, d, ids, maxp);
                for (size_t j = 0; j < r.size(); j++) per[i].push_back({r[j], d[j], ids[j]});
            }
        }
        catch (const std::exception&)
        {
            rc = -5;
        }
    });
    if (rc) return rc;
    H.rad.clear();
    off[0] = 0;
    for (size_t i = 0; i < n; i++)
    {
        off[i + 1] = off[i] + per[i].size();
        H.rad.insert(H.rad.end(), per[i].begin(), per[i].end());
    }
    return 0;
}
void mmo_nn_radius_fetch(void* h, float* oxyz, float* od2, uint64_t* oid)
{
    const Handle& H = *(Handle*)h;
    for (size_t i = 0; i < H.rad.size(); i++)
        oxyz[3 * i] = H.rad[i].p.x, oxyz[3 * i + 1] = H.rad[i].p.y, oxyz[3 * i + 2] = H.rad[i].p.z, od2[i] = H.rad[i].d2, oid[i] = H.rad[i].id;
}
int mmo_likelihood(void* h, const float* xs, const float* ys, const float* zs, size_t n, const double* T, double* out)
{
    Handle& H = *(Handle*)h;
    if (H.prm.decimation == 0) return -1;  // the reference would loop forever
    mrpt::obs::CObservationPointCloud obs;
    auto                              pc = mrpt::maps::CSimplePointsMap::Create();
    for (size_t i = 0; i < n; i++) pc->insertPointFast(xs[i], ys[i], zs[i]);
    obs.pointcloud = pc;
    try
    {
        *out = H.base()->computeObservationLikelihood(obs, T ? mrpt::poses::CPose3D::FromRowMajor12(T) : mrpt::poses::CPose3D());
    }
    catch (const std::exception& e)
    {
        return std::string(e.what()).find("n-search") != std::string::npos ? -5 : -1;
    }
    return 0;
}
void mmo_bbox(void* h, float out[6])
{
    const auto b = ((Handle*)h)->base()->boundingBox();
    out[0] = b.min.x, out[1] = b.min.y, out[2] = b.min.z, out[3] = b.max.x, out[4] = b.max.y, out[5] = b.max.z;
}
uint64_t mmo_num_voxels(void* h) { return rows_of(*(Handle*)h).size(); }
uint64_t mmo_num_points(void* h)
{
    uint64_t n = 0;
    for (auto& r : rows_of(*(Handle*)h)) n += r.pts.size();
    return n;
}
uint32_t mmo_row_capacity(void* h)
{
    const Handle& H = *(Handle*)h;
    return cap_of(H, rows_of(H));
}
void mmo_export(void* h, int32_t* keys, uint32_t* counts, float* pts)
{
    const Handle&  H   = *(Handle*)h;
    const auto     r   = rows_of(H);
    const uint32_t cap = cap_of(H, r);
    for (size_t i = 0; i < r.size(); i++)
    {
        keys[3 * i] = r[i].k[0], keys[3 * i + 1] = r[i].k[1], keys[3 * i + 2] = r[i].k[2];
        counts[i]  = (uint32_t)r[i].pts.size();
        float* row = pts + i * (size_t)cap * 3;
        std::memset(row, 0, sizeof(float) * cap * 3);
        for (size_t j = 0; j < r[i].pts.size(); j++) row[3 * j] = r[i].pts[j].x, row[3 * j + 1] = r[i].pts[j].y, row[3 * j + 2] = r[i].pts[j].z;
    }
}
void mmo_voxel_stats(void* h, float* mean, float* cov)
{  // mean: the reference's VoxelData::mean(); cov: NOT in the reference (extension N4), same formula as the port
    const Handle& H = *(Handle*)h;
    if (H.kind != 1) return;
    const auto r = rows_of(H);
    for (size_t i = 0; i < r.size(); i++)
    {
        const TPoint3Df m = r[i].mean;
        mean[3 * i] = m.x, mean[3 * i + 1] = m.y, mean[3 * i + 2] = m.z;
        float c[6] = {0, 0, 0, 0, 0, 0};
        for (const auto& p : r[i].pts)
        {
            const float dx = p.x - m.x, dy = p.y - m.y, dz = p.z - m.z;
            c[0] += dx * dx, c[1] += dx * dy, c[2] += dx * dz, c[3] += dy * dy, c[4] += dy * dz, c[5] += dz * dz;
        }
        const float k = r[i].pts.size() > 1 ? 1.0f / (float)(r[i].pts.size() - 1) : 0.0f;
        for (int t = 0; t < 6; t++) cov[6 * i + t] = c[t] * k;
    }
}
double mmo_mean_stencil_points(void* h, const float* xs, const float* ys, const float* zs, size_t n)
{
    const Handle& H = *(Handle*)h;
    if (H.kind != 0 || !n) return 0;
    double s = 0;
    for (size_t i = 0; i < n; i++)
    {
        const auto c = H.hv->coordToGlobalIdx({xs[i], ys[i], zs[i]});
        for (int dx = -1; dx <= 1; dx++)
            for (int dy = -1; dy <= 1; dy++)
                for (int dz = -1; dz <= 1; dz++)
                    if (auto* v = H.hv->voxelByGlobalIdxs({c.cx + dx, c.cy + dy, c.cz + dz})) s += v->points().size();
    }
    return s / (double)n;
}
}
