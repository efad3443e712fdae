// GpuMetricMaps.cpp — see GpuMetricMaps.h. Every method is a thin forward to the C ABI
// (include/mola_b200/mmg.h); reference behaviour is cited per method (mola_metric_maps/src/).
#include "GpuMetricMaps.h"

#include <array>
#include <map>

#include <mrpt/core/exceptions.h>
#include <mrpt/obs/CObservation2DRangeScan.h>
#include <mrpt/obs/CObservation3DRangeScan.h>
#include <mrpt/obs/CObservationVelodyneScan.h>

#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

namespace MOLA_B200_NAMESPACE
{
namespace
{
// CPose3D -> row-major R (9) followed by t (3)
void pose12(const mrpt::poses::CPose3D& p, double o[12])
{
#ifdef MOLA_B200_MRPT_SHIM
    std::memcpy(o, p.rot(), 9 * sizeof(double));
    std::memcpy(o + 9, p.trans(), 3 * sizeof(double));
#else
    const auto& R = p.getRotationMatrix();
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) o[3 * i + j] = R(i, j);
    o[9] = p.x(), o[10] = p.y(), o[11] = p.z();
#endif
}
struct PoseArg
{
    double       v[12];
    const double* ptr = nullptr;
    explicit PoseArg(const std::optional<const mrpt::poses::CPose3D>& p)
    {
        if (p) pose12(*p, v), ptr = v;
    }
};
}  // namespace

void GpuPointMapBase::check(int rc)
{
    if (rc == MMG_OK) return;
    THROW_EXCEPTION(std::string("mola_b200: ") + mmg_last_error());
}

GpuPointMapBase::GpuPointMapBase(int kind, float size) : kind_(kind), size_(size)
{
    mmg_default_params(kind, &pushed_);
    pushed_.voxel_size = size;
    check(mmg_create(kind, &pushed_, /*device*/ 0, &h_));  // throws without a GPU: there is no CPU path
}
GpuPointMapBase::~GpuPointMapBase()
{
    if (h_) mmg_destroy(h_);
}
GpuPointMapBase::GpuPointMapBase(const GpuPointMapBase& o) : CMetricMap(o), kind_(o.kind_), size_(o.size_), pushed_(o.pushed_)
{
    check(mmg_clone(o.h_, &h_));  // deep copy, like the reference's value semantics (tests/test-serialization.cpp:35-51)
}
GpuPointMapBase& GpuPointMapBase::operator=(const GpuPointMapBase& o)
{
    if (this == &o) return *this;
    CMetricMap::operator=(o);
    mmg_map* n = nullptr;
    check(mmg_clone(o.h_, &n));
    if (h_) mmg_destroy(h_);
    h_ = n, kind_ = o.kind_, size_ = o.size_, pushed_ = o.pushed_;
    cachedPoints_.reset();
    return *this;
}
GpuPointMapBase::GpuPointMapBase(GpuPointMapBase&& o) noexcept : CMetricMap(o), kind_(o.kind_), size_(o.size_), h_(o.h_), pushed_(o.pushed_) { o.h_ = nullptr; }
GpuPointMapBase& GpuPointMapBase::operator=(GpuPointMapBase&& o) noexcept
{
    if (this == &o) return *this;
    if (h_) mmg_destroy(h_);
    h_ = o.h_, o.h_ = nullptr, kind_ = o.kind_, size_ = o.size_, pushed_ = o.pushed_;
    cachedPoints_.reset();
    return *this;
}

void GpuPointMapBase::syncParams() const
{
    std::lock_guard<std::mutex> lk(paramsMtx_);
    mmg_params p = pushed_;
    fillParams(p);
    p.voxel_size = size_;
    if (std::memcmp(&p, &pushed_, sizeof(p)) != 0)
    {
        check(mmg_set_params(h_, &p));
        pushed_ = p;
    }
}
void GpuPointMapBase::resize(float size)
{  // setVoxelProperties(): always clears, even for an equal size (HashedVoxelPointCloud.cpp:189-200)
    size_ = size;
    syncParams();
    check(mmg_clear(h_));
    check(mmg_invalidate_bbox(h_));
    cachedPoints_.reset();
}
void GpuPointMapBase::internal_clear()
{
    check(mmg_clear(h_));
    cachedPoints_.reset();
}
bool GpuPointMapBase::isEmpty() const
{
    int e = 1;
    check(mmg_is_empty(h_, &e));
    return e != 0;
}
mrpt::math::TBoundingBoxf GpuPointMapBase::boundingBox() const
{
    float b[6];
    check(mmg_bbox(h_, b));
    return mrpt::math::TBoundingBoxf({b[0], b[1], b[2]}, {b[3], b[4], b[5]});
}

void GpuPointMapBase::insertPointCloud(const float* xs, const float* ys, const float* zs, size_t n, const std::optional<const mrpt::poses::CPose3D>& pose,
                                       bool evict)
{
    syncParams();
    const PoseArg P(pose);
    const int     rc = mmg_insert_batch(h_, xs, ys, zs, n, P.ptr, (evict && P.ptr) ? 1 : 0, MMG_MEM_HOST, nullptr);
    if (rc != MMG_ERR_OUT_OF_RANGE) check(rc);  // out-of-range points are skipped, the rest is inserted
    cachedPoints_.reset();
}
void GpuPointMapBase::insertPoint(const mrpt::math::TPoint3Df& pt)
{   // the reference's public insertPoint(): decimating for the voxel maps, a plain append for SparseTrees (.h:224-234)
    syncParams();
    const int rc = mmg_insert_batch(h_, &pt.x, &pt.y, &pt.z, 1, nullptr, MMG_INSERT_NO_CLEARANCE, MMG_MEM_HOST, nullptr);
    if (rc != MMG_ERR_OUT_OF_RANGE) check(rc);
    cachedPoints_.reset();
}

void GpuPointMapBase::insertObservationPoints(const float* xs, const float* ys, const float* zs, size_t n,
                                              const std::optional<const mrpt::poses::CPose3D>& robotPose, bool pointByPoint)
{
    // The reference composes every point with robotPose3D, which is the identity when no robot pose is given
    // (HashedVoxelPointCloud.cpp:297-331); the eviction only runs with a given pose (:300-327).
    syncParams();
    double T[12];
    pose12(robotPose ? mrpt::poses::CPose3D(*robotPose) : mrpt::poses::CPose3D(), T);
    int flags = robotPose ? MMG_INSERT_ROBOT_POSE : 0;
    // SparseTrees: the branches that loop over the public insertPoint() append without the clearance test
    // (SparseTreesPointCloud.cpp:349-356 against :381-383, :400-401); for the voxel maps both forms decimate alike
    if (pointByPoint && kind_ == MMG_SPARSE_TREES) flags |= MMG_INSERT_NO_CLEARANCE;
    const int rc = mmg_insert_batch(h_, xs, ys, zs, n, T, flags, MMG_MEM_HOST, nullptr);
    if (rc != MMG_ERR_OUT_OF_RANGE) check(rc);  // out-of-range points are skipped, the rest is inserted
    cachedPoints_.reset();
}

bool GpuPointMapBase::internal_insertObservation(const mrpt::obs::CObservation& obs, const std::optional<const mrpt::poses::CPose3D>& robotPose)
{
    // HashedVoxelPointCloud.cpp:289-432, SparseVoxelPointCloud.cpp:363-480, SparseTreesPointCloud.cpp:289-410.
    // Every class composes its points with the ROBOT pose only: a 2-D scan's auxiliary map and an unprojected range
    // image already are in the robot frame (they include sensorPose), while points3D, a Velodyne cloud and a
    // CObservationPointCloud are sensor-frame points whose sensorPose the reference ignores on insertion.
    using namespace mrpt::obs;
    // which branches loop over insertPoint() (true) and which call internal_insertPointCloud3D (false)
    const bool trees = kind_ == MMG_SPARSE_TREES;
    if (auto* o = dynamic_cast<const CObservation2DRangeScan*>(&obs); o)
    {
        const auto* scanPoints = o->buildAuxPointsMap<mrpt::maps::CPointsMap>();
        // "if (scanPoints->empty()) return 0;" comes after the eviction in the reference: run it for its side effect
        const auto& xs = scanPoints->getPointsBufferRef_x();
        const auto& ys = scanPoints->getPointsBufferRef_y();
        const auto& zs = scanPoints->getPointsBufferRef_z();
        insertObservationPoints(xs.data(), ys.data(), zs.data(), xs.size(), robotPose, false);
        return !scanPoints->empty();
    }
    if (auto* o = dynamic_cast<const CObservation3DRangeScan*>(&obs); o)
    {
        if (o->hasPoints3D)
        {
            insertObservationPoints(o->points3D_x.data(), o->points3D_y.data(), o->points3D_z.data(), o->points3D_x.size(), robotPose, true);
            return true;
        }
        if (o->hasRangeImage)
        {
            T3DPointsProjectionParams pp;
            pp.takeIntoAccountSensorPoseOnRobot = true;
            mrpt::maps::CSimplePointsMap pointMap;
            const_cast<CObservation3DRangeScan*>(o)->unprojectInto(pointMap, pp);
            const auto& xs = pointMap.getPointsBufferRef_x();
            const auto& ys = pointMap.getPointsBufferRef_y();
            const auto& zs = pointMap.getPointsBufferRef_z();
            insertObservationPoints(xs.data(), ys.data(), zs.data(), xs.size(), robotPose, false);
            return true;
        }
        insertObservationPoints(nullptr, nullptr, nullptr, 0, robotPose, false);  // the eviction has already run there
        return false;
    }
    if (auto* o = dynamic_cast<const CObservationVelodyneScan*>(&obs); o)
    {
        if (!o->point_cloud.size()) const_cast<CObservationVelodyneScan*>(o)->generatePointCloud();
        const auto& pc = o->point_cloud;
        insertObservationPoints(pc.x.data(), pc.y.data(), pc.z.data(), pc.x.size(), robotPose, !trees);
        return true;
    }
    if (auto* o = dynamic_cast<const CObservationPointCloud*>(&obs); o)
    {
        if (!o->pointcloud) insertObservationPoints(nullptr, nullptr, nullptr, 0, robotPose, false);
        ASSERT_(o->pointcloud);
        const auto& xs = o->pointcloud->getPointsBufferRef_x();
        const auto& ys = o->pointcloud->getPointsBufferRef_y();
        const auto& zs = o->pointcloud->getPointsBufferRef_z();
        insertObservationPoints(xs.data(), ys.data(), zs.data(), xs.size(), robotPose, !trees);
        return true;
    }
    insertObservationPoints(nullptr, nullptr, nullptr, 0, robotPose, false);  // unknown class (:426-432): eviction only
    return false;
}

bool GpuPointMapBase::internal_canComputeObservationLikelihood(const mrpt::obs::CObservation& obs) const
{  // HashedVoxelPointCloud.cpp:548-556
    return dynamic_cast<const mrpt::obs::CObservation2DRangeScan*>(&obs) != nullptr ||
           dynamic_cast<const mrpt::obs::CObservationVelodyneScan*>(&obs) != nullptr ||
           dynamic_cast<const mrpt::obs::CObservationPointCloud*>(&obs) != nullptr;
}
double GpuPointMapBase::internal_computeObservationLikelihood(const mrpt::obs::CObservation& obs, const mrpt::poses::CPose3D& takenFrom) const
{  // HashedVoxelPointCloud.cpp:434-507: a 2-D scan's points are in the robot frame (takenFrom alone), the other two
   // classes hold sensor-frame points (takenFrom + sensorPose)
    using namespace mrpt::obs;
    if (isEmpty()) return 0;
    const float *xs = nullptr, *ys = nullptr, *zs = nullptr;
    size_t       n  = 0;
    mrpt::poses::CPose3D pc_in_map = takenFrom;
    if (auto* s = dynamic_cast<const CObservation2DRangeScan*>(&obs); s)
    {
        const auto* scanPoints = s->buildAuxPointsMap<mrpt::maps::CPointsMap>();
        xs = scanPoints->getPointsBufferRef_x().data(), ys = scanPoints->getPointsBufferRef_y().data(), zs = scanPoints->getPointsBufferRef_z().data();
        n  = scanPoints->size();
    }
    else if (auto* v = dynamic_cast<const CObservationVelodyneScan*>(&obs); v)
    {
        if (!v->point_cloud.size()) const_cast<CObservationVelodyneScan*>(v)->generatePointCloud();
        xs = v->point_cloud.x.data(), ys = v->point_cloud.y.data(), zs = v->point_cloud.z.data(), n = v->point_cloud.size();
        pc_in_map = takenFrom + v->sensorPose;
    }
    else if (auto* o = dynamic_cast<const CObservationPointCloud*>(&obs); o)
    {
        ASSERT_(o->pointcloud);
        xs = o->pointcloud->getPointsBufferRef_x().data(), ys = o->pointcloud->getPointsBufferRef_y().data(),
        zs = o->pointcloud->getPointsBufferRef_z().data(), n = o->pointcloud->size();
        pc_in_map = takenFrom + o->sensorPose;
    }
    else
        return 0;
    if (!n) return 0;
    syncParams();
    double T[12], out = 0;
    pose12(pc_in_map, T);
    check(mmg_likelihood(h_, xs, ys, zs, n, T, &out, MMG_MEM_HOST, nullptr));
    return out;
}

// ------------------------------------------------------------------ queries
void GpuPointMapBase::nn_single_search_batch(const float* xs, const float* ys, const float* zs, size_t n, NNBatchResult& out,
                                             const std::optional<const mrpt::poses::CPose3D>& pose) const
{
    syncParams();
    const PoseArg P(pose);
    out.found.assign(n, 0), out.points.resize(n), out.dist_sqr.resize(n), out.ids.resize(n);
    static_assert(sizeof(mrpt::math::TPoint3Df) == 3 * sizeof(float), "TPoint3Df must be 3 packed floats");
    check(mmg_nn_single_batch(h_, xs, ys, zs, n, P.ptr, out.found.data(), reinterpret_cast<float*>(out.points.data()), out.dist_sqr.data(),
                              out.ids.data(), MMG_MEM_HOST, nullptr));
}
void GpuPointMapBase::nn_multiple_search_batch(const float* xs, const float* ys, const float* zs, size_t n, size_t N, NNBatchResult& out,
                                               const std::optional<const mrpt::poses::CPose3D>& pose) const
{
    ASSERT_(N > 0);  // HashedVoxelPointCloud.cpp:702
    syncParams();
    const PoseArg P(pose);
    out.count.assign(n, 0), out.points.resize(n * N), out.dist_sqr.resize(n * N), out.ids.resize(n * N);
    check(mmg_nn_multi_batch(h_, xs, ys, zs, n, P.ptr, N, out.count.data(), reinterpret_cast<float*>(out.points.data()), out.dist_sqr.data(),
                             out.ids.data(), MMG_MEM_HOST, nullptr));
}
void GpuPointMapBase::nn_radius_search_batch(const float* xs, const float* ys, const float* zs, size_t n, float r2, size_t maxPoints, NNBatchResult& out,
                                             const std::optional<const mrpt::poses::CPose3D>& pose) const
{
    syncParams();
    const PoseArg P(pose);
    out.offsets.assign(n + 1, 0);
    uint64_t total = 0;
    check(mmg_nn_radius_batch(h_, xs, ys, zs, n, P.ptr, r2, maxPoints, out.offsets.data(), &total, MMG_MEM_HOST, nullptr));
    out.points.resize(total), out.dist_sqr.resize(total), out.ids.resize(total);
    if (total) check(mmg_nn_radius_fetch(h_, reinterpret_cast<float*>(out.points.data()), out.dist_sqr.data(), out.ids.data(), MMG_MEM_HOST, nullptr));
}

void GpuPointMapBase::match_points_batch(const float* xs, const float* ys, const float* zs, size_t n, float max_dist_sqr, NNBatchResult& out,
                                         const std::optional<const mrpt::poses::CPose3D>& pose) const
{
    syncParams();
    const PoseArg P(pose);
    out.local_idx.resize(n), out.points.resize(n), out.dist_sqr.resize(n), out.ids.resize(n);
    uint64_t m = 0;
    check(mmg_match_batch(h_, xs, ys, zs, n, P.ptr, max_dist_sqr, &m, out.local_idx.data(), reinterpret_cast<float*>(out.points.data()),
                          out.dist_sqr.data(), out.ids.data(), MMG_MEM_HOST, nullptr));
    out.local_idx.resize(m), out.points.resize(m), out.dist_sqr.resize(m), out.ids.resize(m);
}
void GpuPointMapBase::reserve(uint64_t voxels) { check(mmg_reserve(h_, voxels)); }

bool GpuPointMapBase::nn_single_search(const mrpt::math::TPoint3Df& q, mrpt::math::TPoint3Df& result, float& d2, uint64_t& id) const
{  // HashedVoxelPointCloud.cpp:666-679
    NNBatchResult r;
    nn_single_search_batch(&q.x, &q.y, &q.z, 1, r);
    if (!r.found[0]) return false;
    result = r.points[0], d2 = r.dist_sqr[0], id = r.ids[0];
    return true;
}
void GpuPointMapBase::nn_multiple_search(const mrpt::math::TPoint3Df& q, const size_t N, std::vector<mrpt::math::TPoint3Df>& results,
                                         std::vector<float>& d2, std::vector<uint64_t>& ids) const
{  // :681-771; outputs are cleared first (:698-700)
    results.clear(), d2.clear(), ids.clear();
    NNBatchResult r;
    nn_multiple_search_batch(&q.x, &q.y, &q.z, 1, N, r);
    const size_t k = r.count[0];
    results.assign(r.points.begin(), r.points.begin() + k);
    d2.assign(r.dist_sqr.begin(), r.dist_sqr.begin() + k);
    ids.assign(r.ids.begin(), r.ids.begin() + k);
}
void GpuPointMapBase::nn_radius_search(const mrpt::math::TPoint3Df& q, const float r2, std::vector<mrpt::math::TPoint3Df>& results, std::vector<float>& d2,
                                       std::vector<uint64_t>& ids, size_t maxPoints) const
{  // :773-870
    NNBatchResult r;
    nn_radius_search_batch(&q.x, &q.y, &q.z, 1, r2, maxPoints, r);
    results = std::move(r.points), d2 = std::move(r.dist_sqr), ids = std::move(r.ids);
}
// the 2D overloads throw, like the reference (:638-664)
bool GpuPointMapBase::nn_single_search(const mrpt::math::TPoint2Df&, mrpt::math::TPoint2Df&, float&, uint64_t&) const
{
    THROW_EXCEPTION(std::string("Cannot run a 2D search on a ") + mapName());
}
void GpuPointMapBase::nn_multiple_search(const mrpt::math::TPoint2Df&, const size_t, std::vector<mrpt::math::TPoint2Df>&, std::vector<float>&,
                                         std::vector<uint64_t>&) const
{
    THROW_EXCEPTION(std::string("Cannot run a 2D search on a ") + mapName());
}
void GpuPointMapBase::nn_radius_search(const mrpt::math::TPoint2Df&, const float, std::vector<mrpt::math::TPoint2Df>&, std::vector<float>&,
                                       std::vector<uint64_t>&, size_t) const
{
    THROW_EXCEPTION(std::string("Cannot run a 2D search on a ") + mapName());
}

// ----------------------------------------------------------------- contents
size_t GpuPointMapBase::numVoxels() const
{
    uint64_t v = 0;
    check(mmg_num_voxels(h_, &v));
    return v;
}
size_t GpuPointMapBase::numPoints() const
{
    uint64_t v = 0;
    check(mmg_num_points(h_, &v));
    return v;
}
void GpuPointMapBase::exportVoxels(std::vector<int32_t>& keys, std::vector<uint32_t>& counts, std::vector<mrpt::math::TPoint3Df>& points,
                                   uint32_t& cap) const
{
    const size_t V = numVoxels();
    check(mmg_row_capacity(h_, &cap));
    keys.assign(3 * V, 0), counts.assign(V, 0), points.assign(V * cap, mrpt::math::TPoint3Df(0, 0, 0));
    if (V) check(mmg_export(h_, keys.data(), counts.data(), reinterpret_cast<float*>(points.data()), MMG_MEM_HOST, nullptr));
}
void GpuPointMapBase::visitAllPoints(const std::function<void(const mrpt::math::TPoint3Df&)>& f) const
{  // :904-919 (voxel order is unspecified there as well: hash-map iteration)
    std::vector<int32_t>               k;
    std::vector<uint32_t>              c;
    std::vector<mrpt::math::TPoint3Df> p;
    uint32_t                           cap = 0;
    exportVoxels(k, c, p, cap);
    for (size_t v = 0; v < c.size(); v++)
        for (uint32_t j = 0; j < c[v]; j++) f(p[v * cap + j]);
}
const mrpt::maps::CSimplePointsMap* GpuPointMapBase::getAsSimplePointsMap() const
{  // :1084-1096: cached copy of all points
    if (!cachedPoints_)
    {
        cachedPoints_ = mrpt::maps::CSimplePointsMap::Create();
        cachedPoints_->reserve(numPoints());
        visitAllPoints([this](const mrpt::math::TPoint3Df& p) { cachedPoints_->insertPointFast(p.x, p.y, p.z); });
    }
    return cachedPoints_.get();
}
void GpuPointMapBase::saveMetricMapRepresentationToFile(const std::string& prefix) const
{  // :202-215 writes <prefix>.txt with one "x y z" line per point
    std::ofstream f(prefix + ".txt");
    visitAllPoints([&f](const mrpt::math::TPoint3Df& p) { f << p.x << " " << p.y << " " << p.z << "\n"; });
}

// =================================================================== Hashed
IMPLEMENTS_SERIALIZABLE(HashedVoxelPointCloud, CMetricMap, MOLA_B200_NAMESPACE)

HashedVoxelPointCloud::HashedVoxelPointCloud(float voxel_size) : GpuPointMapBase(MMG_HASHED_VOXEL, voxel_size) {}
std::string HashedVoxelPointCloud::asString() const
{
    std::ostringstream s;
    s << "HashedVoxelPointCloud, resolution=" << voxelSize() << " voxels=" << numVoxels() << " points=" << numPoints();
    return s.str();
}
void HashedVoxelPointCloud::fillParams(mmg_params& p) const
{
    p.max_points_per_voxel        = insertionOptions.max_points_per_voxel;
    p.remove_voxels_farther_than  = insertionOptions.remove_voxels_farther_than;
    p.min_distance_between_points = insertionOptions.min_distance_between_points;
    p.sigma_dist                  = likelihoodOptions.sigma_dist;
    p.max_corr_distance           = likelihoodOptions.max_corr_distance;
    p.decimation                  = likelihoodOptions.decimation;
}
// option structs: same keys and stream layout as the reference (.cpp:922-1063)
void HashedVoxelPointCloud::TInsertionOptions::loadFromConfigFile(const mrpt::config::CConfigFileBase& c, const std::string& s)
{
    MRPT_LOAD_CONFIG_VAR(max_points_per_voxel, int, c, s);  // read as int into a uint32 (.cpp:1041)
    MRPT_LOAD_CONFIG_VAR(remove_voxels_farther_than, double, c, s);
    MRPT_LOAD_CONFIG_VAR(min_distance_between_points, float, c, s);
}
void HashedVoxelPointCloud::TInsertionOptions::dumpToTextStream(std::ostream& out) const
{
    out << "\n------ [HashedVoxelPointCloud::TInsertionOptions] ------- \n\n";
    LOADABLEOPTS_DUMP_VAR(max_points_per_voxel, int);
    LOADABLEOPTS_DUMP_VAR(remove_voxels_farther_than, double);
    LOADABLEOPTS_DUMP_VAR(min_distance_between_points, float)
}
void HashedVoxelPointCloud::TInsertionOptions::writeToStream(mrpt::serialization::CArchive& out) const
{
    const int8_t version = 0;
    out << version << max_points_per_voxel << remove_voxels_farther_than << min_distance_between_points;
}
void HashedVoxelPointCloud::TInsertionOptions::readFromStream(mrpt::serialization::CArchive& in)
{
    int8_t version;
    in >> version;
    if (version != 0) MRPT_THROW_UNKNOWN_SERIALIZATION_VERSION(version);
    in >> max_points_per_voxel >> remove_voxels_farther_than >> min_distance_between_points;
}
void HashedVoxelPointCloud::TLikelihoodOptions::loadFromConfigFile(const mrpt::config::CConfigFileBase& c, const std::string& s)
{
    MRPT_LOAD_CONFIG_VAR(sigma_dist, double, c, s);
    MRPT_LOAD_CONFIG_VAR(max_corr_distance, double, c, s);
    MRPT_LOAD_CONFIG_VAR(decimation, int, c, s);
}
void HashedVoxelPointCloud::TLikelihoodOptions::dumpToTextStream(std::ostream& out) const
{
    out << "\n------ [HashedVoxelPointCloud::TLikelihoodOptions] ------- \n\n";
    LOADABLEOPTS_DUMP_VAR(sigma_dist, double);
    LOADABLEOPTS_DUMP_VAR(max_corr_distance, double);
    LOADABLEOPTS_DUMP_VAR(decimation, int);
}
void HashedVoxelPointCloud::TLikelihoodOptions::writeToStream(mrpt::serialization::CArchive& out) const
{
    const int8_t version = 0;
    out << version << sigma_dist << max_corr_distance << decimation;
}
void HashedVoxelPointCloud::TLikelihoodOptions::readFromStream(mrpt::serialization::CArchive& in)
{
    int8_t version;
    in >> version;
    if (version != 0) MRPT_THROW_UNKNOWN_SERIALIZATION_VERSION(version);
    in >> sigma_dist >> max_corr_distance >> decimation;
}
void HashedVoxelPointCloud::TRenderOptions::loadFromConfigFile(const mrpt::config::CConfigFileBase& c, const std::string& s)
{
    MRPT_LOAD_CONFIG_VAR(point_size, float, c, s);
    MRPT_LOAD_CONFIG_VAR(color.R, float, c, s);
    MRPT_LOAD_CONFIG_VAR(color.G, float, c, s);
    MRPT_LOAD_CONFIG_VAR(color.B, float, c, s);
    colormap                    = static_cast<int8_t>(c.read_int(s, "colormap", colormap));
    recolorizeByCoordinateIndex = static_cast<uint8_t>(c.read_int(s, "recolorizeByCoordinateIndex", recolorizeByCoordinateIndex));
}
void HashedVoxelPointCloud::TRenderOptions::dumpToTextStream(std::ostream& out) const
{
    out << "\n------ [HashedVoxelPointCloud::TRenderOptions] ------- \n\n";
    LOADABLEOPTS_DUMP_VAR(point_size, float);
    LOADABLEOPTS_DUMP_VAR(color.R, float);
    LOADABLEOPTS_DUMP_VAR(color.G, float);
    LOADABLEOPTS_DUMP_VAR(color.B, float);
    LOADABLEOPTS_DUMP_VAR(colormap, int);
    LOADABLEOPTS_DUMP_VAR(recolorizeByCoordinateIndex, int);
}
void HashedVoxelPointCloud::TRenderOptions::writeToStream(mrpt::serialization::CArchive& out) const
{
    const int8_t version = 0;
    out << version << point_size << color << colormap << recolorizeByCoordinateIndex;
}
void HashedVoxelPointCloud::TRenderOptions::readFromStream(mrpt::serialization::CArchive& in)
{
    int8_t version;
    in >> version;
    if (version != 0) MRPT_THROW_UNKNOWN_SERIALIZATION_VERSION(version);
    in >> point_size >> color >> colormap >> recolorizeByCoordinateIndex;
}

// Serialization v0, byte-compatible with the reference (.cpp:114-177): voxel size, the three
// option blocks, then per voxel its index triple, a uint16 count and the points.
uint8_t HashedVoxelPointCloud::serializeGetVersion() const { return 0; }
void    HashedVoxelPointCloud::serializeTo(mrpt::serialization::CArchive& out) const
{
    out << voxelSize();
    insertionOptions.writeToStream(out);
    likelihoodOptions.writeToStream(out);
    renderOptions.writeToStream(out);
    std::vector<int32_t>               k;
    std::vector<uint32_t>              c;
    std::vector<mrpt::math::TPoint3Df> p;
    uint32_t                           cap = 0;
    exportVoxels(k, c, p, cap);
    out.WriteAs<uint32_t>(c.size());
    for (size_t v = 0; v < c.size(); v++)
    {
        out << k[3 * v] << k[3 * v + 1] << k[3 * v + 2];
        out.WriteAs<uint16_t>(c[v]);
        for (uint32_t j = 0; j < c[v]; j++) out << p[v * cap + j];
    }
}
void HashedVoxelPointCloud::serializeFrom(mrpt::serialization::CArchive& in, uint8_t version)
{
    if (version != 0) MRPT_THROW_UNKNOWN_SERIALIZATION_VERSION(version);
    float vs;
    in >> vs;
    setVoxelProperties(vs);  // clears
    insertionOptions.readFromStream(in);
    likelihoodOptions.readFromStream(in);
    renderOptions.readFromStream(in);
    // The reference re-files each stored point with VoxelData::insertPoint (no decimation):
    // one batch with the decimation options switched off is the same thing.
    const auto         nVoxels = in.ReadAs<uint32_t>();
    std::vector<float> xs, ys, zs;
    for (uint32_t i = 0; i < nVoxels; i++)
    {
        int32_t cx, cy, cz;
        in >> cx >> cy >> cz;
        const auto n = in.ReadAs<uint16_t>();
        for (uint16_t j = 0; j < n; j++)
        {
            mrpt::math::TPoint3Df pt;
            in >> pt;
            xs.push_back(pt.x), ys.push_back(pt.y), zs.push_back(pt.z);
        }
    }
    const TInsertionOptions keep = insertionOptions;
    insertionOptions.max_points_per_voxel        = 0;
    insertionOptions.min_distance_between_points = 0;
    insertPointCloud(xs.data(), ys.data(), zs.data(), xs.size(), std::nullopt, false);
    insertionOptions = keep;
    check(mmg_invalidate_bbox(h_));  // cached_.reset() (.cpp:176)
}

HashedVoxelPointCloud::TMapDefinition::TMapDefinition() = default;
void HashedVoxelPointCloud::TMapDefinition::loadFromConfigFile_map_specific(const mrpt::config::CConfigFileBase& s, const std::string& sectionPrefix)
{  // .cpp:61-77
    const std::string sSectCreation = sectionPrefix + std::string("_creationOpts");
    MRPT_LOAD_CONFIG_VAR(voxel_size, float, s, sSectCreation);
    ASSERT_(s.sectionExists(sectionPrefix + std::string("_insertionOpts")));
    insertionOpts.loadFromConfigFile(s, sectionPrefix + std::string("_insertionOpts"));
    if (s.sectionExists(sectionPrefix + std::string("_likelihoodOpts"))) likelihoodOpts.loadFromConfigFile(s, sectionPrefix + std::string("_likelihoodOpts"));
    if (s.sectionExists(sectionPrefix + std::string("_renderOpts"))) renderOpts.loadFromConfigFile(s, sectionPrefix + std::string("_renderOpts"));
}
void HashedVoxelPointCloud::TMapDefinition::dumpToTextStream_map_specific(std::ostream& out) const
{
    LOADABLEOPTS_DUMP_VAR(voxel_size, float);
    insertionOpts.dumpToTextStream(out);
    likelihoodOpts.dumpToTextStream(out);
    renderOpts.dumpToTextStream(out);
}
mrpt::maps::CMetricMap::Ptr HashedVoxelPointCloud::internal_CreateFromMapDefinition(const mrpt::maps::TMetricMapInitializer& _def)
{  // .cpp:91-105
    const auto* def = dynamic_cast<const HashedVoxelPointCloud::TMapDefinition*>(&_def);
    ASSERT_(def);
    auto obj               = HashedVoxelPointCloud::Create(def->voxel_size);
    obj->insertionOptions  = def->insertionOpts;
    obj->likelihoodOptions = def->likelihoodOpts;
    obj->renderOptions     = def->renderOpts;
    return obj;
}

// ============================================================== SparseVoxel
IMPLEMENTS_SERIALIZABLE(SparseVoxelPointCloud, CMetricMap, MOLA_B200_NAMESPACE)

SparseVoxelPointCloud::SparseVoxelPointCloud(float voxel_size) : GpuPointMapBase(MMG_SPARSE_VOXEL, voxel_size) {}
std::string SparseVoxelPointCloud::asString() const
{
    std::ostringstream s;
    s << "SparseVoxelPointCloud, resolution=" << voxelSize() << " voxels=" << numVoxels() << " points=" << numPoints();
    return s.str();
}
void SparseVoxelPointCloud::fillParams(mmg_params& p) const
{
    p.max_points_per_voxel = insertionOptions.max_points_per_voxel;
    p.sigma_dist           = likelihoodOptions.sigma_dist;
    p.max_corr_distance    = likelihoodOptions.max_corr_distance;
    p.decimation           = likelihoodOptions.decimation;
    p.match_mean           = likelihoodOptions.match_mean ? 1u : 0u;
}
bool SparseVoxelPointCloud::nn_single_search(const mrpt::math::TPoint3Df& q, mrpt::math::TPoint3Df& r, float& d2, uint64_t& id) const
{  // SparseVoxelPointCloud.cpp:695-719
    if (!enable_knn_extension) THROW_EXCEPTION("n-search not available in this map type");
    return GpuPointMapBase::nn_single_search(q, r, d2, id);
}
void SparseVoxelPointCloud::nn_multiple_search(const mrpt::math::TPoint3Df& q, const size_t N, std::vector<mrpt::math::TPoint3Df>& r,
                                               std::vector<float>& d2, std::vector<uint64_t>& ids) const
{
    if (!enable_knn_extension) THROW_EXCEPTION("n-search not available in this map type");
    GpuPointMapBase::nn_multiple_search(q, N, r, d2, ids);
}
void SparseVoxelPointCloud::voxelStats(std::vector<mrpt::math::TPoint3Df>& mean, std::vector<float>& cov6) const
{
    const size_t V = numVoxels();
    mean.assign(V, mrpt::math::TPoint3Df(0, 0, 0)), cov6.assign(6 * V, 0.f);
    if (V) check(mmg_voxel_stats(h_, reinterpret_cast<float*>(mean.data()), cov6.data(), MMG_MEM_HOST, nullptr));
}
void SparseVoxelPointCloud::TInsertionOptions::loadFromConfigFile(const mrpt::config::CConfigFileBase& c, const std::string& s)
{
    MRPT_LOAD_CONFIG_VAR(max_points_per_voxel, int, c, s);  // SparseVoxelPointCloud.cpp:998
}
void SparseVoxelPointCloud::TInsertionOptions::dumpToTextStream(std::ostream& out) const
{
    out << "\n------ [SparseVoxelPointCloud::TInsertionOptions] ------- \n\n";
    LOADABLEOPTS_DUMP_VAR(max_points_per_voxel, int);
}
void SparseVoxelPointCloud::TLikelihoodOptions::loadFromConfigFile(const mrpt::config::CConfigFileBase& c, const std::string& s)
{
    MRPT_LOAD_CONFIG_VAR(sigma_dist, double, c, s);
    MRPT_LOAD_CONFIG_VAR(max_corr_distance, double, c, s);
    MRPT_LOAD_CONFIG_VAR(decimation, int, c, s);
    MRPT_LOAD_CONFIG_VAR(match_mean, bool, c, s);
}
void SparseVoxelPointCloud::TLikelihoodOptions::dumpToTextStream(std::ostream& out) const
{
    out << "\n------ [SparseVoxelPointCloud::TLikelihoodOptions] ------- \n\n";
    LOADABLEOPTS_DUMP_VAR(sigma_dist, double);
    LOADABLEOPTS_DUMP_VAR(max_corr_distance, double);
    LOADABLEOPTS_DUMP_VAR(decimation, int);
    LOADABLEOPTS_DUMP_VAR(match_mean, bool);
}
// Option blocks: SparseVoxelPointCloud.cpp:880-960.
static void check_opts_version(int8_t version)
{
    if (version != 0) MRPT_THROW_UNKNOWN_SERIALIZATION_VERSION(version);
}
void SparseVoxelPointCloud::TInsertionOptions::writeToStream(mrpt::serialization::CArchive& out) const
{
    const int8_t version = 0;
    out << version << max_points_per_voxel;
}
void SparseVoxelPointCloud::TInsertionOptions::readFromStream(mrpt::serialization::CArchive& in)
{
    int8_t version;
    in >> version;
    check_opts_version(version);
    in >> max_points_per_voxel;
}
void SparseVoxelPointCloud::TLikelihoodOptions::writeToStream(mrpt::serialization::CArchive& out) const
{
    const int8_t version = 0;
    out << version << sigma_dist << max_corr_distance << decimation << match_mean;
}
void SparseVoxelPointCloud::TLikelihoodOptions::readFromStream(mrpt::serialization::CArchive& in)
{
    int8_t version;
    in >> version;
    check_opts_version(version);
    in >> sigma_dist >> max_corr_distance >> decimation >> match_mean;
}
void SparseVoxelPointCloud::TRenderOptions::writeToStream(mrpt::serialization::CArchive& out) const
{
    const int8_t version = 0;
    out << version << point_size << show_mean_only << show_inner_grid_boxes << color << colormap << recolorizeByCoordinateIndex;
}
void SparseVoxelPointCloud::TRenderOptions::readFromStream(mrpt::serialization::CArchive& in)
{
    int8_t version;
    in >> version;
    check_opts_version(version);
    in >> point_size >> show_mean_only >> show_inner_grid_boxes >> color >> colormap >> recolorizeByCoordinateIndex;
}

// Serialization v0 in the reference's layout (SparseVoxelPointCloud.cpp:111-189): inner-grid bit count, voxel size, the
// three option blocks, then per 32^3 block its outer index, the block's points as an embedded CSimplePointsMap object
// (written by MRPT's own operator<<, so the point-map format is whatever the linked MRPT defines) and, for each of the
// 32768 cells, a uint16 count and the uint32 indices of the cell's points inside that block. The reference keeps a
// block's points in insertion order across cells; the device map only knows the order inside a voxel, so blocks are
// written cell by cell (a reader gets the same voxels with the same per-voxel point order). Blocks go out in
// lexicographic (cx, cy, cz) order.
uint8_t SparseVoxelPointCloud::serializeGetVersion() const { return 0; }
void    SparseVoxelPointCloud::serializeTo(mrpt::serialization::CArchive& out) const
{
    constexpr uint32_t INNER_BITS = 5, INNER_MASK = (1u << INNER_BITS) - 1u, CELLS = 1u << (3 * INNER_BITS);
    out << INNER_BITS << voxelSize();
    insertionOptions.writeToStream(out);
    likelihoodOptions.writeToStream(out);
    renderOptions.writeToStream(out);
    std::vector<int32_t>               k;
    std::vector<uint32_t>              c;
    std::vector<mrpt::math::TPoint3Df> p;
    uint32_t                           cap = 0;
    exportVoxels(k, c, p, cap);
    // outer index (g2o: idx & ~31, SparseVoxelPointCloud.h:84-90) -> voxels of the block as (plain inner index, voxel)
    std::map<std::array<int32_t, 3>, std::vector<std::pair<uint32_t, uint32_t>>> blocks;
    for (size_t v = 0; v < c.size(); v++)
    {
        if (!c[v]) continue;
        const int32_t  gx = k[3 * v], gy = k[3 * v + 1], gz = k[3 * v + 2];
        const uint32_t plain = ((uint32_t)gx & INNER_MASK) | (((uint32_t)gy & INNER_MASK) << INNER_BITS) | (((uint32_t)gz & INNER_MASK) << (2 * INNER_BITS));
        blocks[{gx & ~(int32_t)INNER_MASK, gy & ~(int32_t)INNER_MASK, gz & ~(int32_t)INNER_MASK}].emplace_back(plain, (uint32_t)v);
    }
    out.WriteAs<uint32_t>(blocks.size());
    std::vector<uint32_t> first(CELLS), count(CELLS);
    for (auto& kv : blocks)
    {
        out << kv.first[0] << kv.first[1] << kv.first[2];
        std::sort(kv.second.begin(), kv.second.end());
        mrpt::maps::CSimplePointsMap pts;
        std::fill(count.begin(), count.end(), 0u);
        uint32_t next = 0;
        for (const auto& cell : kv.second)
        {
            first[cell.first] = next, count[cell.first] = c[cell.second];
            for (uint32_t j = 0; j < c[cell.second]; j++)
            {
                const auto& q = p[(size_t)cell.second * cap + j];
                pts.insertPointFast(q.x, q.y, q.z);
            }
            next += c[cell.second];
        }
        out << pts;
        for (uint32_t i = 0; i < CELLS; i++)
        {
            out.WriteAs<uint16_t>(count[i]);
            for (uint32_t j = 0; j < count[i]; j++) out.WriteAs<uint32_t>(first[i] + j);
        }
    }
}
void SparseVoxelPointCloud::serializeFrom(mrpt::serialization::CArchive& in, uint8_t version)
{
    if (version != 0) MRPT_THROW_UNKNOWN_SERIALIZATION_VERSION(version);
    constexpr uint32_t INNER_BITS = 5, CELLS = 1u << (3 * INNER_BITS);
    const auto         bits = in.ReadAs<uint32_t>();
    ASSERT_EQUAL_(bits, INNER_BITS);
    float vs;
    in >> vs;
    setVoxelProperties(vs);  // clears
    insertionOptions.readFromStream(in);
    likelihoodOptions.readFromStream(in);
    renderOptions.readFromStream(in);
    // The reference restores each cell's index list verbatim (and leaves the cells' running means at zero, .cpp:166-181);
    // here the cells' points are re-filed in list order with the cap switched off, which gives the same voxels and
    // per-voxel order, with the means recomputed.
    const auto         nGrids = in.ReadAs<uint32_t>();
    std::vector<float> xs, ys, zs;
    for (uint32_t g = 0; g < nGrids; g++)
    {
        int32_t ox, oy, oz;
        in >> ox >> oy >> oz;
        mrpt::maps::CSimplePointsMap pts;
        in >> pts;
        for (uint32_t i = 0; i < CELLS; i++)
        {
            const auto n = in.ReadAs<uint16_t>();
            for (uint16_t j = 0; j < n; j++)
            {
                const auto idx = in.ReadAs<uint32_t>();
                ASSERT_LT_(idx, pts.size());
                float x, y, z;
                pts.getPointFast(idx, x, y, z);
                xs.push_back(x), ys.push_back(y), zs.push_back(z);
            }
        }
    }
    const TInsertionOptions keep = insertionOptions;
    insertionOptions.max_points_per_voxel = 0;
    insertPointCloud(xs.data(), ys.data(), zs.data(), xs.size(), std::nullopt, false);
    insertionOptions = keep;
    check(mmg_invalidate_bbox(h_));  // cached_.reset() (.cpp:188)
}
SparseVoxelPointCloud::TMapDefinition::TMapDefinition() = default;
void SparseVoxelPointCloud::TMapDefinition::loadFromConfigFile_map_specific(const mrpt::config::CConfigFileBase& s, const std::string& sectionPrefix)
{
    const std::string sSectCreation = sectionPrefix + std::string("_creationOpts");
    MRPT_LOAD_CONFIG_VAR(voxel_size, float, s, sSectCreation);
    ASSERT_(s.sectionExists(sectionPrefix + std::string("_insertionOpts")));
    insertionOpts.loadFromConfigFile(s, sectionPrefix + std::string("_insertionOpts"));
    if (s.sectionExists(sectionPrefix + std::string("_likelihoodOpts"))) likelihoodOpts.loadFromConfigFile(s, sectionPrefix + std::string("_likelihoodOpts"));
}
void SparseVoxelPointCloud::TMapDefinition::dumpToTextStream_map_specific(std::ostream& out) const
{
    LOADABLEOPTS_DUMP_VAR(voxel_size, float);
    insertionOpts.dumpToTextStream(out);
    likelihoodOpts.dumpToTextStream(out);
}
mrpt::maps::CMetricMap::Ptr SparseVoxelPointCloud::internal_CreateFromMapDefinition(const mrpt::maps::TMetricMapInitializer& _def)
{
    const auto* def = dynamic_cast<const SparseVoxelPointCloud::TMapDefinition*>(&_def);
    ASSERT_(def);
    auto obj               = SparseVoxelPointCloud::Create(def->voxel_size);
    obj->insertionOptions  = def->insertionOpts;
    obj->likelihoodOptions = def->likelihoodOpts;
    return obj;
}

// ============================================================== SparseTrees
IMPLEMENTS_SERIALIZABLE(SparseTreesPointCloud, CMetricMap, MOLA_B200_NAMESPACE)

SparseTreesPointCloud::SparseTreesPointCloud(float grid_size) : GpuPointMapBase(MMG_SPARSE_TREES, grid_size) {}
std::string SparseTreesPointCloud::asString() const
{
    std::ostringstream s;
    s << "SparseTreesPointCloud, grid_size=" << gridSize() << " blocks=" << numVoxels() << " points=" << numPoints();
    return s.str();
}
void SparseTreesPointCloud::fillParams(mmg_params& p) const
{
    p.minimum_points_clearance   = insertionOptions.minimum_points_clearance;
    p.remove_voxels_farther_than = insertionOptions.remove_submaps_farther_than;
    p.sigma_dist                 = likelihoodOptions.sigma_dist;
    p.max_corr_distance          = likelihoodOptions.max_corr_distance;
    p.decimation                 = likelihoodOptions.decimation;
}
void SparseTreesPointCloud::TInsertionOptions::loadFromConfigFile(const mrpt::config::CConfigFileBase& c, const std::string& s)
{
    MRPT_LOAD_CONFIG_VAR(minimum_points_clearance, double, c, s);
    // read as uint64_t into a float member: fractions are truncated (SparseTreesPointCloud.cpp:893)
    MRPT_LOAD_CONFIG_VAR(remove_submaps_farther_than, uint64_t, c, s);
}
void SparseTreesPointCloud::TInsertionOptions::dumpToTextStream(std::ostream& out) const
{
    out << "\n------ [SparseTreesPointCloud::TInsertionOptions] ------- \n\n";
    LOADABLEOPTS_DUMP_VAR(minimum_points_clearance, double);
    LOADABLEOPTS_DUMP_VAR(remove_submaps_farther_than, double);
}
void SparseTreesPointCloud::TLikelihoodOptions::loadFromConfigFile(const mrpt::config::CConfigFileBase& c, const std::string& s)
{
    MRPT_LOAD_CONFIG_VAR(sigma_dist, double, c, s);
    MRPT_LOAD_CONFIG_VAR(max_corr_distance, double, c, s);
    MRPT_LOAD_CONFIG_VAR(decimation, int, c, s);
}
void SparseTreesPointCloud::TLikelihoodOptions::dumpToTextStream(std::ostream& out) const
{
    out << "\n------ [SparseTreesPointCloud::TLikelihoodOptions] ------- \n\n";
    LOADABLEOPTS_DUMP_VAR(sigma_dist, double);
    LOADABLEOPTS_DUMP_VAR(max_corr_distance, double);
    LOADABLEOPTS_DUMP_VAR(decimation, int);
}
// Option blocks: SparseTreesPointCloud.cpp:770-850.
void SparseTreesPointCloud::TInsertionOptions::writeToStream(mrpt::serialization::CArchive& out) const
{
    const int8_t version = 0;
    out << version << minimum_points_clearance << remove_submaps_farther_than;
}
void SparseTreesPointCloud::TInsertionOptions::readFromStream(mrpt::serialization::CArchive& in)
{
    int8_t version;
    in >> version;
    check_opts_version(version);
    in >> minimum_points_clearance >> remove_submaps_farther_than;
}
void SparseTreesPointCloud::TLikelihoodOptions::writeToStream(mrpt::serialization::CArchive& out) const
{
    const int8_t version = 0;
    out << version << sigma_dist << max_corr_distance << decimation;
}
void SparseTreesPointCloud::TLikelihoodOptions::readFromStream(mrpt::serialization::CArchive& in)
{
    int8_t version;
    in >> version;
    check_opts_version(version);
    in >> sigma_dist >> max_corr_distance >> decimation;
}
void SparseTreesPointCloud::TRenderOptions::writeToStream(mrpt::serialization::CArchive& out) const
{
    const int8_t version = 0;
    out << version << point_size << show_inner_grid_boxes << color << colormap << recolorizeByCoordinateIndex;
}
void SparseTreesPointCloud::TRenderOptions::readFromStream(mrpt::serialization::CArchive& in)
{
    int8_t version;
    in >> version;
    check_opts_version(version);
    in >> point_size >> show_inner_grid_boxes >> color >> colormap >> recolorizeByCoordinateIndex;
}

// Serialization v0 in the reference's layout (SparseTreesPointCloud.cpp:111-167): the id bit count (uint8 20), grid size,
// the three option blocks, then per block its index and its points as an embedded CSimplePointsMap object in
// block-local index order (the ids nn_*_search returns survive the round trip). Blocks go out in lexicographic order.
uint8_t SparseTreesPointCloud::serializeGetVersion() const { return 0; }
void    SparseTreesPointCloud::serializeTo(mrpt::serialization::CArchive& out) const
{
    const uint8_t id_bits = 20;  // GLOBAL_ID_SUBVOXEL_BITCOUNT, SparseTreesPointCloud.h:58
    out << id_bits << gridSize();
    insertionOptions.writeToStream(out);
    likelihoodOptions.writeToStream(out);
    renderOptions.writeToStream(out);
    std::vector<int32_t>               k;
    std::vector<uint32_t>              c;
    std::vector<mrpt::math::TPoint3Df> p;
    uint32_t                           cap = 0;
    exportVoxels(k, c, p, cap);
    std::map<std::array<int32_t, 3>, size_t> blocks;
    for (size_t v = 0; v < c.size(); v++) blocks[{k[3 * v], k[3 * v + 1], k[3 * v + 2]}] = v;
    out.WriteAs<uint32_t>(blocks.size());
    for (const auto& kv : blocks)
    {
        out << kv.first[0] << kv.first[1] << kv.first[2];
        mrpt::maps::CSimplePointsMap pts;
        pts.reserve(c[kv.second]);
        for (uint32_t j = 0; j < c[kv.second]; j++)
        {
            const auto& q = p[kv.second * (size_t)cap + j];
            pts.insertPointFast(q.x, q.y, q.z);
        }
        out << pts;
    }
}
void SparseTreesPointCloud::serializeFrom(mrpt::serialization::CArchive& in, uint8_t version)
{
    if (version != 0) MRPT_THROW_UNKNOWN_SERIALIZATION_VERSION(version);
    const auto bits = in.ReadAs<uint8_t>();
    ASSERT_EQUAL_(bits, 20);
    float gs;
    in >> gs;
    setGridProperties(gs);  // clears
    insertionOptions.readFromStream(in);
    likelihoodOptions.readFromStream(in);
    renderOptions.readFromStream(in);
    // All blocks go in as ONE batch, block after block and each block's points in stored order: the clearance test only
    // looks at the map as it was before the batch (empty here, .cpp:916-959), so every point is accepted, and a block's
    // local indices are assigned in input order, so the ids nn_*_search returns are preserved. The archive's block
    // index is implied by the points themselves (coordToOuterIdx, .h:70-77).
    const auto         nGrids = in.ReadAs<uint32_t>();
    std::vector<float> xs, ys, zs;
    for (uint32_t g = 0; g < nGrids; g++)
    {
        int32_t ox, oy, oz;
        in >> ox >> oy >> oz;
        mrpt::maps::CSimplePointsMap pts;
        in >> pts;
        for (size_t i = 0; i < pts.size(); i++)
        {
            float x, y, z;
            pts.getPointFast(i, x, y, z);
            xs.push_back(x), ys.push_back(y), zs.push_back(z);
        }
    }
    insertPointCloud(xs.data(), ys.data(), zs.data(), xs.size(), std::nullopt, false);
    check(mmg_invalidate_bbox(h_));  // cached_.reset() (.cpp:166)
}
SparseTreesPointCloud::TMapDefinition::TMapDefinition() = default;
void SparseTreesPointCloud::TMapDefinition::loadFromConfigFile_map_specific(const mrpt::config::CConfigFileBase& s, const std::string& sectionPrefix)
{
    const std::string sSectCreation = sectionPrefix + std::string("_creationOpts");
    MRPT_LOAD_CONFIG_VAR(grid_size, float, s, sSectCreation);
    ASSERT_(s.sectionExists(sectionPrefix + std::string("_insertionOpts")));
    insertionOpts.loadFromConfigFile(s, sectionPrefix + std::string("_insertionOpts"));
    if (s.sectionExists(sectionPrefix + std::string("_likelihoodOpts"))) likelihoodOpts.loadFromConfigFile(s, sectionPrefix + std::string("_likelihoodOpts"));
}
void SparseTreesPointCloud::TMapDefinition::dumpToTextStream_map_specific(std::ostream& out) const
{
    LOADABLEOPTS_DUMP_VAR(grid_size, float);
    insertionOpts.dumpToTextStream(out);
    likelihoodOpts.dumpToTextStream(out);
}
mrpt::maps::CMetricMap::Ptr SparseTreesPointCloud::internal_CreateFromMapDefinition(const mrpt::maps::TMetricMapInitializer& _def)
{
    const auto* def = dynamic_cast<const SparseTreesPointCloud::TMapDefinition*>(&_def);
    ASSERT_(def);
    auto obj               = SparseTreesPointCloud::Create(def->grid_size);
    obj->insertionOptions  = def->insertionOpts;
    obj->likelihoodOptions = def->likelihoodOpts;
    return obj;
}
}  // namespace MOLA_B200_NAMESPACE
