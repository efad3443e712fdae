// mrpt_min_part3.h — opengl stubs, observations, maps (see mrpt_min.h)
#pragma once

namespace mrpt
{
namespace system
{
namespace os
{
inline FILE* fopen(const char* f, const char* m) { return ::fopen(f, m); }
inline void  fclose(FILE* f) { ::fclose(f); }
template <typename... A>
inline int fprintf(FILE* f, const char* fmt, A... a)
{
    return ::fprintf(f, fmt, a...);
}
}  // namespace os
}  // namespace system

// -------------------------------------------------------------------- opengl
// Scene-graph objects are CPU-side containers only (no graphics stack here).
namespace opengl
{
class CRenderizable
{
   public:
    using Ptr = std::shared_ptr<CRenderizable>;
    virtual ~CRenderizable() = default;
    void setColor(const mrpt::img::TColorf&) {}
    void setPointSize(float) {}
    void enableColorFromZ(bool) {}
};
class CPointCloud : public CRenderizable
{
   public:
    using Ptr = std::shared_ptr<CPointCloud>;
    static Ptr Create() { return std::make_shared<CPointCloud>(); }
    void       insertPoint(const mrpt::math::TPoint3Df& p) { pts.push_back(p); }
    std::vector<mrpt::math::TPoint3Df> pts;
};
struct TPointXYZfRGBAu8
{
    float   x, y, z;
    uint8_t r = 0, g = 0, b = 0, a = 255;
    TPointXYZfRGBAu8(float X, float Y, float Z, uint8_t R, uint8_t G, uint8_t B, uint8_t A = 255) : x(X), y(Y), z(Z), r(R), g(G), b(B), a(A) {}
};
class CPointCloudColoured : public CRenderizable
{
   public:
    using Ptr = std::shared_ptr<CPointCloudColoured>;
    static Ptr Create() { return std::make_shared<CPointCloudColoured>(); }
    void       insertPoint(const TPointXYZfRGBAu8& p) { pts.push_back(p); }
    void       recolorizeByCoordinate(float, float, int, mrpt::img::TColormap) {}
    std::vector<TPointXYZfRGBAu8> pts;
};
class CBox : public CRenderizable
{
   public:
    using Ptr = std::shared_ptr<CBox>;
    static Ptr Create() { return std::make_shared<CBox>(); }
    template <typename A, typename B>
    void setBoxCorners(const A&, const B&)
    {
    }
    void setWireframe(bool) {}
    void setLineWidth(float) {}
};
class CSetOfObjects : public CRenderizable
{
   public:
    using Ptr = std::shared_ptr<CSetOfObjects>;
    static Ptr Create() { return std::make_shared<CSetOfObjects>(); }
    template <typename P>
    void insert(const P& o)
    {
        objs.push_back(o);
    }
    std::vector<CRenderizable::Ptr> objs;
};
}  // namespace opengl

namespace maps
{
class CPointsMap;
class CSimplePointsMap;
}  // namespace maps

// ----------------------------------------------------------------------- obs
namespace obs
{
class CObservation;  // defined after the maps (the 2-D scan builds a CSimplePointsMap)
}  // namespace obs

// ---------------------------------------------------------------------- maps
namespace maps
{
class NearestNeighborsCapable
{
   public:
    virtual ~NearestNeighborsCapable()                = default;
    [[nodiscard]] virtual bool   nn_has_indices_or_ids() const = 0;
    [[nodiscard]] virtual size_t nn_index_count() const        = 0;
    virtual void                 nn_prepare_for_2d_queries() const {}
    virtual void                 nn_prepare_for_3d_queries() const {}
    [[nodiscard]] virtual bool nn_single_search(const mrpt::math::TPoint3Df& query, mrpt::math::TPoint3Df& result, float& out_dist_sqr,
                                                uint64_t& resultIndexOrIDOrID) const = 0;
    [[nodiscard]] virtual bool nn_single_search(const mrpt::math::TPoint2Df& query, mrpt::math::TPoint2Df& result, float& out_dist_sqr,
                                                uint64_t& resultIndexOrIDOrID) const = 0;
    virtual void nn_multiple_search(const mrpt::math::TPoint3Df& query, const size_t N, std::vector<mrpt::math::TPoint3Df>& results,
                                    std::vector<float>& out_dists_sqr, std::vector<uint64_t>& resultIndicesOrIDs) const = 0;
    virtual void nn_multiple_search(const mrpt::math::TPoint2Df& query, const size_t N, std::vector<mrpt::math::TPoint2Df>& results,
                                    std::vector<float>& out_dists_sqr, std::vector<uint64_t>& resultIndicesOrIDs) const = 0;
    virtual void nn_radius_search(const mrpt::math::TPoint3Df& query, const float search_radius_sqr,
                                  std::vector<mrpt::math::TPoint3Df>& results, std::vector<float>& out_dists_sqr,
                                  std::vector<uint64_t>& resultIndicesOrIDs, size_t maxPoints = 0) const = 0;
    virtual void nn_radius_search(const mrpt::math::TPoint2Df& query, const float search_radius_sqr,
                                  std::vector<mrpt::math::TPoint2Df>& results, std::vector<float>& out_dists_sqr,
                                  std::vector<uint64_t>& resultIndicesOrIDs, size_t maxPoints = 0) const = 0;
};

struct TMetricMapInitializer : public mrpt::config::CLoadableOptions
{
    virtual ~TMetricMapInitializer() = default;
    void loadFromConfigFile(const mrpt::config::CConfigFileBase& s, const std::string& prefix) override
    {
        loadFromConfigFile_map_specific(s, prefix);
    }
    void dumpToTextStream(std::ostream& o) const override { dumpToTextStream_map_specific(o); }
    virtual void loadFromConfigFile_map_specific(const mrpt::config::CConfigFileBase&, const std::string&) = 0;
    virtual void dumpToTextStream_map_specific(std::ostream&) const                                          = 0;
};

class CMetricMap : public mrpt::serialization::CSerializable
{
   public:
    using Ptr = std::shared_ptr<CMetricMap>;
    struct TGenericParams
    {
        bool enableSaveAs3DObject = true;
    } genericMapParams;
    void clear() { internal_clear(); }
    bool insertObservation(const mrpt::obs::CObservation& obs, const std::optional<const mrpt::poses::CPose3D>& robotPose = std::nullopt)
    {
        return internal_insertObservation(obs, robotPose);
    }
    double computeObservationLikelihood(const mrpt::obs::CObservation& obs, const mrpt::poses::CPose3D& takenFrom) const
    {
        return internal_computeObservationLikelihood(obs, takenFrom);
    }
    bool canComputeObservationLikelihood(const mrpt::obs::CObservation& obs) const { return internal_canComputeObservationLikelihood(obs); }
    virtual bool                          isEmpty() const = 0;
    virtual mrpt::math::TBoundingBoxf     boundingBox() const { return {}; }
    virtual std::string                   asString() const = 0;
    virtual void                          getVisualizationInto(mrpt::opengl::CSetOfObjects&) const {}
    virtual void                          saveMetricMapRepresentationToFile(const std::string&) const {}
    virtual const CSimplePointsMap*       getAsSimplePointsMap() const { return nullptr; }

   protected:
    virtual void   internal_clear() = 0;
    virtual bool   internal_insertObservation(const mrpt::obs::CObservation&, const std::optional<const mrpt::poses::CPose3D>&) = 0;
    virtual double internal_computeObservationLikelihood(const mrpt::obs::CObservation&, const mrpt::poses::CPose3D&) const { return 0; }
    virtual bool   internal_canComputeObservationLikelihood(const mrpt::obs::CObservation&) const { return false; }
};

// SoA point store with exact brute-force neighbour queries (nanoflann stand-in).
class CPointsMap : public CMetricMap
{
   public:
    size_t size() const { return x_.size(); }
    bool   empty() const { return x_.empty(); }
    bool   isEmpty() const override { return x_.empty(); }
    std::string asString() const override { return "CPointsMap"; }
    const std::vector<float>& getPointsBufferRef_x() const { return x_; }
    const std::vector<float>& getPointsBufferRef_y() const { return y_; }
    const std::vector<float>& getPointsBufferRef_z() const { return z_; }
    void insertPoint(float x, float y, float z) { x_.push_back(x), y_.push_back(y), z_.push_back(z); }
    void insertPoint(const mrpt::math::TPoint3Df& p) { insertPoint(p.x, p.y, p.z); }
    void insertPointFast(float x, float y, float z) { insertPoint(x, y, z); }
    void getPointFast(size_t i, float& x, float& y, float& z) const { x = x_[i], y = y_[i], z = z_[i]; }
    void getPoint(size_t i, mrpt::math::TPoint3D& p) const { p = {x_[i], y_[i], z_[i]}; }
    void reserve(size_t n) { x_.reserve(n), y_.reserve(n), z_.reserve(n); }
    void nn_prepare_for_2d_queries() const {}
    void nn_prepare_for_3d_queries() const {}

    float d2(size_t i, const mrpt::math::TPoint3Df& q) const
    {  // nanoflann L2_Simple_Adaptor: sum over dims of (a-b)^2, accumulated left to right
        const float dx = q.x - x_[i], dy = q.y - y_[i], dz = q.z - z_[i];
        return dx * dx + dy * dy + dz * dz;
    }
    bool nn_single_search(const mrpt::math::TPoint3Df& q, mrpt::math::TPoint3Df& r, float& od, uint64_t& id) const
    {
        if (x_.empty()) return false;
        size_t bi = 0;
        float  bd = d2(0, q);
        for (size_t i = 1; i < x_.size(); i++)
        {
            const float d = d2(i, q);
            if (d < bd) bd = d, bi = i;
        }
        r  = {x_[bi], y_[bi], z_[bi]};
        od = bd, id = bi;
        return true;
    }
    void sorted_hits(const mrpt::math::TPoint3Df& q, float r2, bool use_r, std::vector<std::pair<float, size_t>>& h) const
    {
        h.clear();
        for (size_t i = 0; i < x_.size(); i++)
        {
            const float d = d2(i, q);
            if (!use_r || d < r2) h.emplace_back(d, i);
        }
        std::sort(h.begin(), h.end());
    }
    void emit(const std::vector<std::pair<float, size_t>>& h, size_t n, std::vector<mrpt::math::TPoint3Df>& r, std::vector<float>& d,
              std::vector<uint64_t>& ids) const
    {
        r.clear(), d.clear(), ids.clear();
        for (size_t i = 0; i < n; i++)
        {
            r.push_back({x_[h[i].second], y_[h[i].second], z_[h[i].second]});
            d.push_back(h[i].first);
            ids.push_back(h[i].second);
        }
    }
    void nn_multiple_search(const mrpt::math::TPoint3Df& q, size_t N, std::vector<mrpt::math::TPoint3Df>& r, std::vector<float>& d,
                            std::vector<uint64_t>& ids) const
    {
        std::vector<std::pair<float, size_t>> h;
        sorted_hits(q, 0, false, h);
        emit(h, std::min(N, h.size()), r, d, ids);
    }
    void nn_radius_search(const mrpt::math::TPoint3Df& q, float r2, std::vector<mrpt::math::TPoint3Df>& r, std::vector<float>& d,
                          std::vector<uint64_t>& ids, size_t maxPoints = 0) const
    {
        std::vector<std::pair<float, size_t>> h;
        sorted_hits(q, r2, true, h);
        emit(h, maxPoints ? std::min(maxPoints, h.size()) : h.size(), r, d, ids);
    }
    // CSerializable: u32 count then x,y,z arrays
    uint8_t serializeGetVersion() const override { return 0; }
    void    serializeTo(mrpt::serialization::CArchive& o) const override
    {
        o.WriteAs<uint32_t>(x_.size());
        if (!x_.empty()) o.write(x_.data(), 4 * x_.size()), o.write(y_.data(), 4 * y_.size()), o.write(z_.data(), 4 * z_.size());
    }
    void serializeFrom(mrpt::serialization::CArchive& in, uint8_t) override
    {
        const uint32_t n = in.ReadAs<uint32_t>();
        x_.resize(n), y_.resize(n), z_.resize(n);
        if (n) in.read(x_.data(), 4 * n), in.read(y_.data(), 4 * n), in.read(z_.data(), 4 * n);
    }

   protected:
    void internal_clear() override { x_.clear(), y_.clear(), z_.clear(); }
    // 2-D range scans only (what the reference's tests insert into a CSimplePointsMap); defined after the observations
    bool internal_insertObservation(const mrpt::obs::CObservation&, const std::optional<const mrpt::poses::CPose3D>&) override;
    std::vector<float> x_, y_, z_;
};
class CSimplePointsMap : public CPointsMap
{
   public:
    using Ptr = std::shared_ptr<CSimplePointsMap>;
    static Ptr  Create() { return std::make_shared<CSimplePointsMap>(); }
    const char* className() const override { return "CSimplePointsMap"; }
};
}  // namespace maps

// ----------------------------------------------------------------------- obs
// Observation classes with the members the map classes touch, shaped like upstream MRPT 2.x: `sensorPose` is a member
// of every DERIVED class (the base only has the virtual getSensorPose()), a 2-D scan carries ranges + validity and
// builds its auxiliary points map in the ROBOT frame (sensor pose applied, like CPointsMap::loadFromRangeScan without a
// robot pose), a 3-D range scan unprojects its depth image with a pinhole model and, when asked, applies the sensor
// pose, a Velodyne scan generates its cloud from (azimuth, elevation, range) returns in the sensor frame. The
// arithmetic of these conversions is this stand-in's own (double composePoint, float sin/cos): unpinned against
// upstream MRPT, but shared by the reference classes and the GPU classes, which is what the parity tests need.
namespace obs
{
class CObservation
{
   public:
    using Ptr = std::shared_ptr<CObservation>;
    virtual ~CObservation()                                     = default;
    virtual void getSensorPose(mrpt::poses::CPose3D& out) const = 0;
};
struct T3DPointsProjectionParams
{
    bool takeIntoAccountSensorPoseOnRobot = false;
};
class CObservationPointCloud : public CObservation
{
   public:
    std::shared_ptr<mrpt::maps::CPointsMap> pointcloud;
    mrpt::poses::CPose3D                    sensorPose;
    void getSensorPose(mrpt::poses::CPose3D& out) const override { out = sensorPose; }
};
class CObservation2DRangeScan : public CObservation
{
   public:
    float                aperture    = 3.14159265358979f;  // field of view, radians
    bool                 rightToLeft = true;               // scan direction (counter-clockwise seen from above)
    float                maxRange    = 80.0f;
    mrpt::poses::CPose3D sensorPose;
    void getSensorPose(mrpt::poses::CPose3D& out) const override { out = sensorPose; }
    void resizeScan(size_t n) { ranges_.assign(n, 0.f), valid_.assign(n, 0), aux_.reset(); }
    size_t getScanSize() const { return ranges_.size(); }
    float  getScanRange(size_t i) const { return ranges_[i]; }
    bool   getScanRangeValidity(size_t i) const { return valid_[i] != 0; }
    void   setScanRange(size_t i, float r) { ranges_[i] = r, aux_.reset(); }
    void   setScanRangeValidity(size_t i, bool v) { valid_[i] = v ? 1 : 0, aux_.reset(); }
    /** The scan as points in the robot frame (cached), CObservation2DRangeScan::buildAuxPointsMap. */
    template <class M>
    const M* buildAuxPointsMap(const void* = nullptr) const
    {
        if (!aux_)
        {
            auto         pm = mrpt::maps::CSimplePointsMap::Create();
            const size_t n  = ranges_.size();
            const float  a0 = rightToLeft ? -0.5f * aperture : 0.5f * aperture;
            const float  da = n > 1 ? (rightToLeft ? aperture : -aperture) / (float)(n - 1) : 0.f;
            for (size_t i = 0; i < n; i++)
            {
                if (!valid_[i]) continue;
                const float a = a0 + da * (float)i, r = ranges_[i];
                pm->insertPoint(sensorPose.composePoint(mrpt::math::TPoint3Df(r * std::cos(a), r * std::sin(a), 0.f)));
            }
            aux_ = pm;
        }
        return static_cast<const M*>(aux_.get());
    }

   private:
    std::vector<float>                                     ranges_;
    std::vector<uint8_t>                                   valid_;
    mutable std::shared_ptr<mrpt::maps::CSimplePointsMap> aux_;
};
class CObservation3DRangeScan : public CObservation
{
   public:
    bool               hasPoints3D = false, hasRangeImage = false;
    std::vector<float> points3D_x, points3D_y, points3D_z;  // sensor frame
    // depth image (metres along the sensor's +X axis, 0 = no return) and its pinhole intrinsics
    size_t             rangeRows = 0, rangeCols = 0;
    std::vector<float> rangeImage;
    struct
    {
        float fx = 1, fy = 1, cx = 0, cy = 0;
    } cameraParams;
    mrpt::poses::CPose3D sensorPose;
    void getSensorPose(mrpt::poses::CPose3D& out) const override { out = sensorPose; }
    /** x = depth, y = (cx - col) * depth / fx, z = (cy - row) * depth / fy; optionally moved to the robot frame. */
    template <class M>
    void unprojectInto(M& dest, const T3DPointsProjectionParams& pp)
    {
        for (size_t r = 0; r < rangeRows; r++)
            for (size_t c = 0; c < rangeCols; c++)
            {
                const float d = rangeImage[r * rangeCols + c];
                if (!(d > 0.f)) continue;
                const mrpt::math::TPoint3Df l(d, (cameraParams.cx - (float)c) * d / cameraParams.fx, (cameraParams.cy - (float)r) * d / cameraParams.fy);
                if (pp.takeIntoAccountSensorPoseOnRobot) dest.insertPoint(sensorPose.composePoint(l));
                else dest.insertPoint(l);
            }
    }
};
class CObservationVelodyneScan : public CObservation
{
   public:
    struct TPointCloud
    {
        std::vector<float> x, y, z;
        size_t             size() const { return x.size(); }
        void               clear() { x.clear(), y.clear(), z.clear(); }
    } point_cloud;
    /** Raw returns (stand-in for scan_packets): azimuth, elevation (radians) and range (metres; <= 0 = none). */
    std::vector<float>   raw_azimuth, raw_elevation, raw_range;
    mrpt::poses::CPose3D sensorPose;
    void getSensorPose(mrpt::poses::CPose3D& out) const override { out = sensorPose; }
    /** Sensor-frame points from the raw returns. */
    void generatePointCloud()
    {
        point_cloud.clear();
        for (size_t i = 0; i < raw_range.size(); i++)
        {
            const float r = raw_range[i];
            if (!(r > 0.f)) continue;
            const float ce = std::cos(raw_elevation[i]);
            point_cloud.x.push_back(r * ce * std::cos(raw_azimuth[i]));
            point_cloud.y.push_back(r * ce * std::sin(raw_azimuth[i]));
            point_cloud.z.push_back(r * std::sin(raw_elevation[i]));
        }
    }
};
}  // namespace obs
inline bool maps::CPointsMap::internal_insertObservation(const mrpt::obs::CObservation& obs, const std::optional<const mrpt::poses::CPose3D>& robotPose)
{
    const auto* s = dynamic_cast<const mrpt::obs::CObservation2DRangeScan*>(&obs);
    if (!s) return false;
    const auto*                aux = s->buildAuxPointsMap<CPointsMap>();
    const mrpt::poses::CPose3D T   = robotPose ? mrpt::poses::CPose3D(*robotPose) : mrpt::poses::CPose3D();
    for (size_t i = 0; i < aux->size(); i++) insertPoint(T.composePoint(mrpt::math::TPoint3Df(aux->x_[i], aux->y_[i], aux->z_[i])));
    return true;
}
}  // namespace mrpt

// Map-definition plumbing (mrpt/maps/metric_map_types.h), reduced to what the classes need.
#define MAP_DEFINITION_START(C)                                                                        \
   public:                                                                                             \
    struct TMapDefinitionBase                                                                          \
    {                                                                                                  \
    };                                                                                                 \
    struct TMapDefinition : public mrpt::maps::TMetricMapInitializer                                   \
    {
#define MAP_DEFINITION_END(C)                                                                          \
    TMapDefinition();                                                                                  \
                                                                                                       \
   protected:                                                                                          \
    void loadFromConfigFile_map_specific(const mrpt::config::CConfigFileBase& source,                  \
                                         const std::string& sectionNamePrefix) override;               \
    void dumpToTextStream_map_specific(std::ostream& out) const override;                              \
    }                                                                                                  \
    ;                                                                                                  \
    static mrpt::maps::CMetricMap::Ptr internal_CreateFromMapDefinition(const mrpt::maps::TMetricMapInitializer& def);
#define MAP_DEFINITION_REGISTER(names, C)
