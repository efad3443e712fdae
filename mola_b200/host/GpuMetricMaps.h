// GpuMetricMaps.h — drop-in C++ host classes for the B200 metric maps.
//
// mola::HashedVoxelPointCloud, mola::SparseVoxelPointCloud and mola::SparseTreesPointCloud with the
// public surface of mola_metric_maps (include/mola_metric_maps/HashedVoxelPointCloud.h:285-364,
// 370-530; SparseVoxelPointCloud.h:300-505; SparseTreesPointCloud.h:240-417): they derive from
// mrpt::maps::CMetricMap and mrpt::maps::NearestNeighborsCapable, keep the public option members
// and their YAML/INI keys, register under the same class names, and forward every operation to
// the C ABI of include/mola_b200/mmg.h. The per-point virtuals are 1-element batches; the *_batch
// methods are what a batched matcher should call once per ICP iteration.
//
// Compiles against MRPT 2.x in a MOLA workspace and against the stand-in headers of shim/mrpt in
// this repository (tests/cpp). No CPU fallback: constructing a map without a CUDA device throws.
#pragma once

#include <mola_b200/mmg.h>
#include <mrpt/config/CConfigFileBase.h>
#include <mrpt/maps/CMetricMap.h>
#include <mrpt/maps/CSimplePointsMap.h>
#include <mrpt/maps/NearestNeighborsCapable.h>
#include <mrpt/math/TBoundingBox.h>
#include <mrpt/math/TPoint3D.h>
#include <mrpt/obs/CObservationPointCloud.h>
#include <mrpt/poses/CPose3D.h>
#include <mrpt/serialization/CArchive.h>

#include <cstdint>
#include <functional>
#include <memory>
#include <mutex>
#include <optional>
#include <string>
#include <vector>

#ifndef MOLA_B200_NAMESPACE
#define MOLA_B200_NAMESPACE mola
#endif

namespace MOLA_B200_NAMESPACE
{
/** Results of a batched query (SoA, caller-owned copies). */
struct NNBatchResult
{
    std::vector<uint8_t>  found;   // nn_single: n
    std::vector<uint32_t> count;   // nn_multiple: n
    std::vector<uint64_t> offsets; // nn_radius: n+1 (CSR)
    std::vector<mrpt::math::TPoint3Df> points;
    std::vector<float>                 dist_sqr;
    std::vector<uint64_t>              ids;
    std::vector<uint32_t>              local_idx;  // match_points_batch: index of the local point of each pairing
};

/** Code shared by the three map classes: owns the mmg_map handle (deep-copied on copy). */
class GpuPointMapBase : public mrpt::maps::CMetricMap, public mrpt::maps::NearestNeighborsCapable
{
   public:
    ~GpuPointMapBase() override;
    GpuPointMapBase(const GpuPointMapBase& o);
    GpuPointMapBase& operator=(const GpuPointMapBase& o);
    GpuPointMapBase(GpuPointMapBase&& o) noexcept;
    GpuPointMapBase& operator=(GpuPointMapBase&& o) noexcept;

    /** Insert one point (1-element batch; use insertPointCloud for scans). SparseTrees: appended without the
     *  clearance test, like the reference's public insertPoint(). */
    void insertPoint(const mrpt::math::TPoint3Df& pt);
    /** Compose with `pose` (identity if empty) and insert in input order; a given pose also
     *  triggers the far-voxel / far-block eviction (HashedVoxelPointCloud.cpp:289-331). */
    void insertPointCloud(const float* xs, const float* ys, const float* zs, size_t n, const std::optional<const mrpt::poses::CPose3D>& pose,
                          bool evict);

    // --- CMetricMap
    bool                                    isEmpty() const override;
    mrpt::math::TBoundingBoxf               boundingBox() const override;
    void                                    getVisualizationInto(mrpt::opengl::CSetOfObjects&) const override {}  // no graphics on this path
    void                                    saveMetricMapRepresentationToFile(const std::string& filNamePrefix) const override;
    const mrpt::maps::CSimplePointsMap*     getAsSimplePointsMap() const override;

    // --- NearestNeighborsCapable
    [[nodiscard]] bool   nn_has_indices_or_ids() const override { return false; }
    [[nodiscard]] size_t nn_index_count() const override { return 0; }
    void                 nn_prepare_for_2d_queries() const override {}
    void                 nn_prepare_for_3d_queries() const override {}
    [[nodiscard]] bool   nn_single_search(const mrpt::math::TPoint3Df& query, mrpt::math::TPoint3Df& result, float& out_dist_sqr,
                                          uint64_t& resultIndexOrID) const override;
    [[nodiscard]] bool   nn_single_search(const mrpt::math::TPoint2Df& query, mrpt::math::TPoint2Df& result, float& out_dist_sqr,
                                          uint64_t& resultIndexOrID) const override;
    void nn_multiple_search(const mrpt::math::TPoint3Df& query, const size_t N, std::vector<mrpt::math::TPoint3Df>& results,
                            std::vector<float>& out_dists_sqr, std::vector<uint64_t>& resultIndicesOrIDs) const override;
    void nn_multiple_search(const mrpt::math::TPoint2Df& query, const size_t N, std::vector<mrpt::math::TPoint2Df>& results,
                            std::vector<float>& out_dists_sqr, std::vector<uint64_t>& resultIndicesOrIDs) const override;
    void nn_radius_search(const mrpt::math::TPoint3Df& query, const float search_radius_sqr, std::vector<mrpt::math::TPoint3Df>& results,
                          std::vector<float>& out_dists_sqr, std::vector<uint64_t>& resultIndicesOrIDs, size_t maxPoints) const override;
    void nn_radius_search(const mrpt::math::TPoint2Df& query, const float search_radius_sqr, std::vector<mrpt::math::TPoint2Df>& results,
                          std::vector<float>& out_dists_sqr, std::vector<uint64_t>& resultIndicesOrIDs, size_t maxPoints) const override;

    // --- batched forms (one C-ABI call each)
    void nn_single_search_batch(const float* xs, const float* ys, const float* zs, size_t n, NNBatchResult& out,
                                const std::optional<const mrpt::poses::CPose3D>& pose = std::nullopt) const;
    void nn_multiple_search_batch(const float* xs, const float* ys, const float* zs, size_t n, size_t N, NNBatchResult& out,
                                  const std::optional<const mrpt::poses::CPose3D>& pose = std::nullopt) const;
    void nn_radius_search_batch(const float* xs, const float* ys, const float* zs, size_t n, float search_radius_sqr, size_t maxPoints,
                                NNBatchResult& out, const std::optional<const mrpt::poses::CPose3D>& pose = std::nullopt) const;

    /** Batched distance-threshold matcher (what mp2p_icp's Matcher_Points_DistanceThreshold does with one nn_single_search
     *  per point): pairings (local_idx, map point, d2, id) with d2 < max_dist_sqr, in local-point order; one C-ABI call. */
    void match_points_batch(const float* xs, const float* ys, const float* zs, size_t n, float max_dist_sqr, NNBatchResult& out,
                            const std::optional<const mrpt::poses::CPose3D>& pose = std::nullopt) const;
    /** voxels_.reserve() of the reference's insert (HashedVoxelPointCloud.cpp:1070) */
    void reserve(uint64_t voxels);

    // --- contents
    void   visitAllPoints(const std::function<void(const mrpt::math::TPoint3Df&)>& f) const;
    size_t numVoxels() const;
    size_t numPoints() const;
    /** (cx,cy,cz) keys, counts and points (V x rowCapacity) of every voxel / block. */
    void exportVoxels(std::vector<int32_t>& keys, std::vector<uint32_t>& counts, std::vector<mrpt::math::TPoint3Df>& points, uint32_t& rowCapacity) const;
    /** The C handle, for callers that work with device pointers (MMG_MEM_DEVICE) directly. */
    mmg_map* handle() const
    {
        syncParams();
        return h_;
    }

   protected:
    GpuPointMapBase(int kind, float size);
    void internal_clear() override;
    bool internal_insertObservation(const mrpt::obs::CObservation& obs, const std::optional<const mrpt::poses::CPose3D>& robotPose) override;
    double internal_computeObservationLikelihood(const mrpt::obs::CObservation& obs, const mrpt::poses::CPose3D& takenFrom) const override;
    bool   internal_canComputeObservationLikelihood(const mrpt::obs::CObservation& obs) const override;
    /** One branch of internal_insertObservation: eviction (with a robot pose), compose with the robot pose, insert.
     *  pointByPoint = the reference loops over insertPoint() there instead of calling internal_insertPointCloud3D. */
    void insertObservationPoints(const float* xs, const float* ys, const float* zs, size_t n,
                                 const std::optional<const mrpt::poses::CPose3D>& robotPose, bool pointByPoint);
    /** Writes the class's current option members into p. */
    virtual void fillParams(mmg_params& p) const = 0;
    virtual const char* mapName() const          = 0;
    void  syncParams() const;
    void  resize(float size);  // clears the map (setVoxelProperties / setGridProperties)
    float size() const { return size_; }
    static void check(int rc);

    int              kind_;
    float            size_;
    mutable mmg_map* h_ = nullptr;
    mutable mmg_params pushed_{};
    mutable std::mutex paramsMtx_;  // const queries from several threads all pass through syncParams()
    mutable std::shared_ptr<mrpt::maps::CSimplePointsMap> cachedPoints_;
};

// =================================================================== Hashed
/** GPU twin of mola::HashedVoxelPointCloud (include/mola_metric_maps/HashedVoxelPointCloud.h). */
class HashedVoxelPointCloud : public GpuPointMapBase
{
    DEFINE_SERIALIZABLE(HashedVoxelPointCloud, MOLA_B200_NAMESPACE)
   public:
    static constexpr std::size_t HARD_MAX_MATCHES         = 3;   // HashedVoxelPointCloud.cpp:44
    static constexpr uint32_t    SSO_MAX_POINTS_PER_VOXEL = 32;  // HashedVoxelPointCloud.h:141-186

    explicit HashedVoxelPointCloud(float voxel_size = 1.00f);
    /** Changes the voxel size and clears the map (HashedVoxelPointCloud.cpp:189-200). */
    void  setVoxelProperties(float voxel_size) { resize(voxel_size); }
    float voxelSize() const { return size(); }
    std::string asString() const override;

    struct TInsertionOptions : public mrpt::config::CLoadableOptions
    {
        void loadFromConfigFile(const mrpt::config::CConfigFileBase& source, const std::string& section) override;
        void dumpToTextStream(std::ostream& out) const override;
        void writeToStream(mrpt::serialization::CArchive& out) const;
        void readFromStream(mrpt::serialization::CArchive& in);
        uint32_t max_points_per_voxel        = 0;
        double   remove_voxels_farther_than  = .0;
        float    min_distance_between_points = .0f;
    };
    struct TLikelihoodOptions : public mrpt::config::CLoadableOptions
    {
        void loadFromConfigFile(const mrpt::config::CConfigFileBase& source, const std::string& section) override;
        void dumpToTextStream(std::ostream& out) const override;
        void writeToStream(mrpt::serialization::CArchive& out) const;
        void readFromStream(mrpt::serialization::CArchive& in);
        double   sigma_dist        = 0.5;
        double   max_corr_distance = 1.0;
        uint32_t decimation        = 10;
    };
    struct TRenderOptions : public mrpt::config::CLoadableOptions
    {
        void loadFromConfigFile(const mrpt::config::CConfigFileBase& source, const std::string& section) override;
        void dumpToTextStream(std::ostream& out) const override;
        void writeToStream(mrpt::serialization::CArchive& out) const;
        void readFromStream(mrpt::serialization::CArchive& in);
        float              point_size = 1.0f;
        mrpt::img::TColorf color{.0f, .0f, 1.0f};
        int8_t             colormap                    = 2;  // mrpt::img::cmHOT
        uint8_t            recolorizeByCoordinateIndex = 2;
    };
    TInsertionOptions  insertionOptions;
    TLikelihoodOptions likelihoodOptions;
    TRenderOptions     renderOptions;

    MAP_DEFINITION_START(HashedVoxelPointCloud)
    float              voxel_size = 1.00f;
    TInsertionOptions  insertionOpts;
    TLikelihoodOptions likelihoodOpts;
    TRenderOptions     renderOpts;
    MAP_DEFINITION_END(HashedVoxelPointCloud)

   protected:
    void        fillParams(mmg_params& p) const override;
    const char* mapName() const override { return "HashedVoxelPointCloud"; }
};

// ============================================================== SparseVoxel
/** GPU twin of mola::SparseVoxelPointCloud. nn_single/multiple_search throw like the reference
 *  (SparseVoxelPointCloud.cpp:695-719) unless enable_knn_extension is set. */
class SparseVoxelPointCloud : public GpuPointMapBase
{
    DEFINE_SERIALIZABLE(SparseVoxelPointCloud, MOLA_B200_NAMESPACE)
   public:
    static constexpr uint32_t HARDLIMIT_MAX_POINTS_PER_VOXEL = 16;  // SparseVoxelPointCloud.h:61

    explicit SparseVoxelPointCloud(float voxel_size = 0.20f);
    void  setVoxelProperties(float voxel_size) { resize(voxel_size); }
    float voxelSize() const { return size(); }
    std::string asString() const override;
    bool        enable_knn_extension = false;

    [[nodiscard]] bool nn_single_search(const mrpt::math::TPoint3Df& query, mrpt::math::TPoint3Df& result, float& out_dist_sqr,
                                        uint64_t& resultIndexOrID) const override;
    void nn_multiple_search(const mrpt::math::TPoint3Df& query, const size_t N, std::vector<mrpt::math::TPoint3Df>& results,
                            std::vector<float>& out_dists_sqr, std::vector<uint64_t>& resultIndicesOrIDs) const override;
    using GpuPointMapBase::nn_multiple_search;
    using GpuPointMapBase::nn_single_search;
    /** Per-voxel running mean (reference) and covariance (extension), in exportVoxels() order. */
    void voxelStats(std::vector<mrpt::math::TPoint3Df>& mean, std::vector<float>& cov6) const;

    struct TInsertionOptions : public mrpt::config::CLoadableOptions
    {
        void loadFromConfigFile(const mrpt::config::CConfigFileBase& source, const std::string& section) override;
        void dumpToTextStream(std::ostream& out) const override;
        void writeToStream(mrpt::serialization::CArchive& out) const;
        void readFromStream(mrpt::serialization::CArchive& in);
        uint32_t max_points_per_voxel = 0;
    };
    struct TLikelihoodOptions : public mrpt::config::CLoadableOptions
    {
        void loadFromConfigFile(const mrpt::config::CConfigFileBase& source, const std::string& section) override;
        void dumpToTextStream(std::ostream& out) const override;
        void writeToStream(mrpt::serialization::CArchive& out) const;
        void readFromStream(mrpt::serialization::CArchive& in);
        double   sigma_dist        = 0.5;
        double   max_corr_distance = 1.0;
        uint32_t decimation        = 10;
        bool     match_mean        = false;
    };
    /** Carried through the archive only (SparseVoxelPointCloud.h:470-494); nothing is rendered here. */
    struct TRenderOptions
    {
        void writeToStream(mrpt::serialization::CArchive& out) const;
        void readFromStream(mrpt::serialization::CArchive& in);
        float              point_size            = 1.0f;
        bool               show_mean_only        = true;
        bool               show_inner_grid_boxes = false;
        mrpt::img::TColorf color{.0f, .0f, 1.0f};
        int8_t             colormap                    = 2;  // mrpt::img::cmHOT
        uint8_t            recolorizeByCoordinateIndex = 2;
    };
    TInsertionOptions  insertionOptions;
    TLikelihoodOptions likelihoodOptions;
    TRenderOptions     renderOptions;

    MAP_DEFINITION_START(SparseVoxelPointCloud)
    float              voxel_size = 0.20f;
    TInsertionOptions  insertionOpts;
    TLikelihoodOptions likelihoodOpts;
    MAP_DEFINITION_END(SparseVoxelPointCloud)

   protected:
    void        fillParams(mmg_params& p) const override;
    const char* mapName() const override { return "SparseVoxelPointCloud"; }
};

// ============================================================== SparseTrees
/** GPU twin of mola::SparseTreesPointCloud (exact NN inside the query's coarse block). */
class SparseTreesPointCloud : public GpuPointMapBase
{
    DEFINE_SERIALIZABLE(SparseTreesPointCloud, MOLA_B200_NAMESPACE)
   public:
    explicit SparseTreesPointCloud(float grid_size = 10.0f);
    void  setGridProperties(float grid_size) { resize(grid_size); }
    float gridSize() const { return size(); }
    std::string asString() const override;

    struct TInsertionOptions : public mrpt::config::CLoadableOptions
    {
        void loadFromConfigFile(const mrpt::config::CConfigFileBase& source, const std::string& section) override;
        void dumpToTextStream(std::ostream& out) const override;
        void writeToStream(mrpt::serialization::CArchive& out) const;
        void readFromStream(mrpt::serialization::CArchive& in);
        float minimum_points_clearance    = 0.20f;
        float remove_submaps_farther_than = .0f;
    };
    struct TLikelihoodOptions : public mrpt::config::CLoadableOptions
    {
        void loadFromConfigFile(const mrpt::config::CConfigFileBase& source, const std::string& section) override;
        void dumpToTextStream(std::ostream& out) const override;
        void writeToStream(mrpt::serialization::CArchive& out) const;
        void readFromStream(mrpt::serialization::CArchive& in);
        double   sigma_dist        = 0.5;
        double   max_corr_distance = 1.0;
        uint32_t decimation        = 10;
    };
    /** Carried through the archive only (SparseTreesPointCloud.h:388-406); nothing is rendered here. */
    struct TRenderOptions
    {
        void writeToStream(mrpt::serialization::CArchive& out) const;
        void readFromStream(mrpt::serialization::CArchive& in);
        float              point_size            = 1.0f;
        bool               show_inner_grid_boxes = false;
        mrpt::img::TColorf color{.0f, .0f, 1.0f};
        int8_t             colormap                    = 2;  // mrpt::img::cmHOT
        uint8_t            recolorizeByCoordinateIndex = 2;
    };
    TInsertionOptions  insertionOptions;
    TLikelihoodOptions likelihoodOptions;
    TRenderOptions     renderOptions;

    MAP_DEFINITION_START(SparseTreesPointCloud)
    float              grid_size = 10.0f;
    TInsertionOptions  insertionOpts;
    TLikelihoodOptions likelihoodOpts;
    MAP_DEFINITION_END(SparseTreesPointCloud)

   protected:
    void        fillParams(mmg_params& p) const override;
    const char* mapName() const override { return "SparseTreesPointCloud"; }
};

/** Registers the three classes with MRPT's class factory (reference: src/register.cpp:41-50). */
void registerGpuMetricMaps();
}  // namespace MOLA_B200_NAMESPACE
