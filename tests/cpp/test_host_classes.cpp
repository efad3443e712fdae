// test_host_classes.cpp — the C++ drop-in classes (mola_b200/host) driven through the
// reference's own virtual interface (CMetricMap / NearestNeighborsCapable, map definitions,
// serialization), checked against the CPU oracle through its C interface (oracle/mmo.h; the
// library path is argv[1]: the reference build oracle/_ref when present, else the port).
// Built and run on the GPU box by tests/test_host_cpp_gpu.py.
#include <dlfcn.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <tuple>

#include <mrpt/obs/CObservation2DRangeScan.h>
#include <mrpt/obs/CObservation3DRangeScan.h>
#include <mrpt/obs/CObservationVelodyneScan.h>

#include <thread>
#include <type_traits>

#include "../../mola_b200/host/GpuMetricMaps.h"
#include "../../oracle/mmo.h"

// copy / move traits of the three GPU classes, as the reference asserts for its own (tests/test-serialization.cpp:35-51)
template <typename T>
struct TraitsTest
{
    static_assert(std::is_move_constructible<T>(), "Can't move construct");
    static_assert(std::is_copy_constructible<T>(), "Can't copy construct");
    static_assert(std::is_move_assignable<T>(), "Can't move assign");
    static_assert(std::is_copy_assignable<T>(), "Can't copy assign");
};
template struct TraitsTest<MOLA_B200_NAMESPACE::HashedVoxelPointCloud>;
template struct TraitsTest<MOLA_B200_NAMESPACE::SparseVoxelPointCloud>;
template struct TraitsTest<MOLA_B200_NAMESPACE::SparseTreesPointCloud>;

namespace mg = MOLA_B200_NAMESPACE;
using mrpt::math::TPoint3Df;

static int g_fail = 0;
static int g_archive_cross_checks = 0;
#define CHECK(c)                                                     \
    do {                                                             \
        if (!(c))                                                    \
        {                                                            \
            std::printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #c); \
            g_fail++;                                                \
        }                                                            \
    } while (0)
#define CHECK_THROWS(expr)                  \
    do {                                    \
        bool t__ = false;                   \
        try                                 \
        {                                   \
            expr;                           \
        }                                   \
        catch (const std::exception&)       \
        {                                   \
            t__ = true;                     \
        }                                   \
        CHECK(t__ && "must throw: " #expr); \
    } while (0)

// ---- the oracle, loaded at run time
struct Oracle
{
    void* h = nullptr;
    decltype(&mmo_create)        create;
    decltype(&mmo_destroy)       destroy;
    decltype(&mmo_insert)        insert;
    decltype(&mmo_nn_single)     nn_single;
    decltype(&mmo_nn_multi)      nn_multi;
    decltype(&mmo_nn_radius)     nn_radius;
    decltype(&mmo_nn_radius_fetch) nn_radius_fetch;
    decltype(&mmo_likelihood)    likelihood;
    decltype(&mmo_bbox)          bbox;
    decltype(&mmo_num_voxels)    num_voxels;
    decltype(&mmo_num_points)    num_points;
    decltype(&mmo_row_capacity)  row_capacity;
    decltype(&mmo_export)        export_;
    // reference library only (nullptr with the port): archives of the reference classes themselves
    size_t (*ref_serialize)(void*, uint8_t*, size_t)        = nullptr;
    int (*ref_deserialize)(void*, const uint8_t*, size_t)   = nullptr;
    void* (*ref_metric_map)(void*)                          = nullptr;  // the reference object as its CMetricMap*
    explicit Oracle(const char* path)
    {
        h = dlopen(path, RTLD_NOW | RTLD_LOCAL);
        if (!h)
        {
            std::printf("cannot load %s: %s\n", path, dlerror());
            std::exit(2);
        }
#define S(f, n) *(void**)(&f) = dlsym(h, n)
        S(create, "mmo_create"), S(destroy, "mmo_destroy"), S(insert, "mmo_insert"), S(nn_single, "mmo_nn_single"), S(nn_multi, "mmo_nn_multi");
        S(nn_radius, "mmo_nn_radius"), S(nn_radius_fetch, "mmo_nn_radius_fetch"), S(likelihood, "mmo_likelihood"), S(bbox, "mmo_bbox");
        S(num_voxels, "mmo_num_voxels"), S(num_points, "mmo_num_points"), S(row_capacity, "mmo_row_capacity"), S(export_, "mmo_export");
        S(ref_serialize, "mmo_ref_serialize"), S(ref_deserialize, "mmo_ref_deserialize"), S(ref_metric_map, "mmo_ref_metric_map");
#undef S
    }
};

// ---- deterministic synthetic "scan": points on a floor and two walls, sensor frame
static uint64_t g_rng = 0x1234567ull;
static float    frand()
{
    g_rng = g_rng * 6364136223846793005ull + 1442695040888963407ull;
    return (float)((g_rng >> 40) & 0xFFFFFF) / (float)0x1000000;
}
static mrpt::maps::CSimplePointsMap::Ptr make_scan(size_t n)
{
    auto pc = mrpt::maps::CSimplePointsMap::Create();
    for (size_t i = 0; i < n; i++)
    {
        const float u = frand(), v = frand(), w = frand();
        if (i % 3 == 0)
            pc->insertPointFast(-15.f + 30.f * u, -6.f + 12.f * v, -1.7f + 0.02f * w);
        else if (i % 3 == 1)
            pc->insertPointFast(-15.f + 30.f * u, 6.f + 0.02f * w, -1.7f + 4.f * v);
        else
            pc->insertPointFast(-15.f + 30.f * u, -6.f - 0.02f * w, -1.7f + 4.f * v);
    }
    return pc;
}
static void pose12(const mrpt::poses::CPose3D& p, double o[12])
{
    std::memcpy(o, p.rot(), 9 * sizeof(double));
    std::memcpy(o + 9, p.trans(), 3 * sizeof(double));
}
static bool same_bits(const float* a, const float* b, size_t n) { return std::memcmp(a, b, n * sizeof(float)) == 0; }

// export of a GPU map, sorted by key like the oracle's
static void sorted_export(const mg::GpuPointMapBase& m, std::vector<int32_t>& keys, std::vector<uint32_t>& counts, std::vector<float>& pts,
                          uint32_t& cap)
{
    std::vector<int32_t>   k;
    std::vector<uint32_t>  c;
    std::vector<TPoint3Df> p;
    m.exportVoxels(k, c, p, cap);
    std::map<std::tuple<int, int, int>, size_t> order;
    for (size_t v = 0; v < c.size(); v++) order[{k[3 * v], k[3 * v + 1], k[3 * v + 2]}] = v;
    keys.clear(), counts.clear(), pts.clear();
    for (auto& kv : order)
    {
        const size_t v = kv.second;
        keys.insert(keys.end(), {k[3 * v], k[3 * v + 1], k[3 * v + 2]});
        counts.push_back(c[v]);
        for (uint32_t j = 0; j < cap; j++) pts.insert(pts.end(), {p[v * cap + j].x, p[v * cap + j].y, p[v * cap + j].z});
    }
}
static void check_same_contents(const mg::GpuPointMapBase& m, Oracle& O, void* o)
{
    std::vector<int32_t>  gk;
    std::vector<uint32_t> gc;
    std::vector<float>    gp;
    uint32_t              cap = 0;
    sorted_export(m, gk, gc, gp, cap);
    const size_t V = O.num_voxels(o);
    CHECK(gc.size() == V);
    CHECK(cap == O.row_capacity(o));
    CHECK(m.numPoints() == O.num_points(o));
    if (gc.size() != V || cap != O.row_capacity(o)) return;
    std::vector<int32_t>  ok(3 * V);
    std::vector<uint32_t> oc(V);
    std::vector<float>    op(V * cap * 3);
    if (V) O.export_(o, ok.data(), oc.data(), op.data());
    CHECK(gk == ok);
    CHECK(gc == oc);
    CHECK(same_bits(gp.data(), op.data(), op.size()));
    float ob[6];
    O.bbox(o, ob);
    const auto bb = m.boundingBox();
    const float gb[6] = {bb.min.x, bb.min.y, bb.min.z, bb.max.x, bb.max.y, bb.max.z};
    CHECK(same_bits(gb, ob, 6));
}

struct Case
{
    int         kind;
    const char* cls;
    float       size;
    mmo_params  prm;  // oracle side
};

// builds the GPU map from a map definition (the path a YAML `class:` entry takes)
static mrpt::maps::CMetricMap::Ptr create_from_definition(const Case& c)
{
    mrpt::config::CConfigFileBase cfg;
    const std::string             pre = "map";
    auto                          num = [](double v) { return std::to_string(v); };
    if (c.kind == 0)
    {
        mg::HashedVoxelPointCloud::TMapDefinition def;
        cfg.write(pre + "_creationOpts", "voxel_size", num(c.size));
        cfg.write(pre + "_insertionOpts", "max_points_per_voxel", num(c.prm.max_points_per_voxel));
        cfg.write(pre + "_insertionOpts", "remove_voxels_farther_than", num(c.prm.remove_voxels_farther_than));
        cfg.write(pre + "_insertionOpts", "min_distance_between_points", num(c.prm.min_distance_between_points));
        cfg.write(pre + "_likelihoodOpts", "decimation", num(c.prm.decimation));
        def.loadFromConfigFile(cfg, pre);
        return mg::HashedVoxelPointCloud::internal_CreateFromMapDefinition(def);
    }
    if (c.kind == 1)
    {
        mg::SparseVoxelPointCloud::TMapDefinition def;
        cfg.write(pre + "_creationOpts", "voxel_size", num(c.size));
        cfg.write(pre + "_insertionOpts", "max_points_per_voxel", num(c.prm.max_points_per_voxel));
        cfg.write(pre + "_likelihoodOpts", "match_mean", c.prm.match_mean ? "true" : "false");
        def.loadFromConfigFile(cfg, pre);
        return mg::SparseVoxelPointCloud::internal_CreateFromMapDefinition(def);
    }
    mg::SparseTreesPointCloud::TMapDefinition def;
    cfg.write(pre + "_creationOpts", "grid_size", num(c.size));
    cfg.write(pre + "_insertionOpts", "minimum_points_clearance", num(c.prm.minimum_points_clearance));
    cfg.write(pre + "_insertionOpts", "remove_submaps_farther_than", num(c.prm.remove_voxels_farther_than));
    def.loadFromConfigFile(cfg, pre);
    return mg::SparseTreesPointCloud::internal_CreateFromMapDefinition(def);
}

static void run_case(Oracle& O, const Case& c)
{
    std::printf("case %s size=%.2f\n", c.cls, c.size);
    auto  mapPtr = create_from_definition(c);
    auto* m      = dynamic_cast<mg::GpuPointMapBase*>(mapPtr.get());
    auto* nn     = dynamic_cast<mrpt::maps::NearestNeighborsCapable*>(mapPtr.get());
    CHECK(m && nn);
    CHECK(std::string(mapPtr->className()).find(c.cls) != std::string::npos);
    void* o = O.create(c.kind, &c.prm);
    CHECK(mapPtr->isEmpty());

    // three scans along x through CMetricMap::insertObservation
    mrpt::obs::CObservationPointCloud obs;
    mrpt::poses::CPose3D              last;
    for (int k = 0; k < 3; k++)
    {
        obs.pointcloud = make_scan(6000);
        const auto pose = mrpt::poses::CPose3D::FromXYZYawPitchRoll(7.0 * k, 0.3 * k, 1.7, 0.05 * k, 0.0, 0.01);
        CHECK(mapPtr->insertObservation(obs, pose));
        double T[12];
        pose12(pose, T);
        O.insert(o, obs.pointcloud->getPointsBufferRef_x().data(), obs.pointcloud->getPointsBufferRef_y().data(),
                 obs.pointcloud->getPointsBufferRef_z().data(), obs.pointcloud->size(), T, 1);
        check_same_contents(*m, O, o);
        last = pose;
    }
    CHECK(!mapPtr->isEmpty());
    CHECK(!nn->nn_has_indices_or_ids() && nn->nn_index_count() == 0);

    // queries: the last scan, shifted, in the map frame
    const auto&        sx = obs.pointcloud->getPointsBufferRef_x();
    const auto&        sy = obs.pointcloud->getPointsBufferRef_y();
    const auto&        sz = obs.pointcloud->getPointsBufferRef_z();
    const size_t       nq = 300;
    std::vector<float> qx(nq), qy(nq), qz(nq);
    for (size_t i = 0; i < nq; i++)
    {
        const auto g = last.composePoint(TPoint3Df(sx[i * 7] + 0.04f, sy[i * 7] - 0.03f, sz[i * 7] + 0.02f));
        qx[i] = g.x, qy[i] = g.y, qz[i] = g.z;
    }
    const bool has_knn = c.kind != 1;
    if (has_knn)
    {
        std::vector<uint8_t>  of(nq);
        std::vector<float>    op(3 * nq), od(nq);
        std::vector<uint64_t> oi(nq);
        CHECK(O.nn_single(o, qx.data(), qy.data(), qz.data(), nq, of.data(), op.data(), od.data(), oi.data(), 1) == 0);
        size_t nfound = 0;
        for (size_t i = 0; i < nq; i++)
        {  // the per-point virtual of NearestNeighborsCapable
            TPoint3Df  r;
            float      d2 = 0;
            uint64_t   id = 0;
            const bool f  = nn->nn_single_search(TPoint3Df(qx[i], qy[i], qz[i]), r, d2, id);
            CHECK(f == (of[i] != 0));
            if (f && of[i])
            {
                CHECK(id == oi[i] && same_bits(&d2, &od[i], 1) && same_bits(&r.x, &op[3 * i], 3));
                nfound++;
            }
        }
        CHECK(nfound > nq / 2);
        // batched form == per-point form
        mg::NNBatchResult b;
        m->nn_single_search_batch(qx.data(), qy.data(), qz.data(), nq, b);
        for (size_t i = 0; i < nq; i++) CHECK(b.found[i] == of[i] && (!of[i] || (b.ids[i] == oi[i] && same_bits(&b.dist_sqr[i], &od[i], 1))));
        {   // batched matcher == per-point results filtered by 'd2 < thr^2', in order
            const float   t2 = 0.04f;
            mg::NNBatchResult pm;
            m->match_points_batch(qx.data(), qy.data(), qz.data(), nq, t2, pm);
            size_t k = 0;
            for (size_t i = 0; i < nq; i++)
                if (of[i] && od[i] < t2)
                {
                    CHECK(k < pm.local_idx.size() && pm.local_idx[k] == i && pm.ids[k] == oi[i] && same_bits(&pm.dist_sqr[k], &od[i], 1));
                    k++;
                }
            CHECK(k == pm.local_idx.size() && k > 0);
        }
        for (size_t K : {size_t(1), size_t(2), size_t(3)})
        {
            std::vector<uint32_t> oc(nq);
            std::vector<float>    kp(3 * nq * K), kd(nq * K);
            std::vector<uint64_t> ki(nq * K);
            O.nn_multi(o, qx.data(), qy.data(), qz.data(), nq, K, oc.data(), kp.data(), kd.data(), ki.data(), 1);
            for (size_t i = 0; i < nq; i += 3)
            {
                std::vector<TPoint3Df> r;
                std::vector<float>     d2;
                std::vector<uint64_t>  ids;
                nn->nn_multiple_search(TPoint3Df(qx[i], qy[i], qz[i]), K, r, d2, ids);
                CHECK(r.size() == oc[i] && d2.size() == oc[i] && ids.size() == oc[i]);
                for (size_t j = 0; j < r.size() && j < oc[i]; j++) CHECK(ids[j] == ki[i * K + j] && same_bits(&d2[j], &kd[i * K + j], 1));
            }
        }
        std::vector<TPoint3Df> r;
        std::vector<float>     d2;
        std::vector<uint64_t>  ids;
        CHECK_THROWS(nn->nn_multiple_search(TPoint3Df(0, 0, 0), 0, r, d2, ids));  // ASSERT_(N > 0)
        // likelihood through CMetricMap::computeObservationLikelihood
        const auto   guess = mrpt::poses::CPose3D::FromXYZYawPitchRoll(14.1, 0.55, 1.7, 0.11, 0.0, 0.01);
        double       T[12], ol = 0;
        pose12(guess, T);
        O.likelihood(o, sx.data(), sy.data(), sz.data(), sx.size(), T, &ol);
        CHECK(mapPtr->canComputeObservationLikelihood(obs));
        const double gl = mapPtr->computeObservationLikelihood(obs, guess);
        CHECK(ol < 0 && std::abs(gl - ol) <= 1e-9 * std::abs(ol));
    }
    else
    {  // the reference throws for this map type
        TPoint3Df r;
        float     d2;
        uint64_t  id;
        CHECK_THROWS((void)nn->nn_single_search(TPoint3Df(0, 0, 0), r, d2, id));
    }
    for (auto [rad, maxp] : {std::pair<float, size_t>{0.3f, 0}, {0.9f, 0}, {0.9f, 2}, {1.4f, 5}})
    {
        std::vector<uint64_t> off(nq + 1);
        O.nn_radius(o, qx.data(), qy.data(), qz.data(), nq, rad * rad, maxp, off.data(), 1);
        std::vector<float>    rp(3 * off[nq]), rd(off[nq]);
        std::vector<uint64_t> ri(off[nq]);
        if (off[nq]) O.nn_radius_fetch(o, rp.data(), rd.data(), ri.data());
        for (size_t i = 0; i < nq; i += 5)
        {
            std::vector<TPoint3Df> r;
            std::vector<float>     d2;
            std::vector<uint64_t>  ids;
            nn->nn_radius_search(TPoint3Df(qx[i], qy[i], qz[i]), rad * rad, r, d2, ids, maxp);
            const size_t cnt = off[i + 1] - off[i];
            CHECK(r.size() == cnt);
            if (r.size() == cnt && cnt)
                CHECK(std::memcmp(ids.data(), &ri[off[i]], cnt * 8) == 0 && same_bits(d2.data(), &rd[off[i]], cnt) && same_bits(&r[0].x, &rp[3 * off[i]], 3 * cnt));
        }
    }
    // the 2D overloads throw (reference :638-664)
    {
        mrpt::math::TPoint2Df q2, r2;
        float                 d;
        uint64_t              id;
        CHECK_THROWS((void)nn->nn_single_search(q2, r2, d, id));
    }
    if (O.ref_serialize && O.ref_deserialize)
    {   // archives cross the boundary both ways: the reference class reads what the GPU class wrote, and vice versa
        mrpt::serialization::CArchive ar;
        ar << *m;
        void* o2 = O.create(c.kind, &c.prm);
        CHECK(O.ref_deserialize(o2, ar.buf.data(), ar.buf.size()) == 0);
        std::vector<int32_t>  gk;
        std::vector<uint32_t> gc;
        std::vector<float>    gp;
        uint32_t              cap = 0;
        sorted_export(*m, gk, gc, gp, cap);
        const size_t V = O.num_voxels(o2);
        CHECK(V == gc.size() && O.num_points(o2) == m->numPoints() && O.row_capacity(o2) == cap);
        if (V == gc.size() && O.row_capacity(o2) == cap)
        {
            std::vector<int32_t>  ok(3 * V);
            std::vector<uint32_t> oc(V);
            std::vector<float>    op(V * cap * 3);
            if (V) O.export_(o2, ok.data(), oc.data(), op.data());
            CHECK(gk == ok && gc == oc && same_bits(gp.data(), op.data(), op.size()));
        }
        O.destroy(o2);
        // the other way round: the reference's archive of its own map, read by the GPU class
        std::vector<uint8_t> bytes(O.ref_serialize(o, nullptr, 0));
        CHECK(O.ref_serialize(o, bytes.data(), bytes.size()) == bytes.size());
        mrpt::serialization::CArchive in;
        in.buf = bytes;
        std::unique_ptr<mg::GpuPointMapBase> back;
        if (c.kind == 0) back.reset(new mg::HashedVoxelPointCloud(1.0f));
        else if (c.kind == 1) back.reset(new mg::SparseVoxelPointCloud(1.0f));
        else back.reset(new mg::SparseTreesPointCloud(1.0f));
        in >> *back;
        CHECK(in.rpos == in.buf.size());
        std::vector<int32_t>  bk;
        std::vector<uint32_t> bc;
        std::vector<float>    bp;
        uint32_t              bcap = 0;
        sorted_export(*back, bk, bc, bp, bcap);
        CHECK(bk == gk && bc == gc && bcap == cap && bp.size() == gp.size() && same_bits(bp.data(), gp.data(), gp.size()));
        g_archive_cross_checks++;
    }
    // value semantics: a copy is deep
    if (c.kind == 0)
    {
        auto*                     h = dynamic_cast<mg::HashedVoxelPointCloud*>(m);
        mg::HashedVoxelPointCloud copy(*h);
        h->clear();
        CHECK(h->isEmpty() && !copy.isEmpty());
        check_same_contents(copy, O, o);
        // serialization v0 round trip (same byte layout as the reference)
        mrpt::serialization::CArchive ar;
        ar << copy;
        mg::HashedVoxelPointCloud back(1.0f);
        ar >> back;
        CHECK(back.voxelSize() == copy.voxelSize());
        CHECK(back.insertionOptions.max_points_per_voxel == copy.insertionOptions.max_points_per_voxel);
        std::vector<int32_t>  k1, k2;
        std::vector<uint32_t> c1, c2;
        std::vector<float>    p1, p2;
        uint32_t              cap1, cap2;
        sorted_export(copy, k1, c1, p1, cap1);
        sorted_export(back, k2, c2, p2, cap2);
        CHECK(k1 == k2 && c1 == c2 && p1.size() == p2.size() && same_bits(p1.data(), p2.data(), p1.size()));
        CHECK(back.getAsSimplePointsMap()->size() == back.numPoints());
    }
    else
    {   // serialization v0 round trip of the two block-structured maps (reference layout: SparseVoxelPointCloud.cpp:111-189,
        // SparseTreesPointCloud.cpp:111-167): same voxels / blocks, same per-voxel (per-block) point order, same options
        mrpt::serialization::CArchive ar;
        ar << *m;
        std::unique_ptr<mg::GpuPointMapBase> back;
        if (c.kind == 1) back.reset(new mg::SparseVoxelPointCloud(1.0f));
        else back.reset(new mg::SparseTreesPointCloud(1.0f));
        ar >> *back;
        CHECK(ar.rpos == ar.buf.size());
        CHECK(back->numPoints() == m->numPoints());
        std::vector<int32_t>  k1, k2;
        std::vector<uint32_t> c1, c2;
        std::vector<float>    p1, p2;
        uint32_t              cap1, cap2;
        sorted_export(*m, k1, c1, p1, cap1);
        sorted_export(*back, k2, c2, p2, cap2);
        CHECK(k1 == k2 && c1 == c2 && cap1 == cap2 && p1.size() == p2.size() && same_bits(p1.data(), p2.data(), p1.size()));
        if (c.kind == 1)
        {
            auto* a = dynamic_cast<mg::SparseVoxelPointCloud*>(m);
            auto* b = dynamic_cast<mg::SparseVoxelPointCloud*>(back.get());
            CHECK(b->voxelSize() == a->voxelSize());
            CHECK(b->insertionOptions.max_points_per_voxel == a->insertionOptions.max_points_per_voxel);
            CHECK(b->likelihoodOptions.match_mean == a->likelihoodOptions.match_mean && b->likelihoodOptions.decimation == a->likelihoodOptions.decimation);
        }
        else
        {
            auto* a = dynamic_cast<mg::SparseTreesPointCloud*>(m);
            auto* b = dynamic_cast<mg::SparseTreesPointCloud*>(back.get());
            CHECK(b->gridSize() == a->gridSize());
            CHECK(b->insertionOptions.minimum_points_clearance == a->insertionOptions.minimum_points_clearance);
            CHECK(b->insertionOptions.remove_submaps_farther_than == a->insertionOptions.remove_submaps_farther_than);
            // block-local point indices survive: the ids nn_single_search returns are the same
            mg::NNBatchResult r1, r2;
            m->nn_single_search_batch(qx.data(), qy.data(), qz.data(), nq, r1);
            back->nn_single_search_batch(qx.data(), qy.data(), qz.data(), nq, r2);
            CHECK(r1.found == r2.found && r1.ids == r2.ids && same_bits(r1.dist_sqr.data(), r2.dist_sqr.data(), nq));
        }
        // radius search (all hits, then the 3 nearest) gives the same answers on the restored map
        for (size_t maxp : {size_t(0), size_t(3)})
            for (size_t i = 0; i < nq; i += 7)
            {
                std::vector<TPoint3Df> ra, rb;
                std::vector<float>     da, db;
                std::vector<uint64_t>  ia, ib;
                nn->nn_radius_search(TPoint3Df(qx[i], qy[i], qz[i]), 0.8f * 0.8f, ra, da, ia, maxp);
                back->nn_radius_search(TPoint3Df(qx[i], qy[i], qz[i]), 0.8f * 0.8f, rb, db, ib, maxp);
                CHECK(ia == ib && da.size() == db.size() && (da.empty() || same_bits(da.data(), db.data(), da.size())));
            }
    }
    O.destroy(o);
}


// =====================================================================================================================
// Every observation class of internal_insertObservation, with a NON-identity sensorPose, through CMetricMap's own
// insertObservation / computeObservationLikelihood on the GPU class and on the reference class side by side
// (HashedVoxelPointCloud.cpp:289-507, SparseVoxelPointCloud.cpp:363-600, SparseTreesPointCloud.cpp:289-528).
static int g_observation_checks = 0;
static const mrpt::poses::CPose3D kSensorPose = mrpt::poses::CPose3D::FromXYZYawPitchRoll(0.30, -0.20, 0.50, 0.20, 0.05, -0.03);

// a 2-D scan of a 12 x 8 m room seen from its middle, a few invalid rays (stands in for MRPT's stock example scan)
static void make_scan2d(mrpt::obs::CObservation2DRangeScan& s, int n = 361, float jitter = 0.f)
{
    s.aperture    = 3.14159265f;
    s.rightToLeft = true;
    s.sensorPose  = kSensorPose;
    s.resizeScan(n);
    for (int i = 0; i < n; i++)
    {
        const float a  = -0.5f * s.aperture + s.aperture * (float)i / (float)(n - 1);
        const float ca = std::cos(a), sa = std::sin(a);
        const float rx = std::fabs(ca) > 1e-3f ? 6.0f / std::fabs(ca) : 1e9f, ry = std::fabs(sa) > 1e-3f ? 4.0f / std::fabs(sa) : 1e9f;
        s.setScanRange(i, std::min(rx, ry) + jitter * (frand() - 0.5f));
        s.setScanRangeValidity(i, i % 17 != 3);
    }
}
struct UnknownObservation : public mrpt::obs::CObservation
{
    void getSensorPose(mrpt::poses::CPose3D& out) const override { out = mrpt::poses::CPose3D(); }
};
static void run_observations(Oracle& O, const Case& c)
{
    using namespace mrpt::obs;
    using mrpt::poses::CPose3D;
    std::printf("observations %s size=%.2f\n", c.cls, c.size);
    auto  mapPtr = create_from_definition(c);
    auto* m      = dynamic_cast<mg::GpuPointMapBase*>(mapPtr.get());
    void* o      = O.create(c.kind, &c.prm);
    auto* ref    = O.ref_metric_map ? static_cast<mrpt::maps::CMetricMap*>(O.ref_metric_map(o)) : nullptr;
    if (!ref)
    {
        std::printf("  (port library loaded: the observation classes need the reference build, skipped)\n");
        O.destroy(o);
        return;
    }
    auto both = [&](const CObservation& obs, const CPose3D* pose, const char* what) {
        const bool a = pose ? mapPtr->insertObservation(obs, *pose) : mapPtr->insertObservation(obs);
        const bool b = pose ? ref->insertObservation(obs, *pose) : ref->insertObservation(obs);
        CHECK(a == b);
        const int before = g_fail;
        check_same_contents(*m, O, o);
        if (g_fail != before) std::printf("   ^ after %s\n", what);
        g_observation_checks++;
        return a;
    };
    auto R = [](int k) { return CPose3D::FromXYZYawPitchRoll(3.0 * k, 0.4 * k, 1.2, 0.1 * k, 0.02, -0.01); };
    // the reference's basic-ops test (tests/test-mola_metric_maps_hashedvoxels.cpp:35-42): empty, insertPoint, not empty
    {
        auto  tmp = create_from_definition(c);
        auto* t   = dynamic_cast<mg::GpuPointMapBase*>(tmp.get());
        CHECK(tmp->isEmpty());
        t->insertPoint({1.0f, 2.0f, 3.0f});
        CHECK(!tmp->isEmpty() && t->numPoints() == 1);
        tmp->clear();
        CHECK(tmp->isEmpty());
    }

    CObservation2DRangeScan scan;
    make_scan2d(scan);
    const CPose3D r0 = R(0), r1 = R(1), r2 = R(2), r3 = R(3);
    CHECK(both(scan, &r0, "2D scan, robot pose"));
    CHECK(both(scan, nullptr, "2D scan, no pose"));
    CObservation2DRangeScan none;
    none.resizeScan(10);  // every ray invalid: empty auxiliary map -> false
    CHECK(!both(none, &r0, "empty 2D scan"));

    CObservation3DRangeScan p3;
    p3.sensorPose  = kSensorPose;
    p3.hasPoints3D = true;
    {
        const auto pc = make_scan(3000);
        p3.points3D_x = pc->getPointsBufferRef_x(), p3.points3D_y = pc->getPointsBufferRef_y(), p3.points3D_z = pc->getPointsBufferRef_z();
    }
    CHECK(both(p3, &r1, "3D scan, points3D"));
    CHECK(both(p3, &r1, "3D scan, points3D again (SparseTrees: no clearance test on this branch)"));
    CObservation3DRangeScan ri;
    ri.sensorPose    = kSensorPose;
    ri.hasRangeImage = true;
    ri.rangeRows = 48, ri.rangeCols = 64;
    ri.cameraParams.fx = ri.cameraParams.fy = 60.f, ri.cameraParams.cx = 32.f, ri.cameraParams.cy = 24.f;
    for (size_t i = 0; i < ri.rangeRows * ri.rangeCols; i++) ri.rangeImage.push_back(i % 11 == 5 ? 0.f : 2.0f + 6.0f * frand());
    CHECK(both(ri, &r2, "3D scan, range image"));
    CObservation3DRangeScan nothing;
    CHECK(!both(nothing, &r2, "3D scan without data"));

    CObservationVelodyneScan vraw;
    vraw.sensorPose = kSensorPose;
    for (int i = 0; i < 20000; i++)
    {
        vraw.raw_azimuth.push_back(6.2831853f * frand()), vraw.raw_elevation.push_back(-0.4f + 0.5f * frand());
        vraw.raw_range.push_back(i % 13 == 7 ? 0.f : 3.0f + 27.0f * frand());
    }
    CHECK(both(vraw, &r3, "Velodyne scan, generatePointCloud()"));
    CHECK(vraw.point_cloud.size() > 15000);
    CObservationVelodyneScan vpc;
    vpc.sensorPose = kSensorPose;
    {
        const auto pc = make_scan(4000);
        vpc.point_cloud.x = pc->getPointsBufferRef_x(), vpc.point_cloud.y = pc->getPointsBufferRef_y(), vpc.point_cloud.z = pc->getPointsBufferRef_z();
    }
    CHECK(both(vpc, nullptr, "Velodyne scan, given cloud, no pose"));

    CObservationPointCloud pcl;
    pcl.sensorPose = kSensorPose;
    pcl.pointcloud = make_scan(5000);
    const CPose3D far = CPose3D::FromXYZYawPitchRoll(40.0, 2.0, 1.2, 0.7, 0.0, 0.0);  // evicts when the case has a far limit
    CHECK(both(pcl, &far, "point cloud, far pose"));
    UnknownObservation unk;
    CHECK(!both(unk, &r0, "unknown observation class (eviction only)"));
    CObservationPointCloud nullpc;
    CHECK_THROWS(mapPtr->insertObservation(nullpc, r0));  // ASSERT_(o.pointcloud)
    CHECK_THROWS(ref->insertObservation(nullpc, r0));
    check_same_contents(*m, O, o);

    // likelihood: 2-D scan = takenFrom alone, the sensor-frame classes = takenFrom + sensorPose
    CObservation2DRangeScan scanq;
    make_scan2d(scanq, 361, 0.04f);
    const CPose3D            guess = CPose3D::FromXYZYawPitchRoll(0.05, -0.02, 1.21, 0.01, 0.02, -0.01);
    const CObservation*      lik[] = {&scanq, &vraw, &pcl, &p3, &unk};
    for (const CObservation* ob : lik)
    {
        CHECK(mapPtr->canComputeObservationLikelihood(*ob) == ref->canComputeObservationLikelihood(*ob));
        if (c.kind == 1 && ref->canComputeObservationLikelihood(*ob))
        {   // nn_single_search throws for this map type, and so does the likelihood built on it
            CHECK_THROWS((void)mapPtr->computeObservationLikelihood(*ob, guess));
            CHECK_THROWS((void)ref->computeObservationLikelihood(*ob, guess));
            continue;
        }
        const double a = mapPtr->computeObservationLikelihood(*ob, guess), b = ref->computeObservationLikelihood(*ob, guess);
        CHECK(std::abs(a - b) <= 1e-9 * std::abs(b));
        g_observation_checks++;
    }
    CHECK(c.kind == 1 || c.prm.remove_voxels_farther_than > 0 || mapPtr->computeObservationLikelihood(scanq, guess) < 0);
    O.destroy(o);
}

// The reference's own test scenarios on the GPU classes (MRPT's stock example scan is not on this image: a synthetic
// 2-D scan takes its place, and the 267-point / 96-voxel constants become "the same numbers as the reference class").
static void replay_reference_tests(Oracle& O)
{
    using namespace mrpt::obs;
    std::printf("reference test scenarios\n");
    auto P = [](float size) {
        mmo_params p{};
        p.voxel_size = size, p.minimum_points_clearance = 0.2f, p.sigma_dist = 0.5, p.max_corr_distance = 1.0, p.decimation = 10;
        return p;
    };
    CObservation2DRangeScan scan2D;
    make_scan2d(scan2D);
    {   // tests/test-mola_metric_maps_hashedvoxels.cpp:44-117
        mg::HashedVoxelPointCloud map(0.2f);
        CHECK(map.insertObservation(scan2D));
        size_t nPts = 0;
        map.visitAllPoints([&nPts](const TPoint3Df&) { nPts++; });
        const mmo_params prm = P(0.2f);
        void*            o   = O.create(0, &prm);
        if (O.ref_metric_map)
        {
            CHECK(static_cast<mrpt::maps::CMetricMap*>(O.ref_metric_map(o))->insertObservation(scan2D));
            CHECK(nPts == O.num_points(o));
            check_same_contents(map, O, o);
        }
        CHECK(nPts > 300 && nPts == scan2D.buildAuxPointsMap<mrpt::maps::CPointsMap>()->size());
        mrpt::maps::CSimplePointsMap refPts, queryPts;
        map.visitAllPoints([&](const TPoint3Df& pt) { refPts.insertPoint(pt); });
        queryPts.insertObservation(scan2D, mrpt::poses::CPose3D::FromXYZYawPitchRoll(-0.05, -0.01, -0.01, 0, 0, 0));
        CHECK(queryPts.size() == nPts);
        const size_t knn = 2;
        size_t       compared = 0;
        for (size_t i = 0; i < queryPts.size(); i++)
        {
            mrpt::math::TPoint3D qd;
            queryPts.getPoint(i, qd);
            const TPoint3Df        query((float)qd.x, (float)qd.y, (float)qd.z);
            std::vector<TPoint3Df> results, gt_results;
            std::vector<float>     dists_sqr, gt_dists_sqr;
            std::vector<uint64_t>  IDs, gt_IDs;
            map.nn_multiple_search(query, knn, results, dists_sqr, IDs);
            refPts.nn_multiple_search(query, knn, gt_results, gt_dists_sqr, gt_IDs);
            CHECK(results.size() == knn);
            for (size_t k = 0; k < results.size(); k++, compared++)
            {
                CHECK(std::fabs(results[k].x - gt_results[k].x) < 1e-3f && std::fabs(results[k].y - gt_results[k].y) < 1e-3f &&
                      std::fabs(results[k].z - gt_results[k].z) < 1e-3f && std::fabs(dists_sqr[k] - gt_dists_sqr[k]) < 1e-3f);
            }
        }
        CHECK(compared == 2 * nPts);
        O.destroy(o);
    }
    {   // tests/test-sparsevoxelpointcloud.cpp:44-93
        static_assert(mg::SparseVoxelPointCloud::HARDLIMIT_MAX_POINTS_PER_VOXEL == 16, "the numbers of this test depend on it");
        mg::SparseVoxelPointCloud map(0.2f);
        CHECK(map.isEmpty());
        CHECK(map.insertObservation(scan2D));
        size_t nPts = 0;
        map.visitAllPoints([&nPts](const TPoint3Df&) { nPts++; });
        CHECK(nPts == map.numPoints() && nPts > 300);
        if (O.ref_metric_map)
        {
            const mmo_params prm = P(0.2f);
            void*            o   = O.create(1, &prm);
            CHECK(static_cast<mrpt::maps::CMetricMap*>(O.ref_metric_map(o))->insertObservation(scan2D));
            CHECK(nPts == O.num_points(o) && map.numVoxels() == O.num_voxels(o));
            check_same_contents(map, O, o);
            O.destroy(o);
        }
    }
    {   // tests/test-serialization.cpp:35-51 at run time: copies are deep, moves leave a destructible object
        mg::SparseTreesPointCloud a(5.0f);
        CHECK(a.insertObservation(scan2D));
        mg::SparseTreesPointCloud b(a), c(1.0f);
        c = a;
        mg::SparseTreesPointCloud d(std::move(a));
        CHECK(b.numPoints() == d.numPoints() && c.numPoints() == d.numPoints() && d.numPoints() > 0);
        b.clear();
        CHECK(b.isEmpty() && !c.isEmpty() && !d.isEmpty());
        c = std::move(d);
        CHECK(!c.isEmpty());
    }
}

// Const queries from several host threads on ONE map (SURVEY 5 / 8(b) "Threading"; the reference's Hashed queries are
// read-only and SparseTrees guards its cache with a shared_mutex, SparseTreesPointCloud.h:164-209,449).
static void run_threads(const Case& c)
{
    std::printf("threads %s\n", c.cls);
    auto  mapPtr = create_from_definition(c);
    auto* m      = dynamic_cast<mg::GpuPointMapBase*>(mapPtr.get());
    mrpt::obs::CObservationPointCloud obs;
    obs.pointcloud = make_scan(30000);
    CHECK(mapPtr->insertObservation(obs, mrpt::poses::CPose3D::FromXYZYawPitchRoll(1.0, 0.5, 1.7, 0.05, 0.0, 0.01)));
    const size_t       nq = 20000;
    std::vector<float> qx(nq), qy(nq), qz(nq);
    for (size_t i = 0; i < nq; i++) qx[i] = -14.f + 30.f * frand(), qy[i] = -6.f + 12.f * frand(), qz[i] = 4.f * frand();
    const int T = 8;
    mg::NNBatchResult s1, sr;
    if (c.kind != 1) m->nn_single_search_batch(qx.data(), qy.data(), qz.data(), nq, s1);
    m->nn_radius_search_batch(qx.data(), qy.data(), qz.data(), nq, 0.6f * 0.6f, 4, sr);
    std::vector<int>         bad(T, 0);
    std::vector<std::thread> th;
    for (int t = 0; t < T; t++)
        th.emplace_back([&, t] {
            try
            {
                for (int rep = 0; rep < 6; rep++)
                {
                    // every thread asks for its own slice, so results that leak between threads show up as mismatches
                    const size_t a = nq * t / T / 2, n = nq / 2;
                    mg::NNBatchResult r;
                    if (c.kind != 1)
                    {
                        m->nn_single_search_batch(qx.data() + a, qy.data() + a, qz.data() + a, n, r);
                        for (size_t i = 0; i < n; i++)
                            if (r.found[i] != s1.found[a + i] || (r.found[i] && (r.ids[i] != s1.ids[a + i] || std::memcmp(&r.dist_sqr[i], &s1.dist_sqr[a + i], 4))))
                                bad[t]++;
                    }
                    mg::NNBatchResult q;
                    m->nn_radius_search_batch(qx.data() + a, qy.data() + a, qz.data() + a, n, 0.6f * 0.6f, 4, q);
                    for (size_t i = 0; i < n; i++)
                    {
                        const size_t k = q.offsets[i + 1] - q.offsets[i], k0 = sr.offsets[a + i + 1] - sr.offsets[a + i];
                        if (k != k0 || (k && (std::memcmp(&q.ids[q.offsets[i]], &sr.ids[sr.offsets[a + i]], 8 * k) ||
                                              std::memcmp(&q.dist_sqr[q.offsets[i]], &sr.dist_sqr[sr.offsets[a + i]], 4 * k))))
                            bad[t]++;
                    }
                }
            }
            catch (const std::exception& e)
            {
                std::printf("  thread %d: %s\n", t, e.what());
                bad[t] = -1;
            }
        });
    for (auto& x : th) x.join();
    for (int t = 0; t < T; t++) CHECK(bad[t] == 0);
}

int main(int argc, char** argv)
{
    if (argc < 2)
    {
        std::printf("usage: %s <path to libmmo_port.so or libmmo_ref.so>\n", argv[0]);
        return 2;
    }
    Oracle O(argv[1]);
    auto   P = [](float size, uint32_t maxppv, double far, float mind, float clear, uint32_t mean) {
        mmo_params p{};
        p.voxel_size = size, p.max_points_per_voxel = maxppv, p.remove_voxels_farther_than = far, p.min_distance_between_points = mind;
        p.minimum_points_clearance = clear, p.sigma_dist = 0.5, p.max_corr_distance = 1.0, p.decimation = 10, p.match_mean = mean;
        return p;
    };
    const Case cases[] = {
        {0, "HashedVoxelPointCloud", 0.5f, P(0.5f, 0, 0, 0, 0.2f, 0)},
        {0, "HashedVoxelPointCloud", 0.4f, P(0.4f, 6, 12.0, 0.05f, 0.2f, 0)},
        {1, "SparseVoxelPointCloud", 0.2f, P(0.2f, 0, 0, 0, 0.2f, 0)},
        {1, "SparseVoxelPointCloud", 0.25f, P(0.25f, 5, 0, 0, 0.2f, 1)},
        {2, "SparseTreesPointCloud", 10.0f, P(10.0f, 0, 0, 0, 0.1f, 0)},
        {2, "SparseTreesPointCloud", 8.0f, P(8.0f, 0, 20.0, 0, 0.05f, 0)},
    };
    try
    {
        for (const Case& c : cases) run_case(O, c);
        for (const Case& c : cases) run_observations(O, c);
        replay_reference_tests(O);
        for (int k : {0, 2, 4}) run_threads(cases[k]);
    }
    catch (const std::exception& e)
    {
        std::printf("EXCEPTION: %s\n", e.what());
        return 1;
    }
    std::printf("archives cross-checked with the reference classes: %d cases\n", g_archive_cross_checks);
    std::printf("observation-class calls compared with the reference classes: %d\n", g_observation_checks);
    std::printf(g_fail ? "%d CHECKS FAILED\n" : "ALL OK\n", g_fail);
    return g_fail ? 1 : 0;
}
