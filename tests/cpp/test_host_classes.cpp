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

#include "../../mola_b200/host/GpuMetricMaps.h"
#include "../../oracle/mmo.h"

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
        S(ref_serialize, "mmo_ref_serialize"), S(ref_deserialize, "mmo_ref_deserialize");
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
    }
    catch (const std::exception& e)
    {
        std::printf("EXCEPTION: %s\n", e.what());
        return 1;
    }
    std::printf("archives cross-checked with the reference classes: %d cases\n", g_archive_cross_checks);
    std::printf(g_fail ? "%d CHECKS FAILED\n" : "ALL OK\n", g_fail);
    return g_fail ? 1 : 0;
}
