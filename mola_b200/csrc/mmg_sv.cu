// mmg_sv.cu — SparseVoxelPointCloud::nn_radius_search on the device
// (reference: mola_metric_maps/src/SparseVoxelPointCloud.cpp:721-798).
//
// Cells coordToGlobalIdx(q - r) .. coordToGlobalIdx(q + r) (round-to-nearest indices) are
// visited cx -> cy -> cz -> slot; a candidate is a stored point, or the voxel's running mean
// when likelihoodOptions.match_mean is set (one candidate per voxel, id = g2plain(cell, 0));
// hits satisfy d2 <= r2. maxPoints == 0: all hits in visiting order. maxPoints != 0: the
// reference keys the candidates by d2 in a std::map<float,...> (equal d2 collapse, the last
// visited wins) and emits the first maxPoints in ascending d2 — see mmg_hits.cuh.
#include "mmg_hits.cuh"
#include "mmg_voxel.cuh"

using namespace mmg;

namespace
{
template <int PASS>
__global__ void __launch_bounds__(128)
    k_radius_sv(VoxelMapDev m, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 n, PoseD pose, int use_pose,
                float inv, float r2, float r, int match_mean, u32* __restrict__ counts, const u64* __restrict__ offsets, float* __restrict__ oxyz,
                float* __restrict__ od2, u64* __restrict__ oid, u32* __restrict__ oq)
{
    typedef SparseVoxelTraits Tr;
    const u32                 q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    float3 p = make_float3(qx[q], qy[q], qz[q]);
    if (use_pose) p = compose_point(pose, p.x, p.y, p.z);
    int  a[3], b[3];
    bool ok = Tr::to_index(__fsub_rn(p.x, r), inv, a[0]) & Tr::to_index(__fsub_rn(p.y, r), inv, a[1]) & Tr::to_index(__fsub_rn(p.z, r), inv, a[2]) &
              Tr::to_index(__fadd_rn(p.x, r), inv, b[0]) & Tr::to_index(__fadd_rn(p.y, r), inv, b[1]) & Tr::to_index(__fadd_rn(p.z, r), inv, b[2]);
    u32 c = 0;
    u64 o = PASS ? offsets[q] : 0;
    if (ok && r2 > 0.f)
    {
        auto emit = [&](float d, float x, float y, float z, u64 id) {
            if (PASS)
            {
                oxyz[3 * (o + c)] = x, oxyz[3 * (o + c) + 1] = y, oxyz[3 * (o + c) + 2] = z;
                od2[o + c] = d, oid[o + c] = id;
                if (oq) oq[o + c] = q;
            }
            c++;
        };
        for (int x = a[0]; x <= b[0]; x++)
            for (int y = a[1]; y <= b[1]; y++)
                for (int z = a[2]; z <= b[2]; z++)
                {
                    if (!index_in_range(x, y, z)) continue;
                    u32 vid, cnt;
                    if (!probe_ro(m.table, m.mask, pack_key(x, y, z), vid, cnt) || cnt == 0) continue;
                    if (match_mean)
                    {
                        const float4 t = __ldg(&m.mean[vid]);
                        const float  d = sqr_dist(t.x, t.y, t.z, p.x, p.y, p.z);
                        if (!(d > r2)) emit(d, t.x, t.y, t.z, plain_id<Tr::SLOT_BITS>(x, y, z, 0));
                    }
                    else
                        for (u32 s = 0; s < cnt; s++)
                        {
                            const float4 t = __ldg(&m.pts[(size_t)vid * Tr::CAP + s]);
                            const float  d = sqr_dist(t.x, t.y, t.z, p.x, p.y, p.z);
                            if (!(d > r2)) emit(d, t.x, t.y, t.z, plain_id<Tr::SLOT_BITS>(x, y, z, (int)s));
                        }
                }
    }
    if (!PASS) counts[q] = c;
}
}  // namespace

namespace
{
constexpr int SV_SPLIT = -1000;  // internal: the piece has more hits than one pass indexes with 32 bits
__global__ void k_add_u64(u64* __restrict__ a, size_t n, u64 v)
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) a[i] += v;
}
}  // namespace

static int sv_radius_once(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, float r2, size_t maxp_,
                          uint64_t* ooff, uint64_t* ototal, int mem, cudaStream_t st);

// A batch whose hit list outgrows one pass (32-bit hit indices in the std::map emulation) is answered in pieces of
// queries, the pieces' results concatenated (round 1 refused such batches with "split the batch").
int sv_radius(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, float r2, size_t maxp, uint64_t* ooff,
              uint64_t* ototal, int mem, cudaStream_t st)
{
    int rc = sv_radius_once(m, xs, ys, zs, n, pose12, r2, maxp, ooff, ototal, mem, st);
    if (rc != SV_SPLIT) return rc;
    struct Piece
    {
        char* blob;
        u64   total;
    };
    std::vector<Piece> pieces;
    auto               cleanup = [&] { for (auto& p : pieces) cudaFree(p.blob); };
    u64                grand = 0;
    size_t             step  = std::max<size_t>(1, n / 2);
    for (size_t a = 0; a < n;)
    {
        const size_t k = std::min(step, n - a);
        uint64_t     t = 0;
        rc             = sv_radius_once(m, xs + a, ys + a, zs + a, k, pose12, r2, maxp, ooff + a, &t, mem, st);
        if (rc == SV_SPLIT && k > 1)
        {
            step = std::max<size_t>(1, k / 2);
            continue;
        }
        if (rc == SV_SPLIT) set_error("nn_radius_search: a single query has more than 2^31 hits"), rc = MMG_ERR_INVALID_ARG;
        if (rc)
        {
            cleanup();
            return rc;
        }
        // offsets of this piece start at the hits before it (entry a + k is rewritten by the next piece's first entry)
        if (mem == MMG_MEM_HOST)
            for (size_t i = 0; i <= k; i++) ooff[a + i] += grand;
        else
            k_add_u64<<<(unsigned)((k + 256) / 256), 256, 0, st>>>((u64*)(ooff + a), k + 1, grand);
        Piece p{nullptr, t};
        if (t)
        {
            if (cudaMalloc(&p.blob, t * 24) != cudaSuccess)
            {
                cudaGetLastError();
                cleanup();
                set_error("cudaMalloc(%llu) failed", (unsigned long long)(t * 24));
                return MMG_ERR_OUT_OF_MEM;
            }
            cudaMemcpyAsync(p.blob, rad_result().id, t * 8, cudaMemcpyDeviceToDevice, st);
            cudaMemcpyAsync(p.blob + t * 8, rad_result().xyz, t * 12, cudaMemcpyDeviceToDevice, st);
            cudaMemcpyAsync(p.blob + t * 20, rad_result().d2, t * 4, cudaMemcpyDeviceToDevice, st);
        }
        pieces.push_back(p);
        grand += t;
        a += k;
    }
    char* hb = (char*)m->scratch_get(S_RHITS, std::max<u64>(grand, 1) * 24);
    if (!hb)
    {
        cleanup();
        return MMG_ERR_OUT_OF_MEM;
    }
    u64 at = 0;
    for (auto& p : pieces)
        if (p.total)
        {
            cudaMemcpyAsync(hb + at * 8, p.blob, p.total * 8, cudaMemcpyDeviceToDevice, st);
            cudaMemcpyAsync(hb + grand * 8 + at * 12, p.blob + p.total * 8, p.total * 12, cudaMemcpyDeviceToDevice, st);
            cudaMemcpyAsync(hb + grand * 20 + at * 4, p.blob + p.total * 20, p.total * 4, cudaMemcpyDeviceToDevice, st);
            at += p.total;
        }
    MMG_CUDA(cudaStreamSynchronize(st));
    cleanup();
    rad_result().total = grand, rad_result().id = (u64*)hb, rad_result().xyz = (float*)(hb + grand * 8), rad_result().d2 = (float*)(hb + grand * 20);
    if (ototal) *ototal = grand;
    return MMG_OK;
}

static int sv_radius_once(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, float r2, size_t maxp_,
                          uint64_t* ooff, uint64_t* ototal, int mem, cudaStream_t st)
{
    const float *dx, *dy, *dz;
    int          rc = stage_in(m, xs, ys, zs, n, mem, st, &dx, &dy, &dz);
    if (rc) return rc;
    u32* counts = (u32*)m->scratch_get(S_RCNT, (n + 1) * sizeof(u32));
    u64* offs   = (u64*)m->scratch_get(S_ROFF, (n + 1 + scan_scratch_elems(n)) * sizeof(u64));
    if (!counts || !offs) return MMG_ERR_OUT_OF_MEM;
    const PoseD    T    = make_pose(pose12);
    const float    r    = r2 > 0 ? std::sqrt(r2) : 0.f;
    const u32      maxp = (u32)std::min<size_t>(maxp_, 0xFFFFFFFFu);
    const int      mean = m->prm.match_mean ? 1 : 0;
    const unsigned g    = (unsigned)((n + 127) / 128);
    u64            total = 0;
    if (n)
    {
        MMG_LAUNCH((k_radius_sv<0>), g, 128, 0, st, m->dev, dx, dy, dz, (u32)n, T, pose12 ? 1 : 0, m->inv, r2, r, mean, counts, (const u64*)nullptr,
                   (float*)nullptr, (float*)nullptr, (u64*)nullptr, (u32*)nullptr);
        exclusive_scan<u32, u64>(counts, offs, n, offs + n, offs + n + 1, st);
        MMG_CUDA(cudaMemcpyAsync(&total, offs + n, sizeof(u64), cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaStreamSynchronize(st));
    }
    else
        MMG_CUDA(cudaMemsetAsync(offs, 0, sizeof(u64), st));
    if (total >= (u64)(unsigned)g_tune[T_SV_SPLIT_HITS]) return SV_SPLIT;
    rad_result().total = total;
    char* hb     = (char*)m->scratch_get(S_RHITS, std::max<u64>(total, 1) * 24);
    if (!hb) return MMG_ERR_OUT_OF_MEM;
    rad_result().id = (u64*)hb, rad_result().xyz = (float*)(hb + total * 8), rad_result().d2 = (float*)(hb + total * 20);
    if (total)
    {
        u32* qid = nullptr;
        if (maxp != 0 && !(qid = (u32*)m->scratch_get(S_QID, total * sizeof(u32)))) return MMG_ERR_OUT_OF_MEM;
        MMG_LAUNCH((k_radius_sv<1>), g, 128, 0, st, m->dev, dx, dy, dz, (u32)n, T, pose12 ? 1 : 0, m->inv, r2, r, mean, counts, offs, rad_result().xyz, rad_result().d2,
                   rad_result().id, qid);
        if (maxp != 0)
        {
            u32* sbuf = (u32*)m->scratch_get(S_HSORT, sort_scratch_words(total) * sizeof(u32));
            u32* work = (u32*)m->scratch_get(S_HWORK, hits_work_words(total, n) * sizeof(u32));
            if (!sbuf || !work) return MMG_ERR_OUT_OF_MEM;
            const float* sxyz = rad_result().xyz;
            const float* sd2  = rad_result().d2;
            const u64*   sid  = rad_result().id;
            if ((rc = hits_map_semantics(m, (u32)n, (u32)total, qid, nullptr, sort_scratch_at(sbuf, total), work, sxyz, sd2, sid, maxp, offs, &total, st)))
                return rc;
        }
    }
    MMG_CUDA(cudaGetLastError());
    MMG_CUDA(cudaMemcpyAsync(ooff, offs, (n + 1) * sizeof(u64), mem == MMG_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, st));
    MMG_CUDA(cudaStreamSynchronize(st));
    if (ototal) *ototal = total;
    return MMG_OK;
}
