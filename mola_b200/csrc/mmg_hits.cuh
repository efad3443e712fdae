// mmg_hits.cuh — post-processing of radius-search hit lists.
//
// SparseVoxelPointCloud::nn_radius_search (maxPoints != 0, SparseVoxelPointCloud.cpp:760-795)
// and SparseTreesPointCloud::nn_radius_search (SparseTreesPointCloud.cpp:645-713) collect their
// candidates in a std::map<float, ...> keyed by the squared distance: equal d2 overwrite each
// other (the LAST candidate in scan order survives) and the output is ascending in d2.
// On the device that is: a stable multi-key radix sort of the hit list by (query, d2), keep the
// last element of every run of equal (query, d2), rank the survivors inside their query, cut at
// maxPoints, and compact. Everything here works on index permutations; the hit payload (point,
// d2, id) is gathered once at the end.
#pragma once
#include "mmg_host.h"
#include "mmg_sort.cuh"

#include <algorithm>

namespace mmg
{
static __global__ void k_hits_gather_d2(const float* __restrict__ d2, const u32* __restrict__ perm, u32 H, u32* __restrict__ key)
{
    const u32 j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < H) key[j] = __float_as_uint(d2[perm ? perm[j] : j]);
}
static __global__ void k_hits_gather_u32(const u32* __restrict__ src, const u32* __restrict__ perm, u32 H, u32* __restrict__ key)
{
    const u32 j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < H) key[j] = src[perm ? perm[j] : j];
}
// keep[t] = last of its run of equal (query, d2); qstart[q] = first sorted position of query q
static __global__ void k_hits_keep(const u32* __restrict__ sq, const u32* __restrict__ perm, const float* __restrict__ d2, u32 H, u32 nq,
                                   u32* __restrict__ keep, u32* __restrict__ qstart)
{
    const u32 t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= H) return;
    const u32 q = sq[t];
    u32       k = 0;
    if (q < nq)
    {
        k = 1;
        if (t + 1 < H && sq[t + 1] == q && d2[perm[t + 1]] == d2[perm[t]]) k = 0;
        if (t == 0 || sq[t - 1] != q) qstart[q] = t;
    }
    keep[t] = k;
}
// survivors with rank < maxp are counted per query
static __global__ void k_hits_count(const u32* __restrict__ sq, const u32* __restrict__ keep, const u32* __restrict__ ks, const u32* __restrict__ qstart,
                                    u32 H, u32 nq, u32 maxp, u32* __restrict__ outcnt)
{
    const u32 t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= H || !keep[t]) return;
    const u32 q    = sq[t];
    const u32 rank = ks[t] - ks[qstart[q]];
    if (maxp == 0 || rank < maxp) atomicAdd(&outcnt[q], 1u);
}
static __global__ void k_hits_compact(const u32* __restrict__ sq, const u32* __restrict__ perm, const u32* __restrict__ keep, const u32* __restrict__ ks,
                                      const u32* __restrict__ qstart, const u64* __restrict__ newoff, u32 H, u32 maxp, const float* __restrict__ sxyz,
                                      const float* __restrict__ sd2, const u64* __restrict__ sid, float* __restrict__ oxyz, float* __restrict__ od2,
                                      u64* __restrict__ oid)
{
    const u32 t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= H || !keep[t]) return;
    const u32 q    = sq[t];
    const u32 rank = ks[t] - ks[qstart[q]];
    if (maxp != 0 && rank >= maxp) return;
    const u64 o = newoff[q] + rank;
    const u32 s = perm[t];
    oxyz[3 * o] = sxyz[3 * (size_t)s], oxyz[3 * o + 1] = sxyz[3 * (size_t)s + 1], oxyz[3 * o + 2] = sxyz[3 * (size_t)s + 2];
    od2[o] = sd2[s];
    oid[o] = sid[s];
}

// Source hits: H entries (sxyz, sd2, sid) with their query in qkey[h] (qkey >= nq drops the hit);
// perm0 (nullable = identity) is the order in which the reference would have visited them.
// Result: m->rad_* point to the compacted hits, doff[nq+1] (device) holds the new CSR offsets,
// *total the new hit count.
inline int hits_map_semantics(mmg_map* m, u32 nq, u32 H, const u32* qkey, const u32* perm0, const SortScratch& ss, u32* work /*hits_work_words(H, nq)*/,
                              const float* sxyz, const float* sd2, const u64* sid, u32 maxp, u64* doff /*nq+1 + scan scratch*/, u64* total,
                              cudaStream_t st)
{
    const unsigned g  = (H + 255) / 256;
    u32 *          key = work, *keep = work + H, *ks = work + 2 * (size_t)H + 1, *qstart = work + 3 * (size_t)H + 2, *outcnt = qstart + nq + 1;
    u32*           bsum32 = outcnt + nq + 1;
    u32 *          sk, *p1, *p2;
    MMG_LAUNCH(k_hits_gather_d2, g, 256, 0, st, sd2, perm0, H, key);
    radix_sort_pairs(key, perm0, H, 0xFFFFFFFFu, ss, &sk, &p1, st);
    // the sorted d2 keys are not needed again: reuse `key` for the query keys
    MMG_LAUNCH(k_hits_gather_u32, g, 256, 0, st, qkey, (const u32*)p1, H, key);
    radix_sort_pairs(key, p1, H, nq, ss, &sk, &p2, st);
    MMG_CUDA(cudaMemsetAsync(qstart, 0, (2 * (size_t)nq + 2) * sizeof(u32), st));
    MMG_LAUNCH(k_hits_keep, g, 256, 0, st, sk, p2, sd2, H, nq, keep, qstart);
    // ks[H] = number of survivors
    exclusive_scan<u32, u32>(keep, ks, H, ks + H, bsum32, st);
    MMG_LAUNCH(k_hits_count, g, 256, 0, st, sk, keep, ks, qstart, H, nq, maxp, outcnt);
    exclusive_scan<u32, u64>(outcnt, doff, nq, doff + nq, doff + nq + 1, st);
    u64 tot = 0;
    MMG_CUDA(cudaMemcpyAsync(&tot, doff + nq, sizeof(u64), cudaMemcpyDeviceToHost, st));
    MMG_CUDA(cudaStreamSynchronize(st));
    char* hb = (char*)m->scratch_get(S_RHITS2, std::max<u64>(tot, 1) * 24);
    if (!hb) return MMG_ERR_OUT_OF_MEM;
    u64*   oid  = (u64*)hb;
    float* oxyz = (float*)(hb + tot * 8);
    float* od2  = (float*)(hb + tot * 20);
    if (tot) MMG_LAUNCH(k_hits_compact, g, 256, 0, st, sk, p2, keep, ks, qstart, doff, H, maxp, sxyz, sd2, sid, oxyz, od2, oid);
    MMG_CUDA(cudaGetLastError());
    rad_result().total = tot, rad_result().id = oid, rad_result().xyz = oxyz, rad_result().d2 = od2;
    *total = tot;
    return MMG_OK;
}
// u32 words of `work` for hits_map_semantics
inline size_t hits_work_words(size_t H, size_t nq) { return 3 * H + 2 * nq + 16 + scan_scratch_elems(H); }
}  // namespace mmg
