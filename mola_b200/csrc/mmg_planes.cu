// mmg_planes.cu — point-to-plane pairing against a SparseVoxelPointCloud (extension).
//
// The reference keeps a running mean per voxel (SparseVoxelPointCloud.cpp:193-204) and nothing else; mp2p_icp's
// point-to-plane matchers (not in the reference tree) pair a scan point with a local plane of the map. The batched form
// here: compose the local point, take the voxel it falls into (mrpt::round index, SparseVoxelPointCloud.h:108-115), and if
// that voxel stores at least `min_points` points, fit the plane through them: centroid and 3x3 scatter matrix of the stored
// points in slot order (fp64), smallest eigenvalue by the closed-form (trigonometric) solution, its eigenvector from the
// largest cross product of two rows of (C - lambda I). A pairing is kept when the point lies within max_dist_sqr of the
// voxel's running mean; pairings leave compacted in local-point order: index, running mean (the reference's, bit-exact),
// unit normal (largest component positive), signed point-to-plane distance, the two smallest eigenvalues (callers
// threshold planarity on them). No reference behaviour exists for this: pinned by a CPU restatement (tests, 1e-5).
#include "mmg_host.h"
#include "mmg_sort.cuh"

#include <cmath>

using namespace mmg;

namespace
{
// smallest and middle eigenvalue of the symmetric matrix {xx, xy, xz, yy, yz, zz}; unit eigenvector of the smallest
__device__ bool plane_eig(const double c[6], double& l0, double& l1, double n[3])
{
    const double xx = c[0], xy = c[1], xz = c[2], yy = c[3], yz = c[4], zz = c[5];
    const double p1 = xy * xy + xz * xz + yz * yz, q = (xx + yy + zz) / 3.0;
    const double p2 = (xx - q) * (xx - q) + (yy - q) * (yy - q) + (zz - q) * (zz - q) + 2.0 * p1;
    if (!(p2 > 0.0)) return false;  // isotropic (or empty) scatter: no plane
    const double p = sqrt(p2 / 6.0), ip = 1.0 / p;
    const double b00 = (xx - q) * ip, b11 = (yy - q) * ip, b22 = (zz - q) * ip, b01 = xy * ip, b02 = xz * ip, b12 = yz * ip;
    double       r = 0.5 * (b00 * (b11 * b22 - b12 * b12) - b01 * (b01 * b22 - b12 * b02) + b02 * (b01 * b12 - b11 * b02));
    r              = fmin(1.0, fmax(-1.0, r));
    const double phi = acos(r) / 3.0;
    const double e2  = q + 2.0 * p * cos(phi);                              // largest
    l0               = q + 2.0 * p * cos(phi + 2.0943951023931954923084);   // smallest
    l1               = 3.0 * q - e2 - l0;
    const double r0[3] = {xx - l0, xy, xz}, r1[3] = {xy, yy - l0, yz}, r2[3] = {xz, yz, zz - l0};
    double       v[3][3] = {{r0[1] * r1[2] - r0[2] * r1[1], r0[2] * r1[0] - r0[0] * r1[2], r0[0] * r1[1] - r0[1] * r1[0]},
                      {r0[1] * r2[2] - r0[2] * r2[1], r0[2] * r2[0] - r0[0] * r2[2], r0[0] * r2[1] - r0[1] * r2[0]},
                      {r1[1] * r2[2] - r1[2] * r2[1], r1[2] * r2[0] - r1[0] * r2[2], r1[0] * r2[1] - r1[1] * r2[0]}};
    int          best = 0;
    double       bn   = -1.0;
    for (int k = 0; k < 3; k++)
    {
        const double nn = v[k][0] * v[k][0] + v[k][1] * v[k][1] + v[k][2] * v[k][2];
        if (nn > bn) bn = nn, best = k;
    }
    if (!(bn > 0.0)) return false;
    const double s = 1.0 / sqrt(bn);
    n[0] = v[best][0] * s, n[1] = v[best][1] * s, n[2] = v[best][2] * s;
    // sign: the component of largest magnitude is positive (first axis on ties)
    int a = 0;
    if (fabs(n[1]) > fabs(n[a])) a = 1;
    if (fabs(n[2]) > fabs(n[a])) a = 2;
    if (n[a] < 0.0) n[0] = -n[0], n[1] = -n[1], n[2] = -n[2];
    return true;
}

__global__ void __launch_bounds__(128)
    k_match_planes(VoxelMapDev m, const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, u32 n, PoseD pose, int use_pose,
                   float inv, float max_d2, u32 min_points, uint8_t* __restrict__ keep, float* __restrict__ res /*n x 10*/)
{
    constexpr int CAP = SparseVoxelTraits::CAP;
    const u32     q   = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    float3 p = make_float3(qx[q], qy[q], qz[q]);
    if (use_pose) p = compose_point(pose, p.x, p.y, p.z);
    int  cx, cy, cz;
    bool k = SparseVoxelTraits::to_index(p.x, inv, cx) & SparseVoxelTraits::to_index(p.y, inv, cy) & SparseVoxelTraits::to_index(p.z, inv, cz);
    u32  vid = 0, cnt = 0;
    k        = k && probe_ro(m.table, m.mask, pack_key(cx, cy, cz), vid, cnt) && cnt >= min_points && cnt >= 3;
    float* o = res + (size_t)q * 10;
    if (k)
    {
        const float4 mu = m.mean[vid];
        k               = sqr_dist(p.x, p.y, p.z, mu.x, mu.y, mu.z) < max_d2;
        if (k)
        {
            const float4* row = m.pts + (size_t)vid * CAP;
            double        sx = 0, sy = 0, sz = 0;
            for (u32 s = 0; s < cnt; s++)
            {
                const float4 t = __ldg(row + s);
                sx += t.x, sy += t.y, sz += t.z;
            }
            const double ic = 1.0 / (double)cnt, gx = sx * ic, gy = sy * ic, gz = sz * ic;
            double       c[6] = {0, 0, 0, 0, 0, 0};
            for (u32 s = 0; s < cnt; s++)
            {
                const float4 t = __ldg(row + s);
                const double dx = t.x - gx, dy = t.y - gy, dz = t.z - gz;
                c[0] += dx * dx, c[1] += dx * dy, c[2] += dx * dz, c[3] += dy * dy, c[4] += dy * dz, c[5] += dz * dz;
            }
            const double ik = 1.0 / (double)(cnt - 1);
            for (int a = 0; a < 6; a++) c[a] *= ik;
            double l0, l1, nv[3];
            k = plane_eig(c, l0, l1, nv);
            if (k)
            {
                o[0] = mu.x, o[1] = mu.y, o[2] = mu.z;
                o[3] = (float)nv[0], o[4] = (float)nv[1], o[5] = (float)nv[2];
                o[6] = (float)(nv[0] * ((double)p.x - gx) + nv[1] * ((double)p.y - gy) + nv[2] * ((double)p.z - gz));
                o[7] = (float)l0, o[8] = (float)l1, o[9] = 0.f;
            }
        }
    }
    keep[q] = k ? 1 : 0;
}
__global__ void k_planes_scatter(const uint8_t* __restrict__ keep, const u32* __restrict__ offs, u32 n, const float* __restrict__ res, u32* __restrict__ oidx,
                                 float* __restrict__ omean, float* __restrict__ onormal, float* __restrict__ odist, float* __restrict__ oeig)
{
    const u32 q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n || !keep[q]) return;
    const size_t o = offs[q];
    const float* r = res + (size_t)q * 10;
    oidx[o]        = q;
    omean[3 * o] = r[0], omean[3 * o + 1] = r[1], omean[3 * o + 2] = r[2];
    onormal[3 * o] = r[3], onormal[3 * o + 1] = r[4], onormal[3 * o + 2] = r[5];
    odist[o]       = r[6];
    oeig[2 * o] = r[7], oeig[2 * o + 1] = r[8];
}
}  // namespace

extern "C" int mmg_match_planes_batch(const mmg_map* cm, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12,
                                      float max_dist_sqr, uint32_t min_points, uint64_t* out_count, uint32_t* out_local_idx, float* out_mean,
                                      float* out_normal, float* out_plane_dist, float* out_eig, int mem, void* stream)
{
    if (!cm || !out_count) return MMG_ERR_INVALID_ARG;
    MMG_CUDA(cudaSetDevice(cm->device));
    if (int rc0 = settle(cm)) return rc0;
    mmg_map* m = const_cast<mmg_map*>(cm);
    *out_count = 0;
    if (m->kind != MMG_SPARSE_VOXEL)
    {
        set_error("point-to-plane pairing needs the per-voxel statistics of a SparseVoxelPointCloud");
        return MMG_ERR_UNSUPPORTED;
    }
    if (n >= (1ull << 31) || (n && (!xs || !ys || !zs || !out_local_idx || !out_mean || !out_normal || !out_plane_dist || !out_eig))) return MMG_ERR_INVALID_ARG;
    if (n == 0) return MMG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    ScratchScope scope(st);
    const float *dx, *dy, *dz;
    int          rc = stage_in(m, xs, ys, zs, n, mem, st, &dx, &dy, &dz);
    if (rc) return rc;
    const size_t dense = n * 40, scanb = (n + 2 + scan_scratch_elems(n)) * sizeof(u32), comp = mem == MMG_MEM_HOST ? n * 40 : 0;
    char*        ob    = (char*)m->scratch_get(S_MATCH, dense + n + 64 + scanb + 64 + comp + 256);
    if (!ob) return MMG_ERR_OUT_OF_MEM;
    float*   res  = (float*)ob;
    uint8_t* keep = (uint8_t*)(ob + dense);
    u32*     offs = (u32*)(ob + ((dense + n + 63) / 64) * 64);
    char*    cb   = (char*)offs + ((scanb + 63) / 64) * 64;
    MMG_LAUNCH(k_match_planes, (unsigned)((n + 127) / 128), 128, 0, st, m->dev, dx, dy, dz, (u32)n, make_pose(pose12), pose12 ? 1 : 0, m->inv, max_dist_sqr,
               (u32)min_points, keep, res);
    exclusive_scan<uint8_t, u32>(keep, offs, n, offs + n, offs + n + 1, st);
    u32*   oi = out_local_idx;
    float *om = out_mean, *on = out_normal, *od = out_plane_dist, *oe = out_eig;
    if (mem == MMG_MEM_HOST) om = (float*)cb, on = om + 3 * n, od = on + 3 * n, oe = od + n, oi = (u32*)(oe + 2 * n);
    MMG_LAUNCH(k_planes_scatter, (unsigned)((n + 255) / 256), 256, 0, st, keep, offs, (u32)n, res, oi, om, on, od, oe);
    MMG_CUDA(cudaGetLastError());
    u32 total = 0;
    MMG_CUDA(cudaMemcpyAsync(&total, offs + n, sizeof(u32), cudaMemcpyDeviceToHost, st));
    MMG_CUDA(cudaStreamSynchronize(st));
    *out_count = total;
    if (mem == MMG_MEM_HOST && total)
    {
        MMG_CUDA(cudaMemcpyAsync(out_mean, om, (size_t)total * 12, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(out_normal, on, (size_t)total * 12, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(out_plane_dist, od, (size_t)total * 4, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(out_eig, oe, (size_t)total * 8, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaMemcpyAsync(out_local_idx, oi, (size_t)total * 4, cudaMemcpyDeviceToHost, st));
        MMG_CUDA(cudaStreamSynchronize(st));
    }
    return MMG_OK;
}
