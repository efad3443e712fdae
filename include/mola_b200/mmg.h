/* mmg.h — C ABI of the B200 metric-map library (libmola_b200_maps.so).
 *
 * This is the drop-in boundary for the hot path of mola_metric_maps: the thin
 * MRPT-facing C++ host classes (mola_b200/host/) and any other FFI (ctypes,
 * cgo, JNI) bind exactly these entry points. Nothing here mentions torch,
 * MRPT or C++ types: plain pointers, sizes and a POD parameter block.
 *
 * Reference interfaces each entry point replaces (paths relative to the
 * reference checkout, mola_metric_maps/):
 *   mmg_create / mmg_set_params   HashedVoxelPointCloud ctor + setVoxelProperties
 *                                 src/HashedVoxelPointCloud.cpp:182-200; option structs
 *                                 include/.../HashedVoxelPointCloud.h:370-427,
 *                                 SparseVoxelPointCloud.h:401-460, SparseTreesPointCloud.h:322-380
 *   mmg_clear                     internal_clear()        src/HashedVoxelPointCloud.cpp:287
 *   mmg_insert_batch              internal_insertObservation + internal_insertPointCloud3D +
 *                                 insertPoint             src/HashedVoxelPointCloud.cpp:289-331,588-626,1065-1082
 *                                 (SparseVoxel: src/SparseVoxelPointCloud.cpp:193-204,638-655;
 *                                  SparseTrees: src/SparseTreesPointCloud.cpp:916-1000)
 *   mmg_nn_single_batch           nn_single_search        src/HashedVoxelPointCloud.cpp:666-679
 *   mmg_nn_multi_batch            nn_multiple_search      src/HashedVoxelPointCloud.cpp:681-771
 *   mmg_nn_radius_batch/_fetch    nn_radius_search        src/HashedVoxelPointCloud.cpp:773-870
 *                                 (SparseVoxel :721-798, SparseTrees :645-713)
 *   mmg_likelihood                internal_computeObservationLikelihoodPointCloud3D  :509-546
 *   mmg_bbox / mmg_invalidate_bbox  boundingBox() :872-902, cached_.reset() :176
 *   mmg_num_voxels / mmg_export   voxels()/visitAllVoxels/visitAllPoints  :904-919, serializeTo :114-134
 *   mmg_voxel_stats               (extension) per-voxel mean + covariance
 *   mmg_comm_* / mmg_bcast        new: replicate a built map over NCCL (no reference equivalent)
 *
 * Conventions
 *   - every function returns MMG_OK (0) or a negative mmg_status; nothing throws
 *     across the boundary; mmg_last_error() gives the text for the calling thread.
 *   - point sets are SoA float arrays xs[], ys[], zs[] (MRPT getPointsBufferRef_x/y/z).
 *   - `mem` says where the caller's in/out buffers live: MMG_MEM_HOST (the library
 *     copies H2D/D2H itself and the call returns with results valid) or
 *     MMG_MEM_DEVICE (pointers are device pointers on the map's device; work is
 *     enqueued on `stream`, results valid once the stream has been synchronised).
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).
 *   - pose12 = row-major 3x3 rotation followed by translation (12 doubles), or NULL
 *     for identity; points are composed in double and rounded to float exactly as
 *     mrpt::poses::CPose3D::composePoint does.
 *   - there is no CPU fallback: without a CUDA device mmg_create fails.
 *   - threading: the const queries (mmg_nn_*_batch, mmg_match*_batch, mmg_likelihood, mmg_export, mmg_voxel_stats,
 *     mmg_num_points) may run on ONE map from several host threads and streams at once, like the reference's
 *     (SparseTreesPointCloud.h:164-209,449): every call takes its scratch from the stream-ordered memory pool, and the
 *     hits of mmg_nn_radius_batch belong to the calling thread until its mmg_nn_radius_fetch. MMG_MEM_HOST calls with
 *     pageable buffers serialise on a per-map mutex. Inserts, clear, set_params, bcast are single-writer: no other call
 *     may touch the map meanwhile.
 *   - a MMG_MEM_DEVICE insert without a min-distance test returns once its kernels are enqueued on `stream`; the
 *     counters it produces are collected by the next call on the map that needs them, and points of such a batch whose
 *     voxel index is out of range are reported (MMG_ERR_OUT_OF_RANGE) by that later call. Work on OTHER streams that
 *     reads the map must be ordered after the insert's stream by the caller.
 */
#ifndef MOLA_B200_MMG_H
#define MOLA_B200_MMG_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mmg_map  mmg_map;  /* opaque map handle (one CUDA device) */
typedef struct mmg_comm mmg_comm; /* opaque NCCL communicator wrapper     */

typedef enum mmg_status
{
    MMG_OK               = 0,
    MMG_ERR_INVALID_ARG  = -1,
    MMG_ERR_CUDA         = -2,
    MMG_ERR_NO_DEVICE    = -3,
    MMG_ERR_OUT_OF_MEM   = -4,
    MMG_ERR_UNSUPPORTED  = -5, /* e.g. SparseVoxel nn_single in reference mode throws */
    MMG_ERR_NCCL         = -6,
    MMG_ERR_OUT_OF_RANGE = -7  /* a voxel index left the +-2^20 range of the packed key */
} mmg_status;

typedef enum mmg_kind
{
    MMG_HASHED_VOXEL = 0, /* mola::HashedVoxelPointCloud */
    MMG_SPARSE_VOXEL = 1, /* mola::SparseVoxelPointCloud */
    MMG_SPARSE_TREES = 2  /* mola::SparseTreesPointCloud */
} mmg_kind;

typedef enum mmg_mem
{
    MMG_MEM_HOST   = 0,
    MMG_MEM_DEVICE = 1
} mmg_mem;

/* One POD block with the union of the YAML keys of the three map classes. */
typedef struct mmg_params
{
    /* creationOpts: voxel_size (Hashed, SparseVoxel) or grid_size (SparseTrees) */
    float voxel_size;
    /* insertionOpts */
    uint32_t max_points_per_voxel;        /* 0 = no limit (hard cap 32 / 16 applies) */
    double   remove_voxels_farther_than;  /* Hashed; SparseTrees: remove_submaps_farther_than */
    float    min_distance_between_points; /* Hashed */
    float    minimum_points_clearance;    /* SparseTrees */
    /* likelihoodOpts */
    double   sigma_dist;
    double   max_corr_distance;
    uint32_t decimation;
    uint32_t match_mean; /* SparseVoxel likelihoodOptions.match_mean (bool) */
} mmg_params;

/* Fills *p with the reference defaults for `kind`. */
int mmg_default_params(int kind, mmg_params* p);

int mmg_create(int kind, const mmg_params* p, int device, mmg_map** out);
int mmg_destroy(mmg_map* m);
int mmg_clear(mmg_map* m);
/* Deep copy (reference classes are copy-constructible). */
int mmg_clone(const mmg_map* m, mmg_map** out);
/* Changes insertion/likelihood options; a changed voxel_size clears the map. */
int mmg_set_params(mmg_map* m, const mmg_params* p);
int mmg_get_params(const mmg_map* m, mmg_params* p);

/* Pre-sizes the hash table (load <= 0.5) and the row store for `voxels` voxels, like the
 * reference's voxels_.reserve() (src/HashedVoxelPointCloud.cpp:1070). Optional: inserts grow both on demand. */
int mmg_reserve(mmg_map* m, uint64_t voxels);

/* Inserts n points in input order (per-voxel decimation identical to calling the
 * reference's insertPoint once per point). `flags` (historically "has_robot_pose"):
 *   MMG_INSERT_ROBOT_POSE    pose12 is the robot pose of internal_insertObservation: run the far-voxel / far-block
 *                            eviction first, with pose12's translation as the sensor position
 *   MMG_INSERT_NO_CLEARANCE  SparseTrees only: append every point without the minimum_points_clearance test, i.e. the
 *                            reference's public insertPoint() (SparseTreesPointCloud.h:224-234) called once per point,
 *                            which is what its CObservation3DRangeScan::points3D branch does (SparseTreesPointCloud.cpp:349-356);
 *                            ignored by the voxel maps, whose insertPoint() is the decimating one */
enum
{
    MMG_INSERT_ROBOT_POSE   = 1,
    MMG_INSERT_NO_CLEARANCE = 2
};
int mmg_insert_batch(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n,
                     const double* pose12, int flags, int mem, void* stream);

/* Several scans in one call: scan s = points [scan_offsets[s], scan_offsets[s+1]) with pose
 * poses12 + 12*s (scan_offsets and poses12 are HOST arrays; the points follow `mem`).
 * Identical to calling mmg_insert_batch once per scan in order (the reference's
 * insertObservation loop, src/HashedVoxelPointCloud.cpp:289-331); when no eviction or
 * min-distance test is configured the scans are ranked in one fused pass on the device. */
int mmg_insert_scans(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n_scans,
                     const uint64_t* scan_offsets /*n_scans+1*/, const double* poses12 /*n_scans*12*/,
                     int has_robot_pose, int mem, void* stream);

/* out_found[n] (0/1), out_xyz[n*3], out_d2[n], out_id[n]. Entries of not-found
 * queries are zero. */
int mmg_nn_single_batch(const mmg_map* m, const float* xs, const float* ys, const float* zs,
                        size_t n, const double* pose12, uint8_t* out_found, float* out_xyz,
                        float* out_d2, uint64_t* out_id, int mem, void* stream);

/* nn_single_search for callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc. This entry point has NO counterpart in the reference.
 * Same results as mmg_nn_single_batch (same kernels), but the pipeline's work area is a static __device__ array
 * of the library (8 MiB + 256 B that every GPU loading the library holds until exit) instead of pool memory, and the call
 * allocates nothing: no cudaMalloc / cudaFree / cudaMallocAsync, no allocating library routine.
 *   - inputs and outputs are DEVICE buffers (there is no `mem` argument); HashedVoxel / SparseVoxel maps only
 *   - n <= MMG_NOALLOC_MAX_QUERIES, else MMG_ERR_INVALID_ARG with a message naming the limit (split the batch)
 *   - every call on a device shares the one array: calls that OVERLAP on the GPU (different streams) race on it;
 *     calls issued on one stream are ordered and safe. */
#define MMG_NOALLOC_MAX_QUERIES 65536
int mmg_nn_single_batch_noalloc(const mmg_map* m, const float* xs, const float* ys, const float* zs,
                                size_t n, const double* pose12, uint8_t* out_found, float* out_xyz,
                                float* out_d2, uint64_t* out_id, void* stream);

/* k nearest per query, k>=1. out_count[n] = matches written for that query
 * (<= min(k, map hard limit)); results of query i start at i*k. */
int mmg_nn_multi_batch(const mmg_map* m, const float* xs, const float* ys, const float* zs,
                       size_t n, const double* pose12, size_t k, uint32_t* out_count,
                       float* out_xyz, float* out_d2, uint64_t* out_id, int mem, void* stream);

/* Radius search, CSR output in two steps: the batch call computes
 * out_offsets[n+1] and keeps the hits inside the map handle; *out_total receives
 * out_offsets[n]. mmg_nn_radius_fetch then copies them to buffers of that size. */
int mmg_nn_radius_batch(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n,
                        const double* pose12, float radius_sqr, size_t max_points,
                        uint64_t* out_offsets, uint64_t* out_total, int mem, void* stream);
int mmg_nn_radius_fetch(mmg_map* m, float* out_xyz, float* out_d2, uint64_t* out_id, int mem,
                        void* stream);

/* Batched point matcher: what mp2p_icp's Matcher_Points_DistanceThreshold does with one
 * nn_single_search call per scan point (the reference's NearestNeighborsCapable contract,
 * include/mola_metric_maps/HashedVoxelPointCloud.h:303-332) as ONE device pass: compose each local point
 * with pose12, nearest map point, keep the pair iff found and d2 < max_dist_sqr (strict), and write the
 * pairings compacted in local-point order. Output arrays are sized for n pairings; *out_count (host)
 * receives the number written. The threshold doubles as the initial search radius of the kernel. */
int mmg_match_batch(const mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n,
                    const double* pose12, float max_dist_sqr, uint64_t* out_count,
                    uint32_t* out_local_idx, float* out_xyz /*map point*/, float* out_d2, uint64_t* out_id,
                    int mem, void* stream);

/* Point-to-plane pairing against a SparseVoxelPointCloud (extension, SURVEY N4: the reference has the per-voxel running
 * mean, SparseVoxelPointCloud.cpp:193-204, and no covariance; mp2p_icp's plane matchers are outside the reference tree).
 * Each local point, composed with pose12, is paired with the voxel it falls into when that voxel stores >= min_points
 * (>= 3) points and the point lies within max_dist_sqr of the voxel's running mean. Per pairing, compacted in local-point
 * order: local index, running mean (3), unit normal of the plane fitted to the voxel's stored points (3; eigenvector of
 * the smallest eigenvalue of their covariance, largest component positive), signed point-to-plane distance, and the two
 * smallest eigenvalues (2) for planarity thresholds. Arrays are sized for n pairings; *out_count (host) = pairings written. */
int mmg_match_planes_batch(const mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n,
                           const double* pose12, float max_dist_sqr, uint32_t min_points, uint64_t* out_count,
                           uint32_t* out_local_idx, float* out_mean, float* out_normal, float* out_plane_dist,
                           float* out_eig, int mem, void* stream);

/* log-likelihood of a scan, see reference :509-546. */
int mmg_likelihood(const mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n,
                   const double* pose12, double* out_loglik, int mem, void* stream);

/* Cached bounding box semantics of the reference (min xyz, max xyz). */
int mmg_bbox(mmg_map* m, float out_min_max[6]);
int mmg_is_empty(const mmg_map* m, int* out_empty);
/* Forgets the cached bounding box (serializeFrom ends with cached_.reset(), HashedVoxelPointCloud.cpp:176). */
int mmg_invalidate_bbox(mmg_map* m);

/* Export: V voxels, each with its index (cx,cy,cz), its point count and its
 * points in slot order (row stride = mmg_row_capacity(), unused slots zero).
 * Order of voxels is unspecified (like the reference's hash-map iteration). */
int mmg_num_voxels(const mmg_map* m, uint64_t* out_v);
int mmg_num_points(const mmg_map* m, uint64_t* out_n);
int mmg_row_capacity(const mmg_map* m, uint32_t* out_cap);
int mmg_export(const mmg_map* m, int32_t* out_keys /*V*3*/, uint32_t* out_counts /*V*/,
               float* out_points /*V*cap*3*/, int mem, void* stream);

/* Extension: per-voxel mean (V*3) and covariance (V*6, xx xy xz yy yz zz)
 * in the same voxel order as mmg_export. */
int mmg_voxel_stats(const mmg_map* m, float* out_mean, float* out_cov, int mem, void* stream);

/* Diagnostics for the profiles: mean/max hash-probe length of the current
 * table, load factor, bytes resident in HBM. */
typedef struct mmg_table_stats
{
    uint64_t capacity, live_voxels, points, hbm_bytes;
    double   load_factor, mean_probe, max_probe;
} mmg_table_stats;
int mmg_get_table_stats(const mmg_map* m, mmg_table_stats* out);

/* Diagnostics: mean number of stored points in the 27-cell stencil of the queries
 * (P-bar of the roofline's algorithmic-bytes formula). Voxel maps only. */
int mmg_mean_stencil_points(const mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n,
                            int mem, void* stream, double* out);

/* Multi-GPU: replicate a built map from `root` to every rank over NCCL. */
int mmg_comm_unique_id(uint8_t out_id[128]);
int mmg_comm_init(const uint8_t id[128], int rank, int nranks, int device, mmg_comm** out);
int mmg_comm_destroy(mmg_comm* c);
int mmg_bcast(mmg_map* m, mmg_comm* c, int root, void* stream);

/* Diagnostics: process-wide A/B switches for the profiles ("nn_kernel", "nn_task_blocks", "nn_min_pipe", "nn_permute",
 * "radius_chunk_hits", "sv_split_hits", "nn_task_minb", "radius_warp_max"; also MMG_TUNE="knob=value,..." in the
 * environment). Results are identical for every setting, only the kernel organisation changes. */
int mmg_set_tuning(const char* key, int value);

/* Number of kernels this library has launched in the calling process. */
uint64_t    mmg_kernel_launch_count(void);
const char* mmg_last_error(void);
const char* mmg_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MOLA_B200_MMG_H */
