/* mmo.h — C interface shared by the two CPU checkers under oracle/:
 *   libmmo_port.so  (oracle/mmo_port.cpp) : our CPU restatement of the reference algorithm
 *   _ref/libmmo_ref.so (oracle/ref_bridge.cpp) : the reference's own .cpp files compiled
 *                     unmodified from /root/reference against the minimal MRPT shim
 *
 * TEST INFRASTRUCTURE ONLY. Nothing in mola_b200/ (the product) may include,
 * link, import or execute anything under oracle/. Only tests/, the smoke check
 * and bench.py's cpu_baseline / --impl reference legs load these libraries, and
 * only as the checker / the timed CPU baseline.
 */
#ifndef MMO_H
#define MMO_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct mmo_params
{
    float    voxel_size; /* grid_size for SparseTrees */
    uint32_t max_points_per_voxel;
    double   remove_voxels_farther_than; /* remove_submaps_farther_than for SparseTrees */
    float    min_distance_between_points;
    float    minimum_points_clearance;
    double   sigma_dist;
    double   max_corr_distance;
    uint32_t decimation;
    uint32_t match_mean;
} mmo_params;

/* kind: 0 Hashed, 1 SparseVoxel, 2 SparseTrees */
const char* mmo_impl_name(void); /* "port" or "reference" */
void*       mmo_create(int kind, const mmo_params* p);
void        mmo_destroy(void* h);
void        mmo_clear(void* h);
void        mmo_set_params(void* h, const mmo_params* p);
int         mmo_is_empty(void* h);

/* pose12: row-major R(3x3) then t(3), or NULL (=> no robot pose given). */
int mmo_insert(void* h, const float* xs, const float* ys, const float* zs, size_t n,
               const double* pose12, int has_robot_pose);

/* All query calls return 0, or -5 if the reference throws for this map type.
 * nthreads>1 shards the (const) query batch over std::threads. */
int mmo_nn_single(void* h, const float* xs, const float* ys, const float* zs, size_t n,
                  uint8_t* found, float* out_xyz, float* out_d2, uint64_t* out_id, int nthreads);
/* Port only: SparseVoxel k-NN, which the reference throws for, answers with the k <= 8 smallest
 * (d2, scan order) of the 27-cell stencil. Checker for the extension; off by default. */
int mmo_enable_knn_extension(void* h, int on);
int mmo_nn_multi(void* h, const float* xs, const float* ys, const float* zs, size_t n, size_t k,
                 uint32_t* out_count, float* out_xyz, float* out_d2, uint64_t* out_id,
                 int nthreads);
int mmo_nn_radius(void* h, const float* xs, const float* ys, const float* zs, size_t n,
                  float radius_sqr, size_t max_points, uint64_t* out_offsets /*n+1*/,
                  int nthreads);
void mmo_nn_radius_fetch(void* h, float* out_xyz, float* out_d2, uint64_t* out_id);

int mmo_likelihood(void* h, const float* xs, const float* ys, const float* zs, size_t n,
                   const double* pose12, double* out);
void mmo_bbox(void* h, float out[6]);

/* Voxels (Hashed, SparseVoxel) or blocks (SparseTrees), sorted by (cx,cy,cz). */
uint64_t mmo_num_voxels(void* h);
uint64_t mmo_num_points(void* h);
uint32_t mmo_row_capacity(void* h); /* 32 / 16 / max block size */
void     mmo_export(void* h, int32_t* keys, uint32_t* counts, float* points);
/* SparseVoxel running mean (reference) and covariance (extension N4), same order. */
void mmo_voxel_stats(void* h, float* out_mean, float* out_cov);
/* Mean number of stored points in the 27-cell stencil of the queries (P-bar of the roofline). */
double mmo_mean_stencil_points(void* h, const float* xs, const float* ys, const float* zs,
                               size_t n);

#ifdef __cplusplus
}
#endif
#endif
