// mmg_host.h — host-side state behind the opaque mmg_map handle.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <mutex>
#include <utility>
#include <vector>

#include "../../include/mola_b200/mmg.h"
#include "mmg_voxel.cuh"
#include "mmg_vm.h"

namespace mmg
{
void set_error(const char* fmt, ...);
#define MMG_CUDA(expr)                                                                         \
    do {                                                                                       \
        cudaError_t e__ = (expr);                                                              \
        if (e__ != cudaSuccess)                                                                \
        {                                                                                      \
            ::mmg::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
            cudaGetLastError();                                                                \
            return e__ == cudaErrorMemoryAllocation ? MMG_ERR_OUT_OF_MEM : MMG_ERR_CUDA;       \
        }                                                                                      \
    } while (0)

enum ScratchSlot
{
    S_IN = 0,
    S_GP,
    S_VID,
    S_SORT,
    S_SEG,
    S_OUT,
    S_RCNT,
    S_ROFF,
    S_RHITS,
    S_RHITS2,
    S_HWORK,
    S_HSORT,
    S_QID,
    S_LIK,
    S_SCANS,
    S_MATCH,
    S_TREES0,
    S_TREES1,
    S_TREES2,
    S_TREES3,
    S_TREES_TODO,
    S_COUNT
};
struct TreesState;

// Scratch of one const query call. While a scope is alive on the calling thread, mmg_map::scratch_get() hands out
// blocks of the stream-ordered pool (cudaMallocAsync on the call's stream) that the scope returns when the call ends,
// instead of the map's own grow-only slots: concurrent queries on one map (several host threads, several streams)
// never share or re-allocate scratch. The hit buffers of a radius search outlive the call (mmg_nn_radius_fetch reads
// them), so they belong to the calling thread instead (RadResult).
struct ScratchScope
{
    explicit ScratchScope(cudaStream_t st);
    ~ScratchScope();
    ScratchScope(const ScratchScope&)            = delete;
    ScratchScope& operator=(const ScratchScope&) = delete;
    void*         get(int slot, size_t bytes);

   private:
    cudaStream_t                          st_;
    ScratchScope*                         prev_;
    std::vector<std::pair<void*, size_t>> slots_;
    std::vector<void*>                    all_;
};
// Result of the calling thread's last mmg_nn_radius_batch (read by its next mmg_nn_radius_fetch).
struct RadResult
{
    const mmg_map* owner = nullptr;
    u64            total = 0;
    u64*           id    = nullptr;
    float *        xyz = nullptr, *d2 = nullptr;
};
RadResult& rad_result();
}  // namespace mmg

struct mmg_map
{
    int        kind = 0, device = 0, num_sms = 148;
    mmg_params prm{};
    float      inv = 1.f;
    // voxel maps
    mmg::VoxelMapDev dev{};
    mmg::u64         cap = 0, rowcap = 0;
    mmg::u32         h_nvox_hi = 0, h_free_n = 0;
    mmg::Counters*   h_ctr = nullptr;  // pinned mirror
    void*            h_stage = nullptr;  // pinned staging block (scan poses + offsets of mmg_insert_scans)
    size_t           h_stage_bytes = 0;
    bool             points_dirty = false;
    // Lazy counters of the cap-only insert path: a device-buffer insert returns without waiting for its kernels; the commit
    // kernel mirrors the counters into h_ctr and ev_ctr marks that moment. h_nvox_hi / h_free_n are exact again after
    // settle(); until then (h_nvox_hi + pending_points) bounds the rows in use from above.
    cudaEvent_t       ev_ctr = nullptr;
    std::atomic<bool> ctr_pending{false};
    mmg::u64          pending_points = 0;
    mmg::u32          deferred_oob   = 0;  // out-of-range points of batches whose counters were read late
    std::mutex        ctr_mtx;
    // SparseTrees
    mmg::TreesState* trees = nullptr;
    // host-buffer pipeline (H2D | kernel | D2H overlapped in chunks), created on first use
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaEvent_t  ev_start = nullptr, ev_in[8] = {}, ev_k[8] = {};
    int          ensure_pipeline();
    // row arrays (pts, cnt, vkey, free_list, fill, mean) that grow in place behind reserved virtual ranges (mmg_vm.h)
    mmg::VmArray vm[6];
    bool         use_vm   = false;
    size_t       vm_limit = 0;
    std::mutex   pipe_mtx;  // the pageable-host-buffer pipeline (streams + events above) serves one call at a time
    // grow-only device scratch
    std::vector<std::pair<void*, size_t>> scratch;
    void*                                 scratch_get(int slot, size_t bytes);
    int row_points() const { return kind == MMG_SPARSE_VOXEL ? 16 : 32; }
};

// SparseVoxel radius search (mmg_sv.cu) and SparseTrees (mmg_trees.cu)
int sv_radius(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, float r2, size_t maxp, uint64_t* ooff,
              uint64_t* ototal, int mem, cudaStream_t st);
int trees_create(mmg_map* m);
int trees_destroy(mmg_map* m);
int trees_clear(mmg_map* m);
int trees_settle(mmg_map* m);
int trees_set_grid(mmg_map* m);
int trees_clone(const mmg_map* src, mmg_map* dst);
int trees_insert(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, int has_robot_pose, int mem,
                 cudaStream_t st);
int trees_nn_single(const mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, size_t stride, const double* pose12,
                    uint8_t* ofound, float* oxyz, float* od2, uint64_t* oid, int mem, cudaStream_t st);
int trees_nn_multi(const mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, size_t k, uint32_t* ocount,
                   float* oxyz, float* od2, uint64_t* oid, int mem, cudaStream_t st);
int trees_radius(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12, float r2, size_t maxp,
                 uint64_t* ooff, uint64_t* ototal, int mem, cudaStream_t st);
int trees_is_empty(const mmg_map* m, int* out);
int trees_num_blocks(const mmg_map* m, uint64_t* out);
int trees_num_points(const mmg_map* m, uint64_t* out);
int trees_row_capacity(const mmg_map* m, uint32_t* out);
int trees_export(const mmg_map* m, int32_t* keys, uint32_t* counts, float* points, int mem, cudaStream_t st);
int trees_bbox_from_points(mmg_map* m);
int trees_table_stats(mmg_map* m, mmg_table_stats* out);

namespace mmg
{
// runtime tuning knobs (mmg_set_tuning): A/B switches for the profiles, defaults = the measured best
enum Tune
{
    T_NN_KERNEL = 0,    // 0 = task-queue pipeline (k_nn_own/_tasks/_finish), 1 = k_nn_single_pruned, 2 = grouped, 3 = simple
    T_NN_TASK_BLOCKS,   // blocks of 256 per SM of k_nn_tasks
    T_NN_MIN_PIPE,      // batches smaller than this take the single-kernel path (one launch, no scratch)
    T_NN_PERMUTE,       // k_nn_single_pruned: scattered tile order on/off
    T_RADIUS_CHUNK_HITS,  // SparseTrees radius search: hits per chunk of queries (bounds the 24-byte-per-hit scratch)
    T_SV_SPLIT_HITS,      // SparseVoxel radius search: hits one pass may index (2^31 - 1); smaller values exercise the splitting
    T_NN_TASK_MINB,       // k_nn_tasks: resident blocks of 256 per SM it is compiled for (5 | 6 | 8)
    T_RADIUS_WARP_MAX,    // SparseTrees radius search: lists up to this many hits are sorted by a warp, longer ones by a block
    T_RADIUS_BLOCK_MID,   // SparseTrees radius search: lists up to this many hits share an SM three blocks at a time
    T_RADIUS_Q16,         // SparseTrees radius search: 2 = one bucket pass on the monotone 16-bit image of d2 + per-bucket insertion sort,
                          // 1 = two radix passes on that image + run fix, 0 = radix on the raw key bits
    T_RADIUS_TOPK,        // SparseTrees radius search with maxPoints <= 32: per-block bounded k-nearest search (1) or the hit-list path (0)
    T_COUNT
};
extern int g_tune[T_COUNT];
int      settle(const mmg_map* m);  // waits for the counters of asynchronous inserts still in flight (no-op otherwise)
int      alloc_table(mmg_map* m, u64 cap, cudaStream_t st);
int      ensure_rows(mmg_map* m, u64 rows, cudaStream_t st);
int      read_counters(mmg_map* m, Counters* c, cudaStream_t st);
int      stage_in(mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, int mem, cudaStream_t st, const float** dx, const float** dy,
                  const float** dz);
PoseD    make_pose(const double* pose12);
unsigned persistent_blocks(const mmg_map* m, size_t work_warps, int warps_per_block, int blocks_per_sm);
}  // namespace mmg
