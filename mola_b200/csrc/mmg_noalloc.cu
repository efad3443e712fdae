// mmg_noalloc.cu — mmg_nn_single_batch_noalloc: nn_single_search for callers that cannot allocate (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the
// reference.
//
// It runs the same three-kernel pipeline as mmg_nn_single_batch (k_nn_own -> k_nn_tasks -> k_nn_finish, mmg_query.cuh), but
// the pipeline's work area (128 B per query: the composed query, its best-so-far key, three task queues; plus the queue
// tails) lives in a file-scope __device__ array instead of the stream-ordered pool: 8 MiB + 256 B that every GPU loading the
// library holds until exit, enough for MMG_NOALLOC_MAX_QUERIES = 65536 queries per call. The entry point calls no
// cudaMalloc / cudaFree / cudaMallocAsync and no allocating library routine.
//
// The array is shared by every call on a device and keeps the last call's contents: the queue tails are cleared at the start
// of each call (everything else is written before it is read), and two calls that overlap on the GPU (different streams,
// or different host threads on different streams) race on it. Calls on one stream are ordered and safe.
#include "../../include/mola_b200/mmg.h"
#include "mmg_host.h"
#include "mmg_query.cuh"

#include <mutex>

using namespace mmg;

namespace
{
constexpr size_t NA_MAX     = MMG_NOALLOC_MAX_QUERIES;
constexpr size_t NA_OFF_KEY = 16 * NA_MAX, NA_OFF_Q = NA_OFF_KEY + 8 * NA_MAX, NA_OFF_CNT = NA_OFF_Q + 104 * NA_MAX;
constexpr size_t NA_BYTES   = NA_OFF_CNT + 256;
constexpr int    NA_MAX_DEVICES = 64;

// In host code this name is only a handle for cudaGetSymbolAddress, never an address to hand to a kernel.
static __device__ __align__(256) unsigned char g_nn_scratch[NA_BYTES];

std::mutex g_na_mtx;
char*      g_na_ptr[NA_MAX_DEVICES] = {};  // the array's device address, per device

// Device address of the scratch array on `device` (which must be the current device); looked up once per device.
int scratch_address(int device, char** out)
{
    if (device < 0 || device >= NA_MAX_DEVICES)
    {
        set_error("mmg_nn_single_batch_noalloc: device ordinal %d not supported (limit %d)", device, NA_MAX_DEVICES);
        return MMG_ERR_INVALID_ARG;
    }
    std::lock_guard<std::mutex> lk(g_na_mtx);
    if (!g_na_ptr[device])
    {
        void*             p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_nn_scratch);
        if (e != cudaSuccess || !p)
        {
            cudaGetLastError();
            set_error("cudaGetSymbolAddress(nn scratch) failed on device %d: %s", device, cudaGetErrorString(e));
            return MMG_ERR_CUDA;
        }
        g_na_ptr[device] = (char*)p;
    }
    *out = g_na_ptr[device];
    return MMG_OK;
}

template <class Tr>
int run(const mmg_map* m, const float* xs, const float* ys, const float* zs, u32 n, const double* pose12, const QueryOut& o, char* wb, cudaStream_t st)
{
    NNWork W;
    W.qpt = (float4*)wb, W.qkey = (u64*)(wb + NA_OFF_KEY), W.count = (u32*)(wb + NA_OFF_CNT);
    W.queue[0] = (u32*)(wb + NA_OFF_Q), W.queue[1] = W.queue[0] + 6 * NA_MAX, W.queue[2] = W.queue[1] + 12 * NA_MAX;
    const PoseD T = make_pose(pose12);
    MMG_CUDA(cudaMemsetAsync(W.count, 0, 16, st));  // queue tails + ticket: the only words that must start at zero
    MMG_LAUNCH((k_nn_own<Tr>), (n + NN_OWN_BLOCK - 1) / NN_OWN_BLOCK, NN_OWN_BLOCK, 0, st, m->dev, xs, ys, zs, n, 1u, T, pose12 ? 1 : 0, m->inv,
               m->prm.voxel_size, INFINITY, W);
    const unsigned tb = (unsigned)std::min<size_t>(((size_t)n * 26 + 255) / 256, (size_t)m->num_sms * g_tune[T_NN_TASK_BLOCKS]);
    MMG_CUDA(launch_pdl(k_nn_tasks<Tr, 6>, tb, 256, st, m->dev, m->inv, m->prm.voxel_size, W));
    MMG_CUDA(launch_pdl(k_nn_finish<Tr>, (n + 255) / 256, 256, st, m->dev, n, m->inv, W, o));
    MMG_CUDA(cudaGetLastError());
    return MMG_OK;
}
}  // namespace

extern "C" int mmg_nn_single_batch_noalloc(const mmg_map* m, const float* xs, const float* ys, const float* zs, size_t n, const double* pose12,
                                           uint8_t* out_found, float* out_xyz, float* out_d2, uint64_t* out_id, void* stream)
{
    if (!m)
    {
        set_error("null map handle");
        return MMG_ERR_INVALID_ARG;
    }
    if (n > NA_MAX)
    {
        set_error("mmg_nn_single_batch_noalloc: %zu queries exceed the static work area (MMG_NOALLOC_MAX_QUERIES = %zu); split the batch or call "
                  "mmg_nn_single_batch",
                  n, NA_MAX);
        return MMG_ERR_INVALID_ARG;
    }
    if (n && (!xs || !ys || !zs || !out_found || !out_xyz || !out_d2 || !out_id)) return MMG_ERR_INVALID_ARG;
    if (m->kind != MMG_HASHED_VOXEL && m->kind != MMG_SPARSE_VOXEL)
    {
        set_error("mmg_nn_single_batch_noalloc: voxel-hashed maps only (HashedVoxelPointCloud, SparseVoxelPointCloud)");
        return MMG_ERR_UNSUPPORTED;
    }
    MMG_CUDA(cudaSetDevice(m->device));
    if (int rc = settle(m)) return rc;  // counters of an asynchronous insert (an event wait, no allocation)
    if (n == 0) return MMG_OK;
    char* wb = nullptr;
    if (int rc = scratch_address(m->device, &wb)) return rc;
    const QueryOut o{out_found, out_xyz, out_d2, (u64*)out_id};
    cudaStream_t   st = (cudaStream_t)stream;
    return m->kind == MMG_HASHED_VOXEL ? run<HashedTraits>(m, xs, ys, zs, (u32)n, pose12, o, wb, st)
                                       : run<SparseVoxelTraits>(m, xs, ys, zs, (u32)n, pose12, o, wb, st);
}
