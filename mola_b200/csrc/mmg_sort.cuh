// mmg_sort.cuh — hand-written device primitives: exclusive scan and a stable LSB
// radix sort of (u32 key, u32 value) pairs. Used to turn "voxel id per point" into
// per-voxel segments in input order (the reference inserts points sequentially,
// HashedVoxelPointCloud.cpp:419-422, so per-voxel order is input order).
#pragma once
#include "mmg_device.cuh"

#include <atomic>

namespace mmg
{
extern std::atomic<unsigned long long> g_launches;  // kernels launched by this library (mmg_kernel_launch_count)
#define MMG_LAUNCH(kernel, grid, block, smem, stream, ...)              \
    do {                                                                \
        kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);     \
        ++::mmg::g_launches;                                            \
    } while (0)

// Programmatic dependent launch (sm_90+): the kernel may be scheduled while its predecessor in the stream is still
// draining; it calls pdl_wait() before touching anything the predecessor wrote, and the predecessor calls pdl_trigger()
// as early as it likes. Hides the launch gap between the short stages of a pipeline.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;"); }
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), unsigned grid, unsigned block, cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = dim3(grid, 1, 1);
    cfg.blockDim           = dim3(block, 1, 1);
    cfg.stream             = st;
    cudaLaunchAttribute at[1];
    at[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs                                        = at;
    cfg.numAttrs                                     = 1;
    ++g_launches;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// ------------------------------------------------------------------ scan
constexpr int SCAN_THREADS = 1024, SCAN_ITEMS = 4, SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T* smem /*33*/, T& block_total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    T         inc  = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        T t = __shfl_up_sync(MMG_FULL, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) smem[warp] = inc;
    __syncthreads();
    if (warp == 0)
    {
        T w = (lane < nw) ? smem[lane] : 0, wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            T t = __shfl_up_sync(MMG_FULL, wi, o);
            if (lane >= o) wi += t;
        }
        if (lane < nw) smem[lane] = wi - w;  // exclusive warp offsets
        if (lane == 31) smem[32] = wi;
    }
    __syncthreads();
    const T r   = smem[warp] + inc - v;
    block_total = smem[32];
    __syncthreads();
    return r;
}

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_reduce(const TIn* __restrict__ in, size_t n, TOut* __restrict__ bsum)
{
    __shared__ TOut sm[33];
    const size_t    base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
    TOut            s    = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++)
        if (base + i < n) s += (TOut)in[base + i];
    TOut tot;
    block_exclusive_scan<TOut>(s, sm, tot);
    if (threadIdx.x == 0) bsum[blockIdx.x] = tot;
}
template <typename TOut>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_bsum(TOut* __restrict__ bsum, size_t nb, TOut* __restrict__ total)
{
    __shared__ TOut sm[33];
    TOut            carry = 0;
    for (size_t b0 = 0; b0 < nb; b0 += SCAN_THREADS)
    {
        const size_t i = b0 + threadIdx.x;
        const TOut   v = (i < nb) ? bsum[i] : 0;
        TOut         tot;
        const TOut   e = block_exclusive_scan<TOut>(v, sm, tot);
        if (i < nb) bsum[i] = carry + e;
        carry += tot;
    }
    if (threadIdx.x == 0 && total) *total = carry;
}
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(SCAN_THREADS)
    k_scan_down(const TIn* __restrict__ in, size_t n, const TOut* __restrict__ bsum, TOut* __restrict__ out)
{
    __shared__ TOut sm[33];
    const size_t    base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
    TOut            v[SCAN_ITEMS], s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++)
    {
        v[i] = (base + i < n) ? (TOut)in[base + i] : 0;
        s += v[i];
    }
    TOut tot;
    TOut e = block_exclusive_scan<TOut>(s, sm, tot) + bsum[blockIdx.x];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++)
    {
        if (base + i < n) out[base + i] = e;
        e += v[i];
    }
}
// out[i] = sum_{j<i} in[j] for i in [0,n); *total (device, may be out+n) = sum of all.
// bsum: scratch of ceil(n/SCAN_TILE) TOut. In-place (in == out, same type) is allowed.
template <typename TIn, typename TOut>
inline void exclusive_scan(const TIn* in, TOut* out, size_t n, TOut* total, TOut* bsum, cudaStream_t st)
{
    if (n == 0)
    {
        if (total) cudaMemsetAsync(total, 0, sizeof(TOut), st);
        return;
    }
    const size_t nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    MMG_LAUNCH((k_scan_reduce<TIn, TOut>), (unsigned)nb, SCAN_THREADS, 0, st, in, n, bsum);
    MMG_LAUNCH((k_scan_bsum<TOut>), 1, SCAN_THREADS, 0, st, bsum, nb, total);
    MMG_LAUNCH((k_scan_down<TIn, TOut>), (unsigned)nb, SCAN_THREADS, 0, st, in, n, bsum, out);
}
inline size_t scan_scratch_elems(size_t n) { return (n + SCAN_TILE - 1) / SCAN_TILE + 1; }

// ------------------------------------------------------------ radix sort
constexpr int RS_THREADS = 256, RS_ITEMS = 16, RS_TILE = RS_THREADS * RS_ITEMS, RS_BITS = 8, RS_RADIX = 1 << RS_BITS;
constexpr int RS_WARPS = RS_THREADS / 32, RS_WCHUNK = 32 * RS_ITEMS;

static __global__ void __launch_bounds__(RS_THREADS)
    k_radix_hist(const u32* __restrict__ keys, size_t n, int shift, u32 clampv, u32* __restrict__ ghist, u32 nblocks)
{
    __shared__ u32 h[RS_RADIX];
    h[threadIdx.x] = 0;
    __syncthreads();
    const size_t base = (size_t)blockIdx.x * RS_TILE;
#pragma unroll 4
    for (int i = 0; i < RS_ITEMS; i++)
    {
        const size_t j = base + (size_t)i * RS_THREADS + threadIdx.x;
        if (j < n)
        {
            const u32 k = min(keys[j], clampv);
            atomicAdd(&h[(k >> shift) & (RS_RADIX - 1)], 1u);
        }
    }
    __syncthreads();
    ghist[(size_t)threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

// Stable scatter. Warp w of the block owns the contiguous chunk
// [tile + w*512, tile + (w+1)*512); it walks it in 16 rounds of 32 consecutive
// items, ranking equal digits with __match_any_sync so input order is preserved.
static __global__ void __launch_bounds__(RS_THREADS)
    k_radix_scatter(const u32* __restrict__ keys_in, const u32* __restrict__ vals_in /*nullable = iota*/, u32* __restrict__ keys_out,
                    u32* __restrict__ vals_out, size_t n, int shift, u32 clampv, const u32* __restrict__ gscan, u32 nblocks)
{
    __shared__ u32 wcnt[RS_WARPS][RS_RADIX];
    const int      lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_WARPS * RS_RADIX; i += RS_THREADS) (&wcnt[0][0])[i] = 0;
    __syncthreads();
    const size_t   wbase = (size_t)blockIdx.x * RS_TILE + (size_t)warp * RS_WCHUNK;
    u32            key[RS_ITEMS];
    unsigned short rank[RS_ITEMS];
#pragma unroll
    for (int r = 0; r < RS_ITEMS; r++)
    {
        const size_t j     = wbase + (size_t)r * 32 + lane;
        const bool   valid = j < n;
        key[r]             = valid ? min(keys_in[j], clampv) : 0xFFFFFFFFu;
        const u32 digit    = (key[r] >> shift) & (RS_RADIX - 1);
        // lanes past the end form their own group (digit 256) so they never disturb ranks
        const u32 peers  = __match_any_sync(MMG_FULL, valid ? digit : (u32)RS_RADIX);
        const int leader = __ffs(peers) - 1;
        u32       base   = 0;
        if (valid && lane == leader)
        {
            base = wcnt[warp][digit];
            wcnt[warp][digit] = base + __popc(peers);
        }
        base    = __shfl_sync(MMG_FULL, base, leader);
        rank[r] = (unsigned short)(base + __popc(peers & ((1u << lane) - 1)));
        __syncwarp();
    }
    __syncthreads();
    {  // thread d turns the per-warp counts of digit d into global output offsets
        const int d   = threadIdx.x;
        u32       run = gscan[(size_t)d * nblocks + blockIdx.x];
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++)
        {
            const u32 c = wcnt[w][d];
            wcnt[w][d]  = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RS_ITEMS; r++)
    {
        const size_t j = wbase + (size_t)r * 32 + lane;
        if (j < n)
        {
            const u32    digit = (key[r] >> shift) & (RS_RADIX - 1);
            const size_t pos   = (size_t)wcnt[warp][digit] + rank[r];
            keys_out[pos]      = key[r];
            vals_out[pos]      = vals_in ? vals_in[j] : (u32)j;
        }
    }
}

// Small inputs (one lidar scan's kept points): all passes of the same stable LSD sort inside ONE launch of one
// 1024-thread block, ping-ponging through the scratch arrays (L2-resident at these sizes). The multi-kernel pipeline
// below costs three to five launches per 8-bit pass, each of them latency-bound at a few ten thousand keys.
constexpr int RS_SMALL_THREADS = 1024, RS_SMALL_WARPS = RS_SMALL_THREADS / 32, RS_SMALL_MAX = 1 << 16;
static __global__ void __launch_bounds__(RS_SMALL_THREADS)
    k_radix_small(u32* k0, const u32* v0 /*nullable = iota*/, u32* k1, u32* va, u32* vb,  // no __restrict__: the arrays swap roles every pass
                  u32 n, int passes, u32 clampv)
{
    __shared__ u32 cnt[RS_SMALL_WARPS][RS_RADIX];
    __shared__ u32 sm[33];
    const int      lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const u32      chunk = ((n + RS_SMALL_WARPS - 1) / RS_SMALL_WARPS + 31u) & ~31u;  // per warp, whole rounds of 32
    const u32      beg = min(n, (u32)warp * chunk), end = min(n, beg + chunk);
    u32 *          kin = k0, *kout = k1;
    const u32*     vin  = v0;
    u32*           vout = (v0 == va) ? vb : va;
    for (int p = 0; p < passes; p++)
    {
        const int shift = p * RS_BITS;
        for (int i = threadIdx.x; i < RS_SMALL_WARPS * RS_RADIX; i += RS_SMALL_THREADS) (&cnt[0][0])[i] = 0;
        __syncthreads();
        for (u32 j0 = beg + lane; j0 < end; j0 += 32 * 8)
        {   // eight independent loads in flight per lane
            u32 k[8];
#pragma unroll
            for (int i = 0; i < 8; i++) k[i] = (j0 + 32 * i < end) ? kin[j0 + 32 * i] : 0u;
#pragma unroll
            for (int i = 0; i < 8; i++)
                if (j0 + 32 * i < end) atomicAdd(&cnt[warp][(min(k[i], clampv) >> shift) & (RS_RADIX - 1)], 1u);
        }
        __syncthreads();
        {  // digit d of warp w starts at (keys with a smaller digit) + (keys with digit d in the warps before w)
            u32 tot = 0;
            if (threadIdx.x < RS_RADIX)
                for (int w = 0; w < RS_SMALL_WARPS; w++) tot += cnt[w][threadIdx.x];
            u32 all;
            u32 run = block_exclusive_scan<u32>(threadIdx.x < RS_RADIX ? tot : 0u, sm, all);
            if (threadIdx.x < RS_RADIX)
                for (int w = 0; w < RS_SMALL_WARPS; w++)
                {
                    const u32 c         = cnt[w][threadIdx.x];
                    cnt[w][threadIdx.x] = run;
                    run += c;
                }
        }
        __syncthreads();
        for (u32 jb = beg; jb < end; jb += 32 * 4)
        {   // four rounds of 32 keys loaded ahead, then ranked round by round (input order)
            u32 kk[4], vv[4];
#pragma unroll
            for (int i = 0; i < 4; i++)
            {
                const u32 j = jb + 32 * i + lane;
                kk[i]       = (j < end) ? kin[j] : 0u;
                vv[i]       = (j < end && vin) ? vin[j] : j;
            }
#pragma unroll
            for (int i = 0; i < 4; i++)
            {
                const u32 j = jb + 32 * i + lane;
                if (jb + 32 * i >= end) break;  // warp-uniform
                const bool valid = j < end;
                const u32  key   = valid ? min(kk[i], clampv) : 0u;
                const u32  digit = (key >> shift) & (RS_RADIX - 1);
                const u32  peers = __match_any_sync(MMG_FULL, valid ? digit : (u32)RS_RADIX);
                const int  leader = __ffs(peers) - 1;
                u32        base  = 0;
                if (valid && lane == leader)
                {
                    base             = cnt[warp][digit];
                    cnt[warp][digit] = base + __popc(peers);
                }
                base = __shfl_sync(MMG_FULL, base, leader);
                if (valid)
                {
                    const u32 pos = base + __popc(peers & ((1u << lane) - 1));
                    kout[pos]     = key;
                    vout[pos]     = vv[i];
                }
                __syncwarp();
            }
        }
        __syncthreads();
        u32* t = kin;
        kin    = kout;
        kout   = t;
        vin    = vout;
        vout   = (vout == va) ? vb : va;
    }
}

struct SortScratch
{
    u32 *keys_b, *vals_a, *vals_b, *ghist, *bsum;  // keys_b/vals_*: n each; ghist: 256*nblocks; bsum: scan scratch
};
inline size_t sort_nblocks(size_t n) { return (n + RS_TILE - 1) / RS_TILE; }

// Sorts keys (clamped to clampv) ascending, stable, carrying a u32 value per key:
// vals_in == nullptr means the original index (iota), otherwise the caller's values
// (e.g. the permutation of a previous sort, which chains several stable sorts into
// one multi-key sort). vals_in may be s.vals_a or s.vals_b (the result of a previous
// call with the same scratch). Returns pointers to the sorted keys / values (inside
// the scratch or `keys`, which is clobbered).
inline void radix_sort_pairs(u32* keys, const u32* vals_in, size_t n, u32 clampv, const SortScratch& s, u32** sorted_keys, u32** sorted_vals,
                             cudaStream_t st)
{
    int bits = 1;
    while (bits < 32 && (clampv >> bits) != 0) bits++;
    const int  passes = (bits + RS_BITS - 1) / RS_BITS;
    const u32  nb     = (u32)sort_nblocks(n);
    u32 *      kin = keys, *kout = s.keys_b;
    const u32* vin  = vals_in;
    u32*       vout = (vals_in == s.vals_a) ? s.vals_b : s.vals_a;
    if (n <= (size_t)RS_SMALL_MAX)
    {
        MMG_LAUNCH(k_radix_small, 1, RS_SMALL_THREADS, 0, st, kin, vin, kout, s.vals_a, s.vals_b, (u32)n, passes, clampv);
        for (int p = 0; p < passes; p++)
        {  // where the kernel's ping-pong ends
            u32* t = kin;
            kin    = kout;
            kout   = t;
            vin    = vout;
            vout   = (vout == s.vals_a) ? s.vals_b : s.vals_a;
        }
        *sorted_keys = kin;
        *sorted_vals = const_cast<u32*>(vin);
        return;
    }
    for (int p = 0; p < passes; p++)
    {
        const int shift = p * RS_BITS;
        MMG_LAUNCH(k_radix_hist, nb, RS_THREADS, 0, st, kin, n, shift, clampv, s.ghist, nb);
        exclusive_scan<u32, u32>(s.ghist, s.ghist, (size_t)RS_RADIX * nb, nullptr, s.bsum, st);
        MMG_LAUNCH(k_radix_scatter, nb, RS_THREADS, 0, st, kin, vin, kout, vout, n, shift, clampv, s.ghist, nb);
        u32* t = kin;
        kin    = kout;
        kout   = t;
        vin    = vout;
        vout   = (vout == s.vals_a) ? s.vals_b : s.vals_a;
    }
    *sorted_keys = kin;
    *sorted_vals = const_cast<u32*>(vin);
}
inline void radix_sort_index(u32* keys, size_t n, u32 clampv, const SortScratch& s, u32** sorted_keys, u32** sorted_idx, cudaStream_t st)
{
    radix_sort_pairs(keys, nullptr, n, clampv, s, sorted_keys, sorted_idx, st);
}
// u32 words a SortScratch needs for n items (keys_b, vals_a, vals_b, ghist, bsum)
inline size_t sort_scratch_words(size_t n)
{
    const size_t nb = sort_nblocks(n);
    return 3 * n + (size_t)RS_RADIX * nb + scan_scratch_elems((size_t)RS_RADIX * nb);
}
inline SortScratch sort_scratch_at(u32* base, size_t n)
{
    const size_t nb = sort_nblocks(n);
    return SortScratch{base, base + n, base + 2 * n, base + 3 * n, base + 3 * n + (size_t)RS_RADIX * nb};
}
}  // namespace mmg
