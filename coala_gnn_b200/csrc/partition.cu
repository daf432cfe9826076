// Kernels of the NCCL all-to-all comparison backend:
//   split_node_list : nccl_split_node_list_kernel (M/cache_kernel.cu:79-91) buckets ids by
//                     id % G with atomicAdd slots, so the order inside a bucket changes run to run.
//                     Here it is a stable partition (count -> scan -> scatter): bucket g holds its
//                     ids in request order, which also makes the owner-side cache state reproducible.
//   map_feat_data   : nccl_gather_feature_kernel + block_memcpy<ulonglong4> (M/cache_kernel.cu:113-137):
//                     out[map[r]] = src[r]; one warp per row, 16-byte vectors.
#include "common.cuh"
#include "../../include/coala_b200.h"

namespace {
constexpr int PT = 1024;  // ids per block

__device__ __forceinline__ int stable_rank_in_block(int g, int G, int (*warp_cnt)[COALA_MAX_GPUS], int* my_warp_base) {
    // rank of this thread among the block's threads with the same bucket; g < 0 = inactive
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int rank = 0;
    for (int j = 0; j < G; j++) {
        unsigned b = __ballot_sync(0xffffffffu, g == j);
        if (g == j) rank = __popc(b & ((1u << lane) - 1));
        if (lane == 0) warp_cnt[w][j] = __popc(b);
    }
    __syncthreads();
    if (threadIdx.x < G) {
        int acc = 0;
        for (int ww = 0; ww < PT / 32; ww++) { int x = warp_cnt[ww][threadIdx.x]; warp_cnt[ww][threadIdx.x] = acc; acc += x; }
        my_warp_base[threadIdx.x] = acc;  // block total for bucket threadIdx.x
    }
    __syncthreads();
    return g >= 0 ? warp_cnt[w][g] + rank : 0;
}

__global__ void __launch_bounds__(PT) part_count_kernel(const int64_t* __restrict__ idx, int64_t n, const int64_t* __restrict__ n_dev, int G,
                                                        int* __restrict__ block_cnt) {
    __shared__ int warp_cnt[PT / 32][COALA_MAX_GPUS];
    __shared__ int tot[COALA_MAX_GPUS];
    if (n_dev) n = min(n, *n_dev);
    const int64_t i = (int64_t)blockIdx.x * PT + threadIdx.x;
    const int g = i < n ? (int)((uint64_t)idx[i] % (uint64_t)G) : -1;
    stable_rank_in_block(g, G, warp_cnt, tot);
    if (threadIdx.x < G) block_cnt[(int64_t)blockIdx.x * G + threadIdx.x] = tot[threadIdx.x];
}

// counter[g * counter_stride] = size of bucket g (counter_stride > 1: the count is the header word of the bucket itself)
__global__ void part_scan_kernel(int* __restrict__ block_cnt, int nblocks, int G, int64_t* __restrict__ counter, int64_t counter_stride) {
    const int g = threadIdx.x;
    if (g >= G) return;
    int acc = 0;
    for (int b = 0; b < nblocks; b++) { int x = block_cnt[(int64_t)b * G + g]; block_cnt[(int64_t)b * G + g] = acc; acc += x; }
    counter[(int64_t)g * counter_stride] = acc;
}

// bucket g's ids start at node + g * node_stride, its positions at map + g * max_sample
__global__ void __launch_bounds__(PT) part_scatter_kernel(const int64_t* __restrict__ idx, int64_t n, const int64_t* __restrict__ n_dev, int G,
                                                          int max_sample, int64_t node_stride, const int* __restrict__ block_base,
                                                          int64_t* __restrict__ node, int64_t* __restrict__ map) {
    __shared__ int warp_cnt[PT / 32][COALA_MAX_GPUS];
    __shared__ int tot[COALA_MAX_GPUS];
    if (n_dev) n = min(n, *n_dev);
    const int64_t i = (int64_t)blockIdx.x * PT + threadIdx.x;
    int64_t id = 0;
    int g = -1;
    if (i < n) { id = idx[i]; g = (int)((uint64_t)id % (uint64_t)G); }
    const int r = stable_rank_in_block(g, G, warp_cnt, tot);
    if (g >= 0) {
        const int64_t k = block_base[(int64_t)blockIdx.x * G + g] + r;
        node[node_stride * g + k] = id;
        map[(int64_t)max_sample * g + k] = i;
    }
}

__global__ void unpermute_rows_kernel(float* __restrict__ out, const float* __restrict__ src, const int64_t* __restrict__ map,
                                      int64_t n, int dim) {
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int lane = threadIdx.x & 31;
    for (int64_t r = warp; r < n; r += nwarps) {
        const int64_t d = map[r];
        if ((dim & 3) == 0) {
            const uint4* s = reinterpret_cast<const uint4*>(src + r * dim);
            uint4* o = reinterpret_cast<uint4*>(out + d * dim);
            for (int v = lane; v < dim / 4; v += 32) st_out_v4(o + v, ld_host_v4(s + v));
        } else {
            for (int c = lane; c < dim; c += 32) out[d * dim + c] = src[r * dim + c];
        }
    }
}
}  // namespace

extern "C" int coala_split_node_list(uint64_t idx_ptr, int64_t n, uint64_t node_ptr, uint64_t map_ptr, uint64_t counter_ptr,
                                     int n_gpus, int max_sample, void* stream) {
    if (n_gpus < 1 || n_gpus > COALA_MAX_GPUS) COALA_FAIL("split_node_list: n_gpus out of range");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int64_t* counter = reinterpret_cast<int64_t*>(counter_ptr);
    if (n <= 0) { COALA_CUDA(cudaMemsetAsync(counter, 0, n_gpus * 8, st)); return 0; }
    const int nblocks = (int)((n + PT - 1) / PT);
    int* block_cnt = nullptr;
    COALA_CUDA(cudaMallocAsync(&block_cnt, (size_t)nblocks * n_gpus * sizeof(int), st));
    const int64_t* idx = reinterpret_cast<const int64_t*>(idx_ptr);
    part_count_kernel<<<nblocks, PT, 0, st>>>(idx, n, nullptr, n_gpus, block_cnt);
    part_scan_kernel<<<1, 32, 0, st>>>(block_cnt, nblocks, n_gpus, counter, 1);
    part_scatter_kernel<<<nblocks, PT, 0, st>>>(idx, n, nullptr, n_gpus, max_sample, max_sample, block_cnt, reinterpret_cast<int64_t*>(node_ptr),
                                                reinterpret_cast<int64_t*>(map_ptr));
    COALA_CUDA(cudaGetLastError());
    COALA_CUDA(cudaFreeAsync(block_cnt, st));
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// coala_split_node_list_static: the same stable partition for callers that are not allowed to allocate.
//
// Who that is: code that drives this library from inside somebody else's memory budget - a host that has handed the
// whole GPU to one pool allocator (a framework's caching allocator, an RMM pool, a MIG slice sized to the byte) and
// forbids any side allocation from a plugin, or a real-time / inference loop in which cudaMalloc / cudaMallocAsync are
// banned because they may take driver locks, grow a pool or synchronise the device at an unpredictable moment.
// coala_split_node_list above needs nblocks * G ints between its three kernels and gets them from cudaMallocAsync;
// this variant keeps that scratch in the library image instead.
//
// The scratch is the file-scope array below: STATIC_MAX_BLOCKS * COALA_MAX_GPUS ints = 128 KiB, which every GPU that
// loads this library carries until the process exits.  It bounds one call at COALA_SPLIT_STATIC_MAX_IDS
// (2 097 152) ids - above the largest request list any configuration in BASELINE.json produces (1 362 944).
//
// `g_part_static_scratch` as written in host code is only the host-side shadow of the symbol, not something a kernel can
// dereference, so the device address is looked up with cudaGetSymbolAddress - once per device, with that device
// current - and cached in s_static_scratch[dev].  That look-up happens in the FIRST call on each device (it may load
// the module on that device), so make that first call during set-up, outside any stream capture.
//
// Every call on a device uses the same array and the array keeps whatever the previous call left in it.  Nothing has
// to start at zero: part_count_kernel writes all nblocks * G words before part_scan_kernel / part_scatter_kernel read
// them, so left-over contents are never observed.  What the sharing does forbid is overlap: two calls whose kernels
// can be on the GPU at the same time (different streams, or two host threads) race on the array and corrupt each
// other's buckets - keep calls on one device stream-ordered one after the other.
namespace {
constexpr int STATIC_MAX_BLOCKS = 2048;                          // * PT ids per call
constexpr int64_t STATIC_MAX_IDS = (int64_t)STATIC_MAX_BLOCKS * PT;  // 2 097 152
constexpr int STATIC_MAX_DEVICES = 64;
static __device__ __align__(256) int g_part_static_scratch[STATIC_MAX_BLOCKS * COALA_MAX_GPUS];
std::mutex s_static_mu;
int* s_static_scratch[STATIC_MAX_DEVICES] = {};  // device address of the array on device d, nullptr = not looked up yet

// device address of g_part_static_scratch on the current device (set-up on first use; no allocation)
int static_scratch_for_current_device(int** out) {
    int dev = -1;
    COALA_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= STATIC_MAX_DEVICES) COALA_FAIL("split_node_list_static: device ordinal out of range");
    std::lock_guard<std::mutex> lk(s_static_mu);
    if (!s_static_scratch[dev]) {
        int* p = nullptr;
        COALA_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_part_static_scratch));
        if (!p) COALA_FAIL("split_node_list_static: cudaGetSymbolAddress returned a null address");
        s_static_scratch[dev] = p;
    }
    *out = s_static_scratch[dev];
    return 0;
}
}  // namespace

extern "C" int64_t coala_split_node_list_static_max_ids(void) { return STATIC_MAX_IDS; }

extern "C" int coala_split_node_list_static(uint64_t idx_ptr, int64_t n, uint64_t node_ptr, uint64_t map_ptr, uint64_t counter_ptr,
                                            int n_gpus, int max_sample, void* stream) {
    if (n_gpus < 1 || n_gpus > COALA_MAX_GPUS) COALA_FAIL("split_node_list_static: n_gpus out of range");
    if (n > STATIC_MAX_IDS)
        COALA_FAIL("split_node_list_static: n = " + std::to_string(n) + " ids exceeds the built-in scratch (at most " +
                   std::to_string(STATIC_MAX_IDS) + " ids per call); use coala_split_node_list or split the list");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int64_t* counter = reinterpret_cast<int64_t*>(counter_ptr);
    int* block_cnt = nullptr;
    if (static_scratch_for_current_device(&block_cnt) != 0) return -1;  // also on the n <= 0 path: set-up is call-count independent
    if (n <= 0) { COALA_CUDA(cudaMemsetAsync(counter, 0, n_gpus * 8, st)); return 0; }
    const int nblocks = (int)((n + PT - 1) / PT);
    const int64_t* idx = reinterpret_cast<const int64_t*>(idx_ptr);
    part_count_kernel<<<nblocks, PT, 0, st>>>(idx, n, nullptr, n_gpus, block_cnt);
    part_scan_kernel<<<1, 32, 0, st>>>(block_cnt, nblocks, n_gpus, counter, 1);
    part_scatter_kernel<<<nblocks, PT, 0, st>>>(idx, n, nullptr, n_gpus, max_sample, max_sample, block_cnt, reinterpret_cast<int64_t*>(node_ptr),
                                                reinterpret_cast<int64_t*>(map_ptr));
    COALA_CUDA(cudaGetLastError());
    return 0;
}

// internal form used by the NCCL exchange (nccl_backend.cu): row count on the device, bucket g = [count | ids...] at
// buckets + g * (1 + max_sample), caller-owned scratch (no allocation on the hot path)
int coala_partition_buckets(const int64_t* idx, int64_t n_upper, const int64_t* n_dev, int64_t* buckets, int64_t* map, int n_gpus,
                            int max_sample, int* block_cnt_scratch, cudaStream_t st) {
    const int nblocks = (int)((n_upper + PT - 1) / PT);
    if (nblocks == 0) {
        COALA_CUDA(cudaMemsetAsync(buckets, 0, (size_t)n_gpus * (1 + (size_t)max_sample) * 8, st));
        return 0;
    }
    part_count_kernel<<<nblocks, PT, 0, st>>>(idx, n_upper, n_dev, n_gpus, block_cnt_scratch);
    part_scan_kernel<<<1, 32, 0, st>>>(block_cnt_scratch, nblocks, n_gpus, buckets, 1 + (int64_t)max_sample);
    part_scatter_kernel<<<nblocks, PT, 0, st>>>(idx, n_upper, n_dev, n_gpus, max_sample, 1 + (int64_t)max_sample, block_cnt_scratch, buckets + 1, map);
    COALA_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int coala_map_feat_data(uint64_t out_ptr, uint64_t src_ptr, uint64_t map_ptr, int64_t n, int feat_dim, void* stream) {
    if (n <= 0) return 0;
    int64_t blocks = (n + 7) / 8;
    if (blocks > coala_sm_count() * 8) blocks = coala_sm_count() * 8;
    unpermute_rows_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<float*>(out_ptr), reinterpret_cast<const float*>(src_ptr), reinterpret_cast<const int64_t*>(map_ptr), n, feat_dim);
    COALA_CUDA(cudaGetLastError());
    return 0;
}
