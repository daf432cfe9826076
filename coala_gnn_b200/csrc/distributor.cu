// Stage 1: colour-affinity seed distribution on the GPU.
// Replaces Node_distributor_pybind (M/node_distributor_pybind.cuh:112-238), whose
// distribute_node_with_affinity (:150-222) is a serial CPU loop.  The result must be bit-exact:
// scores are double, accumulated in k-order with separate multiply and add roundings
// (__dmul_rn / __dadd_rn: no FMA, like the reference built by g++ for x86-64), the maximum is the
// first strict one starting from (-1.0, domain 0), a full bucket scores -1.
//
// Parallel formulation: the only sequential dependence of the greedy loop is "is bucket j full
// yet".  Scores are computed for all (seed, domain) pairs at once; one CTA then walks the seed
// list in tiles: every seed takes its arg-max among the buckets that are not full at the start of
// the tile, a per-bucket prefix count gives its position inside the bucket, and the first seed
// whose bucket reaches capacity ends the tile there (the following seeds are re-evaluated with
// that bucket closed).  At most N/1024 + D tile steps instead of N*D*K serial steps.
#include "common.cuh"
#include "distributor.h"
#include "npy.h"
#include <vector>

namespace {

constexpr int ND_MAX_DOMAINS = 16;
constexpr int ND_TILE = 1024;
constexpr int ND_SLOTS = 8;

struct MetaPtrs {
    const int32_t* p[ND_MAX_DOMAINS];
};

// scores[j*N + i] = sum_k meta_j[topk[c,k]] * score[c,k]   (M/node_distributor_pybind.cuh:172-203)
// Domain-major: consecutive threads of the assign kernel (consecutive seeds) then read consecutive doubles.  With the
// seed-major layout every warp load touched 32 lines and one tile step cost 4 us at D = 8, 10 us at D = 16
// (profiles/r2_launches_distributor_config_shapes.csv: 56 / 290 us per call).
__global__ void nd_score_kernel(const int64_t* __restrict__ items, int64_t N, int D, int K,
                                const int32_t* __restrict__ color, const int32_t* __restrict__ topk,
                                const double* __restrict__ score, MetaPtrs meta, int meta_len,
                                double* __restrict__ scores, int* __restrict__ neg_flag) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= N * D) return;
    int64_t i = t % N;
    int j = (int)(t / N);
    int64_t id = items[i];
    int c = color[id];
    double s = 0.0;
    if (c != 0) {
        const int32_t* tk = topk + (int64_t)(c - 1) * K;
        const double* sc = score + (int64_t)(c - 1) * K;
        const int32_t* m = meta.p[j];
        for (int k = 0; k < K; k++) {
            int nc = tk[k];
            if (nc != 0) {
                int mv = nc < meta_len ? m[nc] : 0;
                if (mv == 0) continue;
                s = __dadd_rn(s, __dmul_rn((double)mv, sc[k]));
            }
        }
    }
    scores[t] = s;
    if (s < 0.0) *neg_flag = 1;  // negative counters: the tiled walk below is not valid, use the serial kernel
}

// One CTA of ND_TILE threads.
__global__ void __launch_bounds__(ND_TILE) nd_assign_kernel(const int64_t* __restrict__ items, int64_t N, int D, int cap,
                                                            int node_id, const double* __restrict__ scores,
                                                            int64_t* __restrict__ out, int* __restrict__ bucket_len_out,
                                                            const int* __restrict__ neg_flag) {
    if (*neg_flag) return;  // nd_assign_serial_kernel takes over
    __shared__ int warp_cnt[ND_TILE / 32][ND_MAX_DOMAINS];
    __shared__ int run[ND_MAX_DOMAINS];  // == bucket_len of the reference loop
    __shared__ int s_p;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid < ND_MAX_DOMAINS) run[tid] = 0;
    __syncthreads();
    int64_t start = 0;
    while (start < N) {
        const int64_t i = start + tid;
        const bool valid = i < N;
        int c = -1;
        if (valid) {
            int best = 0;
            double best_score = -1.0;
            for (int j = 0; j < D; j++) {
                double s = scores[(int64_t)j * N + i];
                if (run[j] == cap) s = -1.0;                 // (:205-206)
                if (s > best_score) { best = j; best_score = s; }  // (:208-211)
            }
            c = best;
        }
        if (tid == 0) s_p = 0x7fffffff;
        int myrank = 0;
        for (int j = 0; j < D; j++) {
            unsigned b = __ballot_sync(0xffffffffu, c == j);
            if (c == j) myrank = __popc(b & ((1u << lane) - 1));
            if (lane == 0) warp_cnt[w][j] = __popc(b);
        }
        __syncthreads();
        if (tid < D) {  // exclusive prefix over the 32 warps for bucket `tid`
            int acc = 0;
            for (int ww = 0; ww < ND_TILE / 32; ww++) { int x = warp_cnt[ww][tid]; warp_cnt[ww][tid] = acc; acc += x; }
        }
        __syncthreads();
        int pos = -1;
        if (valid) {
            pos = run[c] + warp_cnt[w][c] + myrank;
            if (pos + 1 == cap) atomicMin(&s_p, tid);  // this seed fills its bucket
        }
        __syncthreads();
        const int p = s_p;  // tile-relative index of the first seed that closes a bucket
        const bool commit = valid && tid <= p;
        if (commit && c == node_id && pos < cap) out[pos] = items[i];  // (:214-217)
        __syncthreads();  // everyone has read run[] / s_p
        for (int j = 0; j < D; j++) {
            unsigned b = __ballot_sync(0xffffffffu, commit && c == j);
            if (lane == 0 && b) atomicAdd(&run[j], __popc(b));
        }
        __syncthreads();
        start += (p == 0x7fffffff) ? ND_TILE : (p + 1);
    }
    if (tid < D && bucket_len_out) bucket_len_out[tid] = run[tid];
}

// Exact serial restatement (one thread).  Used only when a score is negative — i.e. a cache
// colour counter went below zero — where the reference's "bucket_len == cap" test lets an
// over-full bucket 0 re-open and the tiled walk above does not apply.
__global__ void nd_assign_serial_kernel(const int64_t* __restrict__ items, int64_t N, int D, int cap, int node_id,
                                        const double* __restrict__ scores, int64_t* __restrict__ out,
                                        int* __restrict__ bucket_len_out, const int* __restrict__ neg_flag) {
    if (threadIdx.x != 0 || blockIdx.x != 0 || !*neg_flag) return;
    int len[ND_MAX_DOMAINS];
    for (int j = 0; j < D; j++) len[j] = 0;
    for (int64_t i = 0; i < N; i++) {
        int best = 0;
        double best_score = -1.0;
        for (int j = 0; j < D; j++) {
            double s = scores[(int64_t)j * N + i];
            if (len[j] == cap) s = -1.0;
            if (s > best_score) { best = j; best_score = s; }
        }
        if (best == node_id && len[best] < cap) out[len[best]] = items[i];
        len[best]++;
    }
    if (bucket_len_out) for (int j = 0; j < D; j++) bucket_len_out[j] = len[j];
}

// D == 1: every seed scores >= 0 > -1 until the single bucket is full, and the bucket holds exactly
// N = cap seeds, so the reference loop degenerates to out[i] = item[offset + i] (:214-218).
__global__ void nd_single_domain_kernel(const int64_t* __restrict__ items, int cap, int64_t* __restrict__ out,
                                        int* __restrict__ bucket_len_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < cap) out[i] = items[i];
    if (i == 0 && bucket_len_out) bucket_len_out[0] = cap;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// host object
// ---------------------------------------------------------------------------------------------
struct coala_nd {
    int device = 0;
    uint64_t item_ptr = 0;
    bool items_on_device = false;
    int node_id = 0, batch = 0, local_size = 0, n_domains = 0;
    int cap = 0;          // domain_batch_size  (:146)
    int64_t N = 0;        // global_batch_size  (:147)
    bool has_color = false;
    int64_t n_graph_nodes = 0;
    int num_colors = 0, K = 0, meta_len = 0;
    int64_t* h_color = nullptr;   // pinned host copy, int64 like the reference's color_buffer_ptr
    int32_t* d_color = nullptr;
    int32_t* d_topk = nullptr;
    double* d_score = nullptr;
    int64_t* d_items = nullptr;   // staging for host-resident seed lists
    // scratch of one distribution: slot 0 belongs to the synchronous entry point, the others are handed out
    // to callers that keep several distributions in flight on different streams (one per prefetch slot)
    double* d_scores[ND_SLOTS] = {nullptr};
    int* d_flags[ND_SLOTS] = {nullptr};   // [0] negative-score flag, [1..D] bucket_len
    int32_t* d_meta = nullptr;    // staging [D][meta_len]
    int64_t* d_out = nullptr;
    int h_bucket_len[ND_MAX_DOMAINS] = {0};
    cudaStream_t stream = nullptr;
};

const int32_t* coala_nd_color_dev_i32(coala_nd* h) { return h ? h->d_color : nullptr; }

static int nd_slot_scratch(coala_nd* h, int slot) {
    if (slot < 0 || slot >= ND_SLOTS) COALA_FAIL("distribute: scratch slot must be in [0,8)");
    if (h->d_scores[slot]) return 0;
    COALA_CUDA(cudaMalloc(&h->d_scores[slot], (size_t)h->N * h->n_domains * 8));
    COALA_CUDA(cudaMalloc(&h->d_flags[slot], (1 + ND_MAX_DOMAINS) * sizeof(int)));
    return 0;
}

static int nd_init_tables(coala_nd* h, const int64_t* color, int64_t n_nodes, const int64_t* topk, const double* score,
                          int num_colors, int K) {
    COALA_CUDA(cudaSetDevice(h->device));
    h->has_color = true;
    h->n_graph_nodes = n_nodes;
    h->num_colors = num_colors;       // num_colors = topk_shape[0] (:143)
    h->K = K;
    h->meta_len = num_colors;         // callers size their histograms num_colors (S/Training_node_distributor.py:33)
    h->cap = h->batch * h->local_size;
    h->N = (int64_t)h->cap * h->n_domains;
    COALA_CUDA(cudaHostAlloc(&h->h_color, (size_t)n_nodes * 8, cudaHostAllocMapped));   // (:23)
    memcpy(h->h_color, color, (size_t)n_nodes * 8);
    std::vector<int32_t> c32((size_t)n_nodes), t32((size_t)num_colors * K);
    for (int64_t i = 0; i < n_nodes; i++) {
        if (color[i] < 0 || color[i] > num_colors) COALA_FAIL("color.npy holds a colour id outside [0, num_colors]");
        c32[i] = (int32_t)color[i];
    }
    for (int64_t i = 0; i < (int64_t)num_colors * K; i++) {
        if (topk[i] < 0 || topk[i] > num_colors) COALA_FAIL("topk.npy holds a colour id outside [0, num_colors]");
        t32[i] = (int32_t)topk[i];
    }
    COALA_CUDA(cudaMalloc(&h->d_color, (size_t)n_nodes * 4));
    COALA_CUDA(cudaMemcpy(h->d_color, c32.data(), (size_t)n_nodes * 4, cudaMemcpyHostToDevice));
    COALA_CUDA(cudaMalloc(&h->d_topk, std::max<size_t>(4, (size_t)num_colors * K * 4)));
    COALA_CUDA(cudaMemcpy(h->d_topk, t32.data(), (size_t)num_colors * K * 4, cudaMemcpyHostToDevice));
    COALA_CUDA(cudaMalloc(&h->d_score, std::max<size_t>(8, (size_t)num_colors * K * 8)));
    COALA_CUDA(cudaMemcpy(h->d_score, score, (size_t)num_colors * K * 8, cudaMemcpyHostToDevice));
    COALA_CUDA(cudaMalloc(&h->d_items, (size_t)h->N * 8));
    COALA_CUDA(cudaMalloc(&h->d_meta, (size_t)h->n_domains * (num_colors + 1) * 4));
    COALA_CUDA(cudaMemset(h->d_meta, 0, (size_t)h->n_domains * (num_colors + 1) * 4));
    COALA_CUDA(cudaMalloc(&h->d_out, (size_t)h->cap * 8));
    COALA_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    return nd_slot_scratch(h, 0);
}

static coala_nd* nd_new(uint64_t item_ptr, int node_id, int batch, int local_size, int n_domains, int device) {
    if (n_domains < 1 || n_domains > ND_MAX_DOMAINS) { coala_set_error("distributor: n_domains must be in [1,16]"); return nullptr; }
    if (batch < 1 || local_size < 1 || node_id < 0 || node_id >= n_domains) { coala_set_error("distributor: bad batch/local_size/node_id"); return nullptr; }
    coala_nd* h = new coala_nd();
    h->device = device; h->item_ptr = item_ptr; h->node_id = node_id; h->batch = batch;
    h->local_size = local_size; h->n_domains = n_domains;
    return h;
}

extern "C" coala_nd* coala_nd_create_from_arrays(uint64_t item_ptr, int node_id, int batch, int local_size, int n_domains,
                                                 const int64_t* color, int64_t n_nodes, const int64_t* topk,
                                                 const double* score, int num_colors, int K, int device) {
    coala_nd* h = nd_new(item_ptr, node_id, batch, local_size, n_domains, device);
    if (!h) return nullptr;
    if (nd_init_tables(h, color, n_nodes, topk, score, num_colors, K)) { coala_nd_destroy(h); return nullptr; }
    return h;
}

extern "C" coala_nd* coala_nd_create(uint64_t item_ptr, int node_id, int batch, int local_size, int n_domains,
                                     const char* color_file, const char* topk_file, const char* score_file, int device) {
    try {
        NpyArray color = npy_load(color_file);   // 1-D int64  (:141)
        NpyArray topk = npy_load(topk_file);     // 2-D int64  (:142)
        NpyArray score = npy_load(score_file);   // float64    (:144)
        if (color.descr != "<i8" || color.shape.size() != 1) throw std::runtime_error("color.npy must be a 1-D '<i8' array");
        if (topk.descr != "<i8" || topk.shape.size() != 2) throw std::runtime_error("topk.npy must be a 2-D '<i8' array");
        if (score.descr != "<f8" || score.numel() != topk.numel()) throw std::runtime_error("score.npy must be '<f8' with topk's element count");
        return coala_nd_create_from_arrays(item_ptr, node_id, batch, local_size, n_domains,
                                           static_cast<const int64_t*>(color.data()), color.shape[0],
                                           static_cast<const int64_t*>(topk.data()), static_cast<const double*>(score.data()),
                                           (int)topk.shape[0], (int)topk.shape[1], device);
    } catch (const std::exception& e) {
        coala_set_error(e.what());
        return nullptr;
    }
}

extern "C" coala_nd* coala_nd_create_plain(uint64_t item_ptr, int n_domains) {
    coala_nd* h = new coala_nd();
    h->item_ptr = item_ptr; h->n_domains = n_domains;
    return h;
}

extern "C" void coala_nd_destroy(coala_nd* h) {
    if (!h) return;
    if (h->has_color) cudaSetDevice(h->device);
    if (h->stream) { cudaStreamSynchronize(h->stream); cudaStreamDestroy(h->stream); }
    if (h->h_color) cudaFreeHost(h->h_color);
    cudaFree(h->d_color); cudaFree(h->d_topk); cudaFree(h->d_score); cudaFree(h->d_items);
    for (int k = 0; k < ND_SLOTS; k++) { cudaFree(h->d_scores[k]); cudaFree(h->d_flags[k]); }
    cudaFree(h->d_meta); cudaFree(h->d_out);
    delete h;
}

extern "C" int coala_nd_num_colors(coala_nd* h) { return h->num_colors; }
extern "C" uint64_t coala_nd_color_host_ptr(coala_nd* h) { return (uint64_t)h->h_color; }
extern "C" int64_t coala_nd_num_graph_nodes(coala_nd* h) { return h->n_graph_nodes; }
extern "C" int coala_nd_set_meta_len(coala_nd* h, int len) {
    if (len < 1 || len > h->num_colors + 1) COALA_FAIL("meta_len must be in [1, num_colors+1]");
    h->meta_len = len;
    return 0;
}
extern "C" int coala_nd_last_bucket_len(coala_nd* h, int* out) {
    for (int j = 0; j < h->n_domains; j++) out[j] = h->h_bucket_len[j];
    return 0;
}

static int nd_launch(coala_nd* h, const int64_t* d_items, const MetaPtrs& meta, int64_t* d_out, int slot, cudaStream_t st) {
    if (nd_slot_scratch(h, slot)) return -1;
    double* d_scores = h->d_scores[slot];
    int* d_flags = h->d_flags[slot];
    if (h->n_domains == 1) {
        nd_single_domain_kernel<<<(h->cap + 255) / 256, 256, 0, st>>>(d_items, h->cap, d_out, d_flags + 1);
        COALA_CUDA(cudaGetLastError());
        return 0;
    }
    const int64_t total = h->N * h->n_domains;
    COALA_CUDA(cudaMemsetAsync(d_flags, 0, (1 + ND_MAX_DOMAINS) * sizeof(int), st));
    nd_score_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(d_items, h->N, h->n_domains, h->K, h->d_color, h->d_topk,
                                                                   h->d_score, meta, h->meta_len, d_scores, d_flags);
    COALA_CUDA(cudaGetLastError());
    // both assignment kernels are enqueued; each checks the flag so that no host round trip is needed
    nd_assign_kernel<<<1, ND_TILE, 0, st>>>(d_items, h->N, h->n_domains, h->cap, h->node_id, d_scores, d_out, d_flags + 1, d_flags);
    nd_assign_serial_kernel<<<1, 32, 0, st>>>(d_items, h->N, h->n_domains, h->cap, h->node_id, d_scores, d_out, d_flags + 1, d_flags);
    COALA_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int coala_nd_distribute_async(coala_nd* h, uint64_t offset, uint64_t out_dev_ptr, const uint64_t* meta_dev_ptrs,
                                         int n_meta, void* stream) {
    return coala_nd_distribute_async_slot(h, offset, out_dev_ptr, meta_dev_ptrs, n_meta, 1, stream);
}

extern "C" int coala_nd_distribute_async_slot(coala_nd* h, uint64_t offset, uint64_t out_dev_ptr, const uint64_t* meta_dev_ptrs,
                                              int n_meta, int slot, void* stream) {
    // all pointers are device addresses; the seed list itself must be device-resident too
    if (!h->has_color) COALA_FAIL("Error: Node_distributor_pybind is not created with color information.");
    if (n_meta != h->n_domains) COALA_FAIL("distribute: need one cache-meta pointer per domain");
    COALA_CUDA(cudaSetDevice(h->device));
    MetaPtrs meta;
    for (int j = 0; j < ND_MAX_DOMAINS; j++) meta.p[j] = j < n_meta ? reinterpret_cast<const int32_t*>(meta_dev_ptrs[j]) : nullptr;
    return nd_launch(h, reinterpret_cast<const int64_t*>(h->item_ptr) + offset, meta, reinterpret_cast<int64_t*>(out_dev_ptr),
                     slot, static_cast<cudaStream_t>(stream));
}

extern "C" int coala_nd_distribute(coala_nd* h, uint64_t offset, uint64_t out_ptr, const uint64_t* meta_ptrs, int n_meta) {
    // the reference prints and returns when built without colour tables (:153-156); here it is an error
    if (!h->has_color) COALA_FAIL("Error: Node_distributor_pybind is not created with color information.");
    if (n_meta != h->n_domains) COALA_FAIL("distribute: need one cache-meta pointer per domain");
    COALA_CUDA(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    const int64_t* items = reinterpret_cast<const int64_t*>(h->item_ptr) + offset;
    const int64_t* d_items = items;
    if (!coala_is_device_ptr(items)) {
        COALA_CUDA(cudaMemcpyAsync(h->d_items, items, (size_t)h->N * 8, cudaMemcpyHostToDevice, st));
        d_items = h->d_items;
    }
    MetaPtrs meta;
    for (int j = 0; j < ND_MAX_DOMAINS; j++) meta.p[j] = nullptr;
    for (int j = 0; j < n_meta; j++) {
        const void* m = reinterpret_cast<const void*>(meta_ptrs[j]);
        if (coala_is_device_ptr(m)) meta.p[j] = static_cast<const int32_t*>(m);
        else {
            int32_t* dst = h->d_meta + (size_t)j * (h->num_colors + 1);
            COALA_CUDA(cudaMemcpyAsync(dst, m, (size_t)h->meta_len * 4, cudaMemcpyHostToDevice, st));
            meta.p[j] = dst;
        }
    }
    void* out = reinterpret_cast<void*>(out_ptr);
    const bool out_dev = coala_is_device_ptr(out);
    int64_t* d_out = out_dev ? static_cast<int64_t*>(out) : h->d_out;
    if (nd_launch(h, d_items, meta, d_out, 0, st)) return -1;
    if (!out_dev) COALA_CUDA(cudaMemcpyAsync(out, h->d_out, (size_t)h->cap * 8, cudaMemcpyDeviceToHost, st));
    COALA_CUDA(cudaMemcpyAsync(h->h_bucket_len, h->d_flags[0] + 1, h->n_domains * sizeof(int), cudaMemcpyDeviceToHost, st));
    COALA_CUDA(cudaStreamSynchronize(st));
    return 0;
}
