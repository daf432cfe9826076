// GPU mode of the offline colour-metadata generator.  The host implementation in
// coloring.cpp stays the bit-exact reference mode (M/graph_coloring.cpp); this file is the fast mode:
//
//   * colouring: the seeds are drawn on the host exactly like the reference (same rand() stream, same
//     colour ids — coala_gc_draw_train_seeds), then the traversal runs on the device as a
//     level-synchronous multi-source search over in-neighbours.  The reference pops a LIFO work list
//     (M/graph_coloring.cpp:143-157), so which seed wins a contested node depends on its visit order;
//     here the smallest colour id among the claimants of a level wins (atomicMin), which makes the
//     result independent of thread order.  Level structure kept from the reference: level 1 = seeds
//     plus their uncoloured TRAINING in-neighbours (:125-141), levels 2..max_hop = in-neighbours of the
//     previous level; hop[] = level; the reference's final flush drains an empty list and is omitted.
//   * affinity (cpu_calculate_color_affinity, :254-294): score(c, nc) = sum over edges v<-u with
//     color[v] = c, color[u] = nc of exp(-0.5 hop[u]), top-k per colour, divided by the colour's
//     in-edge count.  The device version counts edges per (c, nc, hop) with a radix sort + run-length
//     encode (integers, order independent), then score = sum_h count_h * exp(-0.5 h) in ascending h with
//     explicitly rounded multiply / add, exp() values taken from the host's libm.  Ties are broken by
//     the smaller neighbour colour (the reference: std::sort over unordered_map order).  Scores agree
//     with the reference's edge-order summation to ~1e-15 relative, not bit for bit; the same loop
//     bounds are kept (colours 1..num_c-1 get a row, :266).
#include "../../include/coala_b200.h"
#include "coloring.h"
#include "common.cuh"
#include <algorithm>
#include <cmath>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_run_length_encode.cuh>

namespace {

constexpr uint32_t NO_CLAIM = 0xFFFFFFFFu;
constexpr int HOP_BITS = 4, COLOR_BITS = 28;  // key = c << 32 | nc << 4 | hop

struct DevBuf {  // frees on scope exit (every early return of the entry points)
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    template <class T> T* as() { return reinterpret_cast<T*>(p); }
};

int stage_graph(coala_gc* g, int device) {
    if (g->gpu_device == device && g->d_indptr) return 0;
    coala_gc_gpu_release(g);
    if (!g->indptr || !g->indices) COALA_FAIL("Graph_Coloring: set_adj_csc first");
    COALA_CUDA(cudaSetDevice(device));
    const uint64_t n = g->num_nodes, e = g->indptr[n];
    COALA_CUDA(cudaMalloc(&g->d_indptr, (n + 1) * 8));
    COALA_CUDA(cudaMalloc(&g->d_indices, (e ? e : 1) * 8));
    COALA_CUDA(cudaMemcpy(g->d_indptr, g->indptr, (n + 1) * 8, cudaMemcpyHostToDevice));
    COALA_CUDA(cudaMemcpy(g->d_indices, g->indices, e * 8, cudaMemcpyHostToDevice));
    g->gpu_device = device;
    return 0;
}

__global__ void gc_seed_kernel(const uint32_t* __restrict__ seed_node, const uint32_t* __restrict__ seed_color, uint32_t ns,
                               uint64_t* color, uint16_t* hop, uint32_t* frontier) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ns) return;
    const uint32_t v = seed_node[i];
    color[v] = seed_color[i];
    hop[v] = 1;
    frontier[i] = v;
}

// warp per frontier node: every uncoloured in-neighbour is claimed with atomicMin(colour); the first
// claimant appends it to `next`
template <bool TRAIN_ONLY>
__global__ void gc_expand_kernel(const uint32_t* frontier, uint32_t nf, const uint64_t* __restrict__ indptr,
                                 const uint64_t* __restrict__ indices, const uint64_t* color, const uint8_t* __restrict__ is_train,
                                 uint32_t* claim, uint32_t* next, uint32_t* n_next) {
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= nf) return;
    const uint32_t v = frontier[w];
    const uint32_t c = (uint32_t)color[v];
    const uint64_t b = indptr[v], e = indptr[v + 1];
    for (uint64_t k = b + lane; k < e; k += 32) {
        const uint64_t u = indices[k];
        if (TRAIN_ONLY && !is_train[u]) continue;
        if (color[u] != 0) continue;
        if (atomicMin(&claim[u], c) == NO_CLAIM) next[atomicAdd(n_next, 1u)] = (uint32_t)u;
    }
}

__global__ void gc_commit_kernel(const uint32_t* __restrict__ nodes, uint32_t n, const uint32_t* __restrict__ claim, uint64_t* color,
                                 uint16_t* hop, uint16_t level) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t u = nodes[i];
    color[u] = claim[u];
    hop[u] = level;
}

// warp per node v: in-edge count of its colour, and one key per edge whose far end has another colour
__global__ void aff_keys_kernel(uint64_t n, const uint64_t* __restrict__ indptr, const uint64_t* __restrict__ indices,
                                const uint64_t* __restrict__ color, const uint16_t* __restrict__ hop, uint64_t c_lo, uint64_t c_hi,
                                unsigned long long* in_edges, uint64_t* keys, unsigned long long cap, unsigned long long* n_keys) {
    const uint64_t v = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (v >= n) return;
    const uint64_t c = color[v];
    if (c < c_lo || c >= c_hi) return;  // this pass's colour range; c_hi <= num_c: the reference's loop stops one colour short (M/graph_coloring.cpp:266)
    const uint64_t b = indptr[v], e = indptr[v + 1];
    if (lane == 0 && e > b) atomicAdd(&in_edges[c], (unsigned long long)(e - b));
    for (uint64_t k0 = b; k0 < e; k0 += 32) {
        const uint64_t k = k0 + lane;
        uint64_t key = 0;
        bool take = false;
        if (k < e) {
            const uint64_t u = indices[k];
            const uint64_t nc = color[u];
            take = nc != 0 && nc != c;
            key = (c << 32) | (nc << HOP_BITS) | (uint64_t)hop[u];
        }
        const unsigned m = __ballot_sync(0xffffffffu, take);
        if (m == 0) continue;
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(n_keys, (unsigned long long)__popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        const unsigned long long pos = base + __popc(m & ((1u << lane) - 1));
        if (take && pos < cap) keys[pos] = key;  // an overflowing pass is detected from the counter and refused
    }
}

// runs are unique (c, nc, hop) triples in ascending order; the first run of every (c, nc) pair sums the pair's score
__global__ void aff_pair_kernel(const uint64_t* __restrict__ run_key, const uint32_t* __restrict__ run_cnt, const int* __restrict__ n_runs,
                                const double* __restrict__ exp_tab, double* pair_score) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n = *n_runs;
    if (i >= n) return;
    const uint64_t pair = run_key[i] >> HOP_BITS;
    if (i > 0 && (run_key[i - 1] >> HOP_BITS) == pair) { pair_score[i] = -1.0; return; }
    double s = 0.0;
    for (int64_t j = i; j < n && (run_key[j] >> HOP_BITS) == pair; j++)
        s = __dadd_rn(s, __dmul_rn((double)run_cnt[j], exp_tab[run_key[j] & ((1u << HOP_BITS) - 1)]));
    pair_score[i] = s;
}

__device__ __forceinline__ int64_t lower_bound_key(const uint64_t* a, int64_t n, uint64_t x) {
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (a[mid] < x) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// (score, nc) order: higher score first, smaller colour on equal scores
__device__ __forceinline__ bool better(double s1, uint64_t c1, double s2, uint64_t c2) { return s1 > s2 || (s1 == s2 && c1 < c2); }

// warp per colour: k rounds of "best entry that comes after the previous pick"
__global__ void aff_topk_kernel(uint64_t c_lo, uint64_t c_hi, int topk, const uint64_t* __restrict__ run_key, const double* __restrict__ pair_score,
                                const int* __restrict__ n_runs, const unsigned long long* __restrict__ in_edges, uint64_t* topk_color,
                                double* topk_aff) {
    const uint64_t c = (((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) + c_lo;
    const unsigned lane = threadIdx.x & 31;
    if (c >= c_hi) return;
    const int64_t n = *n_runs;
    const int64_t lo = lower_bound_key(run_key, n, c << 32), hi = lower_bound_key(run_key, n, (c + 1) << 32);
    double last_s = INFINITY;
    uint64_t last_c = 0;
    const double denom = (double)in_edges[c];
    for (int r = 0; r < topk; r++) {
        double best_s = -1.0;
        uint64_t best_c = ~0ULL;
        for (int64_t j = lo + lane; j < hi; j += 32) {
            const double s = pair_score[j];
            if (s < 0.0) continue;
            const uint64_t nc = (run_key[j] >> HOP_BITS) & ((1ULL << COLOR_BITS) - 1);
            if (r > 0 && !better(last_s, last_c, s, nc)) continue;  // already picked (or equal to the last pick)
            if (best_c == ~0ULL || better(s, nc, best_s, best_c)) { best_s = s; best_c = nc; }
        }
        for (int off = 16; off; off >>= 1) {
            const double os = __shfl_xor_sync(0xffffffffu, best_s, off);
            const uint64_t oc = __shfl_xor_sync(0xffffffffu, best_c, off);
            if (oc != ~0ULL && (best_c == ~0ULL || better(os, oc, best_s, best_c))) { best_s = os; best_c = oc; }
        }
        if (best_c == ~0ULL) break;
        if (lane == 0) {
            topk_color[(c - 1) * topk + r] = best_c;
            if (topk_aff) topk_aff[(c - 1) * topk + r] = best_s / denom;
        }
        last_s = best_s;
        last_c = best_c;
    }
}

inline unsigned blocks_for(uint64_t threads, int tpb) { return (unsigned)((threads + tpb - 1) / tpb); }

}  // namespace

void coala_gc_gpu_release(coala_gc* g) {
    if (g->gpu_device < 0) return;
    cudaSetDevice(g->gpu_device);
    if (g->d_indptr) cudaFree(g->d_indptr);
    if (g->d_indices) cudaFree(g->d_indices);
    g->d_indptr = g->d_indices = nullptr;
    g->gpu_device = -1;
}

extern "C" int coala_gc_gpu_color_graph_optimized(coala_gc* g, uint64_t train_ptr, uint64_t n_train, int device) {
    if (!g || !g->color) COALA_FAIL("Graph_Coloring: set_color_buffer first");
    const uint64_t n = g->num_nodes;
    if (n >= 0xFFFFFFFFull) COALA_FAIL("Graph_Coloring GPU mode: node ids must fit 32 bits");
    if (stage_graph(g, device)) return -1;
    coala_gc_draw_train_seeds(g, reinterpret_cast<const int64_t*>(train_ptr), n_train);
    if (g->next_color >= (1ull << COLOR_BITS)) COALA_FAIL("Graph_Coloring GPU mode: too many colours");
    const uint32_t ns = (uint32_t)g->work[0].size();
    std::vector<uint32_t> h_node(ns), h_col(ns);
    for (uint32_t i = 0; i < ns; i++) { h_node[i] = (uint32_t)g->work[0][i].first; h_col[i] = (uint32_t)g->work[0][i].second; }
    g->work[0].clear();

    DevBuf color, hop, claim, train, fa, fb, cnt, sn, sc;
    COALA_CUDA(cudaMalloc(&color.p, n * 8));
    COALA_CUDA(cudaMalloc(&hop.p, n * 2));
    COALA_CUDA(cudaMalloc(&claim.p, n * 4));
    COALA_CUDA(cudaMalloc(&train.p, n));
    COALA_CUDA(cudaMalloc(&fa.p, (n + 1) * 4));
    COALA_CUDA(cudaMalloc(&fb.p, (n + 1) * 4));
    COALA_CUDA(cudaMalloc(&cnt.p, 4));
    COALA_CUDA(cudaMalloc(&sn.p, (ns + 1) * 4));
    COALA_CUDA(cudaMalloc(&sc.p, (ns + 1) * 4));
    COALA_CUDA(cudaMemcpy(color.p, g->color, n * 8, cudaMemcpyHostToDevice));
    COALA_CUDA(cudaMemcpy(hop.p, g->hop.data(), n * 2, cudaMemcpyHostToDevice));
    COALA_CUDA(cudaMemset(claim.p, 0xFF, n * 4));
    COALA_CUDA(cudaMemcpy(train.p, g->is_train.data(), n, cudaMemcpyHostToDevice));
    COALA_CUDA(cudaMemcpy(sn.p, h_node.data(), ns * 4, cudaMemcpyHostToDevice));
    COALA_CUDA(cudaMemcpy(sc.p, h_col.data(), ns * 4, cudaMemcpyHostToDevice));

    uint64_t colored = ns;
    uint32_t* cur = fa.as<uint32_t>();
    uint32_t* nxt = fb.as<uint32_t>();
    uint32_t nf = ns;
    if (ns && g->max_hop >= 1) {
        gc_seed_kernel<<<blocks_for(ns, 256), 256>>>(sn.as<uint32_t>(), sc.as<uint32_t>(), ns, color.as<uint64_t>(), hop.as<uint16_t>(), cur);
        // level 1: training in-neighbours of the seeds join the seeds' level (appended behind them)
        COALA_CUDA(cudaMemcpy(cnt.p, &ns, 4, cudaMemcpyHostToDevice));
        gc_expand_kernel<true><<<blocks_for((uint64_t)ns * 32, 256), 256>>>(cur, ns, g->d_indptr, g->d_indices, color.as<uint64_t>(),
                                                                          train.as<uint8_t>(), claim.as<uint32_t>(), cur, cnt.as<uint32_t>());
        COALA_CUDA(cudaMemcpy(&nf, cnt.p, 4, cudaMemcpyDeviceToHost));
        if (nf > ns) gc_commit_kernel<<<blocks_for(nf - ns, 256), 256>>>(cur + ns, nf - ns, claim.as<uint32_t>(), color.as<uint64_t>(), hop.as<uint16_t>(), 1);
        colored += nf - ns;
        for (int level = 2; level <= g->max_hop && nf; level++) {
            COALA_CUDA(cudaMemset(cnt.p, 0, 4));
            gc_expand_kernel<false><<<blocks_for((uint64_t)nf * 32, 256), 256>>>(cur, nf, g->d_indptr, g->d_indices, color.as<uint64_t>(),
                                                                               train.as<uint8_t>(), claim.as<uint32_t>(), nxt, cnt.as<uint32_t>());
            COALA_CUDA(cudaMemcpy(&nf, cnt.p, 4, cudaMemcpyDeviceToHost));
            if (nf) gc_commit_kernel<<<blocks_for(nf, 256), 256>>>(nxt, nf, claim.as<uint32_t>(), color.as<uint64_t>(), hop.as<uint16_t>(), (uint16_t)level);
            colored += nf;
            uint32_t* t = cur; cur = nxt; nxt = t;
        }
    }
    COALA_CUDA(cudaGetLastError());
    COALA_CUDA(cudaMemcpy(g->color, color.p, n * 8, cudaMemcpyDeviceToHost));
    COALA_CUDA(cudaMemcpy(g->hop.data(), hop.p, n * 2, cudaMemcpyDeviceToHost));
    g->num_colored += colored;
    return 0;
}

extern "C" int coala_gc_gpu_calculate_color_affinity(coala_gc* g, int device) {
    if (!g || !g->color || !g->topk_color) COALA_FAIL("Graph_Coloring: set_color_buffer / set_topk_color_buffer first");
    if (g->max_hop >= (1 << HOP_BITS)) COALA_FAIL("Graph_Coloring GPU mode: hop does not fit the key");
    if (g->next_color >= (1ull << COLOR_BITS)) COALA_FAIL("Graph_Coloring GPU mode: too many colours");
    if (stage_graph(g, device)) return -1;
    const uint64_t n = g->num_nodes, n_edges = g->indptr[n], num_c = g->next_color - 1;
    if (num_c < 2) return 0;
    const int K = g->topk;
    double h_exp[1 << HOP_BITS];
    for (int h = 0; h < (1 << HOP_BITS); h++) h_exp[h] = std::exp(-0.5 * (double)h);  // score_func (M/graph_coloring.cpp:250-252)

    // The sort runs in passes over disjoint colour ranges so that one pass holds fewer than 2^31 keys (cub's
    // item counts are int) and at most ~2^30 by default; every pass scans the edges once.  COALA_GC_PARTS forces
    // the number of passes (tests use it to exercise the multi-pass path on small graphs).
    uint64_t parts = (n_edges >> 30) + 1;
    if (const char* e = getenv("COALA_GC_PARTS")) parts = strtoull(e, nullptr, 10) > 0 ? strtoull(e, nullptr, 10) : parts;
    if (parts > num_c - 1) parts = num_c - 1;
    // key capacity of a pass: all edges when there is one pass, otherwise twice the even share (colour sizes are skewed)
    uint64_t cap = n_edges ? n_edges : 1;
    if (parts > 1) cap = std::min<uint64_t>(cap, 2 * (n_edges / parts) + (1 << 20));
    if (cap >= 0x7FFFFFFFull) COALA_FAIL("Graph_Coloring GPU mode: a sort pass would exceed 2^31 keys; raise COALA_GC_PARTS");

    DevBuf color, hop, in_edges, keys, keys_sorted, n_keys, run_cnt, n_runs, score, tmp, tmp2, exp_tab, d_topk, d_aff;
    COALA_CUDA(cudaMalloc(&color.p, n * 8));
    COALA_CUDA(cudaMalloc(&hop.p, n * 2));
    COALA_CUDA(cudaMalloc(&in_edges.p, (num_c + 1) * 8));
    COALA_CUDA(cudaMalloc(&keys.p, cap * 8));
    COALA_CUDA(cudaMalloc(&keys_sorted.p, cap * 8));
    COALA_CUDA(cudaMalloc(&run_cnt.p, cap * 4));
    COALA_CUDA(cudaMalloc(&score.p, cap * 8));
    COALA_CUDA(cudaMalloc(&n_keys.p, 8));
    COALA_CUDA(cudaMalloc(&n_runs.p, 4));
    COALA_CUDA(cudaMalloc(&exp_tab.p, sizeof(h_exp)));
    COALA_CUDA(cudaMalloc(&d_topk.p, num_c * K * 8));
    COALA_CUDA(cudaMalloc(&d_aff.p, num_c * K * 8));
    size_t tb = 0, tb2 = 0;  // scratch sized for the largest possible pass
    COALA_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, tb, keys.as<uint64_t>(), keys_sorted.as<uint64_t>(), (int)cap, 0, 32 + COLOR_BITS));
    COALA_CUDA(cub::DeviceRunLengthEncode::Encode(nullptr, tb2, keys_sorted.as<uint64_t>(), keys.as<uint64_t>(), run_cnt.as<uint32_t>(),
                                                  n_runs.as<int>(), (int)cap));
    COALA_CUDA(cudaMalloc(&tmp.p, tb ? tb : 1));
    COALA_CUDA(cudaMalloc(&tmp2.p, tb2 ? tb2 : 1));
    COALA_CUDA(cudaMemcpy(color.p, g->color, n * 8, cudaMemcpyHostToDevice));
    COALA_CUDA(cudaMemcpy(hop.p, g->hop.data(), n * 2, cudaMemcpyHostToDevice));
    COALA_CUDA(cudaMemset(in_edges.p, 0, (num_c + 1) * 8));
    COALA_CUDA(cudaMemcpy(exp_tab.p, h_exp, sizeof(h_exp), cudaMemcpyHostToDevice));
    // rows the kernels do not touch keep the caller's contents, like the reference
    COALA_CUDA(cudaMemcpy(d_topk.p, g->topk_color, num_c * K * 8, cudaMemcpyHostToDevice));
    if (g->topk_aff) COALA_CUDA(cudaMemcpy(d_aff.p, g->topk_aff, num_c * K * 8, cudaMemcpyHostToDevice));

    for (uint64_t part = 0; part < parts; part++) {
        // colours 1 .. num_c-1 split evenly
        const uint64_t c_lo = 1 + (num_c - 1) * part / parts, c_hi = 1 + (num_c - 1) * (part + 1) / parts;
        if (c_hi <= c_lo) continue;
        COALA_CUDA(cudaMemset(n_keys.p, 0, 8));
        aff_keys_kernel<<<blocks_for(n * 32, 256), 256>>>(n, g->d_indptr, g->d_indices, color.as<uint64_t>(), hop.as<uint16_t>(), c_lo, c_hi,
                                                          in_edges.as<unsigned long long>(), keys.as<uint64_t>(), cap,
                                                          n_keys.as<unsigned long long>());
        unsigned long long nk = 0;
        COALA_CUDA(cudaMemcpy(&nk, n_keys.p, 8, cudaMemcpyDeviceToHost));
        if (nk > cap) COALA_FAIL("Graph_Coloring GPU mode: a colour range holds more edges than a sort pass; raise COALA_GC_PARTS");
        if (nk == 0) continue;
        size_t t1 = tb, t2 = tb2;
        COALA_CUDA(cub::DeviceRadixSort::SortKeys(tmp.p, t1, keys.as<uint64_t>(), keys_sorted.as<uint64_t>(), (int)nk, 0, 32 + COLOR_BITS));
        // the unsorted key buffer is free again: it receives the run keys
        COALA_CUDA(cub::DeviceRunLengthEncode::Encode(tmp2.p, t2, keys_sorted.as<uint64_t>(), keys.as<uint64_t>(), run_cnt.as<uint32_t>(),
                                                      n_runs.as<int>(), (int)nk));
        aff_pair_kernel<<<blocks_for(nk, 256), 256>>>(keys.as<uint64_t>(), run_cnt.as<uint32_t>(), n_runs.as<int>(), exp_tab.as<double>(),
                                                      score.as<double>());
        aff_topk_kernel<<<blocks_for((c_hi - c_lo) * 32, 256), 256>>>(c_lo, c_hi, K, keys.as<uint64_t>(), score.as<double>(), n_runs.as<int>(),
                                                                      in_edges.as<unsigned long long>(), d_topk.as<uint64_t>(),
                                                                      g->topk_aff ? d_aff.as<double>() : nullptr);
        COALA_CUDA(cudaGetLastError());
    }
    COALA_CUDA(cudaDeviceSynchronize());
    COALA_CUDA(cudaMemcpy(g->topk_color, d_topk.p, num_c * K * 8, cudaMemcpyDeviceToHost));
    if (g->topk_aff) COALA_CUDA(cudaMemcpy(g->topk_aff, d_aff.p, num_c * K * 8, cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int coala_gc_get_hops(coala_gc* g, uint64_t dst_u16_ptr) {
    if (!g || !dst_u16_ptr) COALA_FAIL("coala_gc_get_hops: null argument");
    memcpy(reinterpret_cast<void*>(dst_u16_ptr), g->hop.data(), g->hop.size() * 2);
    return 0;
}
