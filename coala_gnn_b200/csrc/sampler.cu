// Stage 2: multi-layer uniform neighbour sampling + to_block over a CSC graph that lives in
// pinned host memory (UVA) — the work the reference delegates to DGL's GPU sampler at
// sampler.sample(g, seeds) (S/COALA_GNN_DataLoader.py:162) on tensors that alias pinned host
// memory (M/shared_UVA.cuh:65-100, examples/ssd_gnn_dataloader.py:496-524).
//
// Semantics (oracle/c/sampler.c states them on the CPU, integer-identical):
//   per layer, in reversed fan-out order:  deg <= f -> all in-edges in CSC order;
//   deg > f -> f distinct CSC positions from a partial Fisher-Yates keyed by (rng_key, layer, node),
//   emitted in ascending position;  to_block: dst = frontier in order, src = frontier then unseen
//   sources by first appearance in the edge list.
//
// B200 mapping: every dependent host round trip (indptr pair -> neighbour ids) crosses PCIe, so
// the sampling kernel keeps one warp per frontier node with all of a row's reads issued at once
// and enough warps resident to cover the link latency; everything after the two host reads
// (scan, hash insert, first-occurrence numbering) runs out of HBM/L2.  No host synchronisation
// between layers: frontier sizes stay on the device.
#include "common.cuh"
#include "../../include/coala_b200.h"
#include <vector>

namespace {

constexpr int MAX_LAYERS = 8;
constexpr int MAX_FANOUT = 32;
constexpr uint64_t FY_MUL = 0xD1342543DE82EF95ULL;
constexpr uint64_t LAYER_MUL = 0xA24BAED4963EE407ULL;
constexpr int POS_INF = 0x7fffffff;

struct HashTable {
    unsigned long long* keys;  // 0xFFFF.. = empty
    int* vals;                 // minimum position seen
    uint64_t mask;
};

__device__ __forceinline__ void ht_insert_min(const HashTable& t, int64_t key, int pos) {
    uint64_t h = mix64((uint64_t)key) & t.mask;
    while (true) {
        unsigned long long k = t.keys[h];
        if (k == ~0ULL) {
            unsigned long long old = atomicCAS(t.keys + h, ~0ULL, (unsigned long long)key);
            if (old == ~0ULL || old == (unsigned long long)key) break;
        } else if (k == (unsigned long long)key) {
            break;
        }
        h = (h + 1) & t.mask;
    }
    atomicMin(t.vals + h, pos);
}

__device__ __forceinline__ int ht_lookup(const HashTable& t, int64_t key) {
    uint64_t h = mix64((uint64_t)key) & t.mask;
    while (true) {
        unsigned long long k = t.keys[h];
        if (k == (unsigned long long)key) return t.vals[h];
        if (k == ~0ULL) return POS_INF;
        h = (h + 1) & t.mask;
    }
}

__global__ void ht_reset_kernel(HashTable t) {
    const uint64_t n = t.mask + 1;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        t.keys[i] = ~0ULL;
        t.vals[i] = POS_INF;
    }
}

// ---- K1: one warp per frontier node: indptr pair, picks, neighbour ids into fixed-stride slots
__global__ void __launch_bounds__(256) sample_rows_kernel(const int64_t* __restrict__ indptr, const int64_t* __restrict__ indices,
                                                          const int64_t* __restrict__ frontier, const int64_t* __restrict__ n_front,
                                                          int fanout, const uint64_t* __restrict__ batch_key, int layer, HashTable ht,
                                                          int64_t* __restrict__ slot_src, int64_t* __restrict__ slot_eid,
                                                          int* __restrict__ cnt) {
    const int lane = threadIdx.x & 31;
    const int64_t ns = *n_front;
    const uint64_t layer_key = mix64(*batch_key + (uint64_t)(layer + 1) * LAYER_MUL);
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < ns; i += nwarps) {
        const int64_t v = frontier[i];
        int64_t p = 0;
        if (lane < 2) p = indptr[v + lane];  // both words of the pair in one request
        const int64_t start = __shfl_sync(0xffffffffu, p, 0);
        const int64_t deg = __shfl_sync(0xffffffffu, p, 1) - start;
        if (lane == 0) ht_insert_min(ht, v, (int)i);  // dst nodes take local ids 0..ns-1
        // The host link serves ~154 M independent requests/s whatever their size up to 256 B (measured,
        // scripts/probe_links.py), so a row is fetched as ONE coalesced span (<= 64 neighbours: two
        // 256-B requests at most) issued before the picks are computed, instead of one request per pick.
        constexpr int ROW_SPAN = 64;
        const bool whole_row = deg <= ROW_SPAN;
        int64_t r0 = 0, r1 = 0;
        if (whole_row) {
            if (lane < deg) r0 = indices[start + lane];
            if (lane + 32 < deg) r1 = indices[start + 32 + lane];
        }
        int64_t mypos = -1, myval = 0;
        int k;
        if (deg <= fanout) {
            k = (int)deg;
            if (lane < k) { mypos = lane; myval = r0; }  // fanout <= 32 < ROW_SPAN: the row is in r0
        } else {
            k = fanout;
            // partial Fisher-Yates over the virtual array a[x] = x; the sparse swap map lives one
            // entry per lane (mk -> mv), looked up with a ballot
            const uint64_t row_key = mix64(layer_key ^ mix64((uint64_t)v));
            int64_t myj = 0;
            if (lane < k) {
                uint64_t r = mix64(row_key + (uint64_t)(lane + 1) * FY_MUL);
                myj = lane + (int64_t)__umul64hi(r, (uint64_t)(deg - lane));
            }
            int64_t mk = -1, mv = 0, mypick = 0;
            int mn = 0;
            for (int i2 = 0; i2 < k; i2++) {
                const int64_t j = __shfl_sync(0xffffffffu, myj, i2);
                const unsigned bj = __ballot_sync(0xffffffffu, lane < mn && mk == j);
                const unsigned bi = __ballot_sync(0xffffffffu, lane < mn && mk == (int64_t)i2);
                const int lj = __ffs(bj) - 1, li = __ffs(bi) - 1;
                const int64_t vj_m = __shfl_sync(0xffffffffu, mv, lj < 0 ? 0 : lj);
                const int64_t vi_m = __shfl_sync(0xffffffffu, mv, li < 0 ? 0 : li);
                const int64_t vj = bj ? vj_m : j;
                const int64_t vi = bi ? vi_m : (int64_t)i2;
                if (lane == i2) mypick = vj;
                if (bj) { if (lane == lj) mv = vi; }
                else { if (lane == mn) { mk = j; mv = vi; } mn++; }
            }
            // ascending CSC position: rank = number of smaller picks (all distinct)
            int rank = 0;
            for (int u = 0; u < k; u++) {
                const int64_t pu = __shfl_sync(0xffffffffu, mypick, u);
                rank += (lane < k && pu < mypick) ? 1 : 0;
            }
            // lane `lane` holds pick with rank `rank`; route it so that lane r holds the r-th smallest
            int64_t sorted = 0;
            for (int u = 0; u < k; u++) {
                const int ru = __shfl_sync(0xffffffffu, rank, u);
                const int64_t pu = __shfl_sync(0xffffffffu, mypick, u);
                if (ru == lane) sorted = pu;
            }
            if (lane < k) mypos = sorted;
            if (whole_row) {  // pick the neighbour ids out of the registers that hold the row
                for (int u = 0; u < k; u++) {
                    const int pu = (int)__shfl_sync(0xffffffffu, sorted, u);
                    const int64_t a = __shfl_sync(0xffffffffu, r0, pu & 31);
                    const int64_t b = __shfl_sync(0xffffffffu, r1, pu & 31);
                    if (lane == u) myval = pu < 32 ? a : b;
                }
            } else if (lane < k) {
                myval = indices[start + mypos];
            }
        }
        if (lane < k) {
            slot_src[i * fanout + lane] = myval;
            slot_eid[i * fanout + lane] = start + mypos;  // CSC position; the edge id is looked up on demand (coala_lookup_edge_ids)
        }
        if (lane == 0) cnt[i] = k;
    }
}

// ---- single-CTA exclusive scan of int32.  Thread t owns the contiguous chunk [t*per, (t+1)*per):
// two sweeps over L2-resident data around one block-level scan of the chunk sums (frontier / edge
// counts are <= ~1e6 here, so per stays small and no inter-block carry is needed).  512 threads x
// 32 registers on purpose: the CTA must fit next to the resident gather CTAs of the previous batch
// (a 1024-thread CTA cannot, and would wait for a gather CTA to retire — four times per batch).
constexpr int SCAN_T = 512;
__global__ void __launch_bounds__(SCAN_T) scan_kernel(const int* __restrict__ in, int* __restrict__ out, const int64_t* __restrict__ n_ptr,
                                                      int64_t mul, int64_t* __restrict__ total_out, const int64_t* __restrict__ add_ptr,
                                                      int64_t* __restrict__ total_plus_out) {
    __shared__ int warp_sum[SCAN_T / 32];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int64_t n = *n_ptr * mul;
    const int64_t per = (((n + SCAN_T - 1) / SCAN_T) + 3) & ~(int64_t)3;  // multiple of 4: chunks start 16-B aligned
    const int64_t lo = min(n, (int64_t)tid * per), hi = min(n, lo + per);
    int sum = 0;
    {
        int64_t i = lo;
        for (; i + 4 <= hi; i += 4) {
            const int4 x = *reinterpret_cast<const int4*>(in + i);
            sum += x.x + x.y + x.z + x.w;
        }
        for (; i < hi; i++) sum += in[i];
    }
    int incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += y;
    }
    if (lane == 31) warp_sum[w] = incl;
    __syncthreads();
    if (w == 0) {
        int sw = lane < SCAN_T / 32 ? warp_sum[lane] : 0, si = sw;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, si, d);
            if (lane >= d) si += y;
        }
        if (lane < SCAN_T / 32) warp_sum[lane] = si - sw;  // exclusive warp offsets
    }
    __syncthreads();
    int run = warp_sum[w] + incl - sum;  // exclusive prefix of this thread's chunk
    {
        int64_t i = lo;
        for (; i + 4 <= hi; i += 4) {
            const int4 x = *reinterpret_cast<const int4*>(in + i);
            int4 o;
            o.x = run; o.y = run + x.x; o.z = o.y + x.y; o.w = o.z + x.z;
            run = o.w + x.w;
            *reinterpret_cast<int4*>(out + i) = o;
        }
        for (; i < hi; i++) { const int x = in[i]; out[i] = run; run += x; }
    }
    if (tid == SCAN_T - 1) {  // run == grand total (chunks past n are empty)
        if (total_out) *total_out = run;
        if (total_plus_out) *total_plus_out = run + *add_ptr;
    }
}

// ---- prologue: the only node whose arguments change from batch to batch (seed list, count, RNG key);
// everything after it reads these from device memory, so the rest of the sample is a fixed CUDA graph
__global__ void prologue_kernel(const int64_t* __restrict__ seeds, int64_t ns, uint64_t rng_key, int64_t* __restrict__ frontier0,
                                int64_t* __restrict__ n_front0, uint64_t* __restrict__ key_out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < ns; i += (int64_t)gridDim.x * blockDim.x) frontier0[i] = seeds[i];
    if (blockIdx.x == 0 && threadIdx.x == 0) { *n_front0 = ns; *key_out = rng_key; }
}

// ---- K3: compact the fixed-stride slots into the edge list and insert sources (min position)
__global__ void compact_insert_kernel(const int64_t* __restrict__ slot_src, const int64_t* __restrict__ slot_eid,
                                      const int* __restrict__ cnt, const int* __restrict__ off, const int64_t* __restrict__ n_front,
                                      int fanout, HashTable ht, int64_t* __restrict__ e_src, int64_t* __restrict__ e_dst,
                                      int64_t* __restrict__ e_id) {
    const int64_t ns = *n_front;
    const int64_t total = ns * fanout;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = t / fanout;
        const int j = (int)(t % fanout);
        if (j < cnt[i]) {
            const int64_t e = off[i] + j;
            const int64_t src = slot_src[t];
            e_src[e] = src;
            e_dst[e] = i;
            e_id[e] = slot_eid[t];
            ht_insert_min(ht, src, (int)(ns + e));
        }
    }
}

// ---- on demand: edge ids of the sampled CSC positions (the reference's DGL sampler reads them
// while sampling; the SAGE/GAT consumers never look at them, so they are fetched only when asked for)
__global__ void edge_id_lookup_kernel(const int64_t* __restrict__ pos, const int64_t* __restrict__ n_ptr, int64_t n_host,
                                      const int64_t* __restrict__ edge_ids, int64_t* __restrict__ out) {
    const int64_t n = n_ptr ? *n_ptr : n_host;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t p = pos[e];
        out[e] = edge_ids ? edge_ids[p] : p;
    }
}

// ---- K4: an edge is the first occurrence of its source iff the table's minimum is its own position
__global__ void flag_first_kernel(const int64_t* __restrict__ e_src, const int64_t* __restrict__ n_edges,
                                  const int64_t* __restrict__ n_front, HashTable ht, int* __restrict__ val, int* __restrict__ flag) {
    const int64_t ne = *n_edges, ns = *n_front;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < ne; e += (int64_t)gridDim.x * blockDim.x) {
        const int v = ht_lookup(ht, e_src[e]);
        val[e] = v;
        flag[e] = (int64_t)v == ns + e;
    }
}

// ---- K6: local source ids + the next frontier (dst nodes first, then new sources in first-occurrence order)
__global__ void relabel_kernel(const int64_t* __restrict__ e_src, const int* __restrict__ val, const int* __restrict__ flag,
                               const int* __restrict__ rank, const int64_t* __restrict__ n_edges, const int64_t* __restrict__ n_front,
                               const int64_t* __restrict__ frontier, int64_t* __restrict__ e_src_local,
                               int64_t* __restrict__ next_frontier) {
    const int64_t ne = *n_edges, ns = *n_front;
    const int64_t total = ne > ns ? ne : ns;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        if (t < ns) next_frontier[t] = frontier[t];
        if (t < ne) {
            const int64_t v = val[t];
            e_src_local[t] = v < ns ? v : ns + rank[v - ns];
            if (flag[t]) next_frontier[ns + rank[t]] = e_src[t];
        }
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// host object
// ---------------------------------------------------------------------------------------------
struct coala_sampler {
    int device = 0, L = 0, max_seeds = 0;
    int fanout[MAX_LAYERS];            // sampling order: fanout[0] is applied to the seeds
    const int64_t *indptr = nullptr, *indices = nullptr, *edge_ids = nullptr;
    int64_t num_nodes = 0;
    int64_t ns_max[MAX_LAYERS + 1];    // frontier capacity per sampling round
    int64_t ne_max[MAX_LAYERS];
    int64_t* frontier[MAX_LAYERS + 1] = {nullptr};
    int64_t *e_src[MAX_LAYERS] = {nullptr}, *e_dst[MAX_LAYERS] = {nullptr}, *e_id[MAX_LAYERS] = {nullptr},
            *e_src_local[MAX_LAYERS] = {nullptr};
    int64_t *slot_src = nullptr, *slot_eid = nullptr;
    int *cnt = nullptr, *off = nullptr, *val = nullptr, *flag = nullptr, *rank = nullptr;
    HashTable ht{nullptr, nullptr, 0};
    int64_t* d_counts = nullptr;       // [0..L] frontier sizes, [L+1 .. 2L] edge counts
    int64_t* h_counts = nullptr;       // pinned mirror
    cudaEvent_t done = nullptr;
    int sm_count = 148;
    int64_t* indptr_hbm = nullptr;     // optional HBM copy of indptr (coala_sampler_stage_indptr)
    uint64_t* d_key = nullptr;         // RNG key of the batch in flight
    // the whole sample as one CUDA graph: one launch instead of 16 (every launch is a command fetch over
    // the same host link the path keeps busy); only the prologue node's arguments are patched per batch
    bool use_graph = true;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t graph_exec = nullptr;
    cudaGraphNode_t prologue_node = nullptr;
    cudaStream_t cap_stream = nullptr;
    // diagnostics (COALA_SAMPLER_TRACE=1): one event after every enqueued operation of a sample
    bool trace = false;
    std::vector<cudaEvent_t> tev;
    int tev_used = 0;
    std::vector<double> tsum;
    int tcount = 0;
};

static void trace_mark(coala_sampler* s, cudaStream_t st) {
    if (!s->trace) return;
    if (s->tev_used == (int)s->tev.size()) { cudaEvent_t e; cudaEventCreate(&e); s->tev.push_back(e); }
    cudaEventRecord(s->tev[s->tev_used++], st);
}
static void trace_collect(coala_sampler* s) {  // fold the previous sample's events into the running sums
    if (!s->trace || s->tev_used < 2) return;
    if (cudaEventSynchronize(s->tev[s->tev_used - 1]) != cudaSuccess) return;
    if ((int)s->tsum.size() < s->tev_used - 1) s->tsum.resize(s->tev_used - 1, 0.0);
    for (int i = 0; i + 1 < s->tev_used; i++) { float ms = 0; cudaEventElapsedTime(&ms, s->tev[i], s->tev[i + 1]); s->tsum[i] += ms; }
    s->tcount++;
    s->tev_used = 0;
}
extern "C" int coala_sampler_trace(coala_sampler* s, double* avg_ms, int max_n) {
    trace_collect(s);
    int n = (int)s->tsum.size() < max_n ? (int)s->tsum.size() : max_n;
    for (int i = 0; i < n; i++) avg_ms[i] = s->tcount ? s->tsum[i] / s->tcount : 0.0;
    return n;
}

extern "C" coala_sampler* coala_sampler_create(int device, uint64_t indptr_ptr, uint64_t indices_ptr, uint64_t edge_ids_ptr,
                                               int64_t num_nodes, int max_seeds, const int* fanouts, int n_layers) {
    if (n_layers < 1 || n_layers > MAX_LAYERS) { coala_set_error("sampler: 1..8 layers supported"); return nullptr; }
    for (int l = 0; l < n_layers; l++)
        if (fanouts[l] < 1 || fanouts[l] > MAX_FANOUT) { coala_set_error("sampler: fan-out must be in [1,32]"); return nullptr; }
    if (max_seeds < 1 || !indptr_ptr || !indices_ptr) { coala_set_error("sampler: bad arguments"); return nullptr; }
    COALA_CUDA_NULL(cudaSetDevice(device));
    coala_sampler* s = new coala_sampler();
    s->device = device; s->L = n_layers; s->max_seeds = max_seeds; s->num_nodes = num_nodes;
    s->indptr = reinterpret_cast<const int64_t*>(indptr_ptr);
    s->indices = reinterpret_cast<const int64_t*>(indices_ptr);
    s->edge_ids = reinterpret_cast<const int64_t*>(edge_ids_ptr);
    // the reference samples `for fanout in reversed(fanouts)` (DGL NeighborSampler.sample_blocks)
    for (int l = 0; l < n_layers; l++) s->fanout[l] = fanouts[n_layers - 1 - l];
    s->ns_max[0] = max_seeds;
    int64_t slots_max = 0;
    for (int l = 0; l < n_layers; l++) {
        s->ne_max[l] = s->ns_max[l] * s->fanout[l];
        s->ns_max[l + 1] = s->ns_max[l] + s->ne_max[l];  // = batch * prod(f+1): max_sample_size (S/COALA_GNN_Manager.py:79-81)
        slots_max = std::max(slots_max, s->ne_max[l]);
    }
    if (s->ns_max[n_layers] > 0x3fffffff) { coala_set_error("sampler: sample too large for 32-bit positions"); delete s; return nullptr; }
    for (int l = 0; l <= n_layers; l++) COALA_CUDA_NULL(cudaMalloc(&s->frontier[l], (size_t)s->ns_max[l] * 8));
    for (int l = 0; l < n_layers; l++) {
        COALA_CUDA_NULL(cudaMalloc(&s->e_src[l], (size_t)s->ne_max[l] * 8));
        COALA_CUDA_NULL(cudaMalloc(&s->e_dst[l], (size_t)s->ne_max[l] * 8));
        COALA_CUDA_NULL(cudaMalloc(&s->e_id[l], (size_t)s->ne_max[l] * 8));
        COALA_CUDA_NULL(cudaMalloc(&s->e_src_local[l], (size_t)s->ne_max[l] * 8));
    }
    COALA_CUDA_NULL(cudaMalloc(&s->slot_src, (size_t)slots_max * 8));
    COALA_CUDA_NULL(cudaMalloc(&s->slot_eid, (size_t)slots_max * 8));
    COALA_CUDA_NULL(cudaMalloc(&s->cnt, (size_t)s->ns_max[n_layers - 1] * 4));
    COALA_CUDA_NULL(cudaMalloc(&s->off, (size_t)s->ns_max[n_layers - 1] * 4));
    COALA_CUDA_NULL(cudaMalloc(&s->val, (size_t)slots_max * 4));
    COALA_CUDA_NULL(cudaMalloc(&s->flag, (size_t)slots_max * 4));
    COALA_CUDA_NULL(cudaMalloc(&s->rank, (size_t)slots_max * 4));
    uint64_t cap = 1024;
    while (cap < 2 * (uint64_t)s->ns_max[n_layers]) cap <<= 1;
    s->ht.mask = cap - 1;
    COALA_CUDA_NULL(cudaMalloc(&s->ht.keys, cap * 8));
    COALA_CUDA_NULL(cudaMalloc(&s->ht.vals, cap * 4));
    COALA_CUDA_NULL(cudaMalloc(&s->d_key, 8));
    COALA_CUDA_NULL(cudaMalloc(&s->d_counts, (2 * MAX_LAYERS + 2) * 8));
    COALA_CUDA_NULL(cudaMemset(s->d_counts, 0, (2 * MAX_LAYERS + 2) * 8));
    COALA_CUDA_NULL(cudaHostAlloc(&s->h_counts, (2 * MAX_LAYERS + 2) * 8, cudaHostAllocDefault));
    memset(s->h_counts, 0, (2 * MAX_LAYERS + 2) * 8);
    COALA_CUDA_NULL(cudaEventCreateWithFlags(&s->done, cudaEventDisableTiming));
    cudaDeviceProp prop;
    COALA_CUDA_NULL(cudaGetDeviceProperties(&prop, device));
    s->sm_count = prop.multiProcessorCount;
    if (const char* e = getenv("COALA_SAMPLER_TRACE")) s->trace = atoi(e) != 0;
    if (const char* e = getenv("COALA_SAMPLER_GRAPH")) s->use_graph = atoi(e) != 0;
    if (s->trace) s->use_graph = false;  // per-kernel events only make sense for direct launches
    return s;
}

extern "C" void coala_sampler_destroy(coala_sampler* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    cudaDeviceSynchronize();
    for (int l = 0; l <= s->L; l++) cudaFree(s->frontier[l]);
    for (int l = 0; l < s->L; l++) { cudaFree(s->e_src[l]); cudaFree(s->e_dst[l]); cudaFree(s->e_id[l]); cudaFree(s->e_src_local[l]); }
    cudaFree(s->slot_src); cudaFree(s->slot_eid); cudaFree(s->cnt); cudaFree(s->off); cudaFree(s->val);
    cudaFree(s->indptr_hbm); cudaFree(s->d_key);
    if (s->graph_exec) cudaGraphExecDestroy(s->graph_exec);
    if (s->graph) cudaGraphDestroy(s->graph);
    if (s->cap_stream) cudaStreamDestroy(s->cap_stream);
    cudaFree(s->flag); cudaFree(s->rank); cudaFree(s->ht.keys); cudaFree(s->ht.vals); cudaFree(s->d_counts);
    if (s->h_counts) cudaFreeHost(s->h_counts);
    if (s->done) cudaEventDestroy(s->done);
    delete s;
}

extern "C" int64_t coala_sampler_max_sample(coala_sampler* s) { return s->ns_max[s->L]; }

extern "C" int coala_sampler_stage_indptr(coala_sampler* s) {
    if (s->indptr_hbm) return 0;
    COALA_CUDA(cudaSetDevice(s->device));
    COALA_CUDA(cudaMalloc(&s->indptr_hbm, (size_t)(s->num_nodes + 1) * 8));
    COALA_CUDA(cudaMemcpy(s->indptr_hbm, s->indptr, (size_t)(s->num_nodes + 1) * 8, cudaMemcpyDefault));
    s->indptr = s->indptr_hbm;
    if (s->graph_exec) { cudaGraphExecDestroy(s->graph_exec); s->graph_exec = nullptr; }
    if (s->graph) { cudaGraphDestroy(s->graph); s->graph = nullptr; }
    return 0;
}
extern "C" uint64_t coala_sampler_input_count_dev_ptr(coala_sampler* s) { return (uint64_t)(s->d_counts + s->L); }

static inline unsigned grid_for(int64_t work, int per_block, int cap) {
    int64_t b = (work + per_block - 1) / per_block;
    if (b < 1) b = 1;
    if (b > cap) b = cap;
    return (unsigned)b;
}

// enqueue every operation of one sample on `st` (directly, or while `st` is being captured)
static int enqueue_sample(coala_sampler* s, const int64_t* seeds, int64_t ns, uint64_t rng_key, cudaStream_t st) {
    const int L = s->L;
    int64_t* d_nf = s->d_counts;          // frontier sizes [0..L]
    int64_t* d_ne = s->d_counts + L + 1;  // edge counts [0..L-1]
    trace_mark(s, st);
    // the frontier size of round 0 is the only count the host knows
    prologue_kernel<<<4, 256, 0, st>>>(seeds, ns, rng_key, s->frontier[0], d_nf, s->d_key);
    trace_mark(s, st);
    const int cap = s->sm_count * 8;
    for (int l = 0; l < L; l++) {
        const int f = s->fanout[l];
        HashTable ht = s->ht;
        // table sized for this round only: keeps the reset cheap for the small early rounds
        uint64_t need = 1024;
        while (need < 2 * (uint64_t)s->ns_max[l + 1]) need <<= 1;
        ht.mask = need - 1;
        ht_reset_kernel<<<grid_for((int64_t)need, 1024, cap), 256, 0, st>>>(ht);
        trace_mark(s, st);
        sample_rows_kernel<<<grid_for(s->ns_max[l], 8, cap), 256, 0, st>>>(s->indptr, s->indices, s->frontier[l], d_nf + l, f,
                                                                          s->d_key, l, ht, s->slot_src, s->slot_eid, s->cnt);
        trace_mark(s, st);
        scan_kernel<<<1, SCAN_T, 0, st>>>(s->cnt, s->off, d_nf + l, 1, d_ne + l, nullptr, nullptr);
        trace_mark(s, st);
        compact_insert_kernel<<<grid_for(s->ne_max[l], 256, cap), 256, 0, st>>>(s->slot_src, s->slot_eid, s->cnt, s->off, d_nf + l, f, ht,
                                                                              s->e_src[l], s->e_dst[l], s->e_id[l]);
        trace_mark(s, st);
        flag_first_kernel<<<grid_for(s->ne_max[l], 256, cap), 256, 0, st>>>(s->e_src[l], d_ne + l, d_nf + l, ht, s->val, s->flag);
        trace_mark(s, st);
        scan_kernel<<<1, SCAN_T, 0, st>>>(s->flag, s->rank, d_ne + l, 1, nullptr, d_nf + l, d_nf + l + 1);
        trace_mark(s, st);
        relabel_kernel<<<grid_for(s->ne_max[l], 256, cap), 256, 0, st>>>(s->e_src[l], s->val, s->flag, s->rank, d_ne + l, d_nf + l,
                                                                        s->frontier[l], s->e_src_local[l], s->frontier[l + 1]);
        trace_mark(s, st);
    }
    COALA_CUDA(cudaGetLastError());
    COALA_CUDA(cudaMemcpyAsync(s->h_counts, s->d_counts, (2 * L + 1) * 8, cudaMemcpyDeviceToHost, st));
    trace_mark(s, st);
    return 0;
}

static int build_graph(coala_sampler* s) {
    if (!s->cap_stream) COALA_CUDA(cudaStreamCreateWithFlags(&s->cap_stream, cudaStreamNonBlocking));
    COALA_CUDA(cudaStreamBeginCapture(s->cap_stream, cudaStreamCaptureModeRelaxed));
    int rc = enqueue_sample(s, s->frontier[0], 0, 0, s->cap_stream);  // placeholder arguments, patched per launch
    cudaError_t e = cudaStreamEndCapture(s->cap_stream, &s->graph);
    if (rc) return -1;
    COALA_CUDA(e);
    size_t n_root = 1;
    COALA_CUDA(cudaGraphGetRootNodes(s->graph, &s->prologue_node, &n_root));
    if (n_root != 1) COALA_FAIL("sampler graph: expected a single root node");
    cudaGraphNodeType ty;
    COALA_CUDA(cudaGraphNodeGetType(s->prologue_node, &ty));
    if (ty != cudaGraphNodeTypeKernel) COALA_FAIL("sampler graph: root node is not the prologue kernel");
    COALA_CUDA(cudaGraphInstantiate(&s->graph_exec, s->graph, 0));
    return 0;
}

extern "C" int coala_sampler_sample_async(coala_sampler* s, uint64_t seeds_dev_ptr, int ns, uint64_t rng_key, void* stream) {
    if (ns < 0 || ns > s->max_seeds) COALA_FAIL("sampler: more seeds than max_seeds");
    COALA_CUDA(cudaSetDevice(s->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t* seeds = reinterpret_cast<const int64_t*>(seeds_dev_ptr);
    if (!s->use_graph) {
        trace_collect(s);
        if (enqueue_sample(s, seeds, ns, rng_key, st)) return -1;
    } else {
        if (!s->graph_exec && build_graph(s)) return -1;
        int64_t ns64 = ns;
        int64_t* f0 = s->frontier[0];
        int64_t* nf0 = s->d_counts;
        uint64_t* key = s->d_key;
        void* args[6] = {&seeds, &ns64, &rng_key, &f0, &nf0, &key};
        cudaKernelNodeParams kp;
        memset(&kp, 0, sizeof(kp));
        kp.func = reinterpret_cast<void*>(prologue_kernel);
        kp.gridDim = dim3(4);
        kp.blockDim = dim3(256);
        kp.sharedMemBytes = 0;
        kp.kernelParams = args;
        COALA_CUDA(cudaGraphExecKernelNodeSetParams(s->graph_exec, s->prologue_node, &kp));
        COALA_CUDA(cudaGraphLaunch(s->graph_exec, st));
    }
    COALA_CUDA(cudaEventRecord(s->done, st));
    return 0;
}

extern "C" int coala_sampler_counts(coala_sampler* s, int64_t* counts) {
    COALA_CUDA(cudaEventSynchronize(s->done));
    const int L = s->L;
    // reference block order: blocks[0] is the LAST sampled round
    for (int k = 0; k < L; k++) {
        const int l = L - 1 - k;
        counts[2 * k] = s->h_counts[l + 1];          // n_src
        counts[2 * k + 1] = s->h_counts[L + 1 + l];  // n_edges
    }
    counts[2 * L] = s->h_counts[0];
    return 0;
}

extern "C" int coala_sampler_layer_ptrs(coala_sampler* s, int layer, uint64_t* src_nodes, uint64_t* edge_src, uint64_t* edge_dst,
                                        uint64_t* edge_pos) {
    if (layer < 0 || layer >= s->L) COALA_FAIL("sampler: layer out of range");
    const int l = s->L - 1 - layer;
    if (src_nodes) *src_nodes = (uint64_t)s->frontier[l + 1];
    if (edge_src) *edge_src = (uint64_t)s->e_src_local[l];
    if (edge_dst) *edge_dst = (uint64_t)s->e_dst[l];
    if (edge_pos) *edge_pos = (uint64_t)s->e_id[l];
    return 0;
}

extern "C" int coala_lookup_edge_ids(uint64_t edge_ids_ptr, uint64_t pos_dev_ptr, int64_t n, uint64_t out_dev_ptr, void* stream) {
    if (n <= 0) return 0;
    int64_t blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    edge_id_lookup_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const int64_t*>(pos_dev_ptr), nullptr, n, reinterpret_cast<const int64_t*>(edge_ids_ptr),
        reinterpret_cast<int64_t*>(out_dev_ptr));
    COALA_CUDA(cudaGetLastError());
    return 0;
}
