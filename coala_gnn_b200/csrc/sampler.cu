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
// (ordered compaction, hash insert, first-occurrence numbering) runs out of HBM/L2.  No host
// synchronisation between layers: frontier sizes stay on the device.
//
// Three kernels per round, two of them on the critical path (the whole sample is one CUDA graph of 3 L + 1 nodes):
//   K1 sample_rows       picks + edges written at their final, frontier-ordered positions (single-pass ordered
//                        compaction: decoupled look-back over per-tile edge counts) + hash inserts
//   K2 first_occurrence  first-occurrence flags, their ordered count (look-back again) = ids of the new sources,
//                        next frontier
//   K3 relabel           local source ids of the edges; forked beside the next round's K1
// Hash table and look-back words are versioned instead of cleared, so no reset pass runs between rounds or samples.
#include "common.cuh"
#include "../../include/coala_b200.h"
#include <vector>

namespace {

constexpr int MAX_LAYERS = 8;
constexpr int MAX_FANOUT = 32;
constexpr uint64_t FY_MUL = 0xD1342543DE82EF95ULL;
constexpr uint64_t LAYER_MUL = 0xA24BAED4963EE407ULL;
constexpr int POS_INF = 0x7fffffff;

// ---------------------------------------------------------------------------------------------
// Versioned tables.  Nothing is cleared between uses: every word carries the version of the sampling round that
// wrote it ([ver:24 | payload:40]) and a word of another version reads as empty.  A round's version is
// (sample number mod VER_PERIOD) * MAX_LAYERS + round + 1; the host clears the tables when the sample number wraps.
// ---------------------------------------------------------------------------------------------
constexpr int VER_SHIFT = 40;
constexpr uint64_t PAY_MASK = (1ULL << VER_SHIFT) - 1;
constexpr uint64_t VER_PERIOD = ((1ULL << 24) - 2 * MAX_LAYERS) / MAX_LAYERS;

struct HashTable {
    unsigned long long* keys;  // ver | node id
    unsigned long long* vals;  // ver | (PAY_MASK - smallest position seen): atomicMax keeps the minimum, and any
                               // word of the current version beats every stale one (versions only grow until the wrap)
    uint64_t mask;
};

__device__ __forceinline__ void ht_insert_min(const HashTable& t, uint64_t ver, int64_t key, int pos) {
    const unsigned long long mykey = (ver << VER_SHIFT) | (unsigned long long)key;
    uint64_t h = mix64((uint64_t)key) & t.mask;
    while (true) {
        const unsigned long long k = t.keys[h];
        if (k == mykey) break;
        if ((k >> VER_SHIFT) != ver) {  // empty for this round: claim it
            const unsigned long long old = atomicCAS(t.keys + h, k, mykey);
            if (old == k || old == mykey) break;
            continue;  // somebody else took the slot: look at it again
        }
        h = (h + 1) & t.mask;
    }
    atomicMax(t.vals + h, (ver << VER_SHIFT) | (PAY_MASK - (unsigned long long)pos));
}

__device__ __forceinline__ int ht_lookup(const HashTable& t, uint64_t ver, int64_t key) {
    const unsigned long long mykey = (ver << VER_SHIFT) | (unsigned long long)key;
    uint64_t h = mix64((uint64_t)key) & t.mask;
    while (true) {
        const unsigned long long k = t.keys[h];
        if (k == mykey) return (int)(PAY_MASK - (t.vals[h] & PAY_MASK));
        if ((k >> VER_SHIFT) != ver) return POS_INF;
        h = (h + 1) & t.mask;
    }
}

// ---------------------------------------------------------------------------------------------
// Single-pass ordered compaction (decoupled look-back).  Tiles take their index from a ticket counter, so a tile only
// ever waits for tiles that started before it.  status word = [ver:24 | flag:2 | value:38]; flag 1 = the tile's own
// total, 2 = inclusive prefix up to and including the tile.
// ---------------------------------------------------------------------------------------------
constexpr uint64_t ST_AGG = 1, ST_PREFIX = 2;
__device__ __forceinline__ uint64_t st_pack(uint64_t ver, uint64_t flag, int64_t value) { return (ver << VER_SHIFT) | (flag << 38) | (uint64_t)value; }

// executed by warp 0 of the tile: publishes `agg`, returns the exclusive prefix of the tile (valid in every lane)
__device__ __forceinline__ int64_t lookback_exclusive(uint64_t* status, uint64_t ver, int64_t tile, int64_t agg, int lane) {
    if (tile == 0) {
        if (lane == 0) st_relaxed_u64<false>(status, st_pack(ver, ST_PREFIX, agg));
        return 0;
    }
    if (lane == 0) st_relaxed_u64<false>(status + tile, st_pack(ver, ST_AGG, agg));
    int64_t excl = 0;
    int64_t base = tile - 1;
    while (true) {
        const int64_t idx = base - lane;
        uint64_t w = st_pack(ver, ST_PREFIX, 0);  // below tile 0: an empty prefix
        if (idx >= 0) w = ld_relaxed_u64<false>(status + idx);
        const bool ready = (w >> VER_SHIFT) == ver && ((w >> 38) & 3) != 0;
        const unsigned pre = __ballot_sync(0xffffffffu, ready && ((w >> 38) & 3) == ST_PREFIX);
        const int first = pre ? __ffs(pre) - 1 : 32;  // nearest predecessor that already knows its prefix
        const unsigned need = first == 32 ? 0xffffffffu : ((2u << first) - 1u);
        if (__ballot_sync(0xffffffffu, !ready) & need) continue;  // a tile in between has not published yet
        int64_t c = lane <= first ? (int64_t)(w & ((1ULL << 38) - 1)) : 0;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
        excl += c;
        if (first < 32) break;
        base -= 32;
    }
    if (lane == 0) st_relaxed_u64<false>(status + tile, st_pack(ver, ST_PREFIX, excl + agg));
    return excl;
}

struct RoundCtl {            // device-resident control words of one sampler object
    int64_t counts[2 * MAX_LAYERS + 2];   // [0..L] frontier sizes, [L+1 .. 2L] edge counts
    uint64_t key;                         // RNG key of the batch in flight
    uint64_t ver_base;                    // version of round 0 of the batch in flight, minus 1
    unsigned int ticket[2];               // tile tickets of the two look-back kernels
};

// ---- K1: one warp per frontier node: indptr pair, neighbour row, picks — and the edges written straight to their final
// (frontier order) positions through the look-back, sources and destinations inserted in the round's table.  Round 0
// takes seeds / count / key / version from its (per-launch patched) arguments and publishes them for everything after it.
constexpr int K1_WARPS = 8;
__global__ void __launch_bounds__(K1_WARPS * 32) sample_rows_kernel(
    const int64_t* __restrict__ indptr, const int64_t* __restrict__ indices, const int64_t* __restrict__ frontier, RoundCtl* __restrict__ ctl,
    int layer, int n_layers, int fanout, HashTable ht, uint64_t* __restrict__ status, int64_t* __restrict__ e_src, int64_t* __restrict__ e_dst,
    int64_t* __restrict__ e_id, int64_t* __restrict__ frontier_copy, int64_t ns_arg, uint64_t key_arg, uint64_t ver_base_arg) {
    __shared__ int s_cnt[K1_WARPS], s_excl[K1_WARPS];
    __shared__ int64_t s_tile, s_prefix;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool first_round = layer == 0;
    const int64_t ns = first_round ? ns_arg : ctl->counts[layer];
    const uint64_t bkey = first_round ? key_arg : ctl->key;
    const uint64_t ver = (first_round ? ver_base_arg : ctl->ver_base) + (uint64_t)layer + 1;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        ctl->ticket[1] = 0;  // the other look-back kernel is not running now
        if (first_round) { ctl->counts[0] = ns; ctl->key = bkey; ctl->ver_base = ver_base_arg; }
    }
    if (threadIdx.x == 0) s_tile = (int64_t)atomicAdd(&ctl->ticket[0], 1u);
    __syncthreads();
    const int64_t tile = s_tile;
    const int64_t ntiles = (ns + K1_WARPS - 1) / K1_WARPS;
    if (tile >= ntiles) {
        if (tile == 0 && threadIdx.x == 0) ctl->counts[n_layers + 1 + layer] = 0;  // empty frontier: no edges
        return;
    }
    const uint64_t layer_key = mix64(bkey + (uint64_t)(layer + 1) * LAYER_MUL);
    const int64_t i = tile * K1_WARPS + warp;
    const bool valid = i < ns;
    int64_t v = 0, start = 0, deg = 0;
    if (valid) {
        v = frontier[i];
        int64_t p = 0;
        if (lane < 2) p = indptr[v + lane];  // both words of the pair in one request
        start = __shfl_sync(0xffffffffu, p, 0);
        deg = __shfl_sync(0xffffffffu, p, 1) - start;
        if (first_round && lane == 0) frontier_copy[i] = v;
    }
    const int k = valid ? (int)(deg <= fanout ? deg : fanout) : 0;
    if (lane == 0) s_cnt[warp] = k;
    // The host link serves ~154 M independent requests/s whatever their size up to 256 B (measured,
    // scripts/probe_links.py), so a row is fetched as ONE coalesced span (<= 64 neighbours: two
    // 256-B requests at most) issued before the picks are computed, instead of one request per pick.
    constexpr int ROW_SPAN = 64;
    const bool whole_row = deg <= ROW_SPAN;
    int64_t r0 = 0, r1 = 0;
    if (valid && whole_row) {
        if (lane < deg) r0 = indices[start + lane];
        if (lane + 32 < deg) r1 = indices[start + 32 + lane];
    }
    __syncthreads();
    if (warp == 0) {  // tile total -> look-back -> where this tile's edges start
        const int c = lane < K1_WARPS ? s_cnt[lane] : 0;
        int incl = c;
#pragma unroll
        for (int d = 1; d < K1_WARPS; d <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += y;
        }
        if (lane < K1_WARPS) s_excl[lane] = incl - c;
        const int64_t agg = __shfl_sync(0xffffffffu, incl, K1_WARPS - 1);
        const int64_t excl = lookback_exclusive(status, ver, tile, agg, lane);
        if (lane == 0) {
            s_prefix = excl;
            if (tile == ntiles - 1) ctl->counts[n_layers + 1 + layer] = excl + agg;  // number of edges of this round
        }
    }
    // picks (overlap the look-back of warp 0 in the other warps)
    int64_t mypos = -1, myval = 0;
    if (valid) {
        if (deg <= fanout) {
            if (lane < k) { mypos = lane; myval = r0; }  // fanout <= 32 < ROW_SPAN: the row is in r0
        } else {
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
    }
    __syncthreads();
    if (valid) {
        if (lane == 0) ht_insert_min(ht, ver, v, (int)i);  // dst nodes take local ids 0..ns-1
        if (lane < k) {
            const int64_t e = s_prefix + s_excl[warp] + lane;
            e_src[e] = myval;
            e_dst[e] = i;
            e_id[e] = start + mypos;  // CSC position; the edge id is looked up on demand (coala_lookup_edge_ids)
            ht_insert_min(ht, ver, myval, (int)(ns + e));
        }
    }
}

// ---- K2: an edge is the first occurrence of its source iff the table's minimum is its own position; the ordered count of
// first occurrences (look-back again) numbers the new sources, which are appended to the frontier
constexpr int K2_T = 256, K2_EPT = 4;
__global__ void __launch_bounds__(K2_T) first_occurrence_kernel(const int64_t* __restrict__ e_src, RoundCtl* __restrict__ ctl, int layer, int n_layers,
                                                                HashTable ht, uint64_t* __restrict__ status, int* __restrict__ val,
                                                                int* __restrict__ rank, const int64_t* __restrict__ frontier,
                                                                int64_t* __restrict__ next_frontier) {
    __shared__ int s_warp[K2_T / 32];
    __shared__ int64_t s_tile, s_prefix;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t ns = ctl->counts[layer], ne = ctl->counts[n_layers + 1 + layer];
    const uint64_t ver = ctl->ver_base + (uint64_t)layer + 1;
    if (blockIdx.x == 0 && tid == 0) ctl->ticket[0] = 0;  // the other look-back kernel is not running now
    if (tid == 0) s_tile = (int64_t)atomicAdd(&ctl->ticket[1], 1u);
    __syncthreads();
    const int64_t tile = s_tile;
    constexpr int TILE = K2_T * K2_EPT;
    const int64_t ntiles = (ne + TILE - 1) / TILE;
    const int64_t nworkers = ntiles > 0 ? ntiles : 1;
    if (tile >= nworkers) return;
    {   // dst nodes keep their local ids: frontier[0..ns) is copied in slices, one per tile
        const int64_t per = (ns + nworkers - 1) / nworkers;
        const int64_t lo = tile * per, hi = min(ns, lo + per);
        for (int64_t t = lo + tid; t < hi; t += K2_T) next_frontier[t] = frontier[t];
    }
    if (ntiles == 0) {
        if (tid == 0) ctl->counts[layer + 1] = ns;
        return;
    }
    const int64_t e0 = tile * TILE + (int64_t)tid * K2_EPT;
    int v[K2_EPT], f[K2_EPT];
    int64_t src[K2_EPT];
    int mine = 0;
#pragma unroll
    for (int u = 0; u < K2_EPT; u++) {
        const int64_t e = e0 + u;
        v[u] = 0; f[u] = 0; src[u] = 0;
        if (e < ne) {
            src[u] = e_src[e];
            v[u] = ht_lookup(ht, ver, src[u]);
            f[u] = (int64_t)v[u] == ns + e;
            mine += f[u];
        }
    }
    int incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += y;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const int c = lane < K2_T / 32 ? s_warp[lane] : 0;
        int wi = c;
#pragma unroll
        for (int d = 1; d < K2_T / 32; d <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, wi, d);
            if (lane >= d) wi += y;
        }
        const int64_t agg = __shfl_sync(0xffffffffu, wi, K2_T / 32 - 1);
        if (lane < K2_T / 32) s_warp[lane] = wi - c;  // exclusive warp offsets
        const int64_t excl = lookback_exclusive(status, ver, tile, agg, lane);
        if (lane == 0) {
            s_prefix = excl;
            if (tile == ntiles - 1) ctl->counts[layer + 1] = ns + excl + agg;  // size of the next frontier
        }
    }
    __syncthreads();
    int run = (int)s_prefix + s_warp[warp] + incl - mine;
#pragma unroll
    for (int u = 0; u < K2_EPT; u++) {
        const int64_t e = e0 + u;
        if (e < ne) {
            val[e] = v[u];
            rank[e] = run;
            if (f[u]) { next_frontier[ns + run] = src[u]; run++; }
        }
    }
}

// ---- K3 (off the critical path: runs beside the next round): local source ids of the edges
__global__ void relabel_kernel(const int* __restrict__ val, const int* __restrict__ rank, const RoundCtl* __restrict__ ctl, int layer, int n_layers,
                               int64_t* __restrict__ e_src_local) {
    const int64_t ns = ctl->counts[layer], ne = ctl->counts[n_layers + 1 + layer];
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < ne; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t v = val[t];
        e_src_local[t] = v < ns ? v : ns + rank[v - ns];
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
    int *val[MAX_LAYERS] = {nullptr}, *rank[MAX_LAYERS] = {nullptr};   // per round: K3 of a round runs beside the next round
    HashTable ht{nullptr, nullptr, 0};
    uint64_t ht_cap = 0;
    uint64_t *status_rows = nullptr, *status_edges = nullptr;          // look-back words of K1 / K2
    size_t n_status_rows = 0, n_status_edges = 0;
    RoundCtl* ctl = nullptr;           // device control block
    int64_t* h_counts = nullptr;       // pinned mirror of ctl->counts
    cudaEvent_t done = nullptr;
    int sm_count = 148;
    int64_t* indptr_hbm = nullptr;     // HBM copy of indptr (coala_sampler_stage_indptr), or another sampler's (coala_sampler_share_indptr)
    int64_t* indices_hbm = nullptr;    // HBM copy of indices (coala_sampler_stage_indices), same sharing rule
    bool indptr_owned = false, indices_owned = false;
    uint64_t seq = 0;                  // samples enqueued so far (drives the table versions)
    // the whole sample as one CUDA graph: one launch instead of 7 (every launch is a command fetch over
    // the same host link the path keeps busy); only the first node's arguments are patched per batch
    bool use_graph = true;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t graph_exec = nullptr;
    cudaGraphNode_t first_node = nullptr;
    cudaStream_t cap_stream = nullptr, side_stream = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
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
    if ((uint64_t)num_nodes >= PAY_MASK) { coala_set_error("sampler: node ids must fit 40 bits"); return nullptr; }
    COALA_CUDA_NULL(cudaSetDevice(device));
    coala_sampler* s = new coala_sampler();
    s->device = device; s->L = n_layers; s->max_seeds = max_seeds; s->num_nodes = num_nodes;
    s->indptr = reinterpret_cast<const int64_t*>(indptr_ptr);
    s->indices = reinterpret_cast<const int64_t*>(indices_ptr);
    s->edge_ids = reinterpret_cast<const int64_t*>(edge_ids_ptr);
    // the reference samples `for fanout in reversed(fanouts)` (DGL NeighborSampler.sample_blocks)
    for (int l = 0; l < n_layers; l++) s->fanout[l] = fanouts[n_layers - 1 - l];
    s->ns_max[0] = max_seeds;
    for (int l = 0; l < n_layers; l++) {
        s->ne_max[l] = s->ns_max[l] * s->fanout[l];
        s->ns_max[l + 1] = s->ns_max[l] + s->ne_max[l];  // = batch * prod(f+1): max_sample_size (S/COALA_GNN_Manager.py:79-81)
    }
    if (s->ns_max[n_layers] > 0x3fffffff) { coala_set_error("sampler: sample too large for 32-bit positions"); delete s; return nullptr; }
    for (int l = 0; l <= n_layers; l++) COALA_CUDA_NULL(cudaMalloc(&s->frontier[l], (size_t)s->ns_max[l] * 8));
    for (int l = 0; l < n_layers; l++) {
        COALA_CUDA_NULL(cudaMalloc(&s->e_src[l], (size_t)s->ne_max[l] * 8));
        COALA_CUDA_NULL(cudaMalloc(&s->e_dst[l], (size_t)s->ne_max[l] * 8));
        COALA_CUDA_NULL(cudaMalloc(&s->e_id[l], (size_t)s->ne_max[l] * 8));
        COALA_CUDA_NULL(cudaMalloc(&s->e_src_local[l], (size_t)s->ne_max[l] * 8));
        COALA_CUDA_NULL(cudaMalloc(&s->val[l], (size_t)s->ne_max[l] * 4));
        COALA_CUDA_NULL(cudaMalloc(&s->rank[l], (size_t)s->ne_max[l] * 4));
    }
    uint64_t cap = 1024;
    while (cap < 2 * (uint64_t)s->ns_max[n_layers]) cap <<= 1;
    s->ht.mask = cap - 1;
    s->ht_cap = cap;
    COALA_CUDA_NULL(cudaMalloc(&s->ht.keys, cap * 8));
    COALA_CUDA_NULL(cudaMalloc(&s->ht.vals, cap * 8));
    COALA_CUDA_NULL(cudaMemset(s->ht.keys, 0, cap * 8));   // version 0 = never used
    COALA_CUDA_NULL(cudaMemset(s->ht.vals, 0, cap * 8));
    s->n_status_rows = (size_t)(s->ns_max[n_layers - 1] + K1_WARPS - 1) / K1_WARPS + 1;
    s->n_status_edges = (size_t)(s->ne_max[n_layers - 1] + K2_T * K2_EPT - 1) / (K2_T * K2_EPT) + 1;
    COALA_CUDA_NULL(cudaMalloc(&s->status_rows, s->n_status_rows * 8));
    COALA_CUDA_NULL(cudaMalloc(&s->status_edges, s->n_status_edges * 8));
    COALA_CUDA_NULL(cudaMemset(s->status_rows, 0, s->n_status_rows * 8));
    COALA_CUDA_NULL(cudaMemset(s->status_edges, 0, s->n_status_edges * 8));
    COALA_CUDA_NULL(cudaMalloc(&s->ctl, sizeof(RoundCtl)));
    COALA_CUDA_NULL(cudaMemset(s->ctl, 0, sizeof(RoundCtl)));
    COALA_CUDA_NULL(cudaHostAlloc(&s->h_counts, (2 * MAX_LAYERS + 2) * 8, cudaHostAllocDefault));
    memset(s->h_counts, 0, (2 * MAX_LAYERS + 2) * 8);
    COALA_CUDA_NULL(cudaEventCreateWithFlags(&s->done, cudaEventDisableTiming));
    COALA_CUDA_NULL(cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming));
    COALA_CUDA_NULL(cudaEventCreateWithFlags(&s->ev_join, cudaEventDisableTiming));
    COALA_CUDA_NULL(cudaStreamCreateWithFlags(&s->side_stream, cudaStreamNonBlocking));
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
    for (int l = 0; l < s->L; l++) {
        cudaFree(s->e_src[l]); cudaFree(s->e_dst[l]); cudaFree(s->e_id[l]); cudaFree(s->e_src_local[l]);
        cudaFree(s->val[l]); cudaFree(s->rank[l]);
    }
    if (s->indptr_owned) cudaFree(s->indptr_hbm);
    if (s->indices_owned) cudaFree(s->indices_hbm);
    cudaFree(s->ctl); cudaFree(s->status_rows); cudaFree(s->status_edges);
    if (s->graph_exec) cudaGraphExecDestroy(s->graph_exec);
    if (s->graph) cudaGraphDestroy(s->graph);
    if (s->cap_stream) cudaStreamDestroy(s->cap_stream);
    if (s->side_stream) cudaStreamDestroy(s->side_stream);
    cudaFree(s->ht.keys); cudaFree(s->ht.vals);
    if (s->h_counts) cudaFreeHost(s->h_counts);
    if (s->done) cudaEventDestroy(s->done);
    if (s->ev_fork) cudaEventDestroy(s->ev_fork);
    if (s->ev_join) cudaEventDestroy(s->ev_join);
    delete s;
}

extern "C" int64_t coala_sampler_max_sample(coala_sampler* s) { return s->ns_max[s->L]; }

static void drop_graph(coala_sampler* s) {  // the captured graph holds the old CSC addresses
    if (s->graph_exec) { cudaGraphExecDestroy(s->graph_exec); s->graph_exec = nullptr; }
    if (s->graph) { cudaGraphDestroy(s->graph); s->graph = nullptr; }
}
extern "C" int coala_sampler_stage_indptr(coala_sampler* s) {
    if (s->indptr_hbm) return 0;
    COALA_CUDA(cudaSetDevice(s->device));
    COALA_CUDA(cudaMalloc(&s->indptr_hbm, (size_t)(s->num_nodes + 1) * 8));
    COALA_CUDA(cudaMemcpy(s->indptr_hbm, s->indptr, (size_t)(s->num_nodes + 1) * 8, cudaMemcpyDefault));
    s->indptr = s->indptr_hbm;
    s->indptr_owned = true;
    drop_graph(s);
    return 0;
}
// B200 option: the neighbour lists too (8 bytes per edge: 0.96 GB at igb-medium, 12.9 GB at papers100M, 32 GB at IGB-full, all
// small beside 180 GB).  With both arrays staged a sample never leaves the device.
extern "C" int coala_sampler_stage_indices(coala_sampler* s) {
    if (s->indices_hbm) return 0;
    COALA_CUDA(cudaSetDevice(s->device));
    int64_t nnz = 0;
    COALA_CUDA(cudaMemcpy(&nnz, s->indptr + s->num_nodes, 8, cudaMemcpyDefault));
    if (nnz < 0) COALA_FAIL("stage_indices: indptr[num_nodes] is negative");
    COALA_CUDA(cudaMalloc(&s->indices_hbm, (size_t)std::max<int64_t>(nnz, 1) * 8));
    COALA_CUDA(cudaMemcpy(s->indices_hbm, s->indices, (size_t)nnz * 8, cudaMemcpyDefault));
    s->indices = s->indices_hbm;
    s->indices_owned = true;
    drop_graph(s);
    return 0;
}
// The samplers of one engine walk the same graph: one staged copy serves all of them.  `from` must outlive `s`.
extern "C" int coala_sampler_share_indptr(coala_sampler* s, const coala_sampler* from) {
    if (!from->indptr_hbm || from->num_nodes != s->num_nodes || from->device != s->device || s->indptr_hbm || s->indices_hbm)
        COALA_FAIL("share_indptr: the other sampler has no staged indptr for this graph and device");
    s->indptr_hbm = from->indptr_hbm;
    s->indptr = s->indptr_hbm;
    if (from->indices_hbm) {
        s->indices_hbm = from->indices_hbm;
        s->indices = s->indices_hbm;
    }
    drop_graph(s);
    return 0;
}
extern "C" uint64_t coala_sampler_input_count_dev_ptr(coala_sampler* s) { return (uint64_t)(&s->ctl->counts[s->L]); }

static inline unsigned grid_for(int64_t work, int per_block, int cap) {
    int64_t b = (work + per_block - 1) / per_block;
    if (b < 1) b = 1;
    if (cap > 0 && b > cap) b = cap;
    return (unsigned)b;
}

struct FirstArgs {  // what changes from batch to batch: the arguments of round 0's sampling kernel
    const int64_t* seeds;
    int64_t ns;
    uint64_t key, ver_base;
};

static void first_kernel_params(coala_sampler* s, const FirstArgs& a, void** args, int* layer0, int* nl, int* f0, HashTable* ht0, const int64_t** front,
                                RoundCtl** ctl, uint64_t** status, int64_t** es, int64_t** ed, int64_t** ei, int64_t** fcopy,
                                int64_t* ns, uint64_t* key, uint64_t* vb) {
    *layer0 = 0; *nl = s->L; *f0 = s->fanout[0];
    *ht0 = s->ht;
    uint64_t need = 1024;
    while (need < 2 * (uint64_t)s->ns_max[1]) need <<= 1;
    ht0->mask = need - 1;
    *front = a.seeds; *ctl = s->ctl; *status = s->status_rows;
    *es = s->e_src[0]; *ed = s->e_dst[0]; *ei = s->e_id[0]; *fcopy = s->frontier[0];
    *ns = a.ns; *key = a.key; *vb = a.ver_base;
    args[0] = (void*)&s->indptr; args[1] = (void*)&s->indices; args[2] = front; args[3] = ctl; args[4] = layer0; args[5] = nl; args[6] = f0;
    args[7] = ht0; args[8] = status; args[9] = es; args[10] = ed; args[11] = ei; args[12] = fcopy; args[13] = ns; args[14] = key; args[15] = vb;
}

// enqueue every operation of one sample on `st` (directly, or while `st` is being captured).  K3 of every round and the
// count read-back go to a side stream forked after K2, so the critical path of a round is K1 -> K2.
static int enqueue_sample(coala_sampler* s, const FirstArgs& a, cudaStream_t st) {
    const int L = s->L;
    trace_mark(s, st);
    bool forked = false;
    for (int l = 0; l < L; l++) {
        const int f = s->fanout[l];
        HashTable ht = s->ht;
        // table sized for this round only: keeps the early rounds' probes inside a few cache lines
        uint64_t need = 1024;
        while (need < 2 * (uint64_t)s->ns_max[l + 1]) need <<= 1;
        ht.mask = need - 1;
        const int64_t* front = l == 0 ? a.seeds : s->frontier[l];
        sample_rows_kernel<<<grid_for(s->ns_max[l], K1_WARPS, 0), K1_WARPS * 32, 0, st>>>(
            s->indptr, s->indices, front, s->ctl, l, L, f, ht, s->status_rows, s->e_src[l], s->e_dst[l], s->e_id[l], s->frontier[0],
            l == 0 ? a.ns : 0, l == 0 ? a.key : 0, l == 0 ? a.ver_base : 0);
        trace_mark(s, st);
        first_occurrence_kernel<<<grid_for(s->ne_max[l], K2_T * K2_EPT, 0), K2_T, 0, st>>>(s->e_src[l], s->ctl, l, L, ht, s->status_edges, s->val[l],
                                                                                         s->rank[l], s->frontier[l], s->frontier[l + 1]);
        trace_mark(s, st);
        if (s->trace) {  // direct launches, one stream: keep the per-kernel intervals meaningful
            relabel_kernel<<<grid_for(s->ne_max[l], 256, s->sm_count * 4), 256, 0, st>>>(s->val[l], s->rank[l], s->ctl, l, L, s->e_src_local[l]);
            trace_mark(s, st);
        } else {
            COALA_CUDA(cudaEventRecord(s->ev_fork, st));
            COALA_CUDA(cudaStreamWaitEvent(s->side_stream, s->ev_fork, 0));
            relabel_kernel<<<grid_for(s->ne_max[l], 256, s->sm_count * 4), 256, 0, s->side_stream>>>(s->val[l], s->rank[l], s->ctl, l, L, s->e_src_local[l]);
            forked = true;
        }
    }
    COALA_CUDA(cudaGetLastError());
    if (forked) {
        COALA_CUDA(cudaEventRecord(s->ev_join, s->side_stream));
        COALA_CUDA(cudaStreamWaitEvent(st, s->ev_join, 0));
    }
    COALA_CUDA(cudaMemcpyAsync(s->h_counts, s->ctl->counts, (2 * L + 1) * 8, cudaMemcpyDeviceToHost, st));
    trace_mark(s, st);
    return 0;
}

static int build_graph(coala_sampler* s) {
    if (!s->cap_stream) COALA_CUDA(cudaStreamCreateWithFlags(&s->cap_stream, cudaStreamNonBlocking));
    COALA_CUDA(cudaStreamBeginCapture(s->cap_stream, cudaStreamCaptureModeRelaxed));
    FirstArgs a{s->frontier[0], 0, 0, 0};  // placeholder arguments, patched per launch
    int rc = enqueue_sample(s, a, s->cap_stream);
    cudaError_t e = cudaStreamEndCapture(s->cap_stream, &s->graph);
    if (rc) return -1;
    COALA_CUDA(e);
    size_t n_root = 1;
    COALA_CUDA(cudaGraphGetRootNodes(s->graph, &s->first_node, &n_root));
    if (n_root != 1) COALA_FAIL("sampler graph: expected a single root node");
    cudaGraphNodeType ty;
    COALA_CUDA(cudaGraphNodeGetType(s->first_node, &ty));
    if (ty != cudaGraphNodeTypeKernel) COALA_FAIL("sampler graph: root node is not the first sampling kernel");
    COALA_CUDA(cudaGraphInstantiate(&s->graph_exec, s->graph, 0));
    return 0;
}

extern "C" int coala_sampler_graph_nodes(coala_sampler* s) {
    if (!s->graph) return 0;
    size_t n = 0;
    if (cudaGraphGetNodes(s->graph, nullptr, &n) != cudaSuccess) return -1;
    return (int)n;
}

extern "C" int coala_sampler_sample_async(coala_sampler* s, uint64_t seeds_dev_ptr, int ns, uint64_t rng_key, void* stream) {
    if (ns < 0 || ns > s->max_seeds) COALA_FAIL("sampler: more seeds than max_seeds");
    COALA_CUDA(cudaSetDevice(s->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const uint64_t slot = s->seq % VER_PERIOD;
    if (slot == 0 && s->seq > 0) {  // the versions wrap: forget every older word (once per ~2 M samples)
        COALA_CUDA(cudaMemsetAsync(s->ht.keys, 0, s->ht_cap * 8, st));
        COALA_CUDA(cudaMemsetAsync(s->ht.vals, 0, s->ht_cap * 8, st));
        COALA_CUDA(cudaMemsetAsync(s->status_rows, 0, s->n_status_rows * 8, st));
        COALA_CUDA(cudaMemsetAsync(s->status_edges, 0, s->n_status_edges * 8, st));
    }
    s->seq++;
    FirstArgs a{reinterpret_cast<const int64_t*>(seeds_dev_ptr), (int64_t)ns, rng_key, slot * MAX_LAYERS};
    if (!s->use_graph) {
        trace_collect(s);
        if (enqueue_sample(s, a, st)) return -1;
    } else {
        if (!s->graph_exec && build_graph(s)) return -1;
        void* args[16];
        int layer0, nl, f0;
        HashTable ht0;
        const int64_t* front;
        RoundCtl* ctl;
        uint64_t* status;
        int64_t *es, *ed, *ei, *fcopy, nsv;
        uint64_t key, vb;
        first_kernel_params(s, a, args, &layer0, &nl, &f0, &ht0, &front, &ctl, &status, &es, &ed, &ei, &fcopy, &nsv, &key, &vb);
        cudaKernelNodeParams kp;
        memset(&kp, 0, sizeof(kp));
        kp.func = reinterpret_cast<void*>(sample_rows_kernel);
        kp.gridDim = dim3(grid_for(s->ns_max[0], K1_WARPS, 0));
        kp.blockDim = dim3(K1_WARPS * 32);
        kp.sharedMemBytes = 0;
        kp.kernelParams = args;
        COALA_CUDA(cudaGraphExecKernelNodeSetParams(s->graph_exec, s->first_node, &kp));
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
    if (blocks > coala_sm_count() * 8) blocks = coala_sm_count() * 8;
    edge_id_lookup_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const int64_t*>(pos_dev_ptr), nullptr, n, reinterpret_cast<const int64_t*>(edge_ids_ptr),
        reinterpret_cast<int64_t*>(out_dev_ptr));
    COALA_CUDA(cudaGetLastError());
    return 0;
}
