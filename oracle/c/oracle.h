/* TEST INFRASTRUCTURE ONLY — CPU oracle for the COALA-GNN mini-batch data path.
 *
 * Plain-C restatement of the reference algorithms for the three hot-path stages.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library; the product path (coala_gnn_b200/) never does.
 *
 * Reference citations are into /root/reference/COALA_GNN_Modules/ (M/) and
 * /root/reference/COALA-GNN-Setup/COALA_GNN/ (S/).
 *
 * Parity pins:
 *   distributor : pinned against the reference's own node_distributor_pybind.cuh compiled
 *                 from source (oracle/_ref) and the golden vectors in tests/golden/.
 *   cache model : the reference CUDA cache does not build here (BaM libcxx);
 *                 pinned by the invariant the reference states itself
 *                 (examples/nvshmem_homogeneous_train.py:186-198): out[i] == feat[idx[i]].
 *   sampler     : arithmetic lives in DGL (un-vendored, "version 2.5", README.md:12) —
 *                 PARITY UNPINNED by any reference test; full fan-out is exact by
 *                 construction (CSC slices), sampled fan-out follows the RNG defined here.
 */
#ifndef COALA_ORACLE_H
#define COALA_ORACLE_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- shared deterministic generators (also implemented on the GPU, bit-identical) ---- */
uint64_t orc_mix64(uint64_t z);
/* feature value of (row,col) for a table seed: 24 random mantissa bits -> [0,1) exactly */
float orc_feat_value(uint64_t seed, uint64_t row, uint64_t col);
void orc_fill_features(float* dst, uint64_t seed, uint64_t row0, uint64_t nrows, uint64_t dim, uint64_t stride);

/* ---- stage 1: seed distribution (M/node_distributor_pybind.cuh:150-222) ---- */
void orc_distribute(const int64_t* items, int64_t offset, int node_id, int batch, int local_size,
                    int n_domains, const int64_t* color, const int64_t* topk, const double* score,
                    int K, const int32_t* const* meta, int64_t* out, int* bucket_len_out);

/* ---- stage 3: sequential cache model (M/isolated_cache.h:335-475, M/nvshmem_cache.h:336-480) ---- */
typedef struct orc_cache orc_cache;
/* mode: 0 isolated (set = id % S), 1 partitioned/nccl (owner = id % G, set = (id/G) % S).
 * In mode 1 the object models all G shards of one cache domain. */
orc_cache* orc_cache_create(uint64_t cache_mb, uint32_t page_size, int n_gpus, int mode,
                            const int64_t* color, int num_colors, int store_lines);
void orc_cache_destroy(orc_cache* c);
uint64_t orc_cache_num_sets(const orc_cache* c);
/* process requests in order; feat is the host table with row stride page_size bytes;
 * out rows have stride out_stride floats; copies page_size bytes per row */
void orc_cache_read(orc_cache* c, const int64_t* idx, int64_t n, const float* feat,
                    float* out, int64_t out_stride);
void orc_cache_stats(const orc_cache* c, uint64_t* hit, uint64_t* miss);
/* per-shard views (gpu in [0,G) for mode 1, 0 otherwise) */
const uint64_t* orc_cache_keys(const orc_cache* c, int gpu);
const uint64_t* orc_cache_color_meta(const orc_cache* c, int gpu);
const int32_t* orc_cache_color_counters(const orc_cache* c, int gpu);
const uint32_t* orc_cache_set_cnt(const orc_cache* c, int gpu);

/* ---- stage 2: neighbour sampling + to_block (DGL semantics) ---- */
/* One layer.  seeds[ns] unique.  fanout<0 or >= deg keeps every in-edge in CSC order;
 * otherwise `fanout` distinct edges chosen by the partial Fisher-Yates defined in sampler.c,
 * emitted in ascending CSC position.  Outputs (caller-allocated, capacity ns*max(fanout,..)):
 *   e_src_global[ne], e_dst_local[ne], e_id[ne]; returns ne.                                  */
int64_t orc_sample_layer(const int64_t* indptr, const int64_t* indices, const int64_t* edge_ids,
                         const int64_t* seeds, int64_t ns, int fanout, uint64_t rng_key,
                         int64_t cap, int64_t* e_src_global, int64_t* e_dst_local, int64_t* e_id);
/* to_block: src numbering = seeds first (0..ns-1) then unseen sources by first occurrence.
 * src_nodes capacity ns+ne; returns n_src; e_src_local[ne] written. */
int64_t orc_to_block(const int64_t* seeds, int64_t ns, const int64_t* e_src_global, int64_t ne,
                     int64_t* src_nodes, int64_t* e_src_local);
/* whole multi-layer sample with OpenMP over seeds (the timed CPU baseline):
 * fanouts[L] in reference order (input layer first); layers are sampled in reverse.
 * Returns n_input; input_nodes capacity max_sample.  Per-layer edges are discarded
 * (counted into *total_edges) unless blk_* are non-NULL. */
int64_t orc_sample_blocks(const int64_t* indptr, const int64_t* indices, const int64_t* edge_ids,
                          const int64_t* seeds, int64_t ns, const int* fanouts, int L,
                          uint64_t rng_key, int64_t max_sample, int64_t* input_nodes,
                          int64_t* total_edges, int threads);
void orc_index_select(const float* feat, int64_t dim, const int64_t* idx, int64_t n, float* out, int threads);

#ifdef __cplusplus
}
#endif
#endif
