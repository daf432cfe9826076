/* coala_b200.h — C ABI of libcoala_b200.so: the B200-native (sm_100a) mini-batch data path of
 * COALA-GNN.  Every entry point takes plain pointers and sizes (no torch, no pybind types) and
 * is what a binding for the reference's COALA_GNN_Pybind module would call; the reference
 * interface each one replaces is cited as file:line under /root/reference/COALA_GNN_Modules/
 * (M/) and /root/reference/COALA-GNN-Setup/COALA_GNN/ (S/).
 *
 * Conventions
 *   - handles are opaque pointers; NULL / non-zero return = failure, message via
 *     coala_last_error() (thread-local copy of the last failure in this thread).
 *   - "ptr" arguments typed uint64_t are raw addresses exactly as the reference passes them
 *     from tensor.data_ptr(); unless stated they may be host or device
 *     addresses and are classified with cudaPointerGetAttributes.
 *   - calls without a stream argument are synchronous like the reference's (every cache
 *     method there ends in cudaDeviceSynchronize, M/ssd_gnn_cache.cuh:125,169,266);
 *     *_async variants enqueue on the given cudaStream_t (passed as void*) and return.
 */
#ifndef COALA_B200_H
#define COALA_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define COALA_MAX_GPUS 16
#define COALA_WAYS 32

const char* coala_last_error(void);
int coala_abi_version(void);

/* ------------------------------------------------------------------ shared UVA host memory
 * replaces SharedUVAManager (M/shared_UVA.cuh:26-115; bound at M/COALA_GNN_Pybind.cu:28-32).
 * create!=0: shm_open(O_CREAT)+ftruncate (the local-rank-0 branch, :65-77); create==0: open
 * an existing segment (:78-81).  Both mmap + cudaHostRegister + cudaHostGetDevicePointer
 * (:83-98).  The MPI barrier between the two branches is the caller's (torch.distributed).
 * zero!=0 clears the segment like the reference's memset (:99). */
typedef struct coala_uva coala_uva;
coala_uva* coala_uva_create(const char* shm_path, int64_t bytes, int create, int device, int zero);
uint64_t coala_uva_host_ptr(coala_uva* h);
uint64_t coala_uva_device_ptr(coala_uva* h);
int coala_uva_cleanup(coala_uva* h, int unlink_segment); /* M/shared_UVA.cuh:102-110 */
void coala_uva_destroy(coala_uva* h);

/* ------------------------------------------------------------------ stage 1: seed distributor
 * replaces Node_distributor_pybind (M/node_distributor_pybind.cuh:112-238; bound at
 * M/COALA_GNN_Pybind.cu:53-57). */
typedef struct coala_nd coala_nd;
/* ctor (:138-148): loads color.npy (int64[N]), topk.npy (int64[C,K]), score.npy (f8[C,K]) with
 * the reference's header rules (:37-109), keeps a pinned host copy of the colour table and
 * uploads int32 colour / topk / score tables to `device`.  item_ptr is the caller-owned int64
 * seed list (host or device). */
coala_nd* coala_nd_create(uint64_t item_ptr, int node_id, int batch_size, int local_size, int n_domains,
                          const char* color_file, const char* topk_file, const char* score_file, int device);
/* same, tables given in memory (host pointers) instead of .npy files */
coala_nd* coala_nd_create_from_arrays(uint64_t item_ptr, int node_id, int batch_size, int local_size, int n_domains,
                                      const int64_t* color, int64_t n_nodes, const int64_t* topk,
                                      const double* score, int num_colors, int K, int device);
/* ctor without colour information (:133-135) */
coala_nd* coala_nd_create_plain(uint64_t item_ptr, int n_domains);
/* distribute_node_with_affinity (:150-222): bit-exact bucket assignment on the GPU; writes this
 * domain's batch*local_size seeds (arrival order) to out_ptr. meta_ptrs[n_meta] are the per-domain
 * int32 cache colour histograms (length >= num_colors+1 entries are read up to the largest
 * colour id referenced). */
int coala_nd_distribute(coala_nd* h, uint64_t offset, uint64_t out_ptr, const uint64_t* meta_ptrs, int n_meta);
int coala_nd_distribute_async(coala_nd* h, uint64_t offset, uint64_t out_dev_ptr, const uint64_t* meta_dev_ptrs,
                              int n_meta, void* stream);
/* same with an explicit scratch slot in [0,8): distributions that may overlap on different streams (one per
 * in-flight mini-batch of the prefetch engine) must use different slots; slot 0 is the synchronous entry point's */
int coala_nd_distribute_async_slot(coala_nd* h, uint64_t offset, uint64_t out_dev_ptr, const uint64_t* meta_dev_ptrs,
                                   int n_meta, int slot, void* stream);
/* how many int32 entries each cache-meta histogram holds (default num_colors, the size the
 * reference's callers allocate, S/Training_node_distributor.py:33; colour id == num_colors then
 * reads as 0 instead of one past the end). Set num_colors+1 to cover every colour id. */
int coala_nd_set_meta_len(coala_nd* h, int len);
/* bucket_len[n_domains] after the last synchronous distribute (test hook) */
int coala_nd_last_bucket_len(coala_nd* h, int* out);
int coala_nd_num_colors(coala_nd* h);            /* get_num_colors (:224-226) */
uint64_t coala_nd_color_host_ptr(coala_nd* h);   /* get_color_buffer_ptr (:228-231) */
int64_t coala_nd_num_graph_nodes(coala_nd* h);
void coala_nd_destroy(coala_nd* h);

/* ------------------------------------------------------------------ stage 3: feature cache
 * replaces SSD_GNN_SSD_Controllers (M/ssd_gnn_cache.cuh:10-55), Isolated_Cache (:205-372),
 * SSD_GNN_NVSHMEM_Cache (:58-200) and the kernels of M/cache_kernel.cu. */
/* cache_dim rounding 128/256/512/1024 (M/ssd_gnn_cache.cuh:34-44); -1 when dim > 1024 */
int coala_cache_dim_for(int feat_dim);

typedef struct coala_cache coala_cache;
enum { COALA_CACHE_ISOLATED = 0,    /* private cache, set = id % S            (M/isolated_cache.h:183-188) */
       COALA_CACHE_PARTITIONED = 1, /* owner = id % G, set = (id/G) % S, peers reached by P2P loads
                                       (replaces the NVSHMEM owner-push of M/nvshmem_cache.h:336-480) */
       COALA_CACHE_NCCL_SHARD = 2   /* local shard of an id%G partition, set = (id/G) % S; ids arrive by
                                       all-to-all (M/isolated_cache.h:190-195 with nccl_cache=true) */ };
/* ctor (M/ssd_gnn_cache.cuh:84-109, 227-252): num_sets = cache_mb*2^20/page/32. The distributor
 * supplies the colour table and num_colors (may be NULL: colour tracking off). sim_buf_ptr is the
 * pinned-host feature table (device-visible address, row stride = cache_dim*4 bytes), the
 * reference's `sim_buf`; 0 is rejected (the BaM/NVMe tier is out of scope). */
coala_cache* coala_cache_create(int kind, int device, int feat_dim, coala_nd* nd, int global_rank, int local_rank,
                                int n_gpus, uint64_t cache_mb, uint64_t sim_buf_ptr);
void coala_cache_destroy(coala_cache* c);
uint64_t coala_cache_num_sets(coala_cache* c);
int coala_cache_row_bytes(coala_cache* c);

/* peer mapping for COALA_CACHE_PARTITIONED. One allocation per shard => one 64-byte CUDA IPC
 * handle; exchange them with any transport, then attach every peer's. */
int coala_cache_ipc_handle(coala_cache* c, void* handle64);
int coala_cache_attach_ipc(coala_cache* c, int peer_local_rank, const void* handle64);
/* Default sharing of a PARTITIONED shard (n_gpus > 1): the slab is a cuMemCreate allocation exported as
 * a POSIX file descriptor; send the fd to each peer process (SCM_RIGHTS) and attach it there.  Both
 * sides map it with 2 MB granularity (a cudaIpcOpenMemHandle import of a multi-GB slab is mapped with
 * small pages and loses ~3x on random rows).  The caller owns (closes) the descriptors. */
int coala_cache_is_fd_shared(coala_cache* c);
int coala_cache_export_fd(coala_cache* c);
int coala_cache_attach_fd(coala_cache* c, int peer_local_rank, int fd);
/* same-process attachment (several shards driven from one process, or emulated ranks on one GPU) */
int coala_cache_attach_local(coala_cache* c, int peer_local_rank, coala_cache* peer);

/* Isolated_Cache::read_feature / Isolated_read_feature_kernel (M/ssd_gnn_cache.cuh:255-268,
 * M/cache_kernel.cu:59-77) and, in PARTITIONED mode, the fused replacement of
 * send_requests + read_feature (M/ssd_gnn_cache.cuh:111-174, M/cache_kernel.cu:4-57):
 * out[i, 0:cache_dim] = feature row of idx[i]; out row stride = feat_dim floats. */
int coala_cache_read_feature(coala_cache* c, uint64_t out_ptr, uint64_t idx_ptr, int64_t n);
/* n_dev_ptr (optional, device int64*): the row count is read on the device so the launch can be
 * captured in a CUDA graph; n is then the upper bound used for sizing. */
int coala_cache_read_feature_async(coala_cache* c, uint64_t out_ptr, uint64_t idx_ptr, int64_t n,
                                   uint64_t n_dev_ptr, void* stream);
/* Node ids are bounds-checked on the device (the reference dereferences whatever it is given): an id >= the host
 * table's row count (known from the distributor's colour table, or set here) or >= 2^40 - 1 (the tag's id field) is
 * refused - its output row is zero-filled and counted (set_num_nodes never raises the limit above the colour table's length:
 * the fill path looks the id's colour up too).  The synchronous read_feature calls fail when a launch
 * refused ids; callers of the async form poll coala_cache_refused_ids. */
int coala_cache_set_num_nodes(coala_cache* c, int64_t num_nodes);
int coala_cache_refused_ids(coala_cache* c, uint64_t* count, int reset);
/* same kernel, launched as ONE warp so requests resolve strictly in order (test hook: makes
 * hit/miss counts, tags and colour counters comparable with the sequential oracle) */
int coala_cache_read_feature_serial(coala_cache* c, uint64_t out_ptr, uint64_t idx_ptr, int64_t n);
/* get_cache_data (M/ssd_gnn_cache.cuh:176-186, 270-280): copies num_colors int32 counters */
int coala_cache_get_cache_data(coala_cache* c, uint64_t dst_int32_ptr);
/* same with an explicit entry count (<= num_colors+1) */
int coala_cache_get_cache_data_n(coala_cache* c, uint64_t dst_int32_ptr, int count);
/* device-to-device on `stream` (the counters stay on the GPU for the device-resident distributor) */
int coala_cache_get_cache_data_async(coala_cache* c, uint64_t dst_dev_ptr, int count, void* stream);
uint64_t coala_cache_color_counters_dev_ptr(coala_cache* c);
/* print_stats (M/ssd_gnn_cache.cuh:188-190, M/nvshmem_cache.h:133-142): counters since last reset */
int coala_cache_stats(coala_cache* c, uint64_t* hit, uint64_t* miss, uint64_t* bypass, int reset);
int coala_cache_print_stats(coala_cache* c);
/* test hook: dumps tags (id field, 0xFF.. = empty), colour meta and round-robin pointers */
int coala_cache_export_state(coala_cache* c, uint64_t* keys, uint32_t* color_meta, uint32_t* set_cnt);
/* choose the row-copy engine of the gather: 0 = rows through registers with 16-byte vector loads (default),
 * 1 = hit rows as bulk asynchronous copies (cp.async.bulk, 1-D TMA) line -> shared memory -> output; rows whose
 * byte length / stride is not a multiple of 16 keep engine 0.  COALA_GATHER_ENGINE=1 selects it at create time. */
int coala_cache_set_copy_engine(coala_cache* c, int engine);

/* NCCL all-to-all variant helpers (comparison backend).
 * split_node_list (M/ssd_gnn_cache.cuh:283-295, M/cache_kernel.cu:79-91): deterministic stable
 * partition by id % n_gpus: node[max_sample*g + k], map[...] = original position, counter[g]. */
int coala_split_node_list(uint64_t idx_ptr, int64_t n, uint64_t node_ptr, uint64_t map_ptr, uint64_t counter_ptr,
                          int n_gpus, int max_sample, void* stream);
/* No counterpart in the reference (its split_node_list, cited above, has no allocation-free form).
 * coala_split_node_list_static is coala_split_node_list - same arguments, same buckets, bit for bit - for callers that
 * may not allocate device memory: plugins living inside a host whose GPU memory belongs to one pool allocator (a
 * framework's caching allocator, an RMM pool, a MIG slice budgeted to the byte), and latency-critical loops where
 * cudaMalloc / cudaMallocAsync are banned because they can take driver locks or synchronise the device.  The
 * nblocks x n_gpus block-count scratch the three kernels share lives in a 128 KiB static __device__ array inside the
 * library (held by every GPU that loads it, until exit) instead of coming from cudaMallocAsync; the call performs no
 * cudaMalloc / cudaFree / pool operation of any kind.
 *   limit      n <= coala_split_node_list_static_max_ids() = 2 097 152 ids per call; larger n fails with a message
 *              (nothing is launched) - use coala_split_node_list, or split the list.
 *   set-up     the first call on each device (with that device current, as for every call here) resolves the
 *              array's device address with cudaGetSymbolAddress; make that call before any stream capture.
 *   sharing    all calls on a device use the one array, which keeps the previous call's contents.  No word is read
 *              before this call has written it, so nothing needs clearing - but two calls whose kernels can overlap on
 *              the GPU (different streams / threads) race on it.  Keep them stream-ordered. */
int coala_split_node_list_static(uint64_t idx_ptr, int64_t n, uint64_t node_ptr, uint64_t map_ptr, uint64_t counter_ptr,
                                 int n_gpus, int max_sample, void* stream);
int64_t coala_split_node_list_static_max_ids(void);
/* map_feat_data (M/ssd_gnn_cache.cuh:327-356, M/cache_kernel.cu:113-137): out[map[r]] = src[r] */
int coala_map_feat_data(uint64_t out_ptr, uint64_t src_ptr, uint64_t map_ptr, int64_t n, int feat_dim, void* stream);

/* The whole NCCL variant as one stream-ordered call (replaces the host choreography of S/COALA_GNN_Manager.py:143-211:
 * split_node_list, dist.all_to_all of padded id buckets + all_to_all_single of counts, nccl_get_feature on G streams,
 * G x G serialised dist.send/recv of rows, map_feat_data per peer).  `shard` is this rank's COALA_CACHE_NCCL_SHARD cache.
 * The communicator spans the ranks of one cache domain: rank 0 obtains a 128-byte id with coala_nccl_unique_id, the
 * caller ships it to the other ranks, and every rank calls coala_nccl_create (collective; ncclCommInitRank inside).
 * NCCL is loaded at run time (the libnccl.so.2 already in the process, or COALA_NCCL_LIB).
 * coala_nccl_fetch_async: out[i, 0:feat_dim] = feature row of idx[i]; the row count may live on the device (n_dev_ptr,
 * n_upper then bounds it).  Everything is enqueued on `stream`; the call blocks once, for the 2G bucket counts the row
 * exchange needs on the host.  Collective: every rank of the domain calls it the same number of times. */
typedef struct coala_nccl coala_nccl;
int coala_nccl_unique_id(void* id128);
int coala_nccl_version(void);
coala_nccl* coala_nccl_create(coala_cache* shard, int device, const void* id128, int n_ranks, int rank, int64_t max_sample, int feat_dim);
void coala_nccl_destroy(coala_nccl* x);
int coala_nccl_fetch_async(coala_nccl* x, uint64_t out_ptr, uint64_t idx_ptr, int64_t n_upper, uint64_t n_dev_ptr, void* stream);
int coala_nccl_counters(coala_nccl* x, uint64_t* batches, uint64_t* rows_served, uint64_t* rows_received);

/* ------------------------------------------------------------------ stage 2: UVA neighbour sampler
 * replaces dgl MultiLayerNeighborSampler.sample on the UVA CSC graph (call site
 * S/COALA_GNN_DataLoader.py:162; graph arrangement M/shared_UVA.cuh + examples/ssd_gnn_dataloader.py:496-524). */
typedef struct coala_sampler coala_sampler;
/* CSC arrays are device-visible addresses (pinned host through UVA, or HBM); edge_ids_ptr may be 0
 * (edge id = CSC position). fanouts[L] in the reference's order (input layer first). */
coala_sampler* coala_sampler_create(int device, uint64_t indptr_ptr, uint64_t indices_ptr, uint64_t edge_ids_ptr,
                                    int64_t num_nodes, int max_seeds, const int* fanouts, int n_layers);
void coala_sampler_destroy(coala_sampler* s);
int64_t coala_sampler_max_sample(coala_sampler* s);
/* B200 layout: keep a copy of indptr (8(N+1) bytes: 80 MB at 10 M nodes) in HBM, which removes the
 * first of the two dependent host round trips of every sampled row (the Python layer does this unless
 * COALA_HBM_INDPTR=0).  stage_indices copies the neighbour lists too (8 bytes per edge; opt-in,
 * COALA_HBM_GRAPH=1): a sample then never leaves the device.  Edge ids stay in pinned host memory
 * (they are looked up on demand). */
int coala_sampler_stage_indptr(coala_sampler* s);
int coala_sampler_stage_indices(coala_sampler* s);
/* use the copies another sampler of the same graph and device staged (that sampler must be destroyed last) */
int coala_sampler_share_indptr(coala_sampler* s, const coala_sampler* from);
/* diagnostics (COALA_SAMPLER_TRACE=1 at creation): average milliseconds between consecutive enqueued
 * operations of a sample (per round: sample_rows, first_occurrence, relabel; finally the count read-back).
 * Returns the number of intervals written. */
int coala_sampler_trace(coala_sampler* s, double* avg_ms, int max_n);
/* nodes of the CUDA graph one sample is launched as (0 before the first sample): 3 kernels per round + the count read-back */
int coala_sampler_graph_nodes(coala_sampler* s);
/* enqueue all layers for `ns` seeds (device int64); results stay in the sampler's buffers */
int coala_sampler_sample_async(coala_sampler* s, uint64_t seeds_dev_ptr, int ns, uint64_t rng_key, void* stream);
/* after the stream is synchronised: counts[2*L+1] = n_src(l), n_edges(l) for l = 0..L-1 in
 * reference block order (input layer first), then n_seeds */
int coala_sampler_counts(coala_sampler* s, int64_t* counts);
/* device addresses of layer l (reference block order): src node ids [n_src], edge_src_local [ne],
 * edge_dst_local [ne], CSC positions of the sampled edges [ne]; dst nodes of layer l are the first
 * n_dst src nodes */
int coala_sampler_layer_ptrs(coala_sampler* s, int layer, uint64_t* src_nodes, uint64_t* edge_src, uint64_t* edge_dst,
                             uint64_t* edge_pos);
/* edge ids of sampled CSC positions: out[i] = edge_ids[pos[i]] (edge_ids_ptr == 0: out[i] = pos[i]).
 * DGL's sampler reads the edge-id array while it samples (one more host round trip per picked edge);
 * here that lookup is deferred until a consumer asks for block.edata['_ID']. */
int coala_lookup_edge_ids(uint64_t edge_ids_ptr, uint64_t pos_dev_ptr, int64_t n, uint64_t out_dev_ptr, void* stream);
uint64_t coala_sampler_input_count_dev_ptr(coala_sampler* s); /* device int64: n input nodes */

/* ------------------------------------------------------------------ the iterator's per-batch enqueue sequence
 * replaces the body of COALA_GNN_DataLoader.__next__ (S/COALA_GNN_DataLoader.py:149-167: scheduler.run -> seeds .to(device)
 * -> sampler.sample -> fetch_feature, serial per batch) with one call per batch that enqueues seeds H2D + sample on a
 * slot stream and the cached gather on the gather stream; n_slots batches rotate (n_slots - 1 in flight behind the one
 * being consumed).  The batch's RNG key is coala_batch_key(seed, t, rank) (same arithmetic as oracle.batch_key). */
typedef struct coala_engine coala_engine;
uint64_t coala_batch_key(uint64_t seed, uint64_t batch_idx, int rank);
coala_engine* coala_engine_create(int device, int n_slots, int batch_size, uint64_t seed, int rank, int record_times);
void coala_engine_destroy(coala_engine* e);
/* slot k samples with `sampler` and gathers into out_ptr (device fp32[max_sample, feat_dim], caller-owned) */
int coala_engine_bind_slot(coala_engine* e, int k, coala_sampler* sampler, uint64_t out_ptr);
/* the gather of every slot: the fused cache kernel (nccl == NULL) or the NCCL exchange over its shard */
int coala_engine_bind_gather(coala_engine* e, coala_cache* cache, coala_nccl* nccl, int64_t max_sample);
int coala_engine_set_seed(coala_engine* e, uint64_t seed);
uint64_t coala_engine_slot_stream(coala_engine* e, int k);    /* cudaStream_t of slot k (distribution / refresh go here) */
uint64_t coala_engine_slot_seeds_dev(coala_engine* e, int k); /* device int64[batch_size] staging buffer of slot k */
uint64_t coala_engine_gather_stream(coala_engine* e);
/* everything the consumer enqueued so far on `consumer_stream` precedes any later slot reuse */
int coala_engine_checkpoint(coala_engine* e, void* consumer_stream);
/* two-step launch (device-resident seeds): begin returns the slot index (-1 on failure) after ordering the slot's stream;
 * the caller may enqueue the seed distribution on that stream; sample_gather then enqueues sample + gather.
 * consumer_stream == (void*)~0: no checkpoint (the caller made one for this step already). */
int coala_engine_begin(coala_engine* e, int64_t t, void* consumer_stream);
int coala_engine_sample_gather(coala_engine* e, int64_t t, uint64_t seeds_dev_ptr, int n_seeds);
/* one-step launch from a HOST seed list (the drop-in iterator): pinned staging copy + H2D + sample + gather */
int coala_engine_launch_host(coala_engine* e, int64_t t, uint64_t seeds_host_ptr, int n_seeds, void* consumer_stream);
/* consumer_stream waits for batch t's rows; with counts != NULL the host also waits for the block sizes
 * (counts[2L+1] as coala_sampler_counts) and the stream for the blocks */
int coala_engine_collect(coala_engine* e, int64_t t, int64_t* counts, void* consumer_stream);
int coala_engine_slot_of(coala_engine* e, int64_t t); /* slot holding batch t, -1 if it is not in flight */
/* record_times != 0: mean CUDA-event time of the gather launches and of the sample-stream work (distribution + sampling) */
int coala_engine_times(coala_engine* e, double* gather_ms_mean, uint64_t* n_gather, double* sample_ms_mean, uint64_t* n_sample,
                       int sync, int reset);

/* ------------------------------------------------------------------ offline colour metadata
 * replaces Graph_Coloring (M/graph_coloring.h:15-68, M/graph_coloring.cpp; bound at
 * M/COALA_GNN_Pybind.cu:65-77). Host code, as in the reference. */
typedef struct coala_gc coala_gc;
coala_gc* coala_gc_create(uint64_t num_nodes);
void coala_gc_destroy(coala_gc* g);
void coala_gc_set_adj_csc(coala_gc* g, uint64_t indptr_ptr, uint64_t indices_ptr);
void coala_gc_set_color_buffer(coala_gc* g, uint64_t ptr);
void coala_gc_set_topk_color_buffer(coala_gc* g, uint64_t ptr);
void coala_gc_set_topk_affinity_buffer(coala_gc* g, uint64_t ptr);
void coala_gc_cpu_color_graph(coala_gc* g);
void coala_gc_cpu_color_graph_optimized(coala_gc* g, uint64_t train_ptr, uint64_t n_train);
void coala_gc_cpu_count_nearest_color(coala_gc* g);
void coala_gc_cpu_count_nearest_color_less_memory(coala_gc* g);
void coala_gc_cpu_calculate_color_affinity(coala_gc* g);
uint64_t coala_gc_get_num_color(coala_gc* g);
uint64_t coala_gc_get_num_color_node(coala_gc* g);
/* GPU mode (the cpu_* functions above stay the bit-exact reference mode).
 * gpu_color_graph_optimized: same seeds and colour ids as cpu_color_graph_optimized (M/graph_coloring.cpp:77-92, host
 * rand() stream), traversal as a level-synchronous multi-source search on `device` (smallest colour id wins a contested
 * node; the reference's winner depends on its LIFO visit order, :143-157).
 * gpu_calculate_color_affinity: cpu_calculate_color_affinity (:254-294) for whatever colouring / hops the object holds
 * (from either mode): per-(colour, neighbour colour, hop) edge counts by radix sort, score = sum_h count_h * exp(-0.5 h),
 * top-k by (score desc, colour asc), divided by the colour's in-edge count.  Scores agree with the host mode to ~1e-15
 * relative (summation order), top-k lists agree wherever scores are not tied.  Both return 0 or -1 (coala_last_error). */
int coala_gc_gpu_color_graph_optimized(coala_gc* g, uint64_t train_ptr, uint64_t n_train, int device);
int coala_gc_gpu_calculate_color_affinity(coala_gc* g, int device);
/* test hook: hop (level) per node as uint16, 0 = none */
int coala_gc_get_hops(coala_gc* g, uint64_t dst_u16_ptr);

/* ------------------------------------------------------------------ synthetic data + checks (GPU) */
/* feat[r, c] = hash(seed, row0 + r, c) in [0,1), written through a device-visible address */
int coala_synth_fill_features(uint64_t dst_ptr, uint64_t seed, int64_t row0, int64_t nrows, int dim, int64_t stride_floats,
                              void* stream);
/* counts rows of `rows`[n, stride] whose first `dim` floats differ bitwise from hash(seed, idx[i], c) */
int coala_synth_check_rows(uint64_t rows_ptr, uint64_t idx_ptr, int64_t n, int dim, int64_t stride_floats, uint64_t seed,
                           int64_t* mismatches);
/* streaming-read bandwidth probe of a device-visible range (host link / peer link / HBM peaks) */
int coala_probe_read_bandwidth(uint64_t src_ptr, int64_t bytes, int iters, double* gb_per_s);

/* cudaDeviceEnablePeerAccess(peer) on `device` (already enabled = success); for the link probes */
int coala_enable_peer_access(int device, int peer);

/* row-gather probe: out[i] = table[idx[i]] with no cache protocol - the practical ceiling of the gather's access pattern on
 * whatever link `table` sits behind (local HBM, a peer's HBM, pinned host).  idx holds (iters + 1) * n ids: list 0 is the
 * untimed warm-up launch's, lists 1..iters one per timed launch (no timed launch re-reads rows the warm-up left in L2); rows_in_flight 1|2 per warp; ctas_per_sm CTAs of 256 threads per SM; returns the mean launch time (CUDA events). */
int coala_probe_gather(uint64_t table_ptr, uint64_t idx_ptr, uint64_t out_ptr, int64_t n, int row_bytes, int rows_in_flight,
                       int ctas_per_sm, int iters, double* us_per_launch);
/* the same with the table hash-partitioned over the G GPUs of this process like the cache (owner id % G, row id / G; 4 KB rows)
 * and all G GPUs gathering at once: tables[g] / idx[g] / out[g] are device g's part, id lists ((iters + 1) * n, list 0 = warm-up) and output rows;
 * peer access must be enabled between all pairs; us_per_launch[g] = device g's mean launch time while all run. */
int coala_probe_gather_all(int G, const uint64_t* tables, const uint64_t* idx, const uint64_t* out, int64_t n, int rows_in_flight,
                           int ctas_per_sm, int iters, double* us_per_launch);

/* random-read probe: independent locations per second when each location is read as `span_bytes`
 * (8/32/64/128/256) contiguous bytes by neighbouring lanes; misalign != 0 starts 96 B into a 256-B granule */
int coala_probe_random_read(uint64_t src_ptr, int64_t bytes, int span_bytes, int misalign, int warps, int iters,
                            double* locations_per_s);

#ifdef __cplusplus
}
#endif
#endif /* COALA_B200_H */
