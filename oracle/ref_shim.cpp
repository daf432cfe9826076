// TEST INFRASTRUCTURE ONLY — never linked into or imported by the product path.
//
// Compiles the UNMODIFIED reference sources where they lie under /root/reference
// (REF_DIR is passed by oracle/Makefile) into oracle/_ref/libcoala_ref.so and puts a
// C ABI around the two host-side classes of the hot path that build without
// MPI / NVSHMEM / BaM:
//   * Node_distributor_pybind          (COALA_GNN_Modules/node_distributor_pybind.cuh:112-238)
//   * Graph_Coloring                   (COALA_GNN_Modules/graph_coloring.{h,cpp})
// No reference source is copied into this repository; this file only #includes them.
//
// The reference header calls three CUDA runtime functions just to get pinned host
// memory for the .npy tables; they are replaced by malloc/free below (4-line shim,
// SURVEY.md §8c).  graph_coloring.cpp malloc()s its hop buffer without clearing it
// (graph_coloring.cpp:331); inside a long-lived python process malloc hands back
// dirty memory, so malloc is mapped to calloc for that TU — the value fresh mmap'd
// pages give the reference at real graph sizes.
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#define cuda_err_chk(x) (x)
#define cudaHostAllocMapped 0
static inline int cudaHostAlloc(void** p, size_t n, unsigned) { *p = std::malloc(n); return 0; }
static inline int cudaFreeHost(void* p) { std::free(p); return 0; }

#define COALA_STR2(x) #x
#define COALA_STR(x) COALA_STR2(x)
#include COALA_STR(REF_DIR/node_distributor_pybind.cuh)

static inline void* coala_zero_malloc(size_t n) { return std::calloc(1, n); }
#define malloc(n) coala_zero_malloc(n)
#include COALA_STR(REF_DIR/graph_coloring.cpp)
#undef malloc

extern "C" {

void ref_srand(unsigned s) { srand(s); }

void* ref_nd_create(uint64_t item_ptr, int node_id, int batch, int local_size, int n_domains,
                    const char* color_file, const char* topk_file, const char* score_file) {
    try {
        return new Node_distributor_pybind(item_ptr, node_id, batch, local_size, n_domains,
                                           color_file, topk_file, score_file);
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_nd_create: %s\n", e.what());
        return nullptr;
    }
}
void ref_nd_distribute(void* h, uint64_t offset, uint64_t out_ptr, const uint64_t* meta_ptrs, int n) {
    std::vector<uint64_t> v(meta_ptrs, meta_ptrs + n);
    static_cast<Node_distributor_pybind*>(h)->distribute_node_with_affinity(offset, out_ptr, v);
}
int ref_nd_num_colors(void* h) { return static_cast<Node_distributor_pybind*>(h)->get_num_colors(); }
void ref_nd_destroy(void* h) { delete static_cast<Node_distributor_pybind*>(h); }

void* ref_gc_create(uint64_t n) { return new Graph_Coloring(n); }
void ref_gc_destroy(void* h) { delete static_cast<Graph_Coloring*>(h); }
void ref_gc_set_adj_csc(void* h, uint64_t indptr, uint64_t indices) { static_cast<Graph_Coloring*>(h)->set_adj_csc(indptr, indices); }
void ref_gc_set_color_buffer(void* h, uint64_t p) { static_cast<Graph_Coloring*>(h)->set_color_buffer(p); }
void ref_gc_set_topk_color_buffer(void* h, uint64_t p) { static_cast<Graph_Coloring*>(h)->set_topk_color_buffer(p); }
void ref_gc_set_topk_affinity_buffer(void* h, uint64_t p) { static_cast<Graph_Coloring*>(h)->set_topk_affinity_buffer(p); }
void ref_gc_color_graph(void* h) { static_cast<Graph_Coloring*>(h)->cpu_color_graph(); }
void ref_gc_color_graph_optimized(void* h, uint64_t train_ptr, uint64_t n) { static_cast<Graph_Coloring*>(h)->cpu_color_graph_optimized(train_ptr, n); }
void ref_gc_count_nearest_color(void* h) { static_cast<Graph_Coloring*>(h)->cpu_count_nearest_color(); }
void ref_gc_count_nearest_color_less_memory(void* h) { static_cast<Graph_Coloring*>(h)->cpu_count_nearest_color_less_memory(); }
void ref_gc_calculate_color_affinity(void* h) { static_cast<Graph_Coloring*>(h)->cpu_calculate_color_affinity(); }
uint64_t ref_gc_get_num_color(void* h) { return static_cast<Graph_Coloring*>(h)->get_num_color(); }
uint64_t ref_gc_get_num_color_node(void* h) { return static_cast<Graph_Coloring*>(h)->get_num_color_node(); }

}  // extern "C"
