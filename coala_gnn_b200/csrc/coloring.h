// State of the offline colour-metadata generator (Graph_Coloring, M/graph_coloring.h:15-68), shared by
// the host implementation (coloring.cpp, the bit-exact reference mode) and the GPU mode (coloring_gpu.cu).
#pragma once
#include <cstdint>
#include <utility>
#include <vector>

struct coala_gc {
    uint64_t num_nodes = 0, num_colored = 0, next_color = 1;
    int max_hop = 10;            // global_max_hop (M/graph_coloring.h:19)
    float base_rate = 0.005f;    // global_sampling_rate (:20)
    int topk = 10;               // (:22)
    uint64_t* color = nullptr;
    uint64_t* topk_color = nullptr;
    double* topk_aff = nullptr;
    const uint64_t* indptr = nullptr;
    const uint64_t* indices = nullptr;
    std::vector<uint16_t> hop;
    std::vector<uint8_t> is_train;
    std::vector<std::pair<uint64_t, uint64_t>> work[2];  // (node, colour)
    // GPU mode: HBM copy of the CSC arrays, made on first use and shared by colouring and affinity
    int gpu_device = -1;
    uint64_t* d_indptr = nullptr;
    uint64_t* d_indices = nullptr;
};

// seeds of cpu_color_graph_optimized (M/graph_coloring.cpp:77-92): same rand() stream, same colour ids
void coala_gc_draw_train_seeds(coala_gc* g, const int64_t* train, uint64_t n_train);
// frees the HBM copies (no-op when the GPU mode was never used)
void coala_gc_gpu_release(coala_gc* g);
