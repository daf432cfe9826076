// Offline colour metadata: seeded multi-source traversal colouring + per-colour top-k affinity.
// Replaces Graph_Coloring (M/graph_coloring.h:15-68, M/graph_coloring.cpp), host code there and
// here.  The outputs (color / topk / score) must be bit-identical to the reference's, which pins
// three library behaviours this file therefore shares with it on purpose:
//   * seeds are drawn with glibc rand() (never srand()'ed by the reference => seed 1 stream),
//     compared as float(rand())/RAND_MAX <= rate (M/graph_coloring.cpp:84-90);
//   * traversal is LIFO over two std::vector work lists (pop_back), not a FIFO BFS (:143-157);
//   * ties among a colour's neighbour scores are broken by std::sort (introsort, unstable) over
//     std::unordered_map iteration order (:165-181) — so the same containers are used here.
// Known reference quirks kept: colour ids are 1-based and the loop over colours stops one short,
// so the last colour's topk/score row stays zero (:266); nodes coloured by the final flush carry
// no hop (:158-159) — the hop buffer is zero-initialised here (the reference leaves it
// uninitialised, :331; fresh pages read as 0).
#include "../../include/coala_b200.h"
#include "coloring.h"
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <unordered_map>
#include <utility>
#include <vector>

namespace {

inline void push_in_neighbours(coala_gc* g, int buf, uint64_t node, uint64_t c) {
    for (uint64_t e = g->indptr[node]; e < g->indptr[node + 1]; e++) g->work[buf].emplace_back(g->indices[e], c);
}

// drain a work list; returns through `expand` whether newly coloured nodes push their in-neighbours
template <bool EXPAND, bool SET_HOP>
void drain(coala_gc* g, int cur, int nxt, int hop_value) {
    auto& w = g->work[cur];
    while (!w.empty()) {
        const auto item = w.back();
        w.pop_back();
        if (g->color[item.first] != 0) continue;
        g->color[item.first] = item.second;
        if (SET_HOP) g->hop[item.first] = (uint16_t)hop_value;
        g->num_colored++;
        if (EXPAND) push_in_neighbours(g, nxt, item.first, item.second);
    }
}

template <typename T>
std::vector<std::pair<uint64_t, T>> top_k(const std::unordered_map<uint64_t, T>& m, size_t k) {
    std::vector<std::pair<uint64_t, T>> v(m.begin(), m.end());
    std::sort(v.begin(), v.end(),
              [](const std::pair<uint64_t, T>& a, const std::pair<uint64_t, T>& b) { return a.second > b.second; });
    if (v.size() > k) v.resize(k);
    return v;
}

// nodes grouped by colour, ascending node id inside a colour (what the reference's
// unordered_map<colour, vector<node>> holds, M/graph_coloring.cpp:257-263)
void group_by_color(const coala_gc* g, uint64_t num_c, std::vector<uint64_t>& start, std::vector<uint64_t>& nodes) {
    start.assign(num_c + 2, 0);
    for (uint64_t v = 0; v < g->num_nodes; v++) {
        uint64_t c = g->color[v];
        if (c != 0 && c <= num_c) start[c + 1]++;
    }
    for (uint64_t c = 1; c < start.size(); c++) start[c] += start[c - 1];
    nodes.resize(start.back());
    std::vector<uint64_t> fill(start.begin(), start.end() - 1);
    for (uint64_t v = 0; v < g->num_nodes; v++) {
        uint64_t c = g->color[v];
        if (c != 0 && c <= num_c) nodes[fill[c]++] = v;
    }
}

}  // namespace

void coala_gc_draw_train_seeds(coala_gc* g, const int64_t* train, uint64_t n_train) {
    for (uint64_t t = 0; t < n_train; t++) g->is_train[train[t]] = 1;                        // (:99-105)
    const double frac = std::min(20.0, (double)g->num_nodes / (double)n_train);               // (:77)
    const double rate = g->base_rate * frac;                                                  // (:78) float * double
    for (uint64_t t = 0; t < n_train; t++) {                                                  // (:81-92)
        const int64_t v = train[t];
        if (g->color[v] != 0) continue;
        float samp = static_cast<float>(rand()) / RAND_MAX;
        if (samp <= rate) g->work[0].emplace_back((uint64_t)v, g->next_color++);
    }
}

extern "C" {

coala_gc* coala_gc_create(uint64_t num_nodes) {
    coala_gc* g = new coala_gc();
    g->num_nodes = num_nodes;
    g->hop.assign(num_nodes, 0);
    g->is_train.assign(num_nodes, 0);
    return g;
}
void coala_gc_destroy(coala_gc* g) {
    if (g) coala_gc_gpu_release(g);
    delete g;
}
void coala_gc_set_adj_csc(coala_gc* g, uint64_t indptr, uint64_t indices) {
    g->indptr = reinterpret_cast<const uint64_t*>(indptr);
    g->indices = reinterpret_cast<const uint64_t*>(indices);
}
void coala_gc_set_color_buffer(coala_gc* g, uint64_t p) { g->color = reinterpret_cast<uint64_t*>(p); }
void coala_gc_set_topk_color_buffer(coala_gc* g, uint64_t p) { g->topk_color = reinterpret_cast<uint64_t*>(p); }
void coala_gc_set_topk_affinity_buffer(coala_gc* g, uint64_t p) { g->topk_aff = reinterpret_cast<double*>(p); }
uint64_t coala_gc_get_num_color(coala_gc* g) { return g->next_color == 0 ? 0 : g->next_color - 1; }
uint64_t coala_gc_get_num_color_node(coala_gc* g) { return g->num_colored; }

// cpu_color_graph (M/graph_coloring.cpp:43-73): seeds drawn over all nodes at the base rate
void coala_gc_cpu_color_graph(coala_gc* g) {
    for (uint64_t v = 0; v < g->num_nodes; v++) {
        if (g->color[v] != 0) continue;
        float samp = static_cast<float>(rand()) / RAND_MAX;
        if (samp <= g->base_rate) g->work[0].emplace_back(v, g->next_color++);
    }
    int hop = 0;
    for (; hop < g->max_hop; hop++) drain<true, false>(g, hop % 2, (hop + 1) % 2, 0);
    drain<false, true>(g, (hop + 1) % 2, 0, hop + 2);  // cpu_flush_buffer<true>(next, hop+1) stores hop+1+1
}

// cpu_color_graph_optimized (M/graph_coloring.cpp:108-160)
void coala_gc_cpu_color_graph_optimized(coala_gc* g, uint64_t train_ptr, uint64_t n_train) {
    const int64_t* train = reinterpret_cast<const int64_t*>(train_ptr);
    coala_gc_draw_train_seeds(g, train, n_train);
    int hop = 0;
    for (; hop < g->max_hop; hop++) {
        const int cur = hop % 2, nxt = (hop + 1) % 2;
        if (hop == 0) {  // uncoloured TRAINING in-neighbours of each seed join list 0 with the seed's colour (:125-141)
            const size_t n0 = g->work[0].size();
            for (size_t s = 0; s < n0; s++) {
                const uint64_t node = g->work[0][s].first, c = g->work[0][s].second;
                for (uint64_t e = g->indptr[node]; e < g->indptr[node + 1]; e++) {
                    const uint64_t u = g->indices[e];
                    if (g->is_train[u] && g->color[u] == 0) g->work[0].emplace_back(u, c);
                }
            }
        }
        drain<true, true>(g, cur, nxt, hop + 1);                                              // (:143-157)
    }
    drain<false, false>(g, (hop + 1) % 2, 0, 0);                                              // cpu_flush_buffer<false> (:158-159)
}

// cpu_calculate_color_affinity (M/graph_coloring.cpp:254-294)
void coala_gc_cpu_calculate_color_affinity(coala_gc* g) {
    const uint64_t num_c = g->next_color - 1;
    std::vector<uint64_t> start, nodes;
    group_by_color(g, num_c, start, nodes);
    for (uint64_t c = 0; c < num_c; c++) {  // colour num_c itself is never visited (reference loop bound)
        std::unordered_map<uint64_t, double> acc;
        double in_edges = 0;
        for (uint64_t k = start[c]; k < start[c + 1]; k++) {
            const uint64_t node = nodes[k];
            in_edges += (double)(g->indptr[node + 1] - g->indptr[node]);
            for (uint64_t e = g->indptr[node]; e < g->indptr[node + 1]; e++) {
                const uint64_t u = g->indices[e];
                const uint64_t nc = g->color[u];
                if (nc != 0 && nc != c) acc[nc] += std::exp(-0.5 * (double)g->hop[u]);   // score_func (:250-252)
            }
        }
        if (c == 0) continue;  // colour 0 owns no nodes; the reference computes an empty list for it
        auto best = top_k<double>(acc, (size_t)g->topk);
        for (size_t i = 0; i < best.size(); i++) {
            g->topk_color[(c - 1) * g->topk + i] = best[i].first;
            if (g->topk_aff) g->topk_aff[(c - 1) * g->topk + i] = best[i].second / in_edges;
        }
    }
}

// cpu_count_nearest_color_less_memory (M/graph_coloring.cpp:218-246): edge counts instead of scores
void coala_gc_cpu_count_nearest_color_less_memory(coala_gc* g) {
    const uint64_t num_c = g->next_color - 1;
    std::vector<uint64_t> start, nodes;
    group_by_color(g, num_c, start, nodes);
    for (uint64_t c = 1; c < num_c; c++) {
        std::unordered_map<uint64_t, uint64_t> acc;
        for (uint64_t k = start[c]; k < start[c + 1]; k++)
            for (uint64_t e = g->indptr[nodes[k]]; e < g->indptr[nodes[k] + 1]; e++) {
                const uint64_t nc = g->color[g->indices[e]];
                if (nc != 0 && nc != c) acc[nc] += 1;
            }
        auto best = top_k<uint64_t>(acc, (size_t)g->topk);
        for (size_t i = 0; i < best.size(); i++) g->topk_color[(c - 1) * g->topk + i] = best[i].first;
    }
}

// cpu_count_nearest_color (M/graph_coloring.cpp:184-214): one pass over all nodes into a colour x colour
// map. The reference also accumulates a row for colour 0 and writes it at index (0-1)*topk, out of
// bounds; that row is dropped here.
void coala_gc_cpu_count_nearest_color(coala_gc* g) {
    std::unordered_map<uint64_t, std::unordered_map<uint64_t, uint64_t>> conn;
    for (uint64_t v = 0; v < g->num_nodes; v++) {
        const uint64_t c = g->color[v];
        for (uint64_t e = g->indptr[v]; e < g->indptr[v + 1]; e++) {
            const uint64_t nc = g->color[g->indices[e]];
            if (nc != 0 && nc != c) conn[c][nc] += 1;
        }
    }
    for (const auto& row : conn) {
        if (row.first == 0) continue;
        auto best = top_k<uint64_t>(row.second, (size_t)g->topk);
        for (size_t i = 0; i < best.size(); i++) g->topk_color[(row.first - 1) * g->topk + i] = best[i].first;
    }
}

}  // extern "C"
