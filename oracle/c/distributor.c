/* TEST INFRASTRUCTURE ONLY — see oracle.h.
 * Restates Node_distributor_pybind::distribute_node_with_affinity
 * (M/node_distributor_pybind.cuh:150-222): for each of the batch*local_size*n_domains seeds,
 * in arrival order, score every domain by sum_k meta_j[topk[c,k]] * score[c,k] (double,
 * multiply then add — compiled with -ffp-contract=off so no FMA, like x86-64 g++ -O3
 * without -march), force -1 for a full bucket (:205-206), take the first strict maximum
 * starting from (-1.0, domain 0) (:160-161,:208-211), append to that bucket (:214-218). */
#include "oracle.h"
#include <stdlib.h>

void orc_distribute(const int64_t* items, int64_t offset, int node_id, int batch, int local_size,
                    int n_domains, const int64_t* color, const int64_t* topk, const double* score,
                    int K, const int32_t* const* meta, int64_t* out, int* bucket_len_out) {
    const int cap = batch * local_size;              /* domain_batch_size  (:146) */
    const int64_t N = (int64_t)cap * n_domains;      /* global_batch_size  (:147) */
    int* bucket_len = (int*)calloc((size_t)n_domains, sizeof(int));
    for (int64_t i = 0; i < N; i++) {
        int64_t id = items[offset + i];
        int64_t c = color[id];
        int best = 0;
        double best_score = -1.0;
        for (int j = 0; j < n_domains; j++) {
            double s = 0.0;
            if (c != 0) {
                for (int k = 0; k < K; k++) {
                    int64_t nc = topk[(c - 1) * K + k];
                    double aff = score[(c - 1) * K + k];
                    if (nc != 0) {
                        int32_t m = meta[j][nc];
                        if (m == 0) continue;
                        double prod = (double)m * aff;
                        s = s + prod;
                    }
                }
            }
            if (bucket_len[j] == cap) s = -1.0;
            if (s > best_score) { best = j; best_score = s; }
        }
        /* the reference writes out[bucket_len] unchecked; an over-full bucket 0 (only
         * reachable when every open domain scores < -1, i.e. negative counters) would
         * overrun the caller's buffer there — the oracle drops those writes instead. */
        if (best == node_id && bucket_len[best] < cap) out[bucket_len[best]] = id;
        bucket_len[best] += 1;
    }
    if (bucket_len_out) for (int j = 0; j < n_domains; j++) bucket_len_out[j] = bucket_len[j];
    free(bucket_len);
}
