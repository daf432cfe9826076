/* TEST INFRASTRUCTURE ONLY — see oracle.h. Deterministic generators shared with the GPU
 * (coala_gnn_b200/csrc/common.cuh implements the same integer arithmetic). */
#include "oracle.h"

uint64_t orc_mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

float orc_feat_value(uint64_t seed, uint64_t row, uint64_t col) {
    uint64_t hr = orc_mix64(seed ^ row);
    uint64_t h = orc_mix64(hr + col);
    return (float)(h >> 40) * (1.0f / 16777216.0f); /* 24 bits: exact in fp32 */
}

void orc_fill_features(float* dst, uint64_t seed, uint64_t row0, uint64_t nrows, uint64_t dim, uint64_t stride) {
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < (int64_t)nrows; r++) {
        uint64_t hr = orc_mix64(seed ^ (row0 + (uint64_t)r));
        float* p = dst + (uint64_t)r * stride;
        for (uint64_t c = 0; c < dim; c++) p[c] = (float)(orc_mix64(hr + c) >> 40) * (1.0f / 16777216.0f);
    }
}

void orc_index_select(const float* feat, int64_t dim, const int64_t* idx, int64_t n, float* out, int threads) {
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int64_t i = 0; i < n; i++) {
        const float* s = feat + idx[i] * dim;
        float* d = out + i * dim;
        for (int64_t c = 0; c < dim; c++) d[c] = s[c];
    }
}
