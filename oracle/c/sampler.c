/* TEST INFRASTRUCTURE ONLY — see oracle.h.
 * CPU restatement of what the reference obtains from DGL at its call site
 * sampler.sample(g, seeds) (S/COALA_GNN_DataLoader.py:162; sampler built at
 * examples/sbatch_ssd_gnn_train.py:70-72; graph built from CSC at
 * examples/ssd_gnn_dataloader.py:523-524): per layer, in reversed fan-out order,
 * uniform in-neighbour sampling without replacement on the CSC followed by to_block.
 * DGL (PyPI `dgl`, "version 2.5" per README.md:12, no lock file) is absent from
 * /root/reference and from this image, so this restates DGL's documented behaviour:
 *   - deg <= fanout : every in-edge, CSC order;
 *   - deg  > fanout : `fanout` distinct in-edges, uniform;
 *   - to_block      : dst nodes = seeds in order, src nodes = seeds then unseen sources in
 *                     order of first appearance in the edge list.
 * DGL's RNG stream is internal, so the sampled branch is defined HERE (and implemented
 * identically on the GPU): a partial Fisher-Yates over the row's CSC positions driven by
 * the counter hash below, picks emitted in ascending CSC position. PARITY UNPINNED against
 * DGL itself; distribution tests stand in (tests/test_sampler_oracle.py). */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAX_FANOUT 64

static inline uint64_t mulhi64(uint64_t a, uint64_t b) { return (uint64_t)(((unsigned __int128)a * b) >> 64); }

/* f distinct positions in [0,deg), ascending */
static void pick_positions(uint64_t row_key, int64_t deg, int f, int64_t* picks) {
    int64_t mk[MAX_FANOUT], mv[MAX_FANOUT];
    int mn = 0;
    for (int i = 0; i < f; i++) {
        uint64_t r = orc_mix64(row_key + (uint64_t)(i + 1) * 0xD1342543DE82EF95ULL);
        int64_t j = i + (int64_t)mulhi64(r, (uint64_t)(deg - i));
        int64_t vj = j, vi = i;
        int sj = -1;
        for (int t = 0; t < mn; t++) {
            if (mk[t] == j) { vj = mv[t]; sj = t; }
            if (mk[t] == i) vi = mv[t];
        }
        picks[i] = vj;
        if (sj >= 0) mv[sj] = vi; else { mk[mn] = j; mv[mn] = vi; mn++; }
    }
    for (int a = 1; a < f; a++) { /* insertion sort */
        int64_t x = picks[a]; int b = a - 1;
        while (b >= 0 && picks[b] > x) { picks[b + 1] = picks[b]; b--; }
        picks[b + 1] = x;
    }
}

static inline int row_sample(const int64_t* indptr, int64_t v, int fanout, uint64_t rng_key, int64_t* pos /*cap MAX_FANOUT*/, int64_t* start_out, int64_t* deg_out) {
    int64_t start = indptr[v], deg = indptr[v + 1] - start;
    *start_out = start; *deg_out = deg;
    if (fanout < 0 || deg <= fanout) return -1; /* take all */
    uint64_t row_key = orc_mix64(rng_key ^ orc_mix64((uint64_t)v));
    pick_positions(row_key, deg, fanout, pos);
    return fanout;
}

int64_t orc_sample_layer(const int64_t* indptr, const int64_t* indices, const int64_t* edge_ids,
                         const int64_t* seeds, int64_t ns, int fanout, uint64_t rng_key,
                         int64_t cap, int64_t* e_src, int64_t* e_dst, int64_t* e_id) {
    int64_t ne = 0;
    for (int64_t i = 0; i < ns; i++) {
        int64_t pos[MAX_FANOUT], start, deg;
        int k = row_sample(indptr, seeds[i], fanout, rng_key, pos, &start, &deg);
        if (k < 0) {
            for (int64_t j = 0; j < deg; j++) {
                if (ne >= cap) return -1;
                e_src[ne] = indices[start + j]; e_dst[ne] = i; e_id[ne] = edge_ids ? edge_ids[start + j] : start + j; ne++;
            }
        } else {
            for (int j = 0; j < k; j++) {
                if (ne >= cap) return -1;
                e_src[ne] = indices[start + pos[j]]; e_dst[ne] = i; e_id[ne] = edge_ids ? edge_ids[start + pos[j]] : start + pos[j]; ne++;
            }
        }
    }
    return ne;
}

/* open-addressing first-occurrence map */
typedef struct { int64_t* k; int64_t* v; uint64_t mask; } fmap;
static void fmap_init(fmap* m, int64_t n) {
    uint64_t cap = 16; while ((int64_t)cap < 2 * n + 2) cap <<= 1;
    m->k = (int64_t*)malloc(cap * 8); m->v = (int64_t*)malloc(cap * 8); m->mask = cap - 1;
    memset(m->k, 0xFF, cap * 8);
}
static void fmap_free(fmap* m) { free(m->k); free(m->v); }
static inline int64_t fmap_get_or_put(fmap* m, int64_t key, int64_t val) {
    uint64_t h = orc_mix64((uint64_t)key) & m->mask;
    while (m->k[h] != -1) { if (m->k[h] == key) return m->v[h]; h = (h + 1) & m->mask; }
    m->k[h] = key; m->v[h] = val; return val;
}

int64_t orc_to_block(const int64_t* seeds, int64_t ns, const int64_t* e_src, int64_t ne,
                     int64_t* src_nodes, int64_t* e_src_local) {
    fmap m; fmap_init(&m, ns + ne);
    int64_t n = 0;
    for (int64_t i = 0; i < ns; i++) { fmap_get_or_put(&m, seeds[i], n); src_nodes[n] = seeds[i]; n++; }
    for (int64_t e = 0; e < ne; e++) {
        int64_t id = fmap_get_or_put(&m, e_src[e], n);
        if (id == n) { src_nodes[n] = e_src[e]; n++; }
        e_src_local[e] = id;
    }
    fmap_free(&m);
    return n;
}

int64_t orc_sample_blocks(const int64_t* indptr, const int64_t* indices, const int64_t* edge_ids,
                          const int64_t* seeds, int64_t ns, const int* fanouts, int L,
                          uint64_t rng_key, int64_t max_sample, int64_t* input_nodes,
                          int64_t* total_edges, int threads) {
    (void)edge_ids;
    int64_t* cur = (int64_t*)malloc((size_t)max_sample * 8);
    int64_t* nxt = (int64_t*)malloc((size_t)max_sample * 8);
    memcpy(cur, seeds, (size_t)ns * 8);
    int64_t n_cur = ns, tot = 0;
    for (int l = 0; l < L; l++) {
        int f = fanouts[L - 1 - l];
        uint64_t lkey = orc_mix64(rng_key + (uint64_t)(l + 1) * 0xA24BAED4963EE407ULL);
        int64_t stride = f;
        int64_t* slots = (int64_t*)malloc((size_t)(n_cur * stride) * 8);
        int32_t* cnt = (int32_t*)malloc((size_t)n_cur * 4);
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads)
        for (int64_t i = 0; i < n_cur; i++) {
            int64_t pos[MAX_FANOUT], start, deg;
            int k = row_sample(indptr, cur[i], f, lkey, pos, &start, &deg);
            if (k < 0) { for (int64_t j = 0; j < deg; j++) slots[i * stride + j] = indices[start + j]; cnt[i] = (int32_t)deg; }
            else { for (int j = 0; j < k; j++) slots[i * stride + j] = indices[start + pos[j]]; cnt[i] = k; }
        }
        int64_t ne = 0;
        for (int64_t i = 0; i < n_cur; i++) ne += cnt[i];
        int64_t* e_src = (int64_t*)malloc((size_t)(ne + 1) * 8);
        int64_t* e_loc = (int64_t*)malloc((size_t)(ne + 1) * 8);
        int64_t e = 0;
        for (int64_t i = 0; i < n_cur; i++) for (int j = 0; j < cnt[i]; j++) e_src[e++] = slots[i * stride + j];
        int64_t n_src = orc_to_block(cur, n_cur, e_src, ne, nxt, e_loc);
        tot += ne;
        free(slots); free(cnt); free(e_src); free(e_loc);
        int64_t* t = cur; cur = nxt; nxt = t; n_cur = n_src;
    }
    memcpy(input_nodes, cur, (size_t)n_cur * 8);
    if (total_edges) *total_edges = tot;
    free(cur); free(nxt);
    return n_cur;
}
