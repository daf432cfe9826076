/* TEST INFRASTRUCTURE ONLY — see oracle.h.
 * Sequential model of the reference's 32-way set-associative feature cache, one request at
 * a time in request order (the reference runs one warp per request; with a single warp it
 * degenerates to exactly this):
 *   sizing      M/ssd_gnn_cache.cuh:96-97   num_pages = MB*2^20/page_size, num_sets = pages/32
 *   routing     M/cache_kernel.cu:11,86     owner = id % G
 *   set hash    M/isolated_cache.h:190-202  id % S (isolated) | (id/G) % S (partitioned, nccl)
 *   lookup      M/isolated_cache.h:145-175  linear tag match over 32 ways (keys init 0xFF..)
 *   miss        M/isolated_cache.h:421-450  way = set_cnt[set]++ % 32; counters[color_meta[slot]]--;
 *                                           keys[slot]=id; color_meta[slot]=color[id]; counters[color]++
 *   fill        M/isolated_cache.h:323-331  line = host[id*CL_SIZE .. +CL_SIZE)
 *   output      M/cache_kernel.cu:71-75     out + i*dim  <- CL_SIZE bytes
 * color_counters has num_colors+1 slots here (the reference allocates num_colors and indexes
 * with the raw 1-based colour, one past the end for the last colour — SURVEY §3.2). */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

#define WAYS 32

typedef struct {
    uint64_t* keys;
    uint64_t* color_meta;
    uint32_t* set_cnt;
    int32_t* counters;
    uint8_t* lines;
} shard_t;

struct orc_cache {
    uint64_t num_sets;
    uint32_t page_size;
    int n_gpus, mode, num_colors, store_lines;
    const int64_t* color;
    shard_t* sh;
    int n_shards;
    uint64_t hit, miss;
};

orc_cache* orc_cache_create(uint64_t cache_mb, uint32_t page_size, int n_gpus, int mode,
                            const int64_t* color, int num_colors, int store_lines) {
    orc_cache* c = (orc_cache*)calloc(1, sizeof(*c));
    uint64_t pages = cache_mb * 1024ULL * 1024ULL / page_size;
    c->num_sets = pages / WAYS;
    c->page_size = page_size;
    c->n_gpus = n_gpus;
    c->mode = mode;
    c->color = color;
    c->num_colors = num_colors;
    c->store_lines = store_lines;
    c->n_shards = mode == 1 ? n_gpus : 1;
    c->sh = (shard_t*)calloc((size_t)c->n_shards, sizeof(shard_t));
    uint64_t slots = c->num_sets * WAYS;
    for (int g = 0; g < c->n_shards; g++) {
        shard_t* s = &c->sh[g];
        s->keys = (uint64_t*)malloc(slots * 8);
        memset(s->keys, 0xFF, slots * 8);
        s->color_meta = (uint64_t*)calloc(slots, 8);
        s->set_cnt = (uint32_t*)calloc(c->num_sets, 4);
        s->counters = (int32_t*)calloc((size_t)num_colors + 1, 4);
        s->lines = store_lines ? (uint8_t*)malloc(slots * page_size) : NULL;
    }
    return c;
}

void orc_cache_destroy(orc_cache* c) {
    if (!c) return;
    for (int g = 0; g < c->n_shards; g++) {
        free(c->sh[g].keys); free(c->sh[g].color_meta); free(c->sh[g].set_cnt);
        free(c->sh[g].counters); free(c->sh[g].lines);
    }
    free(c->sh);
    free(c);
}

uint64_t orc_cache_num_sets(const orc_cache* c) { return c->num_sets; }

void orc_cache_read(orc_cache* c, const int64_t* idx, int64_t n, const float* feat,
                    float* out, int64_t out_stride) {
    const uint32_t R = c->page_size;
    for (int64_t i = 0; i < n; i++) {
        uint64_t id = (uint64_t)idx[i];
        int g = 0;
        uint64_t set;
        if (c->mode == 1) { g = (int)(id % (uint64_t)c->n_gpus); set = (id / (uint64_t)c->n_gpus) % c->num_sets; }
        else set = id % c->num_sets;
        shard_t* s = &c->sh[g];
        uint64_t base = set * WAYS;
        int way = -1;
        for (int w = 0; w < WAYS; w++) if (s->keys[base + w] == id) { way = w; break; }
        const uint8_t* host_row = (const uint8_t*)feat + id * (uint64_t)R;
        uint8_t* dst = (uint8_t*)(out + i * out_stride);
        if (way >= 0) {
            c->hit++;
            memcpy(dst, s->lines ? s->lines + (base + way) * R : host_row, R);
        } else {
            c->miss++;
            way = (int)(s->set_cnt[set]++ % WAYS);
            uint64_t slot = base + way;
            s->counters[s->color_meta[slot]] -= 1;
            s->keys[slot] = id;
            int64_t col = c->color ? c->color[id] : 0;
            s->color_meta[slot] = (uint64_t)col;
            s->counters[col] += 1;
            if (s->lines) { memcpy(s->lines + slot * R, host_row, R); memcpy(dst, s->lines + slot * R, R); }
            else memcpy(dst, host_row, R);
        }
    }
}

void orc_cache_stats(const orc_cache* c, uint64_t* hit, uint64_t* miss) { *hit = c->hit; *miss = c->miss; }
const uint64_t* orc_cache_keys(const orc_cache* c, int gpu) { return c->sh[gpu].keys; }
const uint64_t* orc_cache_color_meta(const orc_cache* c, int gpu) { return c->sh[gpu].color_meta; }
const int32_t* orc_cache_color_counters(const orc_cache* c, int gpu) { return c->sh[gpu].counters; }
const uint32_t* orc_cache_set_cnt(const orc_cache* c, int gpu) { return c->sh[gpu].set_cnt; }
