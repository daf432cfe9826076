// Device-side layout and protocol of the feature cache (stage 3).
//
// What the reference does (M/isolated_cache.h:335-475, M/nvshmem_cache.h:336-480): one warp per
// requested node id; set seqlock -> 32-way tag search -> way seqlock -> 4-byte/lane copy ->
// re-check; on a miss round-robin evict, colour bookkeeping, fill the line from the pinned host
// table, re-read the line, copy out.  Peer shards are served by their OWNER (request exchange,
// owner-side kernels, NVSHMEM puts, 4 barriers per batch).
//
// What this build does instead (B200-first):
//   * one fused kernel; the REQUESTER's warp resolves the id itself against whichever shard owns
//     it — its own HBM or a peer's, mapped with CUDA IPC and read over NVLink/NVSwitch with plain
//     loads — so there is no request exchange, no owner-side kernel and no barrier;
//   * tag word = [version:24 | id:40]; an odd version marks a line being refilled.  The version
//     is the line's seqlock (the reference's separate way-lock array is folded into the tag), so
//     a hit costs: 256-B coalesced tag probe -> row -> 8-B re-check;
//   * the per-set word is a real lock (CAS even->odd) whose value/2 is also the round-robin
//     pointer (the reference bumps set_cnt_ exactly once per write section, so the two agree);
//     writers re-probe under the lock, so two GPUs missing on the same id never duplicate a line;
//   * rows move as 16-byte vectors held in registers: a miss forwards the host row to BOTH the
//     line and the output from the same registers (the reference writes the line, re-reads it,
//     then writes the output);
//   * nobody waits unboundedly on another warp/GPU: a busy line or a contended set falls back to
//     reading the (immutable) host row directly — counted as `bypass`.
#pragma once
#include "common.cuh"
#include "../../include/coala_b200.h"

constexpr uint64_t TAG_ID_MASK = (1ULL << 40) - 1;  // empty slot: id field all ones, version 0
constexpr int TAG_VER_SHIFT = 40;

struct ShardView {           // every pointer is valid in the address space of the process using it
    uint64_t* tags;          // [S*32]  version<<40 | id
    uint64_t* set_seq;       // [S]     odd = locked; value>>1 = number of fills = round-robin pointer
    uint32_t* color_meta;    // [S*32]  colour of the resident line (M/isolated_cache.h color_meta_)
    int32_t* color_counters; // [num_colors+1] lines per colour (M/isolated_cache.h color_counters)
    uint4* lines;            // [S*32*R/16]
};

struct CacheParams {
    ShardView shard[COALA_MAX_GPUS];
    int route_gpus;          // id % route_gpus picks the shard (1 = always shard[my_slot])
    int hash_div;            // set = (id / hash_div) % num_sets
    int my_slot;             // index of this GPU's own shard in shard[]
    uint32_t row_bytes;      // R = cache_dim * 4 (CL_SIZE)
    uint64_t num_sets;
    const uint8_t* host_feat;  // pinned host table, row stride R (reference sim_buf)
    const int32_t* color;      // device copy of color[N] (reference reads the pinned int64 array)
    int track_color;
    int out_stride;          // floats between output rows (= feat_dim)
    int out_dim;             // floats copied per output row (min(feat_dim, cache_dim))
    unsigned long long* stats;  // [0]=hit [1]=miss [2]=bypass [3]=ids refused (>= id_limit)
    uint64_t id_limit;       // ids >= id_limit are refused: min(rows of the host table, 2^40 - 1 = the tag's id field)
    uint32_t zero;           // always 0 at run time; unknown to the compiler (see read_line_validated)
    int fenced;              // 1 (default): writers order their stores with fence.acq_rel (PTX-model-exact); 0: by read-back + dependency
};

constexpr int LOCK_TRIES = 64;     // CAS attempts on a set lock before bypassing
constexpr int BUSY_POLLS = 256;    // polls of a line being filled by someone else before bypassing

template <int VPL>
__device__ __forceinline__ void store_out_row(const CacheParams& p, float* out_row, const uint4 (&v)[VPL], unsigned lane) {
    if ((p.out_dim & 3) == 0 && (p.out_stride & 3) == 0) {
        const int nvec = p.out_dim >> 2;
#pragma unroll
        for (int k = 0; k < VPL; k++) {
            int vi = k * 32 + lane;
            if (vi < nvec) st_out_v4(reinterpret_cast<uint4*>(out_row) + vi, v[k]);
        }
    } else {  // feat_dim not a multiple of 4 floats: rows are not 16-B aligned
#pragma unroll
        for (int k = 0; k < VPL; k++) {
            int e = (k * 32 + lane) * 4;
            const uint32_t w[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (e + j < p.out_dim) out_row[e + j] = __uint_as_float(w[j]);
        }
    }
}

template <int VPL>
__device__ __forceinline__ void load_host_row(const CacheParams& p, uint64_t id, uint4 (&v)[VPL], unsigned lane) {
    const uint4* src = reinterpret_cast<const uint4*>(p.host_feat + id * (uint64_t)p.row_bytes);
#pragma unroll
    for (int k = 0; k < VPL; k++) v[k] = ld_host_v4(src + k * 32 + lane);
}

// Optimistic read of `slot` whose tag word was observed as `tw` (even version, id matches).
// Returns true when the registers hold a consistent copy of the line.
//
// Seqlock reader WITHOUT a fence.  A fence here is MEMBAR.ALL.{GPU,SYS} (+ an L1 invalidate for
// every acquire load): on the hit path that is the dominant cost, and over NVLink it drains the SM.
// Instead the three steps are chained by register dependencies, which the hardware cannot reorder:
//   tag probe --(slot address computed from the tag)--> row loads --(address of the re-check is
//   computed from the loaded row words, `& p.zero`)--> tag re-check.
// All of them are strong (.relaxed.gpu/.sys) accesses that resolve at the owner's L2, the single
// point of coherence of that memory.  The re-check is issued after every row word has returned, so
// it observes the tag at a later time than every row read; the writer orders "tag odd -> row
// stores -> tag even" with fences (slow path below).  If both tag observations are equal and
// even, no writer touched the row in between.
template <int VPL, bool SYS>
__device__ __forceinline__ bool read_line_validated(const CacheParams& p, const ShardView& sh, uint64_t slot, uint64_t tw,
                                                    uint32_t row_vecs, uint4 (&v)[VPL], unsigned lane) {
    const uint4* src = sh.lines + slot * row_vecs;
#pragma unroll
    for (int k = 0; k < VPL; k++) v[k] = ld_line_v4<SYS>(src + k * 32 + lane);
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < VPL; k++) acc |= v[k].x | v[k].y | v[k].z | v[k].w;
    const uint64_t t2 = ld_relaxed_u64<SYS>(sh.tags + slot + (acc & p.zero));
    return __all_sync(0xffffffffu, t2 == tw);
}

// One warp resolves one node id. Returns 0 hit, 1 miss (filled), 2 bypass.
// `g`/`set` locate the id (owner shard, set); `t` is this lane's tag of that set, probed earlier
// (prefetched while the previous row was in flight).
template <int VPL, bool SYS>
__device__ __forceinline__ int resolve_row(const CacheParams& p, uint64_t id, int g, uint64_t set, uint64_t t,
                                           float* out_row, unsigned lane) {
    const ShardView& sh = p.shard[g];
    const uint64_t slot0 = set * COALA_WAYS;
    const uint32_t row_vecs = p.row_bytes >> 4;
    uint4 v[VPL];

    // ---- fast path: 32 tags were probed with one coalesced 256-B access
    unsigned m = __ballot_sync(0xffffffffu, (t & TAG_ID_MASK) == id);
    if (m) {
        int way = __ffs(m) - 1;
        uint64_t tw = __shfl_sync(0xffffffffu, t, way);
        if (((tw >> TAG_VER_SHIFT) & 1) == 0) {
            // the row is written out before it is validated (the re-check's latency then overlaps the
            // next row's loads); a failed validation falls through and the slow path rewrites the row
            const uint4* src = sh.lines + (slot0 + way) * row_vecs;
#pragma unroll
            for (int k = 0; k < VPL; k++) v[k] = ld_line_v4<SYS>(src + k * 32 + lane);
            store_out_row<VPL>(p, out_row, v, lane);
            uint32_t acc = 0;
#pragma unroll
            for (int k = 0; k < VPL; k++) acc |= v[k].x | v[k].y | v[k].z | v[k].w;
            const uint64_t t2 = ld_relaxed_u64<SYS>(sh.tags + slot0 + way + (acc & p.zero));
            if (__all_sync(0xffffffffu, t2 == tw)) return 0;
        }
    }

    // ---- the line is being filled right now (id matches, odd version: by another warp of this launch
    // or by a peer): wait for the writer instead of fetching the same row from the host a second time
    if (m) {
        const int way = __ffs(m) - 1;
        const uint64_t slot = slot0 + way;
        uint64_t tw = __shfl_sync(0xffffffffu, t, way);
#pragma unroll 1
        for (int poll = 0; poll < BUSY_POLLS && (tw & TAG_ID_MASK) == id; poll++) {
            if (((tw >> TAG_VER_SHIFT) & 1) == 0) {
                if (read_line_validated<VPL, SYS>(p, sh, slot, tw, row_vecs, v, lane)) {
                    store_out_row<VPL>(p, out_row, v, lane);
                    return 0;
                }
            } else {
                __nanosleep(200);
            }
            tw = ld_relaxed_u64<SYS>(sh.tags + slot);
        }
    }

    // ---- slow path. Start the host read now: it overlaps the lock round trips, and the host
    // table is immutable so the registers are valid whatever happens to the line meanwhile.
    load_host_row<VPL>(p, id, v, lane);

    // Writer-side ordering.  The protocol needs three orders: (a) the lock is taken before the set is re-probed, (b) the
    // busy tag is in place before the unlock and before any line word changes, (c) every line word is in place before the
    // tag turns even again.
    //   p.fenced != 0 (DEFAULT since round 3): PTX acquire on the lock CAS and the re-probe, one fence.acq_rel for (b), one
    //   for (c) - two MEMBAR.{GPU,SYS} per fill (round 2 had three; the release store of the unlock was redundant next to
    //   the fence).  A MEMBAR waits for every access the thread has in flight, the host-row loads over PCIe included.  At
    //   GPU scope (one GPU per cache domain) this costs nothing measurable; at system scope (shards of peers) it costs
    //   19 % on a 3 %-miss launch and 23 % of the N = 2 pipeline (profiles/r3_fenced_vs_readback.txt).
    //   p.fenced == 0 (COALA_CACHE_FENCED=0): the same orders by READ-BACK + register dependency, like the reader side -
    //   all of these are strong accesses that resolve at the owner's L2, the single coherence point of the shard: a strong
    //   load of the word just stored returns only once that store is performed there (same-address order), and an access
    //   whose address depends on the returned value is issued after it.  Two round trips to the owner's L2 per fill, no
    //   MEMBAR.  The PTX memory model does not promise dependency ordering of relaxed accesses: this mode is validated
    //   empirically on B200 only (storm tests), which is why it is opt-in.
    const bool fenced = p.fenced != 0;
    uint64_t s = 0;
    bool locked = false;
    if (lane == 0) {
        unsigned ns = 32;
#pragma unroll 1
        for (int tries = 0; tries < LOCK_TRIES; tries++) {
            s = ld_relaxed_u64<SYS>(sh.set_seq + set);
            if ((s & 1) == 0 && (fenced ? cas_acquire_u64<SYS>(sh.set_seq + set, s, s + 1) : cas_relaxed_u64<SYS>(sh.set_seq + set, s, s + 1)) == s) {
                locked = true;
                break;
            }
            __nanosleep(ns);
            if (ns < 1024) ns <<= 1;
        }
    }
    locked = __shfl_sync(0xffffffffu, (int)locked, 0);
    s = __shfl_sync(0xffffffffu, s, 0);
    if (!locked) {  // contended set: serve from host, do not cache
        store_out_row<VPL>(p, out_row, v, lane);
        return 2;
    }

    // ---- under the set lock: re-probe (someone may have filled it since the optimistic probe).  (a): the probe's
    // address depends on the value the CAS returned.
    t = fenced ? ld_acquire_u64<SYS>(sh.tags + slot0 + lane) : ld_relaxed_u64<SYS>(sh.tags + slot0 + lane + ((uint32_t)s & p.zero));
    m = __ballot_sync(0xffffffffu, (t & TAG_ID_MASK) == id);
    if (m) {
        // present (possibly still being filled by another warp / GPU): nothing to modify,
        // so restore the even value — readers that sampled `s` stay valid, no way is consumed.
        if (lane == 0) {
            if (fenced) st_release_u64<SYS>(sh.set_seq + set, s);
            else st_relaxed_u64<SYS>(sh.set_seq + set, s);
        }
        int way = __ffs(m) - 1;
        const uint64_t slot = slot0 + way;
        uint64_t tw = __shfl_sync(0xffffffffu, t, way);
#pragma unroll 1
        for (int poll = 0; poll < BUSY_POLLS; poll++) {
            if ((tw & TAG_ID_MASK) != id) break;  // evicted again meanwhile
            if (((tw >> TAG_VER_SHIFT) & 1) == 0) {
                if (read_line_validated<VPL, SYS>(p, sh, slot, tw, row_vecs, v, lane)) {
                    store_out_row<VPL>(p, out_row, v, lane);
                    return 0;
                }
            } else {
                __nanosleep(256);
            }
            tw = ld_relaxed_u64<SYS>(sh.tags + slot);
        }
        load_host_row<VPL>(p, id, v, lane);  // registers were reused for the line: host copy is bit-identical
        store_out_row<VPL>(p, out_row, v, lane);
        return 2;
    }

    // ---- miss: round-robin victim (M/isolated_cache.h:198-212: way = set_cnt++ % num_ways)
    const int way = (int)((s >> 1) & (COALA_WAYS - 1));
    const uint64_t slot = slot0 + way;
    const uint64_t told = __shfl_sync(0xffffffffu, t, way);
    if ((told >> TAG_VER_SHIFT) & 1) {  // victim is still being filled by a writer 32 fills ago
        if (lane == 0) {
            if (fenced) st_release_u64<SYS>(sh.set_seq + set, s);
            else st_relaxed_u64<SYS>(sh.set_seq + set, s);
        }
        store_out_row<VPL>(p, out_row, v, lane);
        return 2;
    }
    const uint64_t ver = (told >> TAG_VER_SHIFT) + 1;  // odd: line busy
    uint32_t dep = 0;
    if (lane == 0) {
        const uint64_t busy = (ver << TAG_VER_SHIFT) | id;
        st_relaxed_u64<SYS>(sh.tags + slot, busy);
        if (p.track_color) {  // M/isolated_cache.h:432-446
            // strong accesses: the slot's previous filler may have been another SM or GPU (no acquire invalidates L1 here)
            uint32_t oldc = ld_relaxed_u32<SYS>(sh.color_meta + slot);
            uint32_t newc = (uint32_t)p.color[id];
            st_relaxed_u32<SYS>(sh.color_meta + slot, newc);
            red_add_s32<SYS>(sh.color_counters + oldc, -1);
            red_add_s32<SYS>(sh.color_counters + newc, 1);
        }
        if (fenced) {
            // ONE fence covers both orders of (b): the busy tag (and colour word) precede the unlock - fence + relaxed
            // store is a release pattern - and, through the __syncwarp below, every lane's line stores
            fence_acq_rel<SYS>();
            st_relaxed_u64<SYS>(sh.set_seq + set, s + 2);  // unlock; the line stays busy via its tag
        } else {
            // (b): the busy tag is read back from the owner's L2; the unlock and (through `dep`) every line store depend on it
            // (the colour word of the slot is read back with it: the slot's next filler reads it under the lock)
            const uint64_t tb = ld_relaxed_u64<SYS>(sh.tags + slot);
            const uint32_t cb = p.track_color ? ld_relaxed_u32<SYS>(sh.color_meta + slot) : 0u;
            dep = ((uint32_t)tb | cb) & p.zero;
            st_relaxed_u64<SYS>(sh.set_seq + set, s + 2 + dep);
        }
    }
    dep = __shfl_sync(0xffffffffu, dep, 0);
    __syncwarp();

    // ---- fill: host row -> line and -> output from the same registers
    uint4* dst = sh.lines + slot * row_vecs + dep;
#pragma unroll
    for (int k = 0; k < VPL; k++) st_line_v4<SYS>(dst + k * 32 + lane, v[k]);
    store_out_row<VPL>(p, out_row, v, lane);
    if (fenced) {
        __syncwarp();
        if (lane == 0) {
            fence_acq_rel<SYS>();  // cumulative: the whole warp's line stores precede the publish
            st_relaxed_u64<SYS>(sh.tags + slot, ((ver + 1) << TAG_VER_SHIFT) | id);
        }
    } else {
        // (c): every lane reads its line words back; the publish depends on all of them through the ballot
        uint32_t acc = 0;
#pragma unroll
        for (int k = 0; k < VPL; k++) {
            const uint4 r = ld_line_v4<SYS>(dst + k * 32 + lane);
            acc |= r.x | r.y | r.z | r.w;
        }
        const unsigned all = __ballot_sync(0xffffffffu, (acc & p.zero) != 0);  // 0, known only when every lane's words are back
        if (lane == 0) st_relaxed_u64<SYS>(sh.tags + slot, (((ver + 1) << TAG_VER_SHIFT) | id) + all);
    }
    return 1;
}

// Out-of-line copy of resolve_row for the bulk engine, where it is the rare path (misses, busy lines, failed validations):
// keeps the register file of the bulk kernel's main loop for the pipeline state.
template <int VPL, bool SYS>
__device__ __noinline__ int resolve_row_call(const CacheParams& p, uint64_t id, int g, uint64_t set, uint64_t t, float* out_row, unsigned lane) {
    return resolve_row<VPL, SYS>(p, id, g, set, t, out_row, lane);
}

// An id outside the host table (or wider than the tag's 40-bit id field) must not be dereferenced: its output
// row is zero-filled and the launch counts it in stats[3]; the synchronous entry points turn that into an error.
__device__ __forceinline__ void refuse_row(const CacheParams& p, float* out_row, unsigned lane) {
    for (int e = (int)lane; e < p.out_dim; e += 32) out_row[e] = 0.0f;
}

// Work split: DYNAMIC.  One resident wave of CTAs; a warp's first chunk of GATHER_CH rows is its own index, every further
// one comes from a device-side counter until the list is exhausted (the row count comes from the sampler and never visits
// the host).  A row served from local HBM, a peer's HBM over NVLink or the host (+ the fill protocol) costs about
// 1 : 3 : 10 in warp time; a shared queue keeps every warp busy until the list is empty whatever its rows cost, and
// because the wave is resident from the start the sampler's CTAs (other streams) never displace part of a gather.
//
// Nothing in the steady state waits for the front end: the three dependent steps that open chunk k+1 - the counter
// atomic, the id loads (lane l < GATHER_CH owns row l and does its 64-bit owner/set arithmetic once), the GATHER_CH
// coalesced 256-B tag probes (possibly over NVLink) - are issued BETWEEN the rows of chunk k, one step per row, so each
// has a row's copy time to complete and a warp never holds more than one unprocessed chunk (what it claims ahead is what
// it cannot give back when the list runs dry).  A tag probed one chunk early can be stale by the time it is used; it is
// only a hint (the seqlock re-check / the probe under the set lock decide).
constexpr int GATHER_CH = 3;
constexpr int GS_SHIFT = 28;                       // owner index above the set index in one 32-bit word
constexpr uint32_t GS_SET_MASK = (1u << GS_SHIFT) - 1;
// Register cap of the register-path gather kernel: 2 resident CTAs x 256 threads x 96 = 48 K of the SM's 64 K registers,
// which leaves a sampling CTA of another stream (256 threads x 54) room to be co-resident.
#define GATHER_MAX_REGS 96

struct GatherWork {       // one per in-flight launch (a small rotating pool per cache object)
    unsigned next;        // next dynamic chunk block to hand out (chunk index - number of warps in the grid)
    unsigned done;        // CTAs that have finished (the last one re-arms both words for a later launch)
};

template <bool SYS>
struct RowFeeder {
    const CacheParams& p;
    const int64_t* idx;
    int64_t n;
    GatherWork* wk;
    unsigned lane, n_chunks, per_grab, first_dyn;
    unsigned blk_next = 0, blk_left = 0, raw = 0;
    bool dry = false, claimed = false;
    unsigned cA, cB;                     // chunk being copied, chunk being opened
    uint64_t idA = 0, idB = 0;
    uint32_t gsA = 0, gsB = 0;           // owner << 28 | set  (num_sets < 2^28 is checked when the cache is created)
    uint64_t tA[GATHER_CH], tB[GATHER_CH];

    __device__ __forceinline__ RowFeeder(const CacheParams& p_, const int64_t* idx_, int64_t n_, GatherWork* wk_, unsigned lane_)
        : p(p_), idx(idx_), n(n_), wk(wk_), lane(lane_) {
        n_chunks = (unsigned)((n + GATHER_CH - 1) / GATHER_CH);
        // one counter atomic hands out ~16 KB of rows: 1 chunk of 4 KB rows, 8 chunks of 512-B rows
        per_grab = 4096u / p.row_bytes;
        if (per_grab < 1) per_grab = 1;
        first_dyn = gridDim.x * (blockDim.x >> 5);
    }
    __device__ __forceinline__ void load_ids(unsigned c, uint64_t& id, uint32_t& gs) const {
        id = 0; gs = (uint32_t)p.my_slot << GS_SHIFT;
        const int64_t row = (int64_t)c * GATHER_CH + lane;
        if (c < n_chunks && lane < GATHER_CH && row < n) {
            id = (uint64_t)__ldg(idx + row);
            uint32_t g = (uint32_t)p.my_slot;
            if (p.route_gpus > 1) g = (uint32_t)(id % (uint64_t)p.route_gpus);
            gs = (g << GS_SHIFT) | (uint32_t)((id / (uint64_t)p.hash_div) % p.num_sets);
        }
    }
    __device__ __forceinline__ void probe(unsigned c, uint32_t gs, uint64_t (&t)[GATHER_CH]) const {
#pragma unroll
        for (int i = 0; i < GATHER_CH; i++) {
            const uint32_t x = __shfl_sync(0xffffffffu, gs, i);
            t[i] = 0;
            if (c < n_chunks && (int64_t)c * GATHER_CH + i < n)
                t[i] = ld_relaxed_u64<SYS>(p.shard[x >> GS_SHIFT].tags + (uint64_t)(x & GS_SET_MASK) * COALA_WAYS + lane);
        }
    }
    __device__ __forceinline__ void start() {
        const unsigned w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
        cA = w < n_chunks ? w : n_chunks;
        cB = n_chunks;
        load_ids(cA, idA, gsA);
        probe(cA, gsA, tA);
    }
    __device__ __forceinline__ bool have() const { return cA < n_chunks; }
    // step 1 of opening the next chunk: the counter atomic (only when the block handed out last is used up)
    __device__ __forceinline__ void claim() {
        claimed = false;
        if (blk_left == 0 && !dry) {
            if (lane == 0) raw = atomicAdd(&wk->next, per_grab);
            claimed = true;
        }
    }
    // step 2: the chunk's index is known -> its id loads
    __device__ __forceinline__ void fetch_ids() {
        if (claimed) {
            const unsigned base = __shfl_sync(0xffffffffu, raw, 0);
            if (base >= n_chunks || first_dyn + base >= n_chunks) dry = true;
            else { blk_next = first_dyn + base; blk_left = per_grab; }
        }
        cB = n_chunks;
        if (blk_left) {
            blk_left--;
            const unsigned c = blk_next++;
            if (c < n_chunks) cB = c;
        }
        load_ids(cB, idB, gsB);
    }
    // step 3: its tag probes
    __device__ __forceinline__ void probe_next() { probe(cB, gsB, tB); }
    __device__ __forceinline__ void rotate() {
        cA = cB; idA = idB; gsA = gsB;
#pragma unroll
        for (int i = 0; i < GATHER_CH; i++) tA[i] = tB[i];
    }
    __device__ __forceinline__ int rows_in_A() const {
        const int64_t left = n - (int64_t)cA * GATHER_CH;
        return left < GATHER_CH ? (int)left : GATHER_CH;
    }
};
static_assert(GATHER_CH >= 3, "the three front-end steps of the next chunk are interleaved with the rows of the current one");

// the last CTA out re-arms the work counter for a later launch that gets the same pool entry
__device__ __forceinline__ void gather_work_done(GatherWork* wk) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&wk->done, 1u) == gridDim.x - 1) {
            wk->next = 0;
            wk->done = 0;
            __threadfence();
        }
    }
}

// A hit row whose deferred re-check failed goes back through the full protocol with a fresh probe (rare: out of line).
template <int VPL, bool SYS>
__device__ __noinline__ int redo_row(const CacheParams& p, uint64_t tw, float* out_row, unsigned lane) {
    const uint64_t id = tw & TAG_ID_MASK;
    const int g = p.route_gpus > 1 ? (int)(id % (uint64_t)p.route_gpus) : p.my_slot;
    const uint32_t set = (uint32_t)((id / (uint64_t)p.hash_div) % p.num_sets);
    const uint64_t t = ld_relaxed_u64<SYS>(p.shard[g].tags + (uint64_t)set * COALA_WAYS + lane);
    return resolve_row<VPL, SYS>(p, id, g, set, t, out_row, lane);
}

// Register-path gather.  A hit row costs ONE exposed memory latency: the seqlock re-check of row i (issued once row i's
// words have arrived - its address depends on them) is only looked at after the loads of the NEXT hit row are in flight.
// Until then row i sits in the output unvalidated; if the re-check then fails (a writer refilled the line under the
// copy: rare) the row is redone through resolve_row, which rewrites it.  What is carried across rows is five registers
// (the two tag observations and the row number); id, owner and set of a failed row are recomputed from its tag.
template <int VPL, bool SYS>
__global__ void __maxnreg__(GATHER_MAX_REGS) coala_gather_kernel(const __grid_constant__ CacheParams p, const int64_t* __restrict__ idx,
                                                                 float* __restrict__ out, int64_t n,
                                                                 const int64_t* __restrict__ n_dev, GatherWork* wk) {
    __shared__ int s_stat[3];
    if (threadIdx.x < 3) s_stat[threadIdx.x] = 0;
    __syncthreads();
    if (n_dev) n = min(n, *n_dev);
    const unsigned lane = lane_id();
    const uint32_t row_vecs = p.row_bytes >> 4;
    unsigned n_hit = 0;   // misses, bypasses and refused ids are rare: counted straight into shared / global memory
    uint64_t pend_t2 = 0, pend_tw = 0;   // re-check of the last hit row: observed tag, expected tag
    uint32_t pend_row = 0;
    bool pend = false;
    auto redo = [&](uint64_t tw, uint32_t row) {
        const int res = redo_row<VPL, SYS>(p, tw, out + (int64_t)row * (int64_t)p.out_stride, lane);
        n_hit += (res == 0) - 1;   // it had been counted as a hit
        if (res && lane == 0) atomicAdd(&s_stat[res], 1);
    };
    RowFeeder<SYS> f(p, idx, n, wk, lane);
    f.start();
    while (f.have()) {
        const int cnt = f.rows_in_A();
        const int64_t base = (int64_t)f.cA * GATHER_CH;
#pragma unroll
        for (int i = 0; i < GATHER_CH; i++) {
            if (i == 0) f.claim();
            if (i == 1) f.fetch_ids();
            if (i == 2) f.probe_next();
            if (i < cnt) {
                const uint64_t id = __shfl_sync(0xffffffffu, f.idA, i);
                const uint32_t gs = __shfl_sync(0xffffffffu, f.gsA, i);
                const int g = (int)(gs >> GS_SHIFT);
                const uint32_t set = gs & GS_SET_MASK;
                const uint64_t t = f.tA[i];
                float* out_row = out + (base + i) * (int64_t)p.out_stride;
                if (id >= p.id_limit) {
                    refuse_row(p, out_row, lane);
                    if (lane == 0) atomicAdd(p.stats + 3, 1ULL);
                    continue;
                }
                const unsigned m = __ballot_sync(0xffffffffu, (t & TAG_ID_MASK) == id);
                const int way = m ? __ffs(m) - 1 : 0;
                const uint64_t tw = __shfl_sync(0xffffffffu, t, way);
                if (m && ((tw >> TAG_VER_SHIFT) & 1) == 0) {
                    const ShardView& sh = p.shard[g];
                    const uint64_t slot = (uint64_t)set * COALA_WAYS + way;
                    const uint4* src = sh.lines + slot * row_vecs;
                    uint4 v[VPL];
#pragma unroll
                    for (int k = 0; k < VPL; k++) v[k] = ld_line_v4<SYS>(src + k * 32 + lane);
                    // the previous hit row's re-check was issued before these loads: it is back first
                    const bool bad = pend && !__all_sync(0xffffffffu, pend_t2 == pend_tw);
                    store_out_row<VPL>(p, out_row, v, lane);
                    uint32_t acc = 0;
#pragma unroll
                    for (int k = 0; k < VPL; k++) acc |= v[k].x | v[k].y | v[k].z | v[k].w;
                    const uint64_t t2 = ld_relaxed_u64<SYS>(sh.tags + slot + (acc & p.zero));
                    if (bad) redo(pend_tw, pend_row);
                    pend_t2 = t2; pend_tw = tw; pend_row = (uint32_t)(base + i); pend = true;
                    n_hit++;
                } else {
                    const int res = resolve_row<VPL, SYS>(p, id, g, set, t, out_row, lane);
                    n_hit += res == 0;
                    if (res && lane == 0) atomicAdd(&s_stat[res], 1);
                }
            }
        }
        f.rotate();
    }
    if (pend && !__all_sync(0xffffffffu, pend_t2 == pend_tw)) redo(pend_tw, pend_row);
    if (lane == 0 && n_hit) atomicAdd(&s_stat[0], (int)n_hit);
    __syncthreads();
    if (threadIdx.x < 3 && s_stat[threadIdx.x]) atomicAdd(p.stats + threadIdx.x, (unsigned long long)(long long)s_stat[threadIdx.x]);
    gather_work_done(wk);
}

// ---------------------------------------------------------------------------------------------
// Copy engine 1: hit rows move HBM/peer -> shared memory -> output with bulk asynchronous copies
// (cp.async.bulk, the 1-D TMA path; SASS: UBLKCP) instead of through registers.  One elected lane
// issues a whole row; NBUF (2 or 3) row buffers per warp keep NBUF row loads in flight while the oldest
// one is validated and stored (8-12 KB per warp at R = 4 KB, against 4 KB through registers).  The
// seqlock re-check is ordered behind the data by the mbarrier wait (acquire).  Misses, busy lines and
// failed validations take the register path of resolve_row.
// ---------------------------------------------------------------------------------------------

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

struct BulkPending {   // a hit row whose line is on its way into shared memory
    uint64_t id, tw;
    uint32_t set, row;   // row = index into this launch's id list
    uint32_t meta;       // way | g << 5 | buf << 9  (g < COALA_MAX_GPUS = 16)
};

template <int VPL, bool SYS, int NBUF>
__global__ void __launch_bounds__(256, 2) coala_gather_bulk_kernel(const __grid_constant__ CacheParams p, const int64_t* __restrict__ idx,
                                                                   float* __restrict__ out, int64_t n,
                                                                   const int64_t* __restrict__ n_dev, GatherWork* wk) {
    static_assert(NBUF == 2 || NBUF == 3, "queue below is written for one or two waiting rows");
    extern __shared__ __align__(128) uint8_t dyn_smem[];
    __shared__ uint64_t bars[8][NBUF];
    __shared__ int s_stat[3];
    const unsigned lane = lane_id();
    const int wib = threadIdx.x >> 5;
    uint8_t* mybuf = dyn_smem + (size_t)wib * NBUF * p.row_bytes;
    if (threadIdx.x < 3) s_stat[threadIdx.x] = 0;
    if (lane == 0) {
        for (int b = 0; b < NBUF; b++) mbar_init(&bars[wib][b], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async;" ::: "memory");
    }
    __syncthreads();
    if (n_dev) n = min(n, *n_dev);
    const uint32_t out_bytes = (uint32_t)p.out_dim * 4;
    unsigned n_hit = 0, n_miss = 0, n_byp = 0, n_bad = 0;
    unsigned phase_bits = 0;   // bit b = parity the next wait on buffer b expects
    unsigned cur = 0;          // buffer of the next load
    BulkPending q0, q1;        // q0 = oldest waiting row, q1 = the one behind it (NBUF == 3)
    int waiting = 0;

    auto count = [&](int res) { n_hit += res == 0; n_miss += res == 1; n_byp += res == 2; };
    // finish the oldest waiting row: wait for its line, validate, store from shared memory or fall back
    auto complete_oldest = [&]() {
        const unsigned buf = q0.meta >> 9, way = q0.meta & 31u;
        const int g = (int)((q0.meta >> 5) & 15u);
        mbar_wait(&bars[wib][buf], (phase_bits >> buf) & 1u);
        phase_bits ^= 1u << buf;
        const ShardView& sh = p.shard[g];
        float* out_row = out + (int64_t)q0.row * (int64_t)p.out_stride;
        const uint64_t t2 = ld_relaxed_u64<SYS>(sh.tags + (uint64_t)q0.set * COALA_WAYS + way);
        if (__all_sync(0xffffffffu, t2 == q0.tw)) {  // one decision per warp: the else-branch runs warp-wide collectives
            if (lane == 0) bulk_store(out_row, mybuf + (size_t)buf * p.row_bytes, out_bytes);
            count(0);
        } else {  // the line changed under the copy: redo the row through the register path
            const uint64_t t = ld_relaxed_u64<SYS>(sh.tags + (uint64_t)q0.set * COALA_WAYS + lane);
            count(resolve_row_call<VPL, SYS>(p, q0.id, g, q0.set, t, out_row, lane));
        }
        q0 = q1;
        waiting--;
    };

    RowFeeder<SYS> f(p, idx, n, wk, lane);
    f.start();
    while (f.have()) {
        const int cnt = f.rows_in_A();
        const int64_t base = (int64_t)f.cA * GATHER_CH;
#pragma unroll
        for (int i = 0; i < GATHER_CH; i++) {
            if (i == 0) f.claim();
            if (i == 1) f.fetch_ids();
            if (i == 2) f.probe_next();
            if (i >= cnt) continue;
            const uint64_t id = __shfl_sync(0xffffffffu, f.idA, i);
            const uint32_t gs = __shfl_sync(0xffffffffu, f.gsA, i);
            const int g = (int)(gs >> GS_SHIFT);
            const uint32_t set = gs & GS_SET_MASK;
            const uint64_t t = f.tA[i];
            const unsigned m = __ballot_sync(0xffffffffu, (t & TAG_ID_MASK) == id);
            const unsigned way = m ? __ffs(m) - 1 : 0;
            const uint64_t tw = __shfl_sync(0xffffffffu, t, way);
            if (id >= p.id_limit) {
                refuse_row(p, out + (base + i) * (int64_t)p.out_stride, lane);
                n_bad++;
            } else if (m && ((tw >> TAG_VER_SHIFT) & 1) == 0) {
                // buffer `cur` belonged to a row completed at least one iteration ago
                if (lane == 0) {
                    bulk_wait_read_all();  // ... and the store that drained it has read it
                    mbar_expect_tx(&bars[wib][cur], p.row_bytes);
                    bulk_load(mybuf + (size_t)cur * p.row_bytes,
                              p.shard[g].lines + ((uint64_t)set * COALA_WAYS + way) * (p.row_bytes >> 4), p.row_bytes, &bars[wib][cur]);
                }
                if (waiting == NBUF - 1) complete_oldest();
                BulkPending c;
                c.id = id; c.tw = tw; c.set = set; c.row = (uint32_t)(base + i);
                c.meta = way | ((unsigned)g << 5) | (cur << 9);
                if (waiting == 0) q0 = c; else q1 = c;
                waiting++;
                cur = cur + 1 == NBUF ? 0 : cur + 1;
            } else {
                count(resolve_row_call<VPL, SYS>(p, id, g, set, t, out + (base + i) * (int64_t)p.out_stride, lane));
            }
        }
        f.rotate();
    }
    while (waiting) complete_oldest();
    if (lane == 0) {
        bulk_wait_all();
        if (n_hit) atomicAdd(&s_stat[0], (int)n_hit);
        if (n_miss) atomicAdd(&s_stat[1], (int)n_miss);
        if (n_byp) atomicAdd(&s_stat[2], (int)n_byp);
        if (n_bad) atomicAdd(p.stats + 3, (unsigned long long)n_bad);
    }
    __syncthreads();
    if (threadIdx.x < 3 && s_stat[threadIdx.x]) atomicAdd(p.stats + threadIdx.x, (unsigned long long)(long long)s_stat[threadIdx.x]);
    gather_work_done(wk);
}
