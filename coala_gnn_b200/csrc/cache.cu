// Host side of the feature cache: shard allocation, peer mapping, launches (stage 3).
// Replaces the host classes of M/ssd_gnn_cache.cuh and the handle ctors of
// M/isolated_cache.h:519-636 / M/nvshmem_cache.h:524-641.
#include "cache.cuh"
#include <atomic>
#include "distributor.h"
#include "vmm.h"
#include <unistd.h>
#include <vector>

namespace {

thread_local std::string g_last_error;

struct SlabHeader {  // first 256 bytes of every shard slab: lets a peer verify what it mapped
    uint64_t magic, num_sets, row_bytes, num_colors;
};
constexpr uint64_t SLAB_MAGIC = 0xC0A1A6B200CAC4E0ULL;

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct SlabLayout {
    size_t off_tags, off_seq, off_cmeta, off_ccnt, off_lines, total;
    SlabLayout(uint64_t S, uint32_t R, int C) {
        size_t o = 256;
        off_tags = o;  o = align_up(o + S * COALA_WAYS * 8, 256);
        off_seq = o;   o = align_up(o + S * 8, 256);
        off_cmeta = o; o = align_up(o + S * COALA_WAYS * 4, 256);
        off_ccnt = o;  o = align_up(o + (size_t)(C + 1) * 4, 4096);
        off_lines = o; o = o + S * COALA_WAYS * (size_t)R;
        total = o;
    }
    ShardView view(uint8_t* base) const {
        ShardView v;
        v.tags = reinterpret_cast<uint64_t*>(base + off_tags);
        v.set_seq = reinterpret_cast<uint64_t*>(base + off_seq);
        v.color_meta = reinterpret_cast<uint32_t*>(base + off_cmeta);
        v.color_counters = reinterpret_cast<int32_t*>(base + off_ccnt);
        v.lines = reinterpret_cast<uint4*>(base + off_lines);
        return v;
    }
};

}  // namespace

void coala_set_error(const std::string& msg) { g_last_error = msg; }
extern "C" const char* coala_last_error(void) { return g_last_error.c_str(); }
extern "C" int coala_abi_version(void) { return 1; }

constexpr int GATHER_WORK_POOL = 8;  // launches of one cache object that may be in flight at once (any streams)

struct coala_cache {
    int kind, device, feat_dim, cache_dim, global_rank, local_rank, n_gpus, num_colors;
    uint64_t num_sets;
    uint32_t row_bytes;
    uint8_t* slab = nullptr;
    SlabLayout layout{1, 512, 0};
    CacheParams params;
    unsigned long long* d_stats = nullptr;
    void* ipc_mapped[COALA_MAX_GPUS] = {nullptr};
    bool vmm = false;                    // slab allocated with cuMemCreate (shareable as a POSIX fd)
    VmmRegion region;                    // own slab when vmm
    VmmRegion peer_region[COALA_MAX_GPUS];
    int copy_engine = 0;   // 0 = registers, 1 = bulk asynchronous copies (coala_cache_set_copy_engine / COALA_GATHER_ENGINE)
    int bulk_nbuf = 3;     // row buffers per warp of engine 1 (COALA_GATHER_NBUF = 2|3)
    int ctas_per_sm = 2;   // resident CTAs per SM of the gather's single wave (see launch_gather)
    size_t smem_per_sm = 228 * 1024;
    int sm_count = 148;
    GatherWork* d_work = nullptr;  // pool of device-side work counters: an entry belongs to one launch until that launch has finished
    GatherWork* cur_work = nullptr;                 // entry of the launch being enqueued (valid under work_mu)
    cudaEvent_t work_ev[GATHER_WORK_POOL] = {};     // recorded behind the launch that last used entry k
    bool work_armed[GATHER_WORK_POOL] = {};         // work_ev[k] has been recorded at least once
    unsigned launch_seq = 0;                        // next entry to try (under work_mu)
    uint64_t color_len = 0;                         // entries of the distributor's colour table (0: none): upper bound of id_limit
    std::mutex work_mu;                             // launches of one cache object may come from several host threads
    std::vector<int32_t> h_counters;
};

extern "C" int coala_cache_dim_for(int dim) {
    if (dim <= 0) return -1;
    if (dim <= 128) return 128;
    if (dim <= 256) return 256;
    if (dim <= 512) return 512;
    if (dim <= 1024) return 1024;
    return -1;  // "Only Feature Embedding Size less than 8KB is supported" (M/ssd_gnn_cache.cuh:43-44)
}

static GatherWork* next_work(coala_cache* c) { return c->cur_work; }

// Choose the pool entry of the launch about to be enqueued on `st` (caller holds work_mu).  The kernel's last CTA re-arms
// the entry, so an entry may only be handed out again once the launch that used it has COMPLETED - not merely once eight
// later launches were enqueued: a launch still queued on one stream while another stream starts a later one on the same
// entry would share `next` / `done` with it (chunks skipped, output rows never written).  Each entry therefore carries an
// event recorded right behind its launch; an entry is reused when that event has completed, and when every entry is in
// flight the host waits for the oldest.  While `st` is being captured, events cannot be queried: captured launches rotate
// through the pool as before and must not overlap other launches of the same cache object.
static int acquire_work(coala_cache* c, cudaStream_t st, int* slot, bool* record) {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    COALA_CUDA(cudaStreamIsCapturing(st, &cs));
    if (cs != cudaStreamCaptureStatusNone) {
        *slot = (int)(c->launch_seq++ % GATHER_WORK_POOL);
        *record = false;
        return 0;
    }
    *record = true;
    for (int k = 0; k < GATHER_WORK_POOL; k++) {
        const int i = (int)((c->launch_seq + k) % GATHER_WORK_POOL);
        if (!c->work_armed[i]) { *slot = i; c->launch_seq = i + 1; return 0; }
        const cudaError_t q = cudaEventQuery(c->work_ev[i]);
        if (q == cudaSuccess) { *slot = i; c->launch_seq = i + 1; return 0; }
        if (q != cudaErrorNotReady) COALA_CUDA(q);
        (void)cudaGetLastError();  // "not ready" is an answer, not a failure: keep it out of the launch's error check
    }
    const int i = (int)(c->launch_seq % GATHER_WORK_POOL);  // the oldest
    COALA_CUDA(cudaEventSynchronize(c->work_ev[i]));
    *slot = i;
    c->launch_seq = i + 1;
    return 0;
}

// One resident wave: `ctas_per_sm` CTAs of 8 warps per SM pull chunks of rows from a device-side counter (cache.cuh) until
// the list - whose length may only be known on the device - is exhausted.
//
// SM budget: the gather is bandwidth bound and saturates its link with 16 warps per SM, while the next batch's sampling
// kernels (other streams, higher priority) are latency bound and need resident warps to keep host round trips in flight.
// The residency cap is enforced by reserving (unused) dynamic shared memory, so registers and warp slots stay free for the
// sampler, and because the wave is resident from the start the sampler's CTAs never displace part of a gather (with the
// earlier two-wave grid the second wave queued behind them).  COALA_GATHER_CTAS_PER_SM overrides (1..3).
template <int VPL, bool SYS>
static int launch_gather(coala_cache* c, const int64_t* idx, float* out, int64_t n, const int64_t* n_dev, cudaStream_t st,
                         bool serial) {
    if (n <= 0) return 0;
    if (serial) {
        coala_gather_kernel<VPL, SYS><<<1, 32, 0, st>>>(c->params, idx, out, n, n_dev, next_work(c));
    } else {
        const int threads = 256, wpb = threads / 32;
        const int resident = c->ctas_per_sm;
        // big enough that resident+1 CTAs cannot fit, small enough to leave >= 16 KB per SM for the
        // (1 KB each) driver reservations of the other streams' CTAs
        const size_t smem = (size_t)(c->smem_per_sm / resident) - 16384;
        // the attribute is per (device, instantiation): set it when the value changes, not on every launch
        static std::atomic<int> smem_set[64];
        const bool track = c->device >= 0 && c->device < 64;
        if (!track || smem_set[c->device].load(std::memory_order_relaxed) != (int)smem) {
            COALA_CUDA(cudaFuncSetAttribute(coala_gather_kernel<VPL, SYS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            if (track) smem_set[c->device].store((int)smem, std::memory_order_relaxed);
        }
        int64_t blocks = (int64_t)c->sm_count * resident;
        const int64_t need = (n + (int64_t)wpb * GATHER_CH - 1) / ((int64_t)wpb * GATHER_CH);
        if (blocks > need) blocks = need;
        coala_gather_kernel<VPL, SYS><<<(unsigned)blocks, threads, smem, st>>>(c->params, idx, out, n, n_dev, next_work(c));
    }
    COALA_CUDA(cudaGetLastError());
    return 0;
}

// Copy engine 1 (bulk asynchronous copies through shared memory, see cache.cuh).  NBUF row buffers per warp;
// the dynamic shared memory doubles as the residency cap (2 CTAs per SM).
template <int VPL, bool SYS, int NBUF>
static int launch_gather_bulk(coala_cache* c, const int64_t* idx, float* out, int64_t n, const int64_t* n_dev, cudaStream_t st) {
    if (n <= 0) return 0;
    const int threads = 256, wpb = threads / 32, resident = 2;
    const size_t need_smem = (size_t)wpb * NBUF * c->row_bytes;
    size_t smem = (size_t)(c->smem_per_sm / resident) - 16384;
    if (smem < need_smem) smem = need_smem;
    static std::atomic<int> smem_set[64];
    const bool track = c->device >= 0 && c->device < 64;
    if (!track || smem_set[c->device].load(std::memory_order_relaxed) != (int)smem) {
        COALA_CUDA(cudaFuncSetAttribute(coala_gather_bulk_kernel<VPL, SYS, NBUF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (track) smem_set[c->device].store((int)smem, std::memory_order_relaxed);
    }
    int64_t blocks = (int64_t)c->sm_count * resident;
    const int64_t need = (n + (int64_t)wpb * GATHER_CH - 1) / ((int64_t)wpb * GATHER_CH);
    if (blocks > need) blocks = need;
    coala_gather_bulk_kernel<VPL, SYS, NBUF><<<(unsigned)blocks, threads, smem, st>>>(c->params, idx, out, n, n_dev, next_work(c));
    COALA_CUDA(cudaGetLastError());
    return 0;
}

static int dispatch_gather_locked(coala_cache* c, const int64_t* idx, float* out, int64_t n, const int64_t* n_dev, cudaStream_t st,
                                  bool serial);

static int dispatch_gather(coala_cache* c, const int64_t* idx, float* out, int64_t n, const int64_t* n_dev, cudaStream_t st,
                           bool serial) {
    if (n <= 0) return 0;
    std::lock_guard<std::mutex> lk(c->work_mu);
    int slot = 0;
    bool record = false;
    if (acquire_work(c, st, &slot, &record) != 0) return -1;
    c->cur_work = c->d_work + slot;
    if (dispatch_gather_locked(c, idx, out, n, n_dev, st, serial) != 0) return -1;
    if (record) {
        COALA_CUDA(cudaEventRecord(c->work_ev[slot], st));
        c->work_armed[slot] = true;
    }
    return 0;
}

static int dispatch_gather_locked(coala_cache* c, const int64_t* idx, float* out, int64_t n, const int64_t* n_dev, cudaStream_t st,
                                  bool serial) {
    const bool sys = c->kind == COALA_CACHE_PARTITIONED && c->n_gpus > 1;
    const int vpl = c->row_bytes / 512;
    // bulk copies need 16-byte aligned output rows of a 16-byte multiple
    const bool bulk_ok = c->copy_engine == 1 && !serial && (c->params.out_stride * 4) % 16 == 0 && (c->params.out_dim * 4) % 16 == 0 &&
                         ((uintptr_t)out & 15) == 0;
    if (bulk_ok) {
#define COALA_GOB(V)                                                                                                      \
    if (c->bulk_nbuf == 3)                                                                                                \
        return sys ? launch_gather_bulk<V, true, 3>(c, idx, out, n, n_dev, st) : launch_gather_bulk<V, false, 3>(c, idx, out, n, n_dev, st); \
    return sys ? launch_gather_bulk<V, true, 2>(c, idx, out, n, n_dev, st) : launch_gather_bulk<V, false, 2>(c, idx, out, n, n_dev, st)
        switch (vpl) {
            case 1: COALA_GOB(1);
            case 2: COALA_GOB(2);
            case 4: COALA_GOB(4);
            case 8: COALA_GOB(8);
        }
#undef COALA_GOB
    }
#define COALA_GO(V) \
    return sys ? launch_gather<V, true>(c, idx, out, n, n_dev, st, serial) : launch_gather<V, false>(c, idx, out, n, n_dev, st, serial)
    switch (vpl) {
        case 1: COALA_GO(1);
        case 2: COALA_GO(2);
        case 4: COALA_GO(4);
        case 8: COALA_GO(8);
    }
#undef COALA_GO
    COALA_FAIL("unsupported row size");
}

extern "C" coala_cache* coala_cache_create(int kind, int device, int feat_dim, coala_nd* nd, int global_rank, int local_rank,
                                           int n_gpus, uint64_t cache_mb, uint64_t sim_buf_ptr) {
    if (kind < 0 || kind > 2) { coala_set_error("coala_cache_create: bad kind"); return nullptr; }
    int cache_dim = coala_cache_dim_for(feat_dim);
    if (cache_dim < 0) { coala_set_error("Only Feature Embedding Size less than 8KB is supported"); return nullptr; }
    if (sim_buf_ptr == 0) { coala_set_error("coala_cache_create: sim_buf_ptr == 0 (BaM/NVMe tier is out of scope; pass the pinned host feature table)"); return nullptr; }
    if (n_gpus < 1 || n_gpus > COALA_MAX_GPUS) { coala_set_error("coala_cache_create: n_gpus out of range"); return nullptr; }
    COALA_CUDA_NULL(cudaSetDevice(device));
    coala_cache* c = new coala_cache();
    c->kind = kind; c->device = device; c->feat_dim = feat_dim; c->cache_dim = cache_dim;
    c->global_rank = global_rank; c->local_rank = local_rank; c->n_gpus = n_gpus;
    c->row_bytes = (uint32_t)cache_dim * 4;                                  // page_size (M/ssd_gnn_cache.cuh:47)
    uint64_t num_pages = cache_mb * 1024ULL * 1024ULL / c->row_bytes;        // (:96)
    c->num_sets = num_pages / COALA_WAYS;                                    // (:97)
    if (c->num_sets == 0) { coala_set_error("coala_cache_create: cache smaller than one set"); delete c; return nullptr; }
    if (c->num_sets > GS_SET_MASK) { coala_set_error("coala_cache_create: more than 2^28 sets per shard"); delete c; return nullptr; }
    c->num_colors = nd ? coala_nd_num_colors(nd) : 0;
    c->layout = SlabLayout(c->num_sets, c->row_bytes, c->num_colors);
    // a shard that peers will map is allocated through the VMM API (2 MB-granular mappings on every
    // importer; see vmm.h for the measurement behind this); COALA_CACHE_LEGACY_IPC=1 keeps cudaMalloc + CUDA IPC
    const char* legacy = getenv("COALA_CACHE_LEGACY_IPC");
    c->vmm = kind == COALA_CACHE_PARTITIONED && n_gpus > 1 && !(legacy && atoi(legacy));
    if (c->vmm) {
        std::string err = vmm_alloc(&c->region, c->layout.total, device);
        if (!err.empty()) { coala_set_error("cache slab: " + err); delete c; return nullptr; }
        c->slab = reinterpret_cast<uint8_t*>(c->region.va);
    } else {
        cudaError_t e = cudaMalloc(&c->slab, c->layout.total);
        if (e != cudaSuccess) { coala_set_error(std::string("cudaMalloc(cache slab) failed: ") + cudaGetErrorString(e)); delete c; return nullptr; }
    }
    ShardView own = c->layout.view(c->slab);
    // tags: empty = id field all ones, version 0 (reference keys_ are memset 0xFF, M/isolated_cache.h:546)
    {
        std::vector<uint64_t> init(1 << 16, TAG_ID_MASK);
        size_t ntags = c->num_sets * COALA_WAYS, done = 0;
        // fill by doubling copies on the device
        size_t first = std::min(ntags, init.size());
        COALA_CUDA_NULL(cudaMemcpy(own.tags, init.data(), first * 8, cudaMemcpyHostToDevice));
        done = first;
        while (done < ntags) {
            size_t chunk = std::min(done, ntags - done);
            COALA_CUDA_NULL(cudaMemcpy(own.tags + done, own.tags, chunk * 8, cudaMemcpyDeviceToDevice));
            done += chunk;
        }
    }
    COALA_CUDA_NULL(cudaMemset(own.set_seq, 0, c->num_sets * 8));
    COALA_CUDA_NULL(cudaMemset(own.color_meta, 0, c->num_sets * COALA_WAYS * 4));
    COALA_CUDA_NULL(cudaMemset(own.color_counters, 0, (size_t)(c->num_colors + 1) * 4));
    SlabHeader hdr{SLAB_MAGIC, c->num_sets, c->row_bytes, (uint64_t)c->num_colors};
    COALA_CUDA_NULL(cudaMemcpy(c->slab, &hdr, sizeof(hdr), cudaMemcpyHostToDevice));
    COALA_CUDA_NULL(cudaMalloc(&c->d_stats, 4 * sizeof(unsigned long long)));
    COALA_CUDA_NULL(cudaMemset(c->d_stats, 0, 4 * sizeof(unsigned long long)));
    COALA_CUDA_NULL(cudaMalloc(&c->d_work, GATHER_WORK_POOL * sizeof(GatherWork)));
    COALA_CUDA_NULL(cudaMemset(c->d_work, 0, GATHER_WORK_POOL * sizeof(GatherWork)));
    for (int k = 0; k < GATHER_WORK_POOL; k++) COALA_CUDA_NULL(cudaEventCreateWithFlags(&c->work_ev[k], cudaEventDisableTiming));

    CacheParams& p = c->params;
    memset(&p, 0, sizeof(p));
    p.my_slot = kind == COALA_CACHE_PARTITIONED ? local_rank : 0;
    p.shard[p.my_slot] = own;
    p.route_gpus = kind == COALA_CACHE_PARTITIONED ? n_gpus : 1;
    p.hash_div = kind == COALA_CACHE_ISOLATED ? 1 : n_gpus;
    p.row_bytes = c->row_bytes;
    p.num_sets = c->num_sets;
    p.host_feat = reinterpret_cast<const uint8_t*>(sim_buf_ptr);
    p.color = nd ? coala_nd_color_dev_i32(nd) : nullptr;
    p.track_color = (nd && p.color) ? 1 : 0;                                  // track_color = true (M/ssd_gnn_cache.cuh:67)
    p.out_stride = feat_dim;
    p.out_dim = feat_dim < cache_dim ? feat_dim : cache_dim;
    p.stats = c->d_stats;
    p.fenced = 1;  // PTX-model-exact writer protocol; COALA_CACHE_FENCED=0 opts into the read-back-ordered one (cache.cuh)
    if (const char* e = getenv("COALA_CACHE_FENCED")) p.fenced = atoi(e) != 0;
    p.zero = 0;
    // ids are bounds-checked on the device: against the tag's id field always, against the table's row count as
    // soon as it is known (from the distributor's colour table, or coala_cache_set_num_nodes)
    p.id_limit = TAG_ID_MASK;
    if (nd && coala_nd_num_graph_nodes(nd) > 0 && (uint64_t)coala_nd_num_graph_nodes(nd) < p.id_limit)
        p.id_limit = c->color_len = (uint64_t)coala_nd_num_graph_nodes(nd);
    if (const char* e = getenv("COALA_GATHER_CTAS_PER_SM")) c->ctas_per_sm = atoi(e) >= 1 && atoi(e) <= 3 ? atoi(e) : 2;
    if (const char* e = getenv("COALA_GATHER_ENGINE")) c->copy_engine = atoi(e) == 1 ? 1 : 0;
    if (const char* e = getenv("COALA_GATHER_NBUF")) c->bulk_nbuf = atoi(e) == 2 ? 2 : 3;

    cudaDeviceProp prop;
    COALA_CUDA_NULL(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;
    c->smem_per_sm = prop.sharedMemPerMultiprocessor;
    COALA_CUDA_NULL(cudaDeviceSynchronize());
    return c;
}

extern "C" void coala_cache_destroy(coala_cache* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    for (int g = 0; g < COALA_MAX_GPUS; g++)
        if (c->ipc_mapped[g]) cudaIpcCloseMemHandle(c->ipc_mapped[g]);
    for (int g = 0; g < COALA_MAX_GPUS; g++)
        if (c->peer_region[g].handle) vmm_free(&c->peer_region[g]);
    if (c->vmm) vmm_free(&c->region);
    else if (c->slab) cudaFree(c->slab);
    if (c->d_stats) cudaFree(c->d_stats);
    if (c->d_work) cudaFree(c->d_work);
    for (int k = 0; k < GATHER_WORK_POOL; k++)
        if (c->work_ev[k]) cudaEventDestroy(c->work_ev[k]);
    delete c;
}

extern "C" uint64_t coala_cache_num_sets(coala_cache* c) { return c->num_sets; }
extern "C" int coala_cache_row_bytes(coala_cache* c) { return (int)c->row_bytes; }
extern "C" uint64_t coala_cache_color_counters_dev_ptr(coala_cache* c) {
    return (uint64_t)c->params.shard[c->params.my_slot].color_counters;
}

extern "C" int coala_cache_ipc_handle(coala_cache* c, void* handle64) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    COALA_CUDA(cudaSetDevice(c->device));
    if (c->vmm) COALA_FAIL("ipc_handle: this shard is shared as a file descriptor (coala_cache_export_fd)");
    cudaIpcMemHandle_t h;
    COALA_CUDA(cudaIpcGetMemHandle(&h, c->slab));
    memcpy(handle64, &h, 64);
    return 0;
}

static int attach_base(coala_cache* c, int peer, uint8_t* base) {
    if (c->kind != COALA_CACHE_PARTITIONED) COALA_FAIL("attach: cache is not partitioned");
    if (peer < 0 || peer >= c->n_gpus) COALA_FAIL("attach: peer rank out of range");
    if (peer == c->params.my_slot) return 0;
    SlabHeader hdr;
    COALA_CUDA(cudaMemcpy(&hdr, base, sizeof(hdr), cudaMemcpyDefault));
    if (hdr.magic != SLAB_MAGIC || hdr.num_sets != c->num_sets || hdr.row_bytes != c->row_bytes ||
        hdr.num_colors != (uint64_t)c->num_colors)
        COALA_FAIL("attach: peer shard geometry differs from ours");
    c->params.shard[peer] = c->layout.view(base);
    return 0;
}

extern "C" int coala_cache_attach_ipc(coala_cache* c, int peer, const void* handle64) {
    COALA_CUDA(cudaSetDevice(c->device));
    if (peer == c->params.my_slot) return 0;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void* base = nullptr;
    COALA_CUDA(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
    c->ipc_mapped[peer] = base;
    return attach_base(c, peer, static_cast<uint8_t*>(base));
}

extern "C" int coala_cache_is_fd_shared(coala_cache* c) { return c->vmm ? 1 : 0; }

extern "C" int coala_cache_export_fd(coala_cache* c) {
    if (!c->vmm) { coala_set_error("export_fd: shard was not allocated with the VMM API"); return -1; }
    std::string err;
    int fd = vmm_export_fd(c->region, &err);
    if (fd < 0) coala_set_error(err);
    return fd;
}

extern "C" int coala_cache_attach_fd(coala_cache* c, int peer, int fd) {
    COALA_CUDA(cudaSetDevice(c->device));
    if (peer == c->params.my_slot) return 0;
    if (peer < 0 || peer >= c->n_gpus) COALA_FAIL("attach_fd: peer rank out of range");
    if (!c->vmm) COALA_FAIL("attach_fd: this cache uses CUDA IPC handles");
    std::string err = vmm_import(&c->peer_region[peer], fd, c->region.size, c->device);
    if (!err.empty()) COALA_FAIL("attach_fd: " + err);
    return attach_base(c, peer, reinterpret_cast<uint8_t*>(c->peer_region[peer].va));
}

extern "C" int coala_cache_attach_local(coala_cache* c, int peer, coala_cache* other) {
    COALA_CUDA(cudaSetDevice(c->device));
    if (other->vmm) {  // same process: one more device on the peer's existing mapping
        if (other->device != c->device) {
            std::string err = vmm_grant(other->region, c->device);
            if (!err.empty()) COALA_FAIL("attach_local: " + err);
        }
        return attach_base(c, peer, other->slab);
    }
    if (other->device != c->device) {
        int can = 0;
        COALA_CUDA(cudaDeviceCanAccessPeer(&can, c->device, other->device));
        if (!can) COALA_FAIL("attach_local: no P2P access between the two devices");
        cudaError_t e = cudaDeviceEnablePeerAccess(other->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) COALA_CUDA(e);
        cudaGetLastError();
    }
    return attach_base(c, peer, other->slab);
}

static int check_ready(coala_cache* c) {
    if (c->kind == COALA_CACHE_PARTITIONED)
        for (int g = 0; g < c->n_gpus; g++)
            if (!c->params.shard[g].tags) COALA_FAIL("partitioned cache: peer shard " + std::to_string(g) + " not attached");
    return 0;
}

extern "C" int coala_cache_read_feature_async(coala_cache* c, uint64_t out_ptr, uint64_t idx_ptr, int64_t n,
                                              uint64_t n_dev_ptr, void* stream) {
    if (check_ready(c)) return -1;
    COALA_CUDA(cudaSetDevice(c->device));
    return dispatch_gather(c, reinterpret_cast<const int64_t*>(idx_ptr), reinterpret_cast<float*>(out_ptr), n,
                           reinterpret_cast<const int64_t*>(n_dev_ptr), static_cast<cudaStream_t>(stream), false);
}

static int fail_on_refused_ids(coala_cache* c, const char* what) {
    unsigned long long bad = 0;
    COALA_CUDA(cudaMemcpy(&bad, c->d_stats + 3, sizeof(bad), cudaMemcpyDeviceToHost));
    if (bad == 0) return 0;
    COALA_CUDA(cudaMemset(c->d_stats + 3, 0, sizeof(bad)));
    COALA_FAIL(std::string(what) + ": " + std::to_string(bad) + " node id(s) outside [0, " + std::to_string(c->params.id_limit) +
               ") were refused (their output rows are zero)");
}

extern "C" int coala_cache_set_num_nodes(coala_cache* c, int64_t num_nodes) {
    if (num_nodes <= 0) COALA_FAIL("set_num_nodes: the host table has at least one row");
    uint64_t lim = (uint64_t)num_nodes < TAG_ID_MASK ? (uint64_t)num_nodes : TAG_ID_MASK;
    if (c->color_len && lim > c->color_len) lim = c->color_len;  // the fill path indexes the colour table with the id as well
    c->params.id_limit = lim;
    return 0;
}

extern "C" int coala_cache_refused_ids(coala_cache* c, uint64_t* count, int reset) {
    COALA_CUDA(cudaSetDevice(c->device));
    unsigned long long bad = 0;
    COALA_CUDA(cudaMemcpy(&bad, c->d_stats + 3, sizeof(bad), cudaMemcpyDeviceToHost));
    if (count) *count = bad;
    if (reset && bad) COALA_CUDA(cudaMemset(c->d_stats + 3, 0, sizeof(bad)));
    return 0;
}

extern "C" int coala_cache_read_feature(coala_cache* c, uint64_t out_ptr, uint64_t idx_ptr, int64_t n) {
    if (coala_cache_read_feature_async(c, out_ptr, idx_ptr, n, 0, nullptr)) return -1;
    COALA_CUDA(cudaDeviceSynchronize());  // synchronous like the reference (M/ssd_gnn_cache.cuh:266)
    return fail_on_refused_ids(c, "read_feature");
}

extern "C" int coala_cache_read_feature_serial(coala_cache* c, uint64_t out_ptr, uint64_t idx_ptr, int64_t n) {
    if (check_ready(c)) return -1;
    COALA_CUDA(cudaSetDevice(c->device));
    if (dispatch_gather(c, reinterpret_cast<const int64_t*>(idx_ptr), reinterpret_cast<float*>(out_ptr), n, nullptr, nullptr, true))
        return -1;
    COALA_CUDA(cudaDeviceSynchronize());
    return fail_on_refused_ids(c, "read_feature_serial");
}

extern "C" int coala_cache_get_cache_data(coala_cache* c, uint64_t dst_ptr) {
    // the reference copies num_color int32 out of zero-copy host memory (M/ssd_gnn_cache.cuh:176-186);
    // here the counters live in HBM next to the shard and are fetched with one small copy.
    COALA_CUDA(cudaSetDevice(c->device));
    if (c->num_colors == 0) return 0;
    COALA_CUDA(cudaMemcpy(reinterpret_cast<void*>(dst_ptr), c->params.shard[c->params.my_slot].color_counters,
                          (size_t)c->num_colors * 4, cudaMemcpyDefault));
    return 0;
}

extern "C" int coala_cache_get_cache_data_n(coala_cache* c, uint64_t dst_ptr, int count) {
    COALA_CUDA(cudaSetDevice(c->device));
    if (count < 0 || count > c->num_colors + 1) COALA_FAIL("get_cache_data_n: count out of range");
    if (count == 0) return 0;
    COALA_CUDA(cudaMemcpy(reinterpret_cast<void*>(dst_ptr), c->params.shard[c->params.my_slot].color_counters,
                          (size_t)count * 4, cudaMemcpyDefault));
    return 0;
}

extern "C" int coala_cache_get_cache_data_async(coala_cache* c, uint64_t dst_dev_ptr, int count, void* stream) {
    COALA_CUDA(cudaSetDevice(c->device));
    if (count < 0 || count > c->num_colors + 1) COALA_FAIL("get_cache_data_async: count out of range");
    if (count == 0) return 0;
    COALA_CUDA(cudaMemcpyAsync(reinterpret_cast<void*>(dst_dev_ptr), c->params.shard[c->params.my_slot].color_counters,
                               (size_t)count * 4, cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream)));
    return 0;
}

extern "C" int coala_cache_stats(coala_cache* c, uint64_t* hit, uint64_t* miss, uint64_t* bypass, int reset) {
    COALA_CUDA(cudaSetDevice(c->device));
    unsigned long long h[4];
    COALA_CUDA(cudaMemcpy(h, c->d_stats, sizeof(h), cudaMemcpyDeviceToHost));
    if (hit) *hit = h[0];
    if (miss) *miss = h[1];
    if (bypass) *bypass = h[2];
    if (reset) COALA_CUDA(cudaMemset(c->d_stats, 0, 3 * sizeof(unsigned long long)));  // [3] (refused ids) is read by its own call
    return 0;
}

extern "C" int coala_cache_print_stats(coala_cache* c) {
    uint64_t h, m, b;
    if (coala_cache_stats(c, &h, &m, &b, 1)) return -1;
    // same two lines as NVSHMEM_cache_d_t::print_stats (M/nvshmem_cache.h:133-142); bypasses are misses
    printf("Global Rank: %i Local Rank:%i hit count: %llu miss count: %llu\n", c->global_rank, c->local_rank,
           (unsigned long long)h, (unsigned long long)(m + b));
    printf("Global Rank: %i Local Rank:%i  GPU hit ratio: %f\n", c->global_rank, c->local_rank,
           (float)h / (float)(h + m + b));
    fflush(stdout);
    return 0;
}

extern "C" int coala_cache_export_state(coala_cache* c, uint64_t* keys, uint32_t* color_meta, uint32_t* set_cnt) {
    COALA_CUDA(cudaSetDevice(c->device));
    COALA_CUDA(cudaDeviceSynchronize());
    const ShardView& s = c->params.shard[c->params.my_slot];
    size_t slots = c->num_sets * COALA_WAYS;
    if (keys) {
        COALA_CUDA(cudaMemcpy(keys, s.tags, slots * 8, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < slots; i++) {
            uint64_t id = keys[i] & TAG_ID_MASK;
            keys[i] = id == TAG_ID_MASK ? ~0ULL : id;  // reference encoding of an empty way
        }
    }
    if (color_meta) COALA_CUDA(cudaMemcpy(color_meta, s.color_meta, slots * 4, cudaMemcpyDeviceToHost));
    if (set_cnt) {
        std::vector<uint64_t> seq(c->num_sets);
        COALA_CUDA(cudaMemcpy(seq.data(), s.set_seq, c->num_sets * 8, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < c->num_sets; i++) set_cnt[i] = (uint32_t)(seq[i] >> 1);
    }
    return 0;
}

extern "C" int coala_cache_set_copy_engine(coala_cache* c, int engine) {
    if (engine != 0 && engine != 1) COALA_FAIL("copy engine must be 0 (registers) or 1 (bulk asynchronous copies)");
    c->copy_engine = engine;
    return 0;
}
