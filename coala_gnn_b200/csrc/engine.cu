// Native batch engine: the per-mini-batch enqueue sequence of the data path as ONE C-ABI call.
//
// The reference's iterator (S/COALA_GNN_DataLoader.py:149-167) is serial host code per batch: distributor thread join,
// seeds .to(device), DGL sample (host syncs per layer), fetch_feature (device syncs).  Here a batch is two enqueued
// stages on two CUDA streams (pipeline.py), and at B200 speeds (0.07-0.2 ms per batch once the rows come out of HBM /
// NVLink) the Python that enqueued them — a dozen torch stream/event calls and ctypes hops, ~0.13 ms — had become the
// ceiling of the end-to-end number.  This file is that sequence in C++: per batch the host does one memcpy of the
// seeds into pinned memory and ~12 CUDA runtime calls.
//
//   slot stream k (high priority): [wait consumer, wait previous gather of the slot] seeds H2D -> sample (one CUDA graph)
//   gather stream                : [wait sample, wait consumer] cached gather of the sampler's input-node list
// depth+1 slots rotate; events order slot reuse exactly as pipeline.BatchEngine did.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "../../include/coala_b200.h"

namespace {
struct EngSlot {
    coala_sampler* sampler = nullptr;
    uint64_t out_ptr = 0;
    uint64_t idx_ptr = 0, n_dev_ptr = 0;  // the sampler's input-node list and its device-side count
    int64_t* seeds_dev = nullptr;
    int64_t* seeds_pin = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_sampled = nullptr, ev_gathered = nullptr;
    cudaEvent_t g0 = nullptr, g1 = nullptr, s0 = nullptr, s1 = nullptr;
    bool timed = false, used = false, begun = false;
    int64_t batch = -1;
};
}  // namespace

struct coala_engine {
    int device = 0, n_slots = 0, batch_size = 0;
    uint64_t seed = 0;
    int rank = 0;
    cudaStream_t s_gather = nullptr;
    cudaEvent_t ev_consumed = nullptr;
    bool consumed_recorded = false;
    coala_cache* cache = nullptr;
    coala_nccl* nccl = nullptr;
    int64_t max_sample = 0;
    bool record = false;
    double gather_ms = 0.0, sample_ms = 0.0;
    uint64_t n_gather = 0, n_sample = 0;
    std::vector<EngSlot> slots;
};

extern "C" uint64_t coala_batch_key(uint64_t seed, uint64_t batch_idx, int rank) {
    return mix64(seed ^ mix64(batch_idx * 0x100000001B3ULL + (uint64_t)(int64_t)rank));
}

extern "C" coala_engine* coala_engine_create(int device, int n_slots, int batch_size, uint64_t seed, int rank, int record_times) {
    if (n_slots < 2 || n_slots > 16 || batch_size < 1) { coala_set_error("engine: 2..16 slots, batch_size >= 1"); return nullptr; }
    COALA_CUDA_NULL(cudaSetDevice(device));
    coala_engine* e = new coala_engine();
    e->device = device; e->n_slots = n_slots; e->batch_size = batch_size; e->seed = seed; e->rank = rank; e->record = record_times != 0;
    int lo = 0, hi = 0;
    COALA_CUDA_NULL(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    // COALA_ENGINE_PRIORITY: which side gets the high stream priority - "sample" (default), "gather" or "none"
    const char* pr = getenv("COALA_ENGINE_PRIORITY");
    const int pr_sample = (pr && (!strcmp(pr, "gather") || !strcmp(pr, "none"))) ? lo : hi;
    const int pr_gather = (pr && !strcmp(pr, "gather")) ? hi : lo;
    COALA_CUDA_NULL(cudaStreamCreateWithPriority(&e->s_gather, cudaStreamNonBlocking, pr_gather));
    COALA_CUDA_NULL(cudaEventCreateWithFlags(&e->ev_consumed, cudaEventDisableTiming));
    e->slots.resize(n_slots);
    for (auto& s : e->slots) {
        // sampling kernels are small and latency bound: high priority so that they take SM slots as gather CTAs retire
        COALA_CUDA_NULL(cudaStreamCreateWithPriority(&s.stream, cudaStreamNonBlocking, pr_sample));
        COALA_CUDA_NULL(cudaEventCreateWithFlags(&s.ev_sampled, cudaEventDisableTiming));
        COALA_CUDA_NULL(cudaEventCreateWithFlags(&s.ev_gathered, cudaEventDisableTiming));
        if (e->record) {
            COALA_CUDA_NULL(cudaEventCreate(&s.g0)); COALA_CUDA_NULL(cudaEventCreate(&s.g1));
            COALA_CUDA_NULL(cudaEventCreate(&s.s0)); COALA_CUDA_NULL(cudaEventCreate(&s.s1));
        }
        COALA_CUDA_NULL(cudaMalloc(&s.seeds_dev, (size_t)batch_size * 8));
        COALA_CUDA_NULL(cudaMemset(s.seeds_dev, 0, (size_t)batch_size * 8));
        COALA_CUDA_NULL(cudaHostAlloc(&s.seeds_pin, (size_t)batch_size * 8, cudaHostAllocDefault));
    }
    return e;
}

extern "C" void coala_engine_destroy(coala_engine* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    cudaDeviceSynchronize();
    for (auto& s : e->slots) {
        if (s.stream) cudaStreamDestroy(s.stream);
        for (cudaEvent_t ev : {s.ev_sampled, s.ev_gathered, s.g0, s.g1, s.s0, s.s1})
            if (ev) cudaEventDestroy(ev);
        cudaFree(s.seeds_dev);
        cudaFreeHost(s.seeds_pin);
    }
    if (e->s_gather) cudaStreamDestroy(e->s_gather);
    if (e->ev_consumed) cudaEventDestroy(e->ev_consumed);
    delete e;
}

extern "C" int coala_engine_bind_slot(coala_engine* e, int k, coala_sampler* sampler, uint64_t out_ptr) {
    if (k < 0 || k >= e->n_slots || !sampler || !out_ptr) COALA_FAIL("engine: bad slot binding");
    EngSlot& s = e->slots[k];
    s.sampler = sampler; s.out_ptr = out_ptr;
    if (coala_sampler_layer_ptrs(sampler, 0, &s.idx_ptr, nullptr, nullptr, nullptr)) return -1;
    s.n_dev_ptr = coala_sampler_input_count_dev_ptr(sampler);
    return 0;
}

extern "C" int coala_engine_bind_gather(coala_engine* e, coala_cache* cache, coala_nccl* nccl, int64_t max_sample) {
    if (!cache && !nccl) COALA_FAIL("engine: a cache or an NCCL exchange is needed");
    e->cache = cache; e->nccl = nccl; e->max_sample = max_sample;
    return 0;
}

extern "C" int coala_engine_set_seed(coala_engine* e, uint64_t seed) { e->seed = seed; return 0; }
extern "C" uint64_t coala_engine_slot_stream(coala_engine* e, int k) { return (uint64_t)e->slots[k].stream; }
extern "C" uint64_t coala_engine_slot_seeds_dev(coala_engine* e, int k) { return (uint64_t)e->slots[k].seeds_dev; }
extern "C" uint64_t coala_engine_gather_stream(coala_engine* e) { return (uint64_t)e->s_gather; }

extern "C" int coala_engine_checkpoint(coala_engine* e, void* consumer_stream) {
    COALA_CUDA(cudaEventRecord(e->ev_consumed, static_cast<cudaStream_t>(consumer_stream)));
    e->consumed_recorded = true;
    return 0;
}

static int fold_times(coala_engine* e, EngSlot& s) {
    if (!s.timed) return 0;
    COALA_CUDA(cudaEventSynchronize(s.g1));
    float g = 0.f, t = 0.f;
    COALA_CUDA(cudaEventElapsedTime(&g, s.g0, s.g1));
    COALA_CUDA(cudaEventElapsedTime(&t, s.s0, s.s1));
    e->gather_ms += g; e->n_gather++;
    e->sample_ms += t; e->n_sample++;
    s.timed = false;
    return 0;
}

// Opens batch t on its slot: orders the slot's stream behind the consumer and behind the slot's previous gather (which
// read the sampler's id list).  Returns the slot index, or -1.  What the caller enqueues on the slot's stream between
// begin and sample_gather (the GPU seed distribution, a histogram refresh) runs ahead of the sample.
extern "C" int coala_engine_begin(coala_engine* e, int64_t t, void* consumer_stream) {
    if (cudaSetDevice(e->device) != cudaSuccess) { coala_set_error("engine: cudaSetDevice failed"); return -1; }
    const int k = (int)(t % e->n_slots);
    EngSlot& s = e->slots[k];
    if (!s.sampler) { coala_set_error("engine: slot not bound"); return -1; }
    if (consumer_stream != (void*)~0ULL) {  // ~0: the caller already called coala_engine_checkpoint for this step
        if (coala_engine_checkpoint(e, consumer_stream)) return -1;
    }
    if (e->record && fold_times(e, s)) return -1;
    if (e->consumed_recorded && cudaStreamWaitEvent(s.stream, e->ev_consumed, 0) != cudaSuccess) { coala_set_error("engine: wait failed"); return -1; }
    if (s.used && cudaStreamWaitEvent(s.stream, s.ev_gathered, 0) != cudaSuccess) { coala_set_error("engine: wait failed"); return -1; }
    if (e->record && cudaEventRecord(s.s0, s.stream) != cudaSuccess) { coala_set_error("engine: record failed"); return -1; }
    s.begun = true;
    s.batch = -1;
    return k;
}

// Sample + gather of batch t from a device-resident seed list (valid on the slot's stream).
extern "C" int coala_engine_sample_gather(coala_engine* e, int64_t t, uint64_t seeds_dev_ptr, int n_seeds) {
    const int k = (int)(t % e->n_slots);
    EngSlot& s = e->slots[k];
    if (!s.begun) COALA_FAIL("engine: sample_gather without begin");
    s.begun = false;
    if (coala_sampler_sample_async(s.sampler, seeds_dev_ptr, n_seeds, coala_batch_key(e->seed, (uint64_t)t, e->rank), s.stream)) return -1;
    if (e->record) COALA_CUDA(cudaEventRecord(s.s1, s.stream));
    COALA_CUDA(cudaEventRecord(s.ev_sampled, s.stream));
    COALA_CUDA(cudaStreamWaitEvent(e->s_gather, s.ev_sampled, 0));
    if (e->consumed_recorded) COALA_CUDA(cudaStreamWaitEvent(e->s_gather, e->ev_consumed, 0));  // rows of the batch that used this buffer before are no longer needed
    if (e->record) COALA_CUDA(cudaEventRecord(s.g0, e->s_gather));
    int rc;
    if (e->nccl) rc = coala_nccl_fetch_async(e->nccl, s.out_ptr, s.idx_ptr, e->max_sample, s.n_dev_ptr, e->s_gather);
    else rc = coala_cache_read_feature_async(e->cache, s.out_ptr, s.idx_ptr, e->max_sample, s.n_dev_ptr, e->s_gather);
    if (rc) return rc;
    if (e->record) { COALA_CUDA(cudaEventRecord(s.g1, e->s_gather)); s.timed = true; }
    COALA_CUDA(cudaEventRecord(s.ev_gathered, e->s_gather));
    s.used = true;
    s.batch = t;
    return 0;
}

// Whole launch of batch t from a HOST seed list: pinned staging copy, H2D on the slot's stream, sample, gather.
extern "C" int coala_engine_launch_host(coala_engine* e, int64_t t, uint64_t seeds_host_ptr, int n_seeds, void* consumer_stream) {
    if (n_seeds < 0 || n_seeds > e->batch_size) COALA_FAIL("engine: more seeds than batch_size");
    const int k = coala_engine_begin(e, t, consumer_stream);
    if (k < 0) return -1;
    EngSlot& s = e->slots[k];
    memcpy(s.seeds_pin, reinterpret_cast<const void*>(seeds_host_ptr), (size_t)n_seeds * 8);
    COALA_CUDA(cudaMemcpyAsync(s.seeds_dev, s.seeds_pin, (size_t)n_seeds * 8, cudaMemcpyHostToDevice, s.stream));
    return coala_engine_sample_gather(e, t, (uint64_t)s.seeds_dev, n_seeds);
}

// The consumer's stream waits for batch t's blocks and rows; the host blocks only for the block sizes
// (counts[2L+1], coala_sampler_counts order).  counts may be NULL: device-side wait only.
extern "C" int coala_engine_collect(coala_engine* e, int64_t t, int64_t* counts, void* consumer_stream) {
    const int k = (int)(t % e->n_slots);
    EngSlot& s = e->slots[k];
    if (s.batch != t) COALA_FAIL("engine: batch is not in flight");
    cudaStream_t cs = static_cast<cudaStream_t>(consumer_stream);
    if (counts) {
        COALA_CUDA(cudaStreamWaitEvent(cs, s.ev_sampled, 0));
        if (coala_sampler_counts(s.sampler, counts)) return -1;
    }
    COALA_CUDA(cudaStreamWaitEvent(cs, s.ev_gathered, 0));
    return 0;
}

extern "C" int coala_engine_slot_of(coala_engine* e, int64_t t) {
    const int k = (int)(t % e->n_slots);
    return e->slots[k].batch == t ? k : -1;
}

// Mean kernel-stream times (CUDA events around the gather launch / around distribution + sampling) of the launches
// folded so far; sync != 0 first waits for and folds every launch still in flight.
extern "C" int coala_engine_times(coala_engine* e, double* gather_ms_mean, uint64_t* n_gather, double* sample_ms_mean, uint64_t* n_sample,
                                  int sync, int reset) {
    if (sync)
        for (auto& s : e->slots)
            if (fold_times(e, s)) return -1;
    if (gather_ms_mean) *gather_ms_mean = e->n_gather ? e->gather_ms / (double)e->n_gather : 0.0;
    if (sample_ms_mean) *sample_ms_mean = e->n_sample ? e->sample_ms / (double)e->n_sample : 0.0;
    if (n_gather) *n_gather = e->n_gather;
    if (n_sample) *n_sample = e->n_sample;
    if (reset) { e->gather_ms = e->sample_ms = 0.0; e->n_gather = e->n_sample = 0; }
    return 0;
}
