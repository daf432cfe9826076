// The NCCL all-to-all variant of the cached gather (--cache_backend nccl), kept for comparison with the P2P
// partitioned cache as the reference keeps it next to its NVSHMEM cache (examples/Cache_compare_script.sh:28-35).
//
// Reference (S/COALA_GNN_Manager.py:143-211, M/ssd_gnn_cache.cuh:283-356, M/cache_kernel.cu:79-137): bucket the ids
// by id % G with atomicAdd slots, dist.barrier, all_to_all of G buckets padded to max_sample, all_to_all_single of
// the counts (host-visible), allocate per-peer row tensors, owner-side lookups on G fresh streams, then G x G
// serialised dist.send / dist.recv pairs of rows, then one un-permute kernel per peer.
//
// Here the whole exchange of a mini-batch is enqueued on ONE stream by one call:
//   1 stable partition (count / scan / scatter; bucket g = [count | ids...])            partition.cu
//   2 ONE grouped ncclSend/ncclRecv round for the ids: fixed-size buckets whose header word carries the count, so
//     neither side needs to know a size on the host beforehand
//   3 the 2G counts land in pinned memory and the host waits for them - the only synchronisation: ncclSend/ncclRecv
//     take their element counts from the host, and row blocks are far too large to pad (max_sample x 4 KB per peer)
//   4 the received id lists are packed and looked up in this rank's shard by ONE launch of the gather kernel
//   5 ONE grouped ncclSend/ncclRecv round for the rows, exact counts
//   6 ONE un-permute kernel into the caller's output rows (bucket order -> request order).
// NCCL is resolved at run time (dlopen of the libnccl.so.2 torch ships and has already loaded): the library does not
// depend on it unless this backend is used, and then a missing NCCL is an error, never a fallback.
#include "common.cuh"
#include "../../include/coala_b200.h"
#include <dlfcn.h>
#include <algorithm>

int coala_partition_buckets(const int64_t* idx, int64_t n_upper, const int64_t* n_dev, int64_t* buckets, int64_t* map, int n_gpus,
                            int max_sample, int* block_cnt_scratch, cudaStream_t st);

namespace {

// the slice of nccl.h this file needs (the ABI of these entry points is stable across NCCL 2.x)
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
constexpr int NCCL_INT64 = 4, NCCL_FLOAT32 = 7;

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*CommAbort)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    int (*GetVersion)(int*) = nullptr;
};

NcclApi g_nccl;
std::mutex g_nccl_mu;

int load_nccl() {
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (g_nccl.lib) return 0;
    void* h = nullptr;
    if (const char* e = getenv("COALA_NCCL_LIB")) h = dlopen(e, RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);  // the copy torch has loaded
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) COALA_FAIL(std::string("nccl backend: libnccl.so.2 cannot be loaded (") + dlerror() + "); set COALA_NCCL_LIB");
#define COALA_SYM(field, name)                                                          \
    *reinterpret_cast<void**>(&g_nccl.field) = dlsym(h, name);                          \
    if (!g_nccl.field) { dlclose(h); COALA_FAIL(std::string("nccl backend: symbol ") + name + " missing"); }
    COALA_SYM(GetUniqueId, "ncclGetUniqueId")
    COALA_SYM(CommInitRank, "ncclCommInitRank")
    COALA_SYM(CommDestroy, "ncclCommDestroy")
    COALA_SYM(CommAbort, "ncclCommAbort")
    COALA_SYM(GroupStart, "ncclGroupStart")
    COALA_SYM(GroupEnd, "ncclGroupEnd")
    COALA_SYM(Send, "ncclSend")
    COALA_SYM(Recv, "ncclRecv")
    COALA_SYM(GetErrorString, "ncclGetErrorString")
    COALA_SYM(GetVersion, "ncclGetVersion")
#undef COALA_SYM
    g_nccl.lib = h;
    return 0;
}

#define COALA_NCCL(call)                                                                                        \
    do {                                                                                                        \
        ncclResult_t _r = (call);                                                                               \
        if (_r != 0) COALA_FAIL(std::string(#call) + " failed: " + g_nccl.GetErrorString(_r));                  \
    } while (0)

struct Offsets {
    int64_t v[COALA_MAX_GPUS + 1];
};

// [0..G) = my bucket sizes (what I ask of each owner), [G..2G) = what each peer asks of me
__global__ void nx_counts_kernel(const int64_t* __restrict__ send_b, const int64_t* __restrict__ recv_b, int G, int64_t stride,
                                 int64_t* __restrict__ counts) {
    const int g = threadIdx.x;
    if (g < G) {
        counts[g] = send_b[(int64_t)g * stride];
        counts[G + g] = recv_b[(int64_t)g * stride];
    }
}

// received buckets -> one contiguous id list (peer order), so that the owner-side lookup is a single launch
__global__ void nx_pack_ids_kernel(const int64_t* __restrict__ recv_b, int G, int64_t stride, Offsets off, int64_t* __restrict__ packed) {
    const int64_t total = off.v[G];
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int g = 0;
        while (t >= off.v[g + 1]) g++;
        packed[t] = recv_b[(int64_t)g * stride + 1 + (t - off.v[g])];
    }
}

// out[map[g][k]] = rows[off[g] + k]  (M/cache_kernel.cu:113-137), one warp per row, 16-byte vectors
__global__ void nx_unpermute_kernel(float* __restrict__ out, const float* __restrict__ rows, const int64_t* __restrict__ map, int G,
                                    int max_sample, Offsets off, int dim, int out_stride) {
    const int64_t total = off.v[G];
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int lane = threadIdx.x & 31;
    for (int64_t r = warp; r < total; r += nwarps) {
        int g = 0;
        while (r >= off.v[g + 1]) g++;
        const int64_t d = map[(int64_t)g * max_sample + (r - off.v[g])];
        const float* s = rows + r * dim;
        float* o = out + d * out_stride;
        if ((dim & 3) == 0 && (out_stride & 3) == 0) {
            for (int v = lane; v < dim / 4; v += 32) st_out_v4(reinterpret_cast<uint4*>(o) + v, *(reinterpret_cast<const uint4*>(s) + v));
        } else {
            for (int c = lane; c < dim; c += 32) o[c] = s[c];
        }
    }
}

}  // namespace

struct coala_nccl {
    coala_cache* shard = nullptr;
    ncclComm_t comm = nullptr;
    int device = 0, G = 1, rank = 0, dim = 0, max_sample = 0;
    int64_t stride = 0;              // 1 + max_sample: one bucket
    int64_t served_cap = 0;          // rows this rank can serve per batch
    int64_t *send_b = nullptr, *recv_b = nullptr, *map = nullptr, *packed = nullptr, *d_counts = nullptr, *h_counts = nullptr;
    int* block_cnt = nullptr;
    float *served = nullptr, *rows_in = nullptr;
    uint64_t batches = 0, rows_served = 0, rows_received = 0;
};

extern "C" int coala_nccl_unique_id(void* id128) {
    if (load_nccl()) return -1;
    ncclUniqueId id;
    COALA_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(id128, &id, 128);
    return 0;
}

extern "C" int coala_nccl_version(void) {
    if (load_nccl()) return -1;
    int v = 0;
    g_nccl.GetVersion(&v);
    return v;
}

extern "C" void coala_nccl_destroy(coala_nccl* x) {
    if (!x) return;
    cudaSetDevice(x->device);
    cudaDeviceSynchronize();
    if (x->comm) g_nccl.CommDestroy(x->comm);
    cudaFree(x->send_b); cudaFree(x->recv_b); cudaFree(x->map); cudaFree(x->packed); cudaFree(x->d_counts);
    cudaFree(x->block_cnt); cudaFree(x->served); cudaFree(x->rows_in);
    if (x->h_counts) cudaFreeHost(x->h_counts);
    delete x;
}

extern "C" coala_nccl* coala_nccl_create(coala_cache* shard, int device, const void* id128, int n_ranks, int rank, int64_t max_sample,
                                         int feat_dim) {
    if (!shard || n_ranks < 1 || n_ranks > COALA_MAX_GPUS || rank < 0 || rank >= n_ranks || max_sample < 1 || max_sample > 0x3fffffff) {
        coala_set_error("coala_nccl_create: bad arguments");
        return nullptr;
    }
    if (load_nccl()) return nullptr;
    COALA_CUDA_NULL(cudaSetDevice(device));
    coala_nccl* x = new coala_nccl();
    x->shard = shard; x->device = device; x->G = n_ranks; x->rank = rank; x->dim = feat_dim; x->max_sample = (int)max_sample;
    x->stride = 1 + max_sample;
    // id % G spreads a batch evenly over the owners: a rank serves about as many rows as it asks for (<= max_sample);
    // twice that is allocated up front, a batch that needs more grows the buffers (up to G x max_sample, the true bound)
    x->served_cap = 2 * max_sample;
    const size_t bucket_bytes = (size_t)n_ranks * x->stride * 8;
#define COALA_ALLOC(ptr, bytes)                                                                                 \
    do {                                                                                                        \
        cudaError_t _e = cudaMalloc(&(ptr), (bytes));                                                           \
        if (_e != cudaSuccess) {                                                                                \
            coala_set_error(std::string("coala_nccl_create: cudaMalloc failed: ") + cudaGetErrorString(_e));   \
            coala_nccl_destroy(x);                                                                              \
            return nullptr;                                                                                     \
        }                                                                                                       \
    } while (0)
    COALA_ALLOC(x->send_b, bucket_bytes);
    COALA_ALLOC(x->recv_b, bucket_bytes);
    COALA_ALLOC(x->map, (size_t)n_ranks * max_sample * 8);
    COALA_ALLOC(x->packed, (size_t)x->served_cap * 8);
    COALA_ALLOC(x->d_counts, 2 * COALA_MAX_GPUS * 8);
    COALA_ALLOC(x->block_cnt, ((size_t)(max_sample + 1023) / 1024 + 1) * n_ranks * sizeof(int));
    COALA_ALLOC(x->served, (size_t)x->served_cap * feat_dim * 4);
    COALA_ALLOC(x->rows_in, (size_t)max_sample * feat_dim * 4);
#undef COALA_ALLOC
    COALA_CUDA_NULL(cudaMemset(x->send_b, 0, bucket_bytes));
    COALA_CUDA_NULL(cudaMemset(x->recv_b, 0, bucket_bytes));
    COALA_CUDA_NULL(cudaHostAlloc(&x->h_counts, 2 * COALA_MAX_GPUS * 8, cudaHostAllocDefault));
    if (n_ranks > 1) {
        ncclUniqueId id;
        memcpy(&id, id128, 128);
        ncclResult_t r = g_nccl.CommInitRank(&x->comm, n_ranks, id, rank);  // collective over the cache domain
        if (r != 0) {
            coala_set_error(std::string("ncclCommInitRank failed: ") + g_nccl.GetErrorString(r));
            coala_nccl_destroy(x);
            return nullptr;
        }
    }
    return x;
}

extern "C" int coala_nccl_fetch_async(coala_nccl* x, uint64_t out_ptr, uint64_t idx_ptr, int64_t n_upper, uint64_t n_dev_ptr, void* stream) {
    COALA_CUDA(cudaSetDevice(x->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n_upper > x->max_sample) COALA_FAIL("nccl fetch: more ids than max_sample");
    const int G = x->G, me = x->rank;
    const int64_t stride = x->stride;
    // 1. stable partition by owner (id % G); bucket g = [count | ids in request order]
    if (coala_partition_buckets(reinterpret_cast<const int64_t*>(idx_ptr), n_upper, reinterpret_cast<const int64_t*>(n_dev_ptr), x->send_b, x->map, G,
                                x->max_sample, x->block_cnt, st))
        return -1;
    // 2. bucket sizes first: one grouped exchange of the G header words (8 bytes per peer).  The reference pads every id
    // bucket to max_sample words (S/COALA_GNN_Manager.py:161-165: 2.4 MB per rank per batch at G = 8); knowing the sizes
    // up front, step 4 ships exact-length id lists and the comparison with the P2P backend is about NCCL, not padding.
    if (G > 1) {
        COALA_NCCL(g_nccl.GroupStart());
        for (int g = 0; g < G; g++) {
            if (g == me) continue;
            COALA_NCCL(g_nccl.Send(x->send_b + g * stride, 1, NCCL_INT64, g, x->comm, st));
            COALA_NCCL(g_nccl.Recv(x->recv_b + g * stride, 1, NCCL_INT64, g, x->comm, st));
        }
        COALA_NCCL(g_nccl.GroupEnd());
    }
    COALA_CUDA(cudaMemcpyAsync(x->recv_b + me * stride, x->send_b + me * stride, (size_t)stride * 8, cudaMemcpyDeviceToDevice, st));
    // 3. the counts are the one thing the host has to know (element counts of both exchanges that follow)
    nx_counts_kernel<<<1, 32, 0, st>>>(x->send_b, x->recv_b, G, stride, x->d_counts);
    COALA_CUDA(cudaMemcpyAsync(x->h_counts, x->d_counts, 2 * G * 8, cudaMemcpyDeviceToHost, st));
    COALA_CUDA(cudaStreamSynchronize(st));
    Offsets ask, serve;
    ask.v[0] = serve.v[0] = 0;
    for (int g = 0; g < G; g++) {
        ask.v[g + 1] = ask.v[g] + x->h_counts[g];
        serve.v[g + 1] = serve.v[g] + x->h_counts[G + g];
    }
    // Between here and the row exchange a plain error return would leave every peer blocked for ever in its ncclRecv (they
    // know this rank's bucket sizes already and post the matching receives of ids and rows).  So nothing here fails locally: a batch that concentrates
    // on this owner - legal, the worst case is G x max_sample rows when the ids are skewed so that id % G favours it -
    // grows the serve buffers, and what cannot be recovered (out of memory, a failed cache launch) aborts the communicator,
    // which fails the peers' pending operations instead of hanging them.
    auto abort_domain = [&](const std::string& why) {
        if (x->comm) { g_nccl.CommAbort(x->comm); x->comm = nullptr; }
        coala_set_error("nccl fetch: " + why + "; the domain's communicator was aborted (every rank's exchange fails)");
        return -1;
    };
    if (serve.v[G] > x->served_cap) {
        int64_t cap = std::max<int64_t>(serve.v[G], 2 * x->served_cap);
        cap = std::min<int64_t>(cap, (int64_t)G * x->max_sample);
        float* ns = nullptr;
        int64_t* np = nullptr;
        // the stream was synchronised for the counts: nothing enqueued earlier still uses the old buffers
        cudaFree(x->served); cudaFree(x->packed);
        x->served = nullptr; x->packed = nullptr; x->served_cap = 0;
        if (cudaMalloc(&ns, (size_t)cap * x->dim * 4) != cudaSuccess || cudaMalloc(&np, (size_t)cap * 8) != cudaSuccess) {
            (void)cudaGetLastError();
            cudaFree(ns);
            return abort_domain("out of memory growing the serve buffers to " + std::to_string(cap) + " rows");
        }
        x->served = ns; x->packed = np; x->served_cap = cap;
    }
    // 4. ids: one grouped exchange, exact lengths (bucket g's ids follow its header word)
    if (G > 1) {
        COALA_NCCL(g_nccl.GroupStart());
        for (int g = 0; g < G; g++) {
            if (g == me) continue;
            const int64_t na = ask.v[g + 1] - ask.v[g], nsv = serve.v[g + 1] - serve.v[g];
            if (na > 0) COALA_NCCL(g_nccl.Send(x->send_b + g * stride + 1, (size_t)na, NCCL_INT64, g, x->comm, st));
            if (nsv > 0) COALA_NCCL(g_nccl.Recv(x->recv_b + g * stride + 1, (size_t)nsv, NCCL_INT64, g, x->comm, st));
        }
        COALA_NCCL(g_nccl.GroupEnd());
    }
    // 5. owner side: one list, one launch of the gather kernel against this rank's shard
    if (serve.v[G] > 0) {
        int64_t blocks = (serve.v[G] + 255) / 256;
        if (blocks > coala_sm_count() * 4) blocks = coala_sm_count() * 4;
        nx_pack_ids_kernel<<<(unsigned)blocks, 256, 0, st>>>(x->recv_b, G, stride, serve, x->packed);
        if (cudaError_t le = cudaGetLastError(); le != cudaSuccess) return abort_domain(std::string("pack launch failed: ") + cudaGetErrorString(le));
        if (coala_cache_read_feature_async(x->shard, (uint64_t)x->served, (uint64_t)x->packed, serve.v[G], 0, st))
            return abort_domain(std::string("owner-side cache read failed (") + coala_last_error() + ")");
    }
    // 6. rows: one grouped exchange, exact counts
    const size_t rowf = (size_t)x->dim;
    if (G > 1) {
        COALA_NCCL(g_nccl.GroupStart());
        for (int g = 0; g < G; g++) {
            if (g == me) continue;
            const int64_t ns = serve.v[g + 1] - serve.v[g], nr = ask.v[g + 1] - ask.v[g];
            if (ns > 0) COALA_NCCL(g_nccl.Send(x->served + serve.v[g] * rowf, (size_t)ns * rowf, NCCL_FLOAT32, g, x->comm, st));
            if (nr > 0) COALA_NCCL(g_nccl.Recv(x->rows_in + ask.v[g] * rowf, (size_t)nr * rowf, NCCL_FLOAT32, g, x->comm, st));
        }
        COALA_NCCL(g_nccl.GroupEnd());
    }
    const int64_t n_self = ask.v[me + 1] - ask.v[me];
    if (n_self > 0)
        COALA_CUDA(cudaMemcpyAsync(x->rows_in + ask.v[me] * rowf, x->served + serve.v[me] * rowf, (size_t)n_self * rowf * 4, cudaMemcpyDeviceToDevice, st));
    // 7. bucket order -> request order
    if (ask.v[G] > 0) {
        int64_t blocks = (ask.v[G] + 7) / 8;
        if (blocks > coala_sm_count() * 8) blocks = coala_sm_count() * 8;
        nx_unpermute_kernel<<<(unsigned)blocks, 256, 0, st>>>(reinterpret_cast<float*>(out_ptr), x->rows_in, x->map, G, x->max_sample, ask, x->dim,
                                                             x->dim);
        COALA_CUDA(cudaGetLastError());
    }
    x->batches++;
    x->rows_served += (uint64_t)serve.v[G];
    x->rows_received += (uint64_t)ask.v[G];
    return 0;
}

extern "C" int coala_nccl_counters(coala_nccl* x, uint64_t* batches, uint64_t* rows_served, uint64_t* rows_received) {
    if (batches) *batches = x->batches;
    if (rows_served) *rows_served = x->rows_served;
    if (rows_received) *rows_received = x->rows_received;
    return 0;
}
