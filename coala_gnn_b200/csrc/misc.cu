// Shared UVA host memory (stage-2/3 substrate), synthetic feature tables, link probes and the
// partition / un-permute kernels of the NCCL all-to-all comparison backend.
#include "common.cuh"
#include "../../include/coala_b200.h"
#include <cerrno>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

// ------------------------------------------------------------------------------------------
// SharedUVAManager (M/shared_UVA.cuh:26-115) without MPI: the caller orders create/open.
// ------------------------------------------------------------------------------------------
struct coala_uva {
    std::string path;
    void* host = nullptr;
    void* dev = nullptr;
    int64_t bytes = 0;
    int fd = -1;
    bool creator = false;
    bool registered = false;
};

extern "C" coala_uva* coala_uva_create(const char* shm_path, int64_t bytes, int create, int device, int zero) {
    if (bytes <= 0) { coala_set_error("coala_uva_create: size must be positive"); return nullptr; }
    COALA_CUDA_NULL(cudaSetDevice(device));
    coala_uva* h = new coala_uva();
    h->path = shm_path; h->bytes = bytes; h->creator = create != 0;
    h->fd = create ? shm_open(shm_path, O_CREAT | O_RDWR, S_IRUSR | S_IWUSR)     // (:67)
                   : shm_open(shm_path, O_RDWR, S_IRUSR | S_IWUSR);               // (:80)
    if (h->fd < 0) { coala_set_error(std::string("shm_open failed, Reason=") + strerror(errno)); delete h; return nullptr; }
    if (create && ftruncate(h->fd, bytes) != 0) {
        coala_set_error(std::string("ftruncate failed, Reason=") + strerror(errno));
        close(h->fd); shm_unlink(shm_path); delete h; return nullptr;
    }
    h->host = mmap(nullptr, (size_t)bytes, PROT_READ | PROT_WRITE, MAP_SHARED, h->fd, 0);   // (:83)
    if (h->host == MAP_FAILED) {
        coala_set_error(std::string("mmap failed, Reason= ") + strerror(errno));
        close(h->fd); if (create) shm_unlink(shm_path); delete h; return nullptr;
    }
    if (create && zero) memset(h->host, 0, (size_t)bytes);                                   // (:99)
    cudaError_t e = cudaHostRegister(h->host, (size_t)bytes, cudaHostRegisterDefault);       // (:89)
    if (e != cudaSuccess) {
        coala_set_error(std::string("cudaHostRegister error: ") + cudaGetErrorString(e));
        munmap(h->host, (size_t)bytes); close(h->fd); if (create) shm_unlink(shm_path); delete h; return nullptr;
    }
    h->registered = true;
    e = cudaHostGetDevicePointer(&h->dev, h->host, 0);                                        // (:94)
    if (e != cudaSuccess) {
        coala_set_error(std::string("cudaHostGetDevicePointer error: ") + cudaGetErrorString(e));
        coala_uva_cleanup(h, create); delete h; return nullptr;
    }
    return h;
}
extern "C" uint64_t coala_uva_host_ptr(coala_uva* h) { return (uint64_t)h->host; }
extern "C" uint64_t coala_uva_device_ptr(coala_uva* h) { return (uint64_t)h->dev; }
extern "C" int coala_uva_cleanup(coala_uva* h, int unlink_segment) {
    if (h->host) {
        if (h->registered) cudaHostUnregister(h->host);
        munmap(h->host, (size_t)h->bytes);
        h->host = nullptr; h->dev = nullptr; h->registered = false;
    }
    if (h->fd >= 0) { close(h->fd); h->fd = -1; }
    if (unlink_segment) shm_unlink(h->path.c_str());
    return 0;
}
extern "C" void coala_uva_destroy(coala_uva* h) {
    if (!h) return;
    coala_uva_cleanup(h, h->creator);
    delete h;
}

// ------------------------------------------------------------------------------------------
// synthetic feature table + verification (same integer hash as oracle/c/common.c)
// ------------------------------------------------------------------------------------------
__global__ void synth_fill_kernel(float* __restrict__ dst, uint64_t seed, int64_t row0, int64_t nrows, int dim, int64_t stride) {
    // one warp per row, 4 floats per lane per step
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int lane = threadIdx.x & 31;
    for (int64_t r = warp; r < nrows; r += nwarps) {
        const uint64_t hr = mix64(seed ^ (uint64_t)(row0 + r));
        float* p = dst + r * stride;
        if ((dim & 3) == 0 && (stride & 3) == 0) {
            for (int c = lane * 4; c < dim; c += 128) {
                float4 v = make_float4(feat_from_hash(mix64(hr + c)), feat_from_hash(mix64(hr + c + 1)),
                                       feat_from_hash(mix64(hr + c + 2)), feat_from_hash(mix64(hr + c + 3)));
                *reinterpret_cast<float4*>(p + c) = v;
            }
        } else {
            for (int c = lane; c < dim; c += 32) p[c] = feat_from_hash(mix64(hr + c));
        }
    }
}

extern "C" int coala_synth_fill_features(uint64_t dst_ptr, uint64_t seed, int64_t row0, int64_t nrows, int dim,
                                         int64_t stride_floats, void* stream) {
    if (nrows <= 0) return 0;
    int64_t blocks = (nrows + 7) / 8;
    if (blocks > coala_sm_count() * 16) blocks = coala_sm_count() * 16;
    synth_fill_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(reinterpret_cast<float*>(dst_ptr), seed, row0,
                                                                                   nrows, dim, stride_floats);
    COALA_CUDA(cudaGetLastError());
    if (!stream) COALA_CUDA(cudaDeviceSynchronize());
    return 0;
}

__global__ void synth_check_kernel(const float* __restrict__ rows, const int64_t* __restrict__ idx, int64_t n, int dim,
                                   int64_t stride, uint64_t seed, unsigned long long* __restrict__ bad) {
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int lane = threadIdx.x & 31;
    for (int64_t r = warp; r < n; r += nwarps) {
        const uint64_t hr = mix64(seed ^ (uint64_t)idx[r]);
        const float* p = rows + r * stride;
        bool ok = true;
        for (int c = lane; c < dim; c += 32)
            ok &= __float_as_uint(p[c]) == __float_as_uint(feat_from_hash(mix64(hr + c)));
        if (__ballot_sync(0xffffffffu, !ok) && lane == 0) atomicAdd(bad, 1ULL);
    }
}

extern "C" int coala_synth_check_rows(uint64_t rows_ptr, uint64_t idx_ptr, int64_t n, int dim, int64_t stride_floats,
                                      uint64_t seed, int64_t* mismatches) {
    unsigned long long* d_bad;
    COALA_CUDA(cudaMalloc(&d_bad, 8));
    COALA_CUDA(cudaMemset(d_bad, 0, 8));
    if (n > 0) {
        int64_t blocks = (n + 7) / 8;
        if (blocks > coala_sm_count() * 16) blocks = coala_sm_count() * 16;
        synth_check_kernel<<<(unsigned)blocks, 256>>>(reinterpret_cast<const float*>(rows_ptr), reinterpret_cast<const int64_t*>(idx_ptr),
                                                      n, dim, stride_floats, seed, d_bad);
        COALA_CUDA(cudaGetLastError());
    }
    unsigned long long bad = 0;
    COALA_CUDA(cudaMemcpy(&bad, d_bad, 8, cudaMemcpyDeviceToHost));
    cudaFree(d_bad);
    *mismatches = (int64_t)bad;
    return 0;
}

// ------------------------------------------------------------------------------------------
// streaming-read probe: measures what a kernel can pull through a link (pinned host over PCIe,
// a peer's HBM over NVLink, or local HBM) — the denominators of the per-link rooflines
// ------------------------------------------------------------------------------------------
__global__ void probe_read_kernel(const uint4* __restrict__ src, int64_t nvec, unsigned* __restrict__ sink) {
    unsigned acc = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < nvec; i += 4 * stride) {
        uint4 a = ld_host_v4(src + i), b = ld_host_v4(src + i + stride), c = ld_host_v4(src + i + 2 * stride),
              d = ld_host_v4(src + i + 3 * stride);
        acc ^= a.x ^ b.y ^ c.z ^ d.w;
    }
    for (; i < nvec; i += stride) acc ^= ld_host_v4(src + i).x;
    if (acc == 0x12345677u) *sink = acc;  // keep the loads alive
}

extern "C" int coala_probe_read_bandwidth(uint64_t src_ptr, int64_t bytes, int iters, double* gb_per_s) {
    unsigned* sink;
    COALA_CUDA(cudaMalloc(&sink, 4));
    cudaEvent_t e0, e1;
    COALA_CUDA(cudaEventCreate(&e0));
    COALA_CUDA(cudaEventCreate(&e1));
    const int64_t nvec = bytes / 16;
    probe_read_kernel<<<coala_sm_count() * 8, 256>>>(reinterpret_cast<const uint4*>(src_ptr), nvec, sink);  // warm-up
    COALA_CUDA(cudaEventRecord(e0));
    for (int i = 0; i < iters; i++) probe_read_kernel<<<coala_sm_count() * 8, 256>>>(reinterpret_cast<const uint4*>(src_ptr), nvec, sink);
    COALA_CUDA(cudaEventRecord(e1));
    COALA_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    COALA_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    *gb_per_s = (double)nvec * 16.0 * iters / (ms * 1e-3) / 1e9;
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
    return 0;
}

// classic peer access for plain cudaMalloc buffers of another GPU in this process (the cache shards do not need it:
// they are VMM mappings); used by the link probes of scripts/
extern "C" int coala_enable_peer_access(int device, int peer) {
    COALA_CUDA(cudaSetDevice(device));
    int can = 0;
    COALA_CUDA(cudaDeviceCanAccessPeer(&can, device, peer));
    if (!can) COALA_FAIL("peer access between these devices is not possible");
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return 0; }
    COALA_CUDA(e);
    return 0;
}

// ------------------------------------------------------------------------------------------
// row-gather probe: out[i] = table[idx[i]] for rows of `row_bytes`, NO cache protocol (no tags, no validation) - the
// practical ceiling of the gather's access pattern on a link (local HBM, a peer's HBM, pinned host), against which the
// cache kernel's protocol overhead is read.  One warp per row, RIF rows in flight per warp, 16-byte vectors.
// ------------------------------------------------------------------------------------------
template <int VPL, int RIF>
__global__ void probe_gather_kernel(const uint4* __restrict__ table, const int64_t* __restrict__ idx, uint4* __restrict__ out, int64_t n) {
    const unsigned lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    constexpr int RV = VPL * 32;  // 16-byte vectors per row
    for (int64_t r = w * RIF; r < n; r += warps * RIF) {
        uint4 v[RIF][VPL];
#pragma unroll
        for (int j = 0; j < RIF; j++) {
            if (r + j < n) {
                const uint4* src = table + (uint64_t)__ldg(idx + r + j) * RV;
#pragma unroll
                for (int k = 0; k < VPL; k++) v[j][k] = ld_host_v4(src + k * 32 + lane);
            }
        }
#pragma unroll
        for (int j = 0; j < RIF; j++) {
            if (r + j < n) {
                uint4* dst = out + (uint64_t)(r + j) * RV;
#pragma unroll
                for (int k = 0; k < VPL; k++) st_out_v4(dst + k * 32 + lane, v[j][k]);
            }
        }
    }
}

// Times `iters` launches (each over its own n ids: idx holds iters * n of them, so no launch re-reads rows that an earlier
// one left in L2) with CUDA events; us_per_launch = mean.  rows_in_flight 1|2, ctas_per_sm 1..8 CTAs of 256 threads.
extern "C" int coala_probe_gather(uint64_t table_ptr, uint64_t idx_ptr, uint64_t out_ptr, int64_t n, int row_bytes, int rows_in_flight,
                                  int ctas_per_sm, int iters, double* us_per_launch) {
    if (row_bytes != 512 && row_bytes != 1024 && row_bytes != 2048 && row_bytes != 4096) COALA_FAIL("probe_gather: row_bytes 512..4096");
    if (rows_in_flight != 1 && rows_in_flight != 2) COALA_FAIL("probe_gather: rows_in_flight 1 or 2");
    int dev = 0, sms = 148;
    COALA_CUDA(cudaGetDevice(&dev));
    COALA_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const unsigned grid = (unsigned)(sms * (ctas_per_sm < 1 ? 1 : ctas_per_sm));
    const uint4* t = reinterpret_cast<const uint4*>(table_ptr);
    const int64_t* ix = reinterpret_cast<const int64_t*>(idx_ptr);
    uint4* o = reinterpret_cast<uint4*>(out_ptr);
    cudaEvent_t e0, e1;
    COALA_CUDA(cudaEventCreate(&e0));
    COALA_CUDA(cudaEventCreate(&e1));
    auto go = [&](int it) {
        const int64_t* p = ix + (int64_t)it * n;
#define COALA_PG(V)                                                                 \
    if (rows_in_flight == 2) probe_gather_kernel<V, 2><<<grid, 256>>>(t, p, o, n);  \
    else probe_gather_kernel<V, 1><<<grid, 256>>>(t, p, o, n)
        switch (row_bytes / 512) {
            case 1: COALA_PG(1); break;
            case 2: COALA_PG(2); break;
            case 4: COALA_PG(4); break;
            default: COALA_PG(8); break;
        }
#undef COALA_PG
    };
    go(0);  // warm-up on list 0; the timed launches read lists 1..iters, so none of them re-reads rows the warm-up left in L2
    COALA_CUDA(cudaDeviceSynchronize());
    COALA_CUDA(cudaEventRecord(e0));
    for (int i = 1; i <= iters; i++) go(i);
    COALA_CUDA(cudaEventRecord(e1));
    COALA_CUDA(cudaEventSynchronize(e1));
    COALA_CUDA(cudaGetLastError());
    float ms = 0;
    COALA_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    *us_per_launch = 1e3 * ms / iters;
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}

// ------------------------------------------------------------------------------------------
// random small-read probe: how many independent locations per second a link serves when every
// location is read as `span` contiguous bytes (8, 32, 64, 128 or 256) by neighbouring lanes.  Tells whether
// the link charges per request or per 32-byte sector — the design input of the UVA sampler.
// ------------------------------------------------------------------------------------------
__global__ void probe_random_kernel(const uint64_t* __restrict__ src, uint64_t n_lines, int lanes_per_loc, int iters,
                                    uint64_t seed, int misalign, unsigned long long* __restrict__ sink) {
    const unsigned lane = threadIdx.x & 31;
    const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int locs_per_instr = 32 / lanes_per_loc;
    const int sub = lane / lanes_per_loc, off = lane % lanes_per_loc;
    uint64_t acc = 0;
    for (int it = 0; it < iters; it += 4) {
        uint64_t v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint64_t h = mix64(seed ^ mix64(warp * 1000003ULL + (uint64_t)(it + u) * 64 + sub));
            const uint64_t line = h % n_lines;                       // 256-byte granules
            v[u] = src[line * 32 + (misalign ? 12 : 0) + off];       // misalign: start 96 B into the granule
        }
        acc ^= v[0] ^ v[1] ^ v[2] ^ v[3];
    }
    (void)locs_per_instr;
    if (acc == 0x1234567812345677ULL) *sink = acc;
}

extern "C" int coala_probe_random_read(uint64_t src_ptr, int64_t bytes, int span_bytes, int misalign, int warps, int iters,
                                       double* locations_per_s) {
    if (span_bytes != 8 && span_bytes != 32 && span_bytes != 64 && span_bytes != 128 && span_bytes != 256)
        COALA_FAIL("probe_random_read: span must be 8, 32, 64, 128 or 256");
    const int lanes_per_loc = span_bytes / 8;
    unsigned long long* sink;
    COALA_CUDA(cudaMalloc(&sink, 8));
    cudaEvent_t e0, e1;
    COALA_CUDA(cudaEventCreate(&e0));
    COALA_CUDA(cudaEventCreate(&e1));
    const uint64_t n_lines = (uint64_t)bytes / 256 - 2;
    const int blocks = (warps + 7) / 8;
    probe_random_kernel<<<blocks, 256>>>(reinterpret_cast<const uint64_t*>(src_ptr), n_lines, lanes_per_loc, 4, 1, misalign, sink);
    COALA_CUDA(cudaEventRecord(e0));
    probe_random_kernel<<<blocks, 256>>>(reinterpret_cast<const uint64_t*>(src_ptr), n_lines, lanes_per_loc, iters, 7, misalign, sink);
    COALA_CUDA(cudaEventRecord(e1));
    COALA_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    COALA_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    *locations_per_s = (double)blocks * 8 * (32 / lanes_per_loc) * iters / (ms * 1e-3);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
    return 0;
}

// The same probe with the table hash-partitioned over G GPUs of this process like the cache (owner = id % G, row = id / G) and
// ALL G GPUs gathering at once: the practical ceiling of the N-GPU all-hit regime, where every GPU pulls (G-1)/G of its rows
// over NVLink while serving its peers.  Launches are issued iteration-major so the devices stay in step.
struct ProbeTables { const uint4* t[8]; int G; };

template <int VPL, int RIF>
__global__ void probe_gather_all_kernel(const __grid_constant__ ProbeTables pt, const int64_t* __restrict__ idx, uint4* __restrict__ out, int64_t n) {
    const unsigned lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    constexpr int RV = VPL * 32;
    const uint64_t G = (uint64_t)pt.G;
    for (int64_t r = w * RIF; r < n; r += warps * RIF) {
        uint4 v[RIF][VPL];
#pragma unroll
        for (int j = 0; j < RIF; j++) {
            if (r + j < n) {
                const uint64_t id = (uint64_t)__ldg(idx + r + j);
                const uint4* src = pt.t[id % G] + (id / G) * RV;
#pragma unroll
                for (int k = 0; k < VPL; k++) v[j][k] = ld_host_v4(src + k * 32 + lane);
            }
        }
#pragma unroll
        for (int j = 0; j < RIF; j++) {
            if (r + j < n) {
                uint4* dst = out + (uint64_t)(r + j) * RV;
#pragma unroll
                for (int k = 0; k < VPL; k++) st_out_v4(dst + k * 32 + lane, v[j][k]);
            }
        }
    }
}

// tables[g] = device g's part (peer access between all pairs must be enabled), idx[g] / out[g] = device g's id list
// ((iters + 1) * n ids: list 0 warms up, lists 1..iters are timed) and output rows; us_per_launch[g] = device g's mean launch time while all G run.  4 KB rows only.
extern "C" int coala_probe_gather_all(int G, const uint64_t* tables, const uint64_t* idx, const uint64_t* out, int64_t n, int rows_in_flight,
                                      int ctas_per_sm, int iters, double* us_per_launch) {
    if (G < 1 || G > 8) COALA_FAIL("probe_gather_all: 1..8 GPUs");
    if (rows_in_flight != 1 && rows_in_flight != 2) COALA_FAIL("probe_gather_all: rows_in_flight 1 or 2");
    ProbeTables pt;
    pt.G = G;
    for (int g = 0; g < 8; g++) pt.t[g] = reinterpret_cast<const uint4*>(tables[g < G ? g : 0]);
    int sms = 148;
    COALA_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const unsigned grid = (unsigned)(sms * (ctas_per_sm < 1 ? 1 : ctas_per_sm));
    cudaEvent_t e0[8], e1[8];
    for (int g = 0; g < G; g++) {
        COALA_CUDA(cudaSetDevice(g));
        COALA_CUDA(cudaEventCreate(&e0[g]));
        COALA_CUDA(cudaEventCreate(&e1[g]));
    }
    auto go = [&](int g, int it) {
        const int64_t* p = reinterpret_cast<const int64_t*>(idx[g]) + (int64_t)it * n;
        uint4* o = reinterpret_cast<uint4*>(out[g]);
        if (rows_in_flight == 2) probe_gather_all_kernel<8, 2><<<grid, 256>>>(pt, p, o, n);
        else probe_gather_all_kernel<8, 1><<<grid, 256>>>(pt, p, o, n);
    };
    for (int g = 0; g < G; g++) { COALA_CUDA(cudaSetDevice(g)); go(g, 0); }
    for (int g = 0; g < G; g++) { COALA_CUDA(cudaSetDevice(g)); COALA_CUDA(cudaDeviceSynchronize()); }
    for (int g = 0; g < G; g++) { COALA_CUDA(cudaSetDevice(g)); COALA_CUDA(cudaEventRecord(e0[g])); }
    for (int i = 1; i <= iters; i++)  // list 0 was the warm-up's
        for (int g = 0; g < G; g++) { COALA_CUDA(cudaSetDevice(g)); go(g, i); }
    for (int g = 0; g < G; g++) { COALA_CUDA(cudaSetDevice(g)); COALA_CUDA(cudaEventRecord(e1[g])); }
    for (int g = 0; g < G; g++) {
        COALA_CUDA(cudaSetDevice(g));
        COALA_CUDA(cudaEventSynchronize(e1[g]));
        COALA_CUDA(cudaGetLastError());
        float ms = 0;
        COALA_CUDA(cudaEventElapsedTime(&ms, e0[g], e1[g]));
        us_per_launch[g] = 1e3 * ms / iters;
        cudaEventDestroy(e0[g]); cudaEventDestroy(e1[g]);
    }
    COALA_CUDA(cudaSetDevice(0));
    return 0;
}

