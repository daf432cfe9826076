// Shared helpers for the sm_100a kernels of the COALA-GNN data path.
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

// ---------------------------------------------------------------------------------------
// error plumbing: every C-ABI entry point returns 0 / a handle on success and non-zero / NULL
// on failure with the message retrievable through coala_last_error().  CUDA errors are never
// swallowed (the reference's cuda_err_chk only prints, bam/include/util.h:22-45).
// ---------------------------------------------------------------------------------------
void coala_set_error(const std::string& msg);

#define COALA_CUDA(call)                                                                       \
    do {                                                                                       \
        cudaError_t _e = (call);                                                               \
        if (_e != cudaSuccess) {                                                               \
            coala_set_error(std::string(#call) + " failed: " + cudaGetErrorString(_e) + " (" + \
                            __FILE__ + ":" + std::to_string(__LINE__) + ")");                  \
            return -1;                                                                         \
        }                                                                                      \
    } while (0)

#define COALA_CUDA_NULL(call)                                                                  \
    do {                                                                                       \
        cudaError_t _e = (call);                                                               \
        if (_e != cudaSuccess) {                                                               \
            coala_set_error(std::string(#call) + " failed: " + cudaGetErrorString(_e) + " (" + \
                            __FILE__ + ":" + std::to_string(__LINE__) + ")");                  \
            return nullptr;                                                                    \
        }                                                                                      \
    } while (0)

// SM count of the current device, looked up once per device: grid caps are written as multiples of it
inline int coala_sm_count() {
    static std::atomic<int> cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int v = cached[dev].load(std::memory_order_relaxed);
    if (v <= 0) {
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        cached[dev].store(v, std::memory_order_relaxed);
    }
    return v;
}

#define COALA_FAIL(msg)          \
    do {                         \
        coala_set_error(msg);    \
        return -1;               \
    } while (0)

// ---------------------------------------------------------------------------------------
// deterministic generators — integer-identical to oracle/c/common.c and oracle/c/sampler.c
// ---------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

__device__ __forceinline__ float feat_from_hash(uint64_t h) {
    return (float)(uint32_t)(h >> 40) * (1.0f / 16777216.0f);  // 24 bits: exact
}

// ---------------------------------------------------------------------------------------
// PTX memory-model wrappers.  SCOPE_SYS=true is used whenever a cache shard can be touched
// by a peer GPU over NVLink (partitioned cache); device scope otherwise.
// ---------------------------------------------------------------------------------------
template <bool SYS> __device__ __forceinline__ uint64_t ld_acquire_u64(const uint64_t* p) {
    uint64_t v;
    if (SYS) asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    else asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
template <bool SYS> __device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t* p) {
    uint64_t v;
    if (SYS) asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    else asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
template <bool SYS> __device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
    uint32_t v;
    if (SYS) asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    else asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
template <bool SYS> __device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) {
    if (SYS) asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
    else asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
template <bool SYS> __device__ __forceinline__ void st_relaxed_u64(uint64_t* p, uint64_t v) {
    if (SYS) asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
    else asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
template <bool SYS> __device__ __forceinline__ void st_release_u64(uint64_t* p, uint64_t v) {
    if (SYS) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
    else asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
template <bool SYS> __device__ __forceinline__ uint64_t cas_acquire_u64(uint64_t* p, uint64_t cmp, uint64_t val) {
    uint64_t old;
    if (SYS) asm volatile("atom.acquire.sys.global.cas.b64 %0, [%1], %2, %3;" : "=l"(old) : "l"(p), "l"(cmp), "l"(val) : "memory");
    else asm volatile("atom.acquire.gpu.global.cas.b64 %0, [%1], %2, %3;" : "=l"(old) : "l"(p), "l"(cmp), "l"(val) : "memory");
    return old;
}
template <bool SYS> __device__ __forceinline__ uint64_t cas_relaxed_u64(uint64_t* p, uint64_t cmp, uint64_t val) {
    uint64_t old;
    if (SYS) asm volatile("atom.relaxed.sys.global.cas.b64 %0, [%1], %2, %3;" : "=l"(old) : "l"(p), "l"(cmp), "l"(val) : "memory");
    else asm volatile("atom.relaxed.gpu.global.cas.b64 %0, [%1], %2, %3;" : "=l"(old) : "l"(p), "l"(cmp), "l"(val) : "memory");
    return old;
}
template <bool SYS> __device__ __forceinline__ void red_add_s32(int* p, int v) {
    if (SYS) asm volatile("red.relaxed.sys.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
    else asm volatile("red.relaxed.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
template <bool SYS> __device__ __forceinline__ void fence_acq_rel() {
    if (SYS) asm volatile("fence.acq_rel.sys;" ::: "memory");
    else asm volatile("fence.acq_rel.gpu;" ::: "memory");
}
// 16-byte row-data accessors.  Cache lines are mutable while the kernel runs (another warp,
// possibly on another GPU, may be refilling the slot), so they are read/written with strong
// relaxed accesses that resolve at L2; the seqlock in the tag word validates them.
template <bool SYS> __device__ __forceinline__ uint4 ld_line_v4(const uint4* p) {
    uint4 v;
    if (SYS) asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    else asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
template <bool SYS> __device__ __forceinline__ void st_line_v4(uint4* p, const uint4& v) {
    if (SYS) asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    else asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// immutable pinned-host feature rows: read-only path, do not pollute L1
__device__ __forceinline__ uint4 ld_host_v4(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
// gathered output rows: written once, consumed by the next kernel
__device__ __forceinline__ void st_out_v4(uint4* p, const uint4& v) {
    asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }

// pointer classification for the drop-in API (callers pass raw integers that may be host or
// device addresses)
static inline bool coala_is_device_ptr(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}
