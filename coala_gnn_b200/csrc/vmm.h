// Shareable device allocations through the CUDA virtual-memory-management driver API.
//
// Why not cudaMalloc + cudaIpcOpenMemHandle: measured on B200 (gpurun_out r1_n2_small_32g vs
// r1_n2_small_c) a 32 GB cache slab imported with cudaIpcOpenMemHandle serves random
// 4 KB rows at ~350 GB/s, the same slab mapped in-process at ~930 GB/s — the IPC import is mapped with
// a page size whose TLB reach does not cover tens of GB.  cuMemCreate + cuMemMap gives an explicit
// 2 MB-granular mapping on both the owner and every importer; the handle crosses processes as a
// POSIX file descriptor (sent over a unix socket by the Python side).
//
// The driver entry points are looked up at run time (cudaGetDriverEntryPoint) so the library keeps
// loading on machines without libcuda (the build container), where only the symbol table is checked.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <string>

struct VmmApi {
    CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
    CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
    bool ok = false;
    std::string err;
};

const VmmApi& vmm_api();  // resolved once; .ok == false (with .err) when the driver is unavailable

struct VmmRegion {           // one mapping of one physical allocation in this process
    CUmemGenericAllocationHandle handle = 0;
    CUdeviceptr va = 0;
    size_t size = 0;         // rounded up to the allocation granularity
    bool mapped = false;
};

// allocate `bytes` on `device`, map it read/write for `device`; returns "" or an error message
std::string vmm_alloc(VmmRegion* r, size_t bytes, int device);
// POSIX fd for the allocation (caller closes it); -1 on failure
int vmm_export_fd(const VmmRegion& r, std::string* err);
// import an fd exported by a peer process/device and map it read/write for `device`
std::string vmm_import(VmmRegion* r, int fd, size_t bytes, int device);
// let one more device of THIS process access an existing mapping (same-process peers)
std::string vmm_grant(const VmmRegion& r, int device);
void vmm_free(VmmRegion* r);
