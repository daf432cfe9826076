// See vmm.h.
#include "vmm.h"
#include <mutex>
#include <unistd.h>

namespace {

template <class F>
bool resolve(const char* name, F* out, std::string* err) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
        cudaGetLastError();
        *err = std::string("driver entry point ") + name + " unavailable";
        return false;
    }
    *out = reinterpret_cast<F>(fn);
    return true;
}

std::string cu_err(const char* what, CUresult r) { return std::string(what) + " failed (CUresult " + std::to_string((int)r) + ")"; }

CUmemAllocationProp prop_for(int device) {
    CUmemAllocationProp p = {};
    p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    p.location.id = device;
    p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return p;
}

std::string map_rw(VmmRegion* r, int device) {
    const VmmApi& a = vmm_api();
    CUresult rc = a.MemAddressReserve(&r->va, r->size, (size_t)2 << 20, 0, 0);
    if (rc != CUDA_SUCCESS) return cu_err("cuMemAddressReserve", rc);
    rc = a.MemMap(r->va, r->size, 0, r->handle, 0);
    if (rc != CUDA_SUCCESS) { a.MemAddressFree(r->va, r->size); r->va = 0; return cu_err("cuMemMap", rc); }
    r->mapped = true;
    return vmm_grant(*r, device);
}

}  // namespace

const VmmApi& vmm_api() {
    static VmmApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        cudaFree(nullptr);  // make sure the runtime (and with it the driver) is initialised
        std::string e;
        bool ok = resolve("cuMemGetAllocationGranularity", &api.MemGetAllocationGranularity, &e) &&
                  resolve("cuMemCreate", &api.MemCreate, &e) && resolve("cuMemRelease", &api.MemRelease, &e) &&
                  resolve("cuMemAddressReserve", &api.MemAddressReserve, &e) && resolve("cuMemAddressFree", &api.MemAddressFree, &e) &&
                  resolve("cuMemMap", &api.MemMap, &e) && resolve("cuMemUnmap", &api.MemUnmap, &e) &&
                  resolve("cuMemSetAccess", &api.MemSetAccess, &e) &&
                  resolve("cuMemExportToShareableHandle", &api.MemExportToShareableHandle, &e) &&
                  resolve("cuMemImportFromShareableHandle", &api.MemImportFromShareableHandle, &e);
        api.ok = ok;
        api.err = e;
    });
    return api;
}

std::string vmm_alloc(VmmRegion* r, size_t bytes, int device) {
    const VmmApi& a = vmm_api();
    if (!a.ok) return a.err;
    CUmemAllocationProp p = prop_for(device);
    size_t gran = 0;
    CUresult rc = a.MemGetAllocationGranularity(&gran, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
    if (rc != CUDA_SUCCESS || gran == 0) return cu_err("cuMemGetAllocationGranularity", rc);
    r->size = (bytes + gran - 1) / gran * gran;
    rc = a.MemCreate(&r->handle, r->size, &p, 0);
    if (rc != CUDA_SUCCESS) return cu_err("cuMemCreate", rc);
    std::string e = map_rw(r, device);
    if (!e.empty()) { vmm_free(r); return e; }
    return "";
}

int vmm_export_fd(const VmmRegion& r, std::string* err) {
    const VmmApi& a = vmm_api();
    int fd = -1;
    CUresult rc = a.MemExportToShareableHandle(&fd, r.handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (rc != CUDA_SUCCESS) { if (err) *err = cu_err("cuMemExportToShareableHandle", rc); return -1; }
    return fd;
}

std::string vmm_import(VmmRegion* r, int fd, size_t bytes, int device) {
    const VmmApi& a = vmm_api();
    if (!a.ok) return a.err;
    CUresult rc = a.MemImportFromShareableHandle(&r->handle, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (rc != CUDA_SUCCESS) return cu_err("cuMemImportFromShareableHandle", rc);
    r->size = bytes;  // the exporter's rounded size
    std::string e = map_rw(r, device);
    if (!e.empty()) { vmm_free(r); return e; }
    return "";
}

std::string vmm_grant(const VmmRegion& r, int device) {
    const VmmApi& a = vmm_api();
    CUmemAccessDesc d = {};
    d.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    d.location.id = device;
    d.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CUresult rc = a.MemSetAccess(r.va, r.size, &d, 1);
    if (rc != CUDA_SUCCESS) return cu_err("cuMemSetAccess", rc);
    return "";
}

void vmm_free(VmmRegion* r) {
    const VmmApi& a = vmm_api();
    if (!a.ok) return;
    if (r->mapped) { a.MemUnmap(r->va, r->size); r->mapped = false; }
    if (r->va) { a.MemAddressFree(r->va, r->size); r->va = 0; }
    if (r->handle) { a.MemRelease(r->handle); r->handle = 0; }
}
