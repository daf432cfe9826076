// COALA_GNN_Pybind as a compiled pybind11 module over the C ABI of libcoala_b200.so (include/coala_b200.h).
//
// This is the binding a maintainer of the reference would put in place of COALA_GNN_Modules/COALA_GNN_Pybind.cu
// (:27-79): the same seven classes, method names, positional arities and raw-integer-pointer convention, linked
// against -lcoala_b200 instead of libnvm / NVSHMEM / MPI.  Built by coala_gnn_b200/build.py into
// COALA_GNN_Pybind/_native*.so; `COALA_PYBIND_NATIVE=1` makes the COALA_GNN package use it instead of the ctypes
// binding (coala_gnn_b200/pybind.py).  No torch type crosses this file.  Methods the reference does not have are
// additions of the kind SURVEY.md 8(b) allows (stream-taking forms, statistics, shard descriptors) and are marked (+).
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>
#include <cuda_runtime_api.h>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <tuple>
#include <unordered_map>
#include <vector>
#include "coala_b200.h"

namespace py = pybind11;

namespace {

void ck(int rc) {
    if (rc) throw std::runtime_error(coala_last_error());
}
template <class T> T* ckh(T* h) {
    if (!h) throw std::runtime_error(coala_last_error());
    return h;
}
int cur_dev() {
    int d = 0;
    cudaGetDevice(&d);
    return d;
}
void* as_stream(uintptr_t s) { return reinterpret_cast<void*>(s); }

// M/shared_UVA.cuh:26-115.  No MPI here: the 5th argument, an MPI_Comm* address in the reference, says whether this
// process creates the segment (non-zero: the local-rank-0 branch, :65-77) or opens it (:78-81); the caller orders the
// two with its own barrier (COALA_GNN.Shared_UVA_Tensor_Manager does).
struct SharedUVAManager {
    coala_uva* h;
    std::string path;
    bool creator;
    SharedUVAManager(const std::string& shm_path, int64_t bytes, int /*node_id*/, int64_t /*global_comm_ptr*/, int64_t creates,
                     int device, bool zero)
        : h(ckh(coala_uva_create(shm_path.c_str(), bytes, creates != 0, device < 0 ? cur_dev() : device, zero ? 1 : 0))),
          path(shm_path), creator(creates != 0) {}
    ~SharedUVAManager() { if (h) coala_uva_destroy(h); }
    uintptr_t get_host_ptr() { return coala_uva_host_ptr(h); }
    uintptr_t get_device_ptr() { return coala_uva_device_ptr(h); }
    void cleanup() { ck(coala_uva_cleanup(h, creator ? 1 : 0)); }
};

// M/ssd_gnn_cache.cuh:10-55: only the geometry survives (the NVMe controllers are out of scope).
struct SSD_GNN_SSD_Controllers {
    uint32_t n_ctrls, page_size;
    int cudaDevice, dim, cache_dim;
    SSD_GNN_SSD_Controllers(uint32_t num_ctrls, uint32_t /*page_size*/, uint64_t, uint64_t, uint32_t device, int feat_dim, bool sim)
        : n_ctrls(num_ctrls), cudaDevice((int)device), dim(feat_dim) {
        if (!sim) throw std::runtime_error("SSD_GNN_SSD_Controllers: the BaM/NVMe tier is out of scope; run with --feat_cpu (sim=True)");
        cache_dim = coala_cache_dim_for(feat_dim);
        if (cache_dim < 0) throw std::runtime_error("Only Feature Embedding Size less than 8KB is supported\n");  // (:43-44)
        page_size = (uint32_t)cache_dim * 4;                                                                     // (:47)
    }
};

// M/node_distributor_pybind.cuh:112-238
struct Node_distributor_pybind {
    coala_nd* h;
    Node_distributor_pybind(uint64_t item_ptr, int n_domains) : h(ckh(coala_nd_create_plain(item_ptr, n_domains))) {}
    Node_distributor_pybind(uint64_t item_ptr, int node_id, int batch, int local_size, int n_domains, const std::string& color_file,
                            const std::string& topk_file, const std::string& score_file, int device)
        : h(ckh(coala_nd_create(item_ptr, node_id, batch, local_size, n_domains, color_file.c_str(), topk_file.c_str(),
                                score_file.c_str(), device < 0 ? cur_dev() : device))) {}
    ~Node_distributor_pybind() { if (h) coala_nd_destroy(h); }
    void distribute_node_with_affinity(uint64_t offset, uint64_t out_ptr, std::vector<uint64_t> meta) {
        py::gil_scoped_release r;  // the reference holds the GIL here although it runs on a thread (S/COALA_GNN_DataLoader.py:32-33)
        ck(coala_nd_distribute(h, offset, out_ptr, meta.data(), (int)meta.size()));
    }
    void distribute_async(uint64_t offset, uint64_t out_dev_ptr, std::vector<uint64_t> meta, uintptr_t stream, int slot) {
        ck(coala_nd_distribute_async_slot(h, offset, out_dev_ptr, meta.data(), (int)meta.size(), slot, as_stream(stream)));
    }
    int get_num_colors() { return coala_nd_num_colors(h); }
    uint64_t get_color_buffer_ptr() { return coala_nd_color_host_ptr(h); }
    void set_meta_len(int n) { ck(coala_nd_set_meta_len(h, n)); }
};

// shared by the two cache front ends (M/ssd_gnn_cache.cuh:58-200, 205-372)
struct CacheBase {
    coala_cache* h = nullptr;
    int dim = 0, num_gpus = 1;
    CacheBase(int kind, const SSD_GNN_SSD_Controllers& c, Node_distributor_pybind* nd, int global_rank, int n_gpus, uint64_t cache_mb,
              uint64_t sim_buf_ptr, int local_slot)
        : dim(c.dim), num_gpus(n_gpus) {
        // the distributor lends its colour table: it must outlive the cache (M/ssd_gnn_cache.cuh:91); keep_alive in the binding
        h = ckh(coala_cache_create(kind, c.cudaDevice, c.dim, nd ? nd->h : nullptr, global_rank, local_slot < 0 ? global_rank % n_gpus : local_slot,
                                   n_gpus, cache_mb, sim_buf_ptr));
    }
    ~CacheBase() { if (h) coala_cache_destroy(h); }
    void get_cache_data(int64_t dst) { ck(coala_cache_get_cache_data(h, (uint64_t)dst)); }
    void get_cache_data_n(int64_t dst, int count) { ck(coala_cache_get_cache_data_n(h, (uint64_t)dst, count)); }
    void get_cache_data_async(uint64_t dst_dev, int count, uintptr_t stream) { ck(coala_cache_get_cache_data_async(h, dst_dev, count, as_stream(stream))); }
    void print_stats() { ck(coala_cache_print_stats(h)); }
    std::tuple<uint64_t, uint64_t, uint64_t> stats(bool reset) {
        uint64_t a = 0, b = 0, c = 0;
        ck(coala_cache_stats(h, &a, &b, &c, reset ? 1 : 0));
        return {a, b, c};
    }
    uint64_t num_sets() { return coala_cache_num_sets(h); }
    void set_copy_engine(int e) { ck(coala_cache_set_copy_engine(h, e)); }
    void set_num_nodes(int64_t n) { ck(coala_cache_set_num_nodes(h, n)); }
    uint64_t refused_ids(bool reset) {
        uint64_t n = 0;
        ck(coala_cache_refused_ids(h, &n, reset ? 1 : 0));
        return n;
    }
    uint64_t color_counters_dev_ptr() { return coala_cache_color_counters_dev_ptr(h); }
    uintptr_t handle() { return reinterpret_cast<uintptr_t>(h); }
    void read_feature_async(uint64_t out, uint64_t idx, int64_t n, uint64_t n_dev_ptr, uintptr_t stream) {
        ck(coala_cache_read_feature_async(h, out, idx, n, n_dev_ptr, as_stream(stream)));
    }
};

struct Isolated_Cache : CacheBase {
    Isolated_Cache(SSD_GNN_SSD_Controllers c, Node_distributor_pybind& nd, int global_rank, int n_gpus, uint64_t cache_mb, uint64_t sim_buf_ptr,
                   int kind, int local_slot)
        : CacheBase(kind, c, &nd, global_rank, n_gpus, cache_mb, sim_buf_ptr, local_slot) {}
    void read_feature(uint64_t out, uint64_t idx, int64_t n) {
        py::gil_scoped_release r;
        ck(coala_cache_read_feature(h, out, idx, n));
    }
    void split_node_list(uint64_t idx, int64_t n, uint64_t node, uint64_t map, uint64_t counter, int G, int max_sample) {
        ck(coala_split_node_list(idx, n, node, map, counter, G, max_sample, nullptr));
        cudaDeviceSynchronize();  // (M/ssd_gnn_cache.cuh:294)
    }
    void nccl_get_feature(std::vector<uint64_t> idx, std::vector<uint64_t> out, std::vector<uint64_t> sizes, int G, int /*max_sample*/) {
        for (int g = 0; g < G; g++)
            if (sizes[g]) ck(coala_cache_read_feature_async(h, out[g], idx[g], (int64_t)sizes[g], 0, nullptr));
        cudaDeviceSynchronize();  // (:321)
    }
    void map_feat_data(uint64_t out, std::vector<uint64_t> src, uint64_t map, std::vector<uint64_t> sizes, int G, int max_sample) {
        for (int g = 0; g < G; g++)
            if (sizes[g]) ck(coala_map_feat_data(out, src[g], map + 8ull * (uint64_t)max_sample * g, (int64_t)sizes[g], dim, nullptr));
        cudaDeviceSynchronize();  // (:351)
    }
};

struct SSD_GNN_NVSHMEM_Cache : CacheBase {
    uint64_t pending_idx = 0;
    int64_t pending_n = -1;
    SSD_GNN_NVSHMEM_Cache(SSD_GNN_SSD_Controllers c, Node_distributor_pybind& nd, int global_rank, int n_gpus, uint64_t cache_mb,
                          uint64_t sim_buf_ptr, int local_slot)
        : CacheBase(COALA_CACHE_PARTITIONED, c, &nd, global_rank, n_gpus, cache_mb, sim_buf_ptr, local_slot) {}
    // no request exchange: the requester's warps pull rows straight out of the owner's shard (M/ssd_gnn_cache.cuh:111-174)
    void send_requests(uint64_t idx, int64_t n, uint64_t /*req_ptr*/, int /*max_index*/) { pending_idx = idx; pending_n = n; }
    void read_feature(uint64_t out, uint64_t /*req_ptr*/, int64_t /*max_index*/) {
        if (pending_n < 0) throw std::runtime_error("read_feature called before send_requests");
        const int64_t n = pending_n;
        pending_n = -1;
        py::gil_scoped_release r;
        ck(coala_cache_read_feature(h, out, pending_idx, n));
    }
    // (+) peers: each shard is a VMM allocation shared as a POSIX descriptor (or a CUDA IPC handle with COALA_CACHE_LEGACY_IPC=1)
    bool is_fd_shared() { return coala_cache_is_fd_shared(h) != 0; }
    int export_fd() {
        int fd = coala_cache_export_fd(h);
        if (fd < 0) throw std::runtime_error(coala_last_error());
        return fd;
    }
    void attach_fd(int peer, int fd) { ck(coala_cache_attach_fd(h, peer, fd)); }
    py::bytes ipc_handle() {
        char buf[64];
        ck(coala_cache_ipc_handle(h, buf));
        return py::bytes(buf, 64);
    }
    void attach_ipc(int peer, const std::string& handle) {
        if (handle.size() != 64) throw std::runtime_error("attach_ipc: a CUDA IPC handle is 64 bytes");
        ck(coala_cache_attach_ipc(h, peer, handle.data()));
    }
    void attach_local(int peer, SSD_GNN_NVSHMEM_Cache& other) { ck(coala_cache_attach_local(h, peer, other.h)); }
};

// M/nvshmem_manager.cuh:9-56: there is no NVSHMEM and no symmetric heap (rows are pulled by the requester, nobody
// writes into a peer's batch buffer), so allocate() is plain device memory with 64-bit sizes.
struct NVSHMEM_Manager {
    std::unordered_map<uint64_t, void*> bufs;
    NVSHMEM_Manager(int64_t /*local_comm_ptr*/, int /*local_rank*/) {}
    ~NVSHMEM_Manager() { finalize(); }
    uint64_t allocate(int64_t bytes) {
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, (size_t)bytes);
        if (e != cudaSuccess) throw std::runtime_error(std::string("NVSHMEM_Manager.allocate: ") + cudaGetErrorString(e));
        bufs[(uint64_t)p] = p;
        return (uint64_t)p;
    }
    void free(uint64_t ptr) {
        auto it = bufs.find(ptr);
        if (it != bufs.end()) { cudaFree(it->second); bufs.erase(it); }
    }
    void finalize() {
        for (auto& kv : bufs) cudaFree(kv.second);
        bufs.clear();
    }
};

// M/graph_coloring.h:15-68
struct Graph_Coloring {
    coala_gc* h;
    explicit Graph_Coloring(uint64_t num_nodes) : h(ckh(coala_gc_create(num_nodes))) {}
    ~Graph_Coloring() { if (h) coala_gc_destroy(h); }
};

}  // namespace

PYBIND11_MODULE(_native, m) {
    m.doc() = "COALA_GNN_Pybind over libcoala_b200.so (B200-native data path); same classes as COALA_GNN_Modules/COALA_GNN_Pybind.cu:27-79";
    m.attr("ISOLATED") = (int)COALA_CACHE_ISOLATED;
    m.attr("PARTITIONED") = (int)COALA_CACHE_PARTITIONED;
    m.attr("NCCL_SHARD") = (int)COALA_CACHE_NCCL_SHARD;
    m.def("abi_version", &coala_abi_version);

    py::class_<SharedUVAManager>(m, "SharedUVAManager")
        .def(py::init<const std::string&, int64_t, int, int64_t, int64_t, int, bool>(), py::arg("shm_path"), py::arg("shm_size"),
             py::arg("node_id") = 0, py::arg("global_comm_ptr") = 0, py::arg("local_comm_ptr") = 1, py::arg("device") = -1, py::arg("zero") = true)
        .def("get_host_ptr", &SharedUVAManager::get_host_ptr)
        .def("get_device_ptr", &SharedUVAManager::get_device_ptr)
        .def("cleanup", &SharedUVAManager::cleanup)
        .def_readonly("_path", &SharedUVAManager::path);

    py::class_<SSD_GNN_SSD_Controllers>(m, "SSD_GNN_SSD_Controllers")
        .def(py::init<uint32_t, uint32_t, uint64_t, uint64_t, uint32_t, int, bool>())
        .def_readonly("cudaDevice", &SSD_GNN_SSD_Controllers::cudaDevice)
        .def_readonly("dim", &SSD_GNN_SSD_Controllers::dim)
        .def_readonly("cache_dim", &SSD_GNN_SSD_Controllers::cache_dim)
        .def_readonly("page_size", &SSD_GNN_SSD_Controllers::page_size);

    py::class_<Node_distributor_pybind>(m, "Node_distributor_pybind")
        .def(py::init<uint64_t, int>())
        .def(py::init<uint64_t, int, int, int, int, const std::string&, const std::string&, const std::string&, int>(), py::arg("item_ptr"),
             py::arg("node_id"), py::arg("batch_size"), py::arg("local_size"), py::arg("num_nodes"), py::arg("color_file"), py::arg("topk_file"),
             py::arg("score_file"), py::arg("device") = -1)
        .def("distribute_node_with_affinity", &Node_distributor_pybind::distribute_node_with_affinity)
        .def("get_num_colors", &Node_distributor_pybind::get_num_colors)
        .def("get_color_buffer_ptr", &Node_distributor_pybind::get_color_buffer_ptr)
        .def("set_meta_len", &Node_distributor_pybind::set_meta_len)                                                          // (+)
        .def("distribute_async", &Node_distributor_pybind::distribute_async, py::arg("offset"), py::arg("out_dev_ptr"),      // (+)
             py::arg("meta_dev_ptrs"), py::arg("stream") = 0, py::arg("slot") = 1);

#define COALA_CACHE_COMMON(cls)                                                                                              \
    .def("get_cache_data", &cls::get_cache_data)                                                                             \
    .def("get_cache_data", &cls::get_cache_data_n)                                                             /* (+) */     \
    .def("get_cache_data_async", &cls::get_cache_data_async, py::arg("dst_dev_ptr"), py::arg("count"), py::arg("stream") = 0) /* (+) */ \
    .def("print_stats", &cls::print_stats)                                                                                   \
    .def("stats", &cls::stats, py::arg("reset") = false)                                                       /* (+) */     \
    .def("num_sets", &cls::num_sets)                                                                           /* (+) */     \
    .def("set_copy_engine", &cls::set_copy_engine)                                                             /* (+) */     \
    .def("set_num_nodes", &cls::set_num_nodes)                                                                 /* (+) */     \
    .def("refused_ids", &cls::refused_ids, py::arg("reset") = true)                                            /* (+) */     \
    .def("color_counters_dev_ptr", &cls::color_counters_dev_ptr)                                               /* (+) */     \
    .def("handle", &cls::handle)                                                 /* (+) the C-ABI handle, for coala_nccl_create */ \
    .def("read_feature_async", &cls::read_feature_async, py::arg("out_ptr"), py::arg("idx_ptr"), py::arg("n"),               \
         py::arg("n_dev_ptr") = 0, py::arg("stream") = 0)                                                      /* (+) */     \
    .def_readonly("num_gpus", &cls::num_gpus)

    py::class_<Isolated_Cache>(m, "Isolated_Cache")
        .def(py::init<SSD_GNN_SSD_Controllers, Node_distributor_pybind&, int, int, uint64_t, uint64_t, int, int>(), py::arg("controllers"),
             py::arg("node_distributor"), py::arg("global_rank"), py::arg("n_gpus"), py::arg("cache_size"), py::arg("sim_buf_ptr"),
             py::arg("kind") = (int)COALA_CACHE_ISOLATED, py::arg("local_slot") = -1, py::keep_alive<1, 3>())
        .def("read_feature", &Isolated_Cache::read_feature)
        .def("split_node_list", &Isolated_Cache::split_node_list)
        .def("nccl_get_feature", &Isolated_Cache::nccl_get_feature)
        .def("map_feat_data", &Isolated_Cache::map_feat_data)
        COALA_CACHE_COMMON(Isolated_Cache);

    py::class_<SSD_GNN_NVSHMEM_Cache>(m, "SSD_GNN_NVSHMEM_Cache")
        .def(py::init<SSD_GNN_SSD_Controllers, Node_distributor_pybind&, int, int, uint64_t, uint64_t, int>(), py::arg("controllers"),
             py::arg("node_distributor"), py::arg("global_rank"), py::arg("n_gpus"), py::arg("cache_size"), py::arg("sim_buf_ptr"),
             py::arg("local_slot") = -1, py::keep_alive<1, 3>())
        .def("send_requests", &SSD_GNN_NVSHMEM_Cache::send_requests)
        .def("read_feature", &SSD_GNN_NVSHMEM_Cache::read_feature)
        .def("is_fd_shared", &SSD_GNN_NVSHMEM_Cache::is_fd_shared)                                             // (+)
        .def("export_fd", &SSD_GNN_NVSHMEM_Cache::export_fd)                                                   // (+)
        .def("attach_fd", &SSD_GNN_NVSHMEM_Cache::attach_fd)                                                   // (+)
        .def("ipc_handle", &SSD_GNN_NVSHMEM_Cache::ipc_handle)                                                 // (+)
        .def("attach_ipc", &SSD_GNN_NVSHMEM_Cache::attach_ipc)                                                 // (+)
        .def("attach_local", &SSD_GNN_NVSHMEM_Cache::attach_local, py::keep_alive<1, 3>())                     // (+)
        COALA_CACHE_COMMON(SSD_GNN_NVSHMEM_Cache);
#undef COALA_CACHE_COMMON

    py::class_<NVSHMEM_Manager>(m, "NVSHMEM_Manager")
        .def(py::init<int64_t, int>())
        .def("allocate", &NVSHMEM_Manager::allocate)
        .def("free", &NVSHMEM_Manager::free)
        .def("finalize", &NVSHMEM_Manager::finalize);

    py::class_<Graph_Coloring>(m, "Graph_Coloring")
        .def(py::init<uint64_t>())
        .def("cpu_color_graph", [](Graph_Coloring& g) { py::gil_scoped_release r; coala_gc_cpu_color_graph(g.h); })
        .def("cpu_color_graph_optimized", [](Graph_Coloring& g, uint64_t train, uint64_t n) { py::gil_scoped_release r; coala_gc_cpu_color_graph_optimized(g.h, train, n); })
        .def("cpu_count_nearest_color", [](Graph_Coloring& g) { coala_gc_cpu_count_nearest_color(g.h); })
        .def("cpu_count_nearest_color_less_memory", [](Graph_Coloring& g) { coala_gc_cpu_count_nearest_color_less_memory(g.h); })
        .def("cpu_calculate_color_affinity", [](Graph_Coloring& g) { py::gil_scoped_release r; coala_gc_cpu_calculate_color_affinity(g.h); })
        .def("set_color_buffer", [](Graph_Coloring& g, uint64_t p) { coala_gc_set_color_buffer(g.h, p); })
        .def("set_topk_color_buffer", [](Graph_Coloring& g, uint64_t p) { coala_gc_set_topk_color_buffer(g.h, p); })
        .def("set_topk_affinity_buffer", [](Graph_Coloring& g, uint64_t p) { coala_gc_set_topk_affinity_buffer(g.h, p); })
        .def("set_adj_csc", [](Graph_Coloring& g, uint64_t a, uint64_t b) { coala_gc_set_adj_csc(g.h, a, b); })
        .def("get_num_color_node", [](Graph_Coloring& g) { return coala_gc_get_num_color_node(g.h); })
        .def("get_num_color", [](Graph_Coloring& g) { return coala_gc_get_num_color(g.h); })
        .def("gpu_color_graph_optimized", [](Graph_Coloring& g, uint64_t train, uint64_t n, int dev) { ck(coala_gc_gpu_color_graph_optimized(g.h, train, n, dev)); },  // (+)
             py::arg("train_ptr"), py::arg("n_train"), py::arg("device") = 0)
        .def("gpu_calculate_color_affinity", [](Graph_Coloring& g, int dev) { ck(coala_gc_gpu_calculate_color_affinity(g.h, dev)); }, py::arg("device") = 0);  // (+)
}
