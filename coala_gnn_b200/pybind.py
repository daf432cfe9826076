"""Drop-in for the reference's native module ``COALA_GNN_Pybind`` (M/COALA_GNN_Pybind.cu:27-79).

Same seven class names, method names, positional arities and raw-integer-pointer convention
(SURVEY.md §8b); every method forwards to the C ABI of libcoala_b200.so.  Deviations, all of
the "allowed" kind listed in SURVEY.md §8b:
  * MPI communicator arguments are ignored or replaced by a torch.distributed process group
    (no MPI in this stack);
  * CUDA / argument errors raise RuntimeError instead of being printed and ignored;
  * NVSHMEM_Manager.allocate takes 64-bit sizes (the reference truncates to int).
"""
import ctypes as C

from ._lib import lib, check, check_handle, CoalaError

__all__ = [
    "SharedUVAManager", "SSD_GNN_SSD_Controllers", "SSD_GNN_NVSHMEM_Cache", "Isolated_Cache",
    "Node_distributor_pybind", "NVSHMEM_Manager", "Graph_Coloring", "tensor_from_ptr", "NcclExchange", "connect_partitioned",
]

ISOLATED, PARTITIONED, NCCL_SHARD = 0, 1, 2


class _CudaArray:
    """Minimal __cuda_array_interface__ carrier so torch can view memory it does not own
    (the reference uses cupy.cuda.UnownedMemory for this, S/Shared_Tensor.py:146-150)."""

    def __init__(self, ptr, shape, typestr, owner=None):
        self.__cuda_array_interface__ = {"shape": tuple(int(s) for s in shape), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 2, "strides": None}
        self._owner = owner


_TYPESTR = {"float32": "<f4", "int64": "<i8", "int32": "<i4", "uint8": "|u1", "float64": "<f8"}


def tensor_from_ptr(ptr, shape, dtype, device, owner=None):
    """torch tensor aliasing device-visible memory at `ptr` (device memory or UVA-mapped pinned host)."""
    import torch
    name = str(dtype).replace("torch.", "").replace("<class 'numpy.", "").replace("'>", "")
    t = torch.as_tensor(_CudaArray(ptr, shape, _TYPESTR[name], owner), device=device)
    t._coala_owner = owner
    return t


class SharedUVAManager:
    """M/shared_UVA.cuh:26-115. (shm_path, bytes, node_id, global_comm, local_comm): the comm
    arguments carry no MPI here; `local_comm` may be an object with .rank()/.barrier() (see
    COALA_GNN.MPI_Comm_Manager.local_comm) — rank 0 creates the segment, the others open it
    after the barrier, as the reference does with MPI_Barrier (:65-81)."""

    def __init__(self, shm_path, shm_size, node_id=0, global_comm_ptr=0, local_comm_ptr=0, device=None, zero=True):
        import torch
        comm = local_comm_ptr if hasattr(local_comm_ptr, "barrier") else None
        rank = comm.rank() if comm else 0
        dev = torch.cuda.current_device() if device is None else device
        self._path = shm_path.encode()
        if rank == 0:
            self._h = check_handle(lib().coala_uva_create(self._path, int(shm_size), 1, dev, 1 if zero else 0), "SharedUVAManager")
            if comm:
                comm.barrier()
        else:
            comm.barrier()
            self._h = check_handle(lib().coala_uva_create(self._path, int(shm_size), 0, dev, 0), "SharedUVAManager")
        self._creator = rank == 0
        self.size = int(shm_size)

    def get_host_ptr(self):
        return int(lib().coala_uva_host_ptr(self._h))

    def get_device_ptr(self):
        return int(lib().coala_uva_device_ptr(self._h))

    def cleanup(self):
        if self._h:
            lib().coala_uva_cleanup(self._h, 1 if self._creator else 0)

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().coala_uva_destroy(self._h)
                self._h = None
        except Exception:
            pass


class SSD_GNN_SSD_Controllers:
    """M/ssd_gnn_cache.cuh:10-55: only the geometry survives (the NVMe controllers are out of scope)."""

    def __init__(self, num_ctrls, page_size, n_elems, read_off, device_id, feat_dim, sim):
        if not sim:
            raise RuntimeError("SSD_GNN_SSD_Controllers: the BaM/NVMe tier is out of scope; run with --feat_cpu (sim=True)")
        cd = lib().coala_cache_dim_for(int(feat_dim))
        if cd < 0:
            raise RuntimeError("Only Feature Embedding Size less than 8KB is supported\n")  # (:43-44)
        self.n_ctrls, self.num_elements, self.offset = num_ctrls, n_elems, read_off
        self.cudaDevice, self.dim, self.cache_dim = int(device_id), int(feat_dim), cd
        self.page_size = cd * 4  # (:47)
        self.SSD_SIM = True


class Node_distributor_pybind:
    """M/node_distributor_pybind.cuh:112-238."""

    def __init__(self, item_ptr, *args, device=None):
        import torch
        self._h = None
        if len(args) == 1:  # (ptr, n_nodes) — no colour information (:133-135)
            self._h = check_handle(lib().coala_nd_create_plain(int(item_ptr), int(args[0])), "Node_distributor_pybind")
            return
        node_id, batch, local_size, n_domains, color_file, topk_file, score_file = args
        dev = (torch.cuda.current_device() if device is None else device)
        self._h = check_handle(lib().coala_nd_create(int(item_ptr), int(node_id), int(batch), int(local_size), int(n_domains),
                                                     str(color_file).encode(), str(topk_file).encode(),
                                                     str(score_file).encode(), dev), "Node_distributor_pybind")

    @classmethod
    def from_arrays(cls, item_ptr, node_id, batch, local_size, n_domains, color, topk, score, device=0):
        import numpy as np
        self = cls.__new__(cls)
        color = np.ascontiguousarray(color, dtype=np.int64)
        topk = np.ascontiguousarray(topk, dtype=np.int64)
        score = np.ascontiguousarray(score, dtype=np.float64)
        self._h = check_handle(lib().coala_nd_create_from_arrays(
            int(item_ptr), node_id, batch, local_size, n_domains, color.ctypes.data, len(color), topk.ctypes.data,
            score.ctypes.data, topk.shape[0], topk.shape[1], device), "Node_distributor_pybind")
        return self

    def distribute_node_with_affinity(self, offset, out_ptr, meta_ptrs):
        arr = (C.c_uint64 * len(meta_ptrs))(*[int(p) for p in meta_ptrs])
        check(lib().coala_nd_distribute(self._h, int(offset), int(out_ptr), arr, len(meta_ptrs)), "distribute_node_with_affinity")

    def distribute_async(self, offset, out_dev_ptr, meta_dev_ptrs, stream=0, slot=1):
        """Device-resident form: every pointer is a device address.  Distributions that may overlap on different
        streams must use different scratch slots (1..7; 0 belongs to distribute_node_with_affinity)."""
        arr = (C.c_uint64 * len(meta_dev_ptrs))(*[int(p) for p in meta_dev_ptrs])
        check(lib().coala_nd_distribute_async_slot(self._h, int(offset), int(out_dev_ptr), arr, len(meta_dev_ptrs), int(slot), stream),
              "distribute_async")

    def get_num_colors(self):
        return int(lib().coala_nd_num_colors(self._h))

    def get_color_buffer_ptr(self):
        return int(lib().coala_nd_color_host_ptr(self._h))

    def set_meta_len(self, n):
        check(lib().coala_nd_set_meta_len(self._h, int(n)), "set_meta_len")

    def last_bucket_len(self, n_domains):
        out = (C.c_int * n_domains)()
        lib().coala_nd_last_bucket_len(self._h, out)
        return list(out)

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().coala_nd_destroy(self._h)
                self._h = None
        except Exception:
            pass


class _CacheBase:
    _kind = ISOLATED

    def __init__(self, controllers, node_distributor, global_rank, n_gpus, cache_size, sim_buf_ptr, kind=None, local_slot=-1):
        self._h = None
        self._ctrl = controllers
        self._nd = node_distributor  # borrowed colour table: the distributor must outlive the cache (M/ssd_gnn_cache.cuh:91)
        self.num_gpus = int(n_gpus)
        self.local_rank = controllers.cudaDevice
        self.kind = self._kind if kind is None else kind
        nd_h = node_distributor._h if node_distributor is not None else None
        slot = int(local_slot) if int(local_slot) >= 0 else self._slot()
        self._h = check_handle(lib().coala_cache_create(self.kind, controllers.cudaDevice, controllers.dim, nd_h, int(global_rank),
                                                        slot, self.num_gpus, int(cache_size), int(sim_buf_ptr)),
                               type(self).__name__)

    def _slot(self):
        return self.local_rank

    def get_cache_data(self, dst_ptr, count=None):
        if count is None:
            check(lib().coala_cache_get_cache_data(self._h, int(dst_ptr)), "get_cache_data")
        else:
            check(lib().coala_cache_get_cache_data_n(self._h, int(dst_ptr), int(count)), "get_cache_data")

    def get_cache_data_async(self, dst_dev_ptr, count, stream=0):
        """device-to-device copy of the colour counters on `stream` (they never visit the host)"""
        check(lib().coala_cache_get_cache_data_async(self._h, int(dst_dev_ptr), int(count), stream), "get_cache_data_async")

    def color_counters_dev_ptr(self):
        return int(lib().coala_cache_color_counters_dev_ptr(self._h))

    def print_stats(self):
        check(lib().coala_cache_print_stats(self._h), "print_stats")

    def stats(self, reset=False):
        h, m, b = C.c_uint64(), C.c_uint64(), C.c_uint64()
        check(lib().coala_cache_stats(self._h, C.byref(h), C.byref(m), C.byref(b), int(reset)), "stats")
        return h.value, m.value, b.value

    def num_sets(self):
        return int(lib().coala_cache_num_sets(self._h))

    def set_num_nodes(self, n):
        """Row count of the host feature table: ids >= n are refused on the device instead of dereferenced."""
        check(lib().coala_cache_set_num_nodes(self._h, int(n)), "set_num_nodes")

    def refused_ids(self, reset=True):
        """Ids refused (out of range) by launches since the last reset; the async gather cannot raise."""
        n = C.c_uint64()
        check(lib().coala_cache_refused_ids(self._h, C.byref(n), int(reset)), "refused_ids")
        return n.value

    def export_state(self):
        import numpy as np
        n = self.num_sets()
        keys = np.zeros(n * 32, dtype=np.uint64)
        cmeta = np.zeros(n * 32, dtype=np.uint32)
        scnt = np.zeros(n, dtype=np.uint32)
        check(lib().coala_cache_export_state(self._h, keys.ctypes.data, cmeta.ctypes.data, scnt.ctypes.data), "export_state")
        return keys, cmeta, scnt

    def read_feature_async(self, out_ptr, idx_ptr, n, n_dev_ptr=0, stream=0):
        check(lib().coala_cache_read_feature_async(self._h, int(out_ptr), int(idx_ptr), int(n), int(n_dev_ptr), stream),
              "read_feature_async")

    def set_copy_engine(self, engine):
        """0 = rows through registers (16-byte vector loads), 1 = bulk asynchronous copies through shared memory."""
        check(lib().coala_cache_set_copy_engine(self._h, int(engine)), "set_copy_engine")

    def read_feature_serial(self, out_ptr, idx_ptr, n):
        check(lib().coala_cache_read_feature_serial(self._h, int(out_ptr), int(idx_ptr), int(n)), "read_feature_serial")

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().coala_cache_destroy(self._h)
                self._h = None
        except Exception:
            pass


class Isolated_Cache(_CacheBase):
    """M/ssd_gnn_cache.cuh:205-372. `isolated` backend: private cache (set = id % S).  For the
    `nccl` backend construct with kind=NCCL_SHARD (set = (id/G) % S, M/isolated_cache.h:190-195)."""
    _kind = ISOLATED

    def _slot(self):
        return self.local_rank

    def read_feature(self, out_ptr, idx_ptr, n):
        check(lib().coala_cache_read_feature(self._h, int(out_ptr), int(idx_ptr), int(n)), "read_feature")

    def split_node_list(self, idx_ptr, n, node_ptr, map_ptr, counter_ptr, local_size, max_sample):
        import torch
        check(lib().coala_split_node_list(int(idx_ptr), int(n), int(node_ptr), int(map_ptr), int(counter_ptr), int(local_size),
                                          int(max_sample), None), "split_node_list")
        torch.cuda.synchronize()  # (M/ssd_gnn_cache.cuh:294)

    def nccl_get_feature(self, idx_ptrs, out_ptrs, sizes, local_size, max_sample):
        """Owner-side lookups of the ids received from each peer (M/ssd_gnn_cache.cuh:297-325)."""
        import torch
        for i in range(local_size):
            if int(sizes[i]) > 0:
                self.read_feature_async(out_ptrs[i], idx_ptrs[i], int(sizes[i]))
        torch.cuda.synchronize()

    def map_feat_data(self, out_ptr, src_ptrs, map_ptr, sizes, local_size, max_sample):
        import torch
        dim = self._ctrl.dim
        for i in range(local_size):
            if int(sizes[i]) > 0:
                check(lib().coala_map_feat_data(int(out_ptr), int(src_ptrs[i]), int(map_ptr) + 8 * i * int(max_sample),
                                                int(sizes[i]), dim, None), "map_feat_data")
        torch.cuda.synchronize()


class SSD_GNN_NVSHMEM_Cache(_CacheBase):
    """M/ssd_gnn_cache.cuh:58-200. The partitioned multi-GPU cache; peers are reached with P2P loads
    over NVLink instead of NVSHMEM puts, so send_requests only records the request list and
    read_feature runs the single fused gather kernel."""
    _kind = PARTITIONED

    def __init__(self, controllers, node_distributor, global_rank, n_gpus, cache_size, sim_buf_ptr, local_slot=-1):
        super().__init__(controllers, node_distributor, global_rank, n_gpus, cache_size, sim_buf_ptr, local_slot=local_slot)
        self._pending = None
        self._peers = []

    def _slot(self):
        return NVSHMEM_Manager.current_rank if NVSHMEM_Manager.current_rank is not None else self.local_rank

    def ipc_handle(self):
        buf = C.create_string_buffer(64)
        check(lib().coala_cache_ipc_handle(self._h, buf), "ipc_handle")
        return buf.raw

    def export_fd(self):
        fd = lib().coala_cache_export_fd(self._h)
        if fd < 0:
            raise CoalaError("export_fd: " + lib().coala_last_error().decode())
        return fd

    def attach_fd(self, peer, fd):
        check(lib().coala_cache_attach_fd(self._h, int(peer), int(fd)), "attach_fd")

    def attach_ipc(self, peer, handle):
        check(lib().coala_cache_attach_ipc(self._h, int(peer), handle), "attach_ipc")

    def is_fd_shared(self):
        return bool(lib().coala_cache_is_fd_shared(self._h))

    def connect(self, group):
        connect_partitioned(self, self.num_gpus, group)

    def attach_local(self, peer_rank, peer_cache):
        self._peers.append(peer_cache)
        check(lib().coala_cache_attach_local(self._h, int(peer_rank), peer_cache._h), "attach_local")

    def send_requests(self, idx_ptr, n, req_ptr, max_index):
        self._pending = (int(idx_ptr), int(n))

    def read_feature(self, out_ptr, req_ptr, max_index):
        if self._pending is None:
            raise CoalaError("read_feature called before send_requests")
        idx_ptr, n = self._pending
        self._pending = None
        check(lib().coala_cache_read_feature(self._h, int(out_ptr), idx_ptr, n), "read_feature")


class NcclExchange:
    """The NCCL all-to-all variant as one stream-ordered call (csrc/nccl_backend.cu; replaces the host choreography of
    S/COALA_GNN_Manager.py:143-211).  `cache` is this rank's NCCL_SHARD cache (either binding); `group` a process group
    of the cache domain used once, to ship the 128-byte NCCL id from the domain's first rank."""

    def __init__(self, cache, device, group, n_ranks, rank, max_sample, feat_dim):
        self._h = None
        self._cache = cache  # the exchange looks rows up in this shard: keep it alive
        uid = C.create_string_buffer(128)
        if n_ranks > 1:
            import torch.distributed as dist
            box = [None]
            if rank == 0:
                check(lib().coala_nccl_unique_id(uid), "nccl unique id")
                box[0] = uid.raw
            dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0), group=group)
            uid = C.create_string_buffer(box[0], 128)
        handle = cache._h if hasattr(cache, "_h") else cache.handle()
        self._h = check_handle(lib().coala_nccl_create(handle, int(device), uid, int(n_ranks), int(rank), int(max_sample), int(feat_dim)),
                               "NcclExchange")

    def fetch_async(self, out_ptr, idx_ptr, n_upper, n_dev_ptr=0, stream=0):
        check(lib().coala_nccl_fetch_async(self._h, int(out_ptr), int(idx_ptr), int(n_upper), int(n_dev_ptr), stream), "nccl fetch")

    def counters(self):
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        check(lib().coala_nccl_counters(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return {"batches": a.value, "rows_served": b.value, "rows_received": c.value}

    @staticmethod
    def version():
        return int(lib().coala_nccl_version())

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().coala_nccl_destroy(self._h)
                self._h = None
        except Exception:
            pass


FD_EXCHANGE_TIMEOUT_S = 120.0


def connect_partitioned(cache, num_gpus, group):
    """Collective over the cache domain: every rank maps every peer's shard.  Shards are shared as POSIX file
    descriptors (VMM allocations) passed over unix sockets (SCM_RIGHTS); with COALA_CACHE_LEGACY_IPC=1 as 64-byte
    CUDA IPC handles.  `cache` is any binding of the C ABI with is_fd_shared / export_fd / attach_fd / ipc_handle /
    attach_ipc (the ctypes classes of this file or the compiled COALA_GNN_Pybind module).

    A received descriptor maps the peer's shard read-write, so the exchange is closed to other local processes: each
    rank listens on a socket inside its own mode-0700 temporary directory (only this uid can reach it), learns its
    peers' paths and pids over the process group, accepts only connections whose SO_PEERCRED (pid, uid) belongs to a
    peer of the domain, and every blocking call carries a timeout that names what did not arrive."""
    import os
    import shutil
    import socket
    import struct
    import tempfile
    import torch.distributed as dist
    me = dist.get_rank(group)
    if not cache.is_fd_shared():
        handles = [None] * num_gpus
        dist.all_gather_object(handles, cache.ipc_handle(), group=group)
        for r, h in enumerate(handles):
            if r != me:
                cache.attach_ipc(r, h)
        dist.barrier(group=group)
        return
    my_fd = cache.export_fd()
    tmpdir = tempfile.mkdtemp(prefix="coala_b200_fd_")  # mode 0700
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    conns, served = {}, []
    try:
        path = os.path.join(tmpdir, "s")
        srv.bind(path)
        srv.listen(num_gpus)
        srv.settimeout(FD_EXCHANGE_TIMEOUT_S)
        peers = [None] * num_gpus
        dist.all_gather_object(peers, (path, os.getpid(), os.getuid()), group=group)  # also: every rank is listening
        allowed = {(pid, uid) for r, (_p, pid, uid) in enumerate(peers) if r != me}
        for r in range(num_gpus):
            if r != me:
                c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                c.settimeout(FD_EXCHANGE_TIMEOUT_S)
                c.connect(peers[r][0])  # completes from the listen backlog, no accept needed yet
                c.sendall(bytes([me]))
                conns[r] = c
        n_served = 0
        while n_served < num_gpus - 1:
            try:
                c, _addr = srv.accept()
            except socket.timeout:
                raise CoalaError("shard exchange: rank %d served %d of %d peers within %.0f s" % (me, n_served, num_gpus - 1, FD_EXCHANGE_TIMEOUT_S))
            c.settimeout(FD_EXCHANGE_TIMEOUT_S)
            pid, uid, _gid = struct.unpack("3i", c.getsockopt(socket.SOL_SOCKET, socket.SO_PEERCRED, struct.calcsize("3i")))
            if (pid, uid) not in allowed:
                c.close()  # not a rank of this cache domain: no descriptor, and it does not use up a slot
                continue
            allowed.discard((pid, uid))
            c.recv(1)
            socket.send_fds(c, [b"f"], [my_fd])
            served.append(c)
            n_served += 1
        for r, c in conns.items():
            try:
                _msg, fds, _flags, _addr = socket.recv_fds(c, 1, 1)
            except socket.timeout:
                raise CoalaError("shard exchange: rank %d received nothing from peer %d within %.0f s" % (me, r, FD_EXCHANGE_TIMEOUT_S))
            if not fds:
                raise CoalaError("shard exchange: peer %d closed the connection without sending its shard" % r)
            try:
                cache.attach_fd(r, fds[0])
            finally:
                os.close(fds[0])
        dist.barrier(group=group)  # every rank has imported every shard
    finally:
        for c in list(conns.values()) + served + [srv]:
            c.close()
        os.close(my_fd)
        shutil.rmtree(tmpdir, ignore_errors=True)


class NVSHMEM_Manager:
    """M/nvshmem_manager.cuh:9-56. There is no NVSHMEM: the "symmetric heap" is plain device memory
    (peers never write into it — rows are pulled by the requester), and the object only remembers
    the cache domain's process group for the IPC handle exchange."""
    current_group = None
    current_rank = None

    def __init__(self, local_comm_ptr, local_rank):
        self._bufs = {}
        group = getattr(local_comm_ptr, "group", None)
        NVSHMEM_Manager.current_group = group
        NVSHMEM_Manager.current_rank = int(local_rank)

    def allocate(self, size):
        import torch
        t = torch.empty(int(size), dtype=torch.uint8, device="cuda")
        self._bufs[t.data_ptr()] = t
        return t.data_ptr()

    def free(self, ptr):
        self._bufs.pop(int(ptr), None)

    def finalize(self):
        self._bufs.clear()
        NVSHMEM_Manager.current_group = None
        NVSHMEM_Manager.current_rank = None


class Graph_Coloring:
    """M/graph_coloring.h:15-68 (bound at M/COALA_GNN_Pybind.cu:65-77)."""

    def __init__(self, num_nodes):
        self._h = check_handle(lib().coala_gc_create(int(num_nodes)), "Graph_Coloring")

    def set_adj_csc(self, indptr_ptr, indices_ptr):
        lib().coala_gc_set_adj_csc(self._h, int(indptr_ptr), int(indices_ptr))

    def set_color_buffer(self, p):
        lib().coala_gc_set_color_buffer(self._h, int(p))

    def set_topk_color_buffer(self, p):
        lib().coala_gc_set_topk_color_buffer(self._h, int(p))

    def set_topk_affinity_buffer(self, p):
        lib().coala_gc_set_topk_affinity_buffer(self._h, int(p))

    def cpu_color_graph(self):
        lib().coala_gc_cpu_color_graph(self._h)

    def cpu_color_graph_optimized(self, train_ptr, n_train):
        lib().coala_gc_cpu_color_graph_optimized(self._h, int(train_ptr), int(n_train))

    def cpu_count_nearest_color(self):
        lib().coala_gc_cpu_count_nearest_color(self._h)

    def cpu_count_nearest_color_less_memory(self):
        lib().coala_gc_cpu_count_nearest_color_less_memory(self._h)

    def cpu_calculate_color_affinity(self):
        lib().coala_gc_cpu_calculate_color_affinity(self._h)

    # GPU mode (not in the reference): same seeds / colour ids, level-synchronous traversal and sort-based
    # affinity on the device; see include/coala_b200.h
    def gpu_color_graph_optimized(self, train_ptr, n_train, device=0):
        check(lib().coala_gc_gpu_color_graph_optimized(self._h, int(train_ptr), int(n_train), int(device)), "gpu_color_graph_optimized")

    def gpu_calculate_color_affinity(self, device=0):
        check(lib().coala_gc_gpu_calculate_color_affinity(self._h, int(device)), "gpu_calculate_color_affinity")

    def get_hops(self, num_nodes):
        import numpy as np
        hop = np.zeros(int(num_nodes), dtype=np.uint16)
        check(lib().coala_gc_get_hops(self._h, hop.ctypes.data), "get_hops")
        return hop

    def get_num_color(self):
        return int(lib().coala_gc_get_num_color(self._h))

    def get_num_color_node(self):
        return int(lib().coala_gc_get_num_color_node(self._h))

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().coala_gc_destroy(self._h)
                self._h = None
        except Exception:
            pass
