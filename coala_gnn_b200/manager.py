"""Mirror of S/COALA_GNN_Manager.py: backend dispatch for the cached feature gather.

Backends (``--cache_backend``):
  nvshmem | p2p : partitioned cache, one fused requester-pull kernel over P2P-mapped shards
                  (replaces send_requests/read_feature + NVSHMEM puts + 4 barriers, :124-135)
  isolated      : private cache per GPU (:137-141)
  nccl          : id%G shards served by their owner, ids/rows exchanged with all-to-all-v
                  (:143-211; exact counts instead of max_sample padding, one grouped exchange
                  instead of O(G^2) serialized send/recv)
Output rows land in a persistent double buffer (no per-batch torch.zeros of 151 MB, :138,146).
"""
import os
import time

import torch
import torch.distributed as dist

from . import binding
from .pybind import connect_partitioned, NcclExchange


class COALA_GNN_Manager(object):
    def __init__(self, node_distributor, num_ssds, page_size, num_elems, ssd_read_offset, cache_size, batch_size, fan_out,
                 dim, MPI_comm_manager, device, cache_backend="nvshmem", sim_buf=None):
        self.node_distributor = node_distributor
        self.device = device
        self.cache_backend = "nvshmem" if cache_backend == "p2p" else cache_backend
        self.MPI_comm_manager = MPI_comm_manager
        self.dim = dim
        self.sim_buf = sim_buf
        self.is_simulation = sim_buf is not None
        self.aggregation_timer = 0.0
        comm = MPI_comm_manager

        B = binding.mod()  # ctypes classes, or the compiled COALA_GNN_Pybind module (COALA_PYBIND_NATIVE=1): same names
        self.SSD_Controllers = B.SSD_GNN_SSD_Controllers(num_ssds, page_size, num_elems, ssd_read_offset, comm.device_index, dim,
                                                         self.is_simulation)
        self.max_sample_size = batch_size
        for i in fan_out:
            self.max_sample_size *= (int(i) + 1)  # (:79-81)
        sim_ptr = self.sim_buf.data_ptr() if self.is_simulation else 0
        nd = self.node_distributor.distribute_manager

        if self.cache_backend == "nvshmem":
            self.nvshmem_manager = B.NVSHMEM_Manager(0 if binding.is_native() else comm.local_comm, comm.local_rank)
            self.COALA_GNN_Cache = B.SSD_GNN_NVSHMEM_Cache(self.SSD_Controllers, nd, comm.global_rank, comm.local_size, cache_size, sim_ptr,
                                                           local_slot=comm.local_rank)
            if comm.local_size > 1:  # every rank of the domain maps every peer's shard (collective)
                connect_partitioned(self.COALA_GNN_Cache, comm.local_size, comm.local_gloo_gather)
        elif self.cache_backend == "isolated":
            self.COALA_GNN_Cache = B.Isolated_Cache(self.SSD_Controllers, nd, comm.global_rank, comm.local_size, cache_size, sim_ptr)
        elif self.cache_backend == "nccl":
            self.COALA_GNN_Cache = B.Isolated_Cache(self.SSD_Controllers, nd, comm.global_rank, comm.local_size, cache_size, sim_ptr,
                                                    kind=B.NCCL_SHARD)
            G, ms = comm.local_size, self.max_sample_size
            # the exchange as one stream-ordered native call (ncclSend/ncclRecv groups); COALA_NCCL_LEGACY=1 keeps the
            # reference-shaped host choreography over torch.distributed (split_node_list / nccl_get_feature / map_feat_data)
            self.nccl_legacy = os.environ.get("COALA_NCCL_LEGACY", "0") == "1"
            self.nccl = None
            if self.nccl_legacy:
                self.nccl_node_tensor = torch.zeros(ms * G, dtype=torch.int64, device=device)
                self.nccl_map_tensor = torch.zeros(ms * G, dtype=torch.int64, device=device)
                self.nccl_counter_tensor = torch.zeros(G, dtype=torch.int64, device=device)
            else:
                self.nccl = NcclExchange(self.COALA_GNN_Cache, comm.device_index, comm.local_gloo_gather, G, comm.local_rank, ms, dim)
        else:
            raise ValueError(f"Unsupported cache backend: {cache_backend}")
        # Device-side id bounds check: the host table's row count is known here (the reference dereferences any id it is
        # given).  Without it an id beyond the table - but below 2^40 - would be read from whatever follows the table.
        if self.is_simulation and getattr(self.sim_buf, "dim", lambda: 0)() == 2:
            self.COALA_GNN_Cache.set_num_nodes(int(self.sim_buf.shape[0]))
        self._out = [torch.empty((self.max_sample_size, dim), dtype=torch.float32, device=device) for _ in range(2)]
        self._flip = 0

    def out_buffers(self, n):
        """Persistent output row buffers (one per in-flight batch of the prefetch engine)."""
        while len(self._out) < n:
            self._out.append(torch.empty((self.max_sample_size, self.dim), dtype=torch.float32, device=self.device))
        return self._out[:n]

    def _next_out(self, n):
        self._flip = (self._flip + 1) % 2
        return self._out[self._flip][:n]

    def gather_async(self, out_ptr, idx_ptr, n_upper, n_dev_ptr, stream):
        """Enqueue the cached gather of a device-resident id list (length possibly known only on the device) on `stream`:
        the fused requester-pull kernel (partitioned / isolated), or the NCCL exchange (which blocks once for its counts)."""
        if self.cache_backend == "nccl" and self.nccl is not None:
            self.nccl.fetch_async(out_ptr, idx_ptr, n_upper, n_dev_ptr, stream)
        else:
            self.COALA_GNN_Cache.read_feature_async(out_ptr, idx_ptr, n_upper, n_dev_ptr, stream)

    def native_handles(self):
        """(coala_cache*, coala_nccl* | None) as integers, for the native batch engine (either binding)."""
        c = self.COALA_GNN_Cache
        ch = c._h if hasattr(c, "_h") else c.handle()
        nh = None
        if self.cache_backend == "nccl" and self.nccl is not None:
            nh = self.nccl._h
        return ch, nh

    @property
    def engine_capable(self):
        """Can the prefetch engine drive this backend (everything enqueued on a stream by gather_async)?"""
        return self.cache_backend in ("nvshmem", "isolated") or (self.cache_backend == "nccl" and self.nccl is not None)

    def fetch_feature_async(self, index, n_upper, n_dev_ptr, stream):
        """Enqueue the gather for a device-resident id list whose length is only known on the device.
        Returns the full-capacity output buffer."""
        out = self._next_out(self.max_sample_size)
        self.gather_async(out.data_ptr(), index, n_upper, n_dev_ptr, stream)
        return out

    def fetch_feature(self, batch):
        index = batch[0].to(self.device)
        index_size = len(index)
        fetch_start = time.time()
        if self.cache_backend in ("nvshmem", "isolated"):
            return_torch = self._next_out(index_size)
            if self.cache_backend == "nvshmem":
                self.COALA_GNN_Cache.send_requests(index.data_ptr(), index_size, 0, self.max_sample_size)
                self.COALA_GNN_Cache.read_feature(return_torch.data_ptr(), 0, self.max_sample_size)
            else:
                self.COALA_GNN_Cache.read_feature(return_torch.data_ptr(), index.data_ptr(), index_size)
        elif self.nccl is not None:
            return_torch = self._next_out(index_size)
            self.nccl.fetch_async(return_torch.data_ptr(), index.data_ptr(), index_size, 0, torch.cuda.current_stream().cuda_stream)
            torch.cuda.current_stream().synchronize()  # synchronous like the reference's fetch
        else:
            return_torch = self._fetch_nccl(index, index_size)
        self.aggregation_timer += (time.time() - fetch_start)
        return (*batch, return_torch)

    def _fetch_nccl(self, index, n):
        comm = self.MPI_comm_manager
        G, ms, dim = comm.local_size, self.max_sample_size, self.dim
        cache = self.COALA_GNN_Cache
        return_torch = self._next_out(n)
        cache.split_node_list(index.data_ptr(), n, self.nccl_node_tensor.data_ptr(), self.nccl_map_tensor.data_ptr(),
                              self.nccl_counter_tensor.data_ptr(), G, ms)
        send_counts = self.nccl_counter_tensor.tolist()
        if G == 1:
            recv_counts = send_counts
            recv_ids = self.nccl_node_tensor[:send_counts[0]]
        else:
            group = comm.nccl_cache_gather
            gathered = torch.empty(G, dtype=torch.int64, device=self.device)
            dist.all_to_all_single(gathered, self.nccl_counter_tensor, group=group)
            recv_counts = gathered.tolist()
            send_ids = torch.cat([self.nccl_node_tensor[g * ms: g * ms + send_counts[g]] for g in range(G)])
            recv_ids = torch.empty(sum(recv_counts), dtype=torch.int64, device=self.device)
            dist.all_to_all_single(recv_ids, send_ids, output_split_sizes=recv_counts, input_split_sizes=send_counts, group=group)
        # owner side: look the received ids up in the local shard (the reference passes the SEND counts here,
        # S/COALA_GNN_Manager.py:184-189 — the received counts are the right ones)
        served = torch.empty((sum(recv_counts), dim), dtype=torch.float32, device=self.device)
        if sum(recv_counts) > 0:
            cache.read_feature(served.data_ptr(), recv_ids.data_ptr(), sum(recv_counts))
        if G == 1:
            rows = served
        else:
            rows = torch.empty((n, dim), dtype=torch.float32, device=self.device)
            dist.all_to_all_single(rows, served, output_split_sizes=send_counts, input_split_sizes=recv_counts, group=group)
        off = 0
        ptrs = []
        for g in range(G):
            ptrs.append(rows.data_ptr() + off * dim * 4)
            off += send_counts[g]
        cache.map_feat_data(return_torch.data_ptr(), ptrs, self.nccl_map_tensor.data_ptr(), send_counts, G, ms)
        return return_torch

    def get_cache_data(self, ptr):
        self.COALA_GNN_Cache.get_cache_data(ptr)

    def check_refused_ids(self):
        """The asynchronous gather zero-fills the row of an id outside the feature table and counts it (the synchronous
        calls fail on the spot).  Surface that count: raises if any id was refused since the last check."""
        n = int(self.COALA_GNN_Cache.refused_ids(True))
        if n:
            raise RuntimeError("feature gather refused %d node id(s) outside the feature table (rows were zero-filled)" % n)

    def print_stats(self):
        self.COALA_GNN_Cache.print_stats()
        self.check_refused_ids()

    def get_aggregate_time(self):
        return self.aggregation_timer

    def __del__(self):
        if getattr(self, "cache_backend", None) == "nvshmem" and hasattr(self, "nvshmem_manager"):
            self.nvshmem_manager.finalize()
