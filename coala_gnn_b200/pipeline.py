"""Two-stream execution of the mini-batch path (B200-first addition behind the drop-in loader).

The reference serialises, per batch: CPU distributor (thread) -> gloo broadcast -> seeds H2D ->
DGL sampling with host syncs -> gather (S/COALA_GNN_DataLoader.py:149-167).  Here a mini-batch is
two enqueued stages on two CUDA streams and nothing returns to the host between them:

  stream `sample` : [seed distribution ->] multi-layer UVA sampling + to_block   (host-link LATENCY bound)
  stream `gather` : cached feature gather of the sampler's input-node list        (host-link / NVLink / HBM BANDWIDTH bound)

Batch t+1 is launched while batch t is being gathered, so the sampler's dependent host round
trips hide behind the gather's bulk traffic (they load different resources of the same links).
Two slots (sampler buffers, seed buffer, output rows) alternate; events order slot reuse:
a slot is rewritten only after its previous gather finished and the consumer moved on.

`BatchEngine` is shared by the drop-in iterator (`COALA_GNN_DataLoader`, seeds come from the host
scheduler) and by `DevicePipeline` (seeds come from the GPU distributor, no host data at all).
"""
import ctypes as C

import torch
import torch.distributed as dist

from ._lib import lib, check, check_handle
from .pybind import tensor_from_ptr
from .sampler import MultiLayerNeighborSampler, batch_key  # noqa: F401  (batch_key re-exported for callers)


class _Slot(object):
    """Python face of one native engine slot: its sampler object, output rows, stream and seed staging buffer."""

    def __init__(self, k, fanouts, graph, batch_size, out, device, hbm_indptr=None, indptr_from=None, hbm_graph=None):
        self.k = k
        self.sampler = MultiLayerNeighborSampler(fanouts, hbm_indptr=hbm_indptr, indptr_from=indptr_from, hbm_graph=hbm_graph)
        self.sampler._ensure(graph, batch_size)
        self.sampler._last_g = graph
        self.out = out
        self.device = device
        self.stream = None       # torch view of the slot's native stream (distribution / refresh are enqueued there)
        self.seeds = None        # device int64[batch_size] staging buffer (host seed lists land here)
        self.batch = -1
        self._last = (0, 0)      # (device pointer, count) of the seed list the last sample of this slot was drawn from
        self.meta_used = None    # DevicePipeline: the domain histograms its distribution read

    @property
    def last_seeds(self):
        return tensor_from_ptr(self._last[0], (self._last[1],), torch.int64, self.device)


class BatchEngine(object):
    """launch*(t, ...) enqueues batch t; collect(t) / wait(t) make the current stream wait for it.  `depth` batches can be
    in flight behind the one being consumed (depth + 1 slots).  Each slot samples on its own stream: a sample is a dependent
    chain of small kernels and host round trips that leaves the GPU and the host link mostly idle, so several of them overlap
    almost perfectly; gathers run in batch order on one stream.  Measured at igb-small N=1 (sampler-bound): depth 1 / 2 / 3 =
    7.7 K / 11.3 K / 13.4 K batches/s.  The enqueue sequence itself is native (csrc/engine.cu): one C call per batch."""

    def __init__(self, graph, sampler, manager, batch_size, device, record_gather_events=False, depth=None):
        self.g, self.mgr, self.device = graph, manager, device
        self._seed, self.rank = int(sampler.seed), int(sampler.rank)
        self.fanouts = list(sampler.fanouts)
        self.batch_size = int(batch_size)
        import os
        if depth is None:
            depth = int(os.environ.get("COALA_PREFETCH_DEPTH", "3"))
        self.depth = max(1, int(depth))
        self.nslots = self.depth + 1
        dev_index = torch.device(device).index
        if dev_index is None:
            dev_index = torch.cuda.current_device()
        self._lib = lib()
        self._dev_index = dev_index
        self._h = check_handle(self._lib.coala_engine_create(dev_index, self.nslots, self.batch_size, self._seed & ((1 << 64) - 1), self.rank,
                                                             1 if record_gather_events else 0), "engine")
        outs = manager.out_buffers(self.nslots)
        self.slots = []
        for k in range(self.nslots):  # slot 0 stages indptr in HBM (unless switched off), the others use its copy
            self.slots.append(_Slot(k, self.fanouts, graph, batch_size, outs[k], device, getattr(sampler, "hbm_indptr", None),
                                    self.slots[0].sampler if k else None, getattr(sampler, "hbm_graph", None)))
        for s in self.slots:
            check(self._lib.coala_engine_bind_slot(self._h, s.k, s.sampler._h, s.out.data_ptr()), "engine slot")
            s.stream = torch.cuda.ExternalStream(int(self._lib.coala_engine_slot_stream(self._h, s.k)), device=device)
            s.seeds = tensor_from_ptr(int(self._lib.coala_engine_slot_seeds_dev(self._h, s.k)), (self.batch_size,), torch.int64, device, owner=self)
        cache_h, nccl_h = manager.native_handles()
        check(self._lib.coala_engine_bind_gather(self._h, cache_h, nccl_h, manager.max_sample_size), "engine gather")
        self.s_gather = torch.cuda.ExternalStream(int(self._lib.coala_engine_gather_stream(self._h)), device=device)
        self.record = record_gather_events
        self.launched = 0
        self._counts = (C.c_int64 * (2 * len(self.fanouts) + 1))()
        self._NOCP = C.c_void_p((1 << 64) - 1)

    def _cur_stream(self):
        """raw handle of torch's current stream on this device (torch.cuda.current_stream() costs ~3 us of Python per call)"""
        return torch._C._cuda_getCurrentRawStream(self._dev_index)

    @property
    def seed(self):
        return self._seed

    @seed.setter
    def seed(self, v):
        self._seed = int(v)
        check(self._lib.coala_engine_set_seed(self._h, self._seed & ((1 << 64) - 1)))

    def consumer_checkpoint(self):
        """Everything the consumer enqueued so far on the current stream precedes any slot reuse."""
        check(self._lib.coala_engine_checkpoint(self._h, self._cur_stream()), "engine checkpoint")

    def launch_host(self, t, seeds_host_ptr, n):
        """Batch t from a host seed list (the drop-in iterator): pinned staging + H2D + sample + gather, one native call."""
        check(self._lib.coala_engine_launch_host(self._h, t, seeds_host_ptr, n, self._NOCP), "engine launch")
        slot = self.slots[t % self.nslots]
        slot.batch, slot._last = t, (slot.seeds.data_ptr(), n)
        self.launched = t + 1
        return slot

    def launch(self, t, seed_fn):
        """Batch t from a device-resident seed list: seed_fn(slot, slot.stream) enqueues whatever produces the seeds on the slot's
        stream (the GPU distributor, a histogram refresh) and returns (device pointer, count) or an int64 device tensor."""
        k = self._lib.coala_engine_begin(self._h, t, self._NOCP)
        if k < 0:
            check(-1, "engine begin")
        slot = self.slots[k]
        seeds = seed_fn(slot, slot.stream)
        ptr, n = seeds if isinstance(seeds, tuple) else (seeds.data_ptr(), seeds.numel())
        check(self._lib.coala_engine_sample_gather(self._h, t, ptr, n), "engine sample/gather")
        slot.batch, slot._last = t, (ptr, n)
        self.launched = t + 1
        return slot

    def slot_of(self, t):
        slot = self.slots[t % self.nslots]
        assert slot.batch == t, "batch %d is not in flight" % t
        return slot

    def wait(self, t):
        """The current stream waits for the rows of batch t (no host wait)."""
        check(self._lib.coala_engine_collect(self._h, t, None, self._cur_stream()), "engine wait")
        return self.slots[t % self.nslots]

    def collect(self, t):
        """The current stream waits for blocks and rows of batch t; the host waits for the block sizes. Returns (slot, counts)."""
        check(self._lib.coala_engine_collect(self._h, t, self._counts, self._cur_stream()), "engine collect")
        return self.slots[t % self.nslots], self._counts

    def _times(self, reset=False):
        g, s, ng, ns = C.c_double(), C.c_double(), C.c_uint64(), C.c_uint64()
        check(self._lib.coala_engine_times(self._h, C.byref(g), C.byref(ng), C.byref(s), C.byref(ns), 1, 1 if reset else 0), "engine times")
        return g.value, ng.value, s.value, ns.value

    def reset_times(self):
        torch.cuda.synchronize()
        self._times(reset=True)

    def gather_ms(self):
        """[mean CUDA-event time of the gather launches since reset_times()] (empty when nothing was recorded)."""
        g, ng, _, _ = self._times()
        return [g] if ng else []

    def sample_ms(self):
        _, _, s, ns = self._times()
        return [s] if ns else []

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.coala_engine_destroy(self._h)
                self._h = None
        except Exception:
            pass


class DevicePipeline(object):
    """Device-resident form of the path for D x L cache domains: every rank runs the (deterministic) GPU
    distributor on the same inputs - the step's global seed slice and the D domain histograms - and takes
    its domain's bucket and its own slice of it, so there is no seed broadcast (S/Shared_Tensor.py:102-103).
    The cache colour histograms never leave the device: every `refresh_counter` batches each rank's counters
    are all-gathered over NCCL and summed per domain (the reference: gloo all_reduce inside the domain, then
    all_gather across the domain masters, on CPU tensors, S/Shared_Tensor.py:96-100).  step() returns the rows
    of batch t (valid on the current stream) while the next batches are already in flight."""

    def __init__(self, loader, record_gather_events=False, prefetch=True, seed=None, depth=None):
        self.loader = loader
        self.nd = loader.node_distributor
        self.comm = self.nd.comm_manager
        self.mgr = loader.COALA_GNN_Manager
        self.g = loader.g
        self.device = loader.device
        assert self.nd._dev_index is not None, "seed list must be device resident"
        self.B, self.L, self.D = self.nd.batch_size, self.comm.local_size, self.comm.num_master_process
        self.engine = BatchEngine(self.g, loader.sampler, self.mgr, self.B, self.device, record_gather_events, depth=depth if prefetch else 1)
        if seed is not None:
            self.engine.seed = int(seed)
        ns = self.engine.nslots
        if ns + 1 > 8 and self.nd.parsing_method == "node_color":
            raise ValueError("the distributor has 7 scratch slots for in-flight distributions; prefetch depth must be <= 6")
        self.domain_seeds = [torch.zeros(self.B * self.L, dtype=torch.int64, device=self.device) for _ in range(ns)]
        ml = self.nd.meta_len
        # ring of histogram sets: a refresh writes the next entry and later launches read it, so a distribution
        # still running on another slot's stream keeps the entry it was launched with (at most `ns` launches are
        # in flight, hence ns + 2 entries even with refresh_counter = 1)
        self.meta_ring = [torch.zeros((self.D, ml), dtype=torch.int32, device=self.device) for _ in range(ns + 2)]
        self.meta_local = [torch.zeros(ml, dtype=torch.int32, device=self.device) for _ in range(ns + 2)]
        self.meta_world = [torch.zeros((self.comm.global_size, ml), dtype=torch.int32, device=self.device) for _ in range(ns + 2)] \
            if self.D > 1 else None
        self.meta_cur = 0
        self._meta_ptrs = [[m[d].data_ptr() for d in range(self.D)] for m in self.meta_ring]
        self._dev_index_ptr = self.nd._dev_index.data_ptr()
        doms = self.comm.global_dist_local_rank_list if self.comm.global_size > 1 else [[0]]
        self._dom_ranks = [torch.tensor(r, dtype=torch.int64, device=self.device) for r in doms]
        self._dom_blocked = all(list(r) == list(range(d * self.L, (d + 1) * self.L)) for d, r in enumerate(doms))
        self.step_idx = 0
        self.refresh = loader.refresh_counter
        self.prefetch = prefetch
        if self.nd.parsing_method != "node_color":
            n_dist = 0
        else:
            n_dist = 1 if self.D == 1 else 3  # pass-through | score + tiled assign + (idle) serial assign
        self.kernels_per_step = n_dist + 3 * len(self.engine.fanouts) + 1  # + three kernels per sampling round, gather

    def steps_per_epoch(self):
        return len(self.nd.index_tensor) // self.nd.global_batch_size - 1

    def seed_offset(self, t):
        return (t % self.steps_per_epoch()) * self.nd.global_batch_size

    def _seed_fn(self, t):
        def fn(slot, stream):
            offset = self.seed_offset(t)
            if t % self.refresh == 0 and t > 0:
                with torch.cuda.stream(stream):  # torch ops (NCCL, reductions) of the refresh go to the slot's stream
                    self.refresh_meta(stream)
            slot.meta_used = self.meta_ring[self.meta_cur]
            if self.nd.parsing_method == "node_color":
                dom = self.domain_seeds[slot.k]
                self.nd.distribute_manager.distribute_async(offset, dom.data_ptr(), self._meta_ptrs[self.meta_cur], stream.cuda_stream, slot=1 + slot.k)
                return dom.data_ptr() + 8 * self.comm.local_rank * self.B, self.B
            lo = offset + self.comm.master_process_index * self.B * self.L + self.comm.local_rank * self.B  # (S/Training_node_distributor.py:41-47)
            return self._dev_index_ptr + 8 * lo, self.B
        return fn

    def refresh_meta(self, stream):
        """Colour histograms of every domain's cache -> the next ring entry, int32[D, meta_len] (on `stream`)."""
        cache = self.mgr.COALA_GNN_Cache
        nxt = (self.meta_cur + 1) % len(self.meta_ring)
        local, buf = self.meta_local[nxt], self.meta_ring[nxt]
        cache.get_cache_data_async(local.data_ptr(), self.nd.meta_len, stream.cuda_stream)
        if self.D == 1:
            if self.L > 1:
                dist.all_reduce(local, group=self.comm.nccl_cache_gather)  # runs on the current (= sample) stream
            buf[0].copy_(local)
        else:
            world = self.meta_world[nxt]
            dist.all_gather_into_tensor(world, local)  # every rank's counters; NCCL, ordered after `stream`
            if self._dom_blocked:
                torch.sum(world.view(self.D, self.L, -1), dim=1, dtype=torch.int32, out=buf)
            else:
                for d, ranks in enumerate(self._dom_ranks):
                    torch.sum(world.index_select(0, ranks), dim=0, dtype=torch.int32, out=buf[d])
        self.meta_cur = nxt

    @property
    def meta(self):
        """The histogram set new launches read (int32[D, meta_len]); D == 1 callers use meta[0]."""
        return self.meta_ring[self.meta_cur]

    def step(self):
        t = self.step_idx
        eng = self.engine
        eng.consumer_checkpoint()
        if eng.launched <= t:
            eng.launch(t, self._seed_fn(t))
        if self.prefetch:
            while eng.launched <= t + eng.depth:
                eng.launch(eng.launched, self._seed_fn(eng.launched))
        slot = eng.wait(t)
        self.step_idx += 1
        return slot.out

    def current_sampler(self):
        """Sampler object holding the blocks of the batch returned by the last step()."""
        return self.engine.slot_of(self.step_idx - 1).sampler

    def reset(self):  # epochs wrap inside _seed_fn; kept for API compatibility
        pass

    def gather_ms(self):
        return self.engine.gather_ms()

    def __del__(self):
        try:
            torch.cuda.synchronize()
        except Exception:
            pass
