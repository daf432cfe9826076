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
import torch
import torch.distributed as dist

from ._lib import lib, check
from .sampler import MultiLayerNeighborSampler, batch_key


class _Slot(object):
    def __init__(self, k, fanouts, batch_size, out, device, hbm_indptr=None):
        self.k = k
        self.sampler = MultiLayerNeighborSampler(fanouts, hbm_indptr=hbm_indptr)
        self.seeds = torch.zeros(batch_size, dtype=torch.int64, device=device)
        self.seeds_pin = torch.zeros(batch_size, dtype=torch.int64).pin_memory()
        self.out = out
        # sampling kernels are small and latency bound: high priority so that they take SM slots as gather CTAs retire
        self.stream = torch.cuda.Stream(device=device, priority=-1)
        self.ev_sampled = torch.cuda.Event()
        self.ev_gathered = torch.cuda.Event()
        self.t0 = torch.cuda.Event(enable_timing=True)
        self.t1 = torch.cuda.Event(enable_timing=True)
        self.used = False
        self.batch = -1
        self.last_seeds = None   # device tensor the last sample of this slot was drawn from
        self.meta_used = None    # DevicePipeline: the domain histograms its distribution read


class BatchEngine(object):
    """launch(t, seed_fn) enqueues batch t; wait(t) makes the current stream wait for its rows.
    `depth` batches can be in flight behind the one being consumed (depth + 1 slots).  Each slot samples on
    its own stream: a sample is a dependent chain of small kernels and host round trips that leaves the GPU
    and the host link mostly idle, so two of them overlap almost perfectly; gathers run in batch order on
    one stream.  Measured at igb-small N=1 (sampler-bound): depth 1 / 2 / 3 = 7.7 K / 11.3 K / 13.4 K batches/s."""

    def __init__(self, graph, sampler, manager, batch_size, device, record_gather_events=False, depth=None):
        self.g, self.mgr, self.device = graph, manager, device
        self.seed, self.rank = sampler.seed, sampler.rank
        self.fanouts = list(sampler.fanouts)
        import os
        if depth is None:
            depth = int(os.environ.get("COALA_PREFETCH_DEPTH", "3"))
        self.depth = max(1, int(depth))
        self.nslots = self.depth + 1
        self.s_gather = torch.cuda.Stream(device=device)
        self.ev_consumed = torch.cuda.Event()
        outs = manager.out_buffers(self.nslots)
        self.slots = [_Slot(k, self.fanouts, batch_size, outs[k], device, getattr(sampler, "hbm_indptr", None)) for k in range(self.nslots)]
        for s in self.slots:
            s.sampler._ensure(graph, batch_size)
        self.record = record_gather_events
        self.gather_times = []  # (start, end) event pairs of finished launches
        self.sample_times = []  # same for distribution + sampling (the other stream)
        self.launched = 0

    def consumer_checkpoint(self):
        """Everything the consumer enqueued so far on the current stream precedes any slot reuse."""
        self.ev_consumed.record(torch.cuda.current_stream())

    def launch(self, t, seed_fn):
        slot = self.slots[t % self.nslots]
        ss, sg = slot.stream, self.s_gather
        ss.wait_event(self.ev_consumed)
        if slot.used:
            ss.wait_event(slot.ev_gathered)  # the previous gather of this slot read the sampler's id list
        with torch.cuda.stream(ss):
            if self.record:
                s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s0.record(ss)
            seeds = seed_fn(slot, ss)
            slot.last_seeds = seeds
            slot.sampler.sample_async(self.g, seeds, batch_key(self.seed, t, self.rank), ss.cuda_stream)
            if self.record:
                s1.record(ss)
                self.sample_times.append((s0, s1))
            slot.ev_sampled.record(ss)
        sg.wait_event(slot.ev_sampled)
        sg.wait_event(self.ev_consumed)      # the rows of the batch that used this buffer before are no longer needed
        idx_ptr, n_ptr = slot.sampler.input_nodes_dev_ptrs()
        with torch.cuda.stream(sg):
            if self.record:
                if slot.used:
                    self.gather_times.append((slot.t0, slot.t1))
                    slot.t0, slot.t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                slot.t0.record(sg)
            self.mgr.gather_async(slot.out.data_ptr(), idx_ptr, self.mgr.max_sample_size, n_ptr, sg.cuda_stream)
            if self.record:
                slot.t1.record(sg)
            slot.ev_gathered.record(sg)
        slot.used, slot.batch = True, t
        self.launched = t + 1
        return slot

    def slot_of(self, t):
        slot = self.slots[t % self.nslots]
        assert slot.batch == t, "batch %d is not in flight" % t
        return slot

    def wait(self, t):
        slot = self.slot_of(t)
        torch.cuda.current_stream().wait_event(slot.ev_gathered)
        return slot

    def gather_ms(self):
        torch.cuda.synchronize()
        return [a.elapsed_time(b) for a, b in self.gather_times]

    def sample_ms(self):
        torch.cuda.synchronize()
        return [a.elapsed_time(b) for a, b in self.sample_times]


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
                self.refresh_meta(stream)
            slot.meta_used = self.meta_ring[self.meta_cur]
            if self.nd.parsing_method == "node_color":
                dom = self.domain_seeds[slot.k]
                meta = slot.meta_used
                self.nd.distribute_manager.distribute_async(offset, dom.data_ptr(), [meta[d].data_ptr() for d in range(self.D)],
                                                            stream.cuda_stream, slot=1 + slot.k)
                return dom[self.comm.local_rank * self.B:(self.comm.local_rank + 1) * self.B]
            lo = offset + self.comm.master_process_index * self.B * self.L + self.comm.local_rank * self.B  # (S/Training_node_distributor.py:41-47)
            return self.nd._dev_index[lo:lo + self.B]
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
