"""Device-resident form of the mini-batch path (B200-first addition next to the drop-in loader).

The reference serialises, per batch: CPU distributor (thread) -> gloo broadcast -> seeds H2D ->
DGL sampling with host syncs -> gather (S/COALA_GNN_DataLoader.py:149-167).  Here the three stages
are enqueued back to back on one CUDA stream and nothing returns to the host between them:
  * every rank of a cache domain runs the (deterministic) GPU distributor on the same inputs and
    takes its own slice, so there is no seed broadcast at all;
  * the sampler reads its seeds from the distributor's device output, the gather reads the
    sampler's input-node list and its length from device memory;
  * cache colour histograms are summed across the domain with one small NCCL all-reduce every
    `refresh_counter` batches and stay on the device.
Results of step t are valid after `sync()`; buffers are double-buffered across steps."""
import ctypes as C

import torch
import torch.distributed as dist

from ._lib import lib, check


class DevicePipeline(object):
    def __init__(self, loader, record_gather_events=False):
        self.loader = loader
        self.nd = loader.node_distributor
        self.comm = self.nd.comm_manager
        self.mgr = loader.COALA_GNN_Manager
        self.sampler = loader.sampler
        self.g = loader.g
        self.device = loader.device
        assert self.nd._dev_index is not None, "seed list must be device resident"
        if self.comm.num_master_process != 1:
            raise NotImplementedError("DevicePipeline drives one cache domain; use the drop-in loader for D > 1")
        self.B, self.L = self.nd.batch_size, self.comm.local_size
        self.out_seeds = [torch.zeros(self.B * self.L, dtype=torch.int64, device=self.device) for _ in range(2)]
        self.meta = torch.zeros(self.nd.meta_len, dtype=torch.int32, device=self.device)
        self.step_idx = 0
        self.offset = 0
        self.refresh = loader.refresh_counter
        self.events = [] if record_gather_events else None
        self.kernels_per_step = 3 + 1 + 7 * len(self.sampler.fanouts) + 1  # distributor, set_count, sampler rounds, gather
        self.sampler._ensure(self.g, self.B)

    def refresh_meta(self):
        """Colour histogram of the whole domain's cache -> device int32[meta_len]."""
        cache = self.mgr.COALA_GNN_Cache
        check(lib().coala_cache_get_cache_data_n(cache._h, self.meta.data_ptr(), self.nd.meta_len), "get_cache_data")
        if self.L > 1:
            dist.all_reduce(self.meta, group=self.comm.nccl_cache_gather)

    def step(self):
        stream = torch.cuda.current_stream().cuda_stream
        if self.step_idx % self.refresh == 0 and self.step_idx > 0:
            self.refresh_meta()
        k = self.step_idx & 1
        seeds_all = self.out_seeds[k]
        if self.nd.parsing_method == "node_color":
            self.nd.distribute_manager.distribute_async(self.offset, seeds_all.data_ptr(), [self.meta.data_ptr()], stream)
            seeds = seeds_all[self.comm.local_rank * self.B:(self.comm.local_rank + 1) * self.B]
        else:
            lo = self.offset + self.comm.local_rank * self.B
            seeds = self.nd._dev_index[lo:lo + self.B]
        self.offset += self.nd.global_batch_size
        self.sampler.sample_async(self.g, seeds, None, stream)
        idx_ptr, n_ptr = self.sampler.input_nodes_dev_ptrs()
        if self.events is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        out = self.mgr.fetch_feature_async(idx_ptr, self.mgr.max_sample_size, n_ptr, stream)
        if self.events is not None:
            e1.record()
            self.events.append((e0, e1))
        self.step_idx += 1
        return out

    def reset(self):
        self.offset = 0

    def steps_per_epoch(self):
        return len(self.nd.index_tensor) // self.nd.global_batch_size - 1

    def gather_ms(self):
        return [a.elapsed_time(b) for a, b in self.events]
