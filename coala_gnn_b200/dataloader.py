"""Mirror of S/COALA_GNN_DataLoader.py: the iterator that glues seed distribution -> neighbour
sampling -> cached feature gather, yielding (input_nodes, output_nodes, blocks, feat)."""
import threading

import torch

from .manager import COALA_GNN_Manager


class COALA_GNN_Node_Distribution_Scheduler(object):
    """S/COALA_GNN_DataLoader.py:8-75.  Same interface (run(is_last) -> this rank's seeds of the next batch,
    cache colour metadata refreshed every `refresh_counter` batches into a double buffer, gloo collectives on
    a background thread), different plumbing:
      * the reference distributes on the domain master (a CPU loop on a thread) and gloo-broadcasts the
        domain's seeds to its ranks every batch.  The GPU distributor is deterministic and takes tens of
        microseconds, so every rank runs it on the same inputs and slices its own seeds: no per-batch
        collective, no per-batch thread.  For that the refreshed metadata is made identical on all ranks
        of the domain (MPI_Comm_Manager.gather_cache_meta broadcasts what the masters gathered);
      * the one-batch-ahead overlap the reference gets from its thread comes from the loader's prefetch
        (run() for batch t+1 is called while batch t is in flight)."""

    def __init__(self, node_distributor, ssd_gnn_manager, refresh_counter=8):
        self.node_distributor = node_distributor
        self.ssd_gnn_manager = ssd_gnn_manager
        self.metadata_reuse_counter = 0
        self.refresh_counter = refresh_counter
        self.cache_color_gathered_header = 0
        self.cache_meta_gather_thread = None
        self.cache_meta_tensor = torch.zeros(self.node_distributor.meta_len, dtype=torch.int32)

    def run(self, is_last: bool):
        nd = self.node_distributor
        comm = nd.comm_manager
        distributed_node_index = nd.parse_domain_training_nodes(self.cache_color_gathered_header)

        if self.metadata_reuse_counter == self.refresh_counter:  # (:46-59)
            self.metadata_reuse_counter = 0
            if self.cache_meta_gather_thread is not None:
                self.cache_meta_gather_thread.join()
                nd.cache_color_db_header = int((nd.cache_color_db_header + 1) % 2)
            self.ssd_gnn_manager.COALA_GNN_Cache.get_cache_data(self.cache_meta_tensor.data_ptr(), nd.meta_len)
            self.cache_color_gathered_header = int((nd.cache_color_db_header + 1) % 2)
            self.cache_meta_gather_thread = threading.Thread(target=nd.gather_cache_meta, args=(self.cache_meta_tensor,))
            self.cache_meta_gather_thread.start()

        self.metadata_reuse_counter += 1
        lr = comm.local_rank
        return distributed_node_index[(lr * nd.batch_size):((lr + 1) * nd.batch_size)]  # (:71-75)

    def drain(self):
        if self.cache_meta_gather_thread is not None:
            self.cache_meta_gather_thread.join()
        self.cache_meta_gather_thread = None


class SSD_INFO(object):
    def __init__(self, num_ssds, page_size, num_elems, ssd_read_offset):
        self.num_ssds = num_ssds
        self.num_elems = num_elems
        self.ssd_read_offset = ssd_read_offset
        self.page_size = page_size


class COALA_GNN_DataLoader(object):
    """S/COALA_GNN_DataLoader.py:92-174 — same constructor arguments and iterator contract."""

    def __init__(self, SSD_info, node_distributor, graph, graph_sampler, batch_size, dim, fan_out, cache_size, device,
                 refresh_counter=10, cache_backend="nvshmem", sim_buf=None, shuffle=False):
        self.refresh_counter = refresh_counter
        self.sampler = graph_sampler
        self.batch_size = batch_size
        self.g = graph
        self.SSD_info = SSD_info
        self.cache_backend = cache_backend
        self.node_distributor = node_distributor
        self.device = device
        self.COALA_GNN_Manager = COALA_GNN_Manager(
            node_distributor=node_distributor, page_size=SSD_info.page_size, num_ssds=SSD_info.num_ssds,
            num_elems=SSD_info.num_elems, ssd_read_offset=SSD_info.ssd_read_offset, cache_size=cache_size,
            batch_size=batch_size, fan_out=fan_out, dim=dim, MPI_comm_manager=node_distributor.comm_manager,
            device=device, cache_backend=cache_backend, sim_buf=sim_buf)
        self.scheduler = COALA_GNN_Node_Distribution_Scheduler(node_distributor=self.node_distributor,
                                                               ssd_gnn_manager=self.COALA_GNN_Manager,
                                                               refresh_counter=self.refresh_counter)
        self.counter = 0
        self.index_len = len(self.node_distributor.index_tensor)
        self.total_count = int(self.index_len / self.node_distributor.global_batch_size) - 1  # (:139-141)
        # fused path: sampler output feeds the gather on the device without a host round trip, and
        # batch t+1 is sampled while batch t is gathered / consumed (pipeline.BatchEngine)
        self.fused = self.COALA_GNN_Manager.engine_capable and hasattr(self.sampler, "sample_async") and torch.cuda.is_available()
        self.engine = None
        if self.fused:
            from .pipeline import BatchEngine
            self.engine = BatchEngine(graph, graph_sampler, self.COALA_GNN_Manager, batch_size, device)
        self._epoch_base = 0  # engine batch index of this epoch's first batch
        self._launched = 0
        self.copy_blocks = False  # True: every batch owns its id/edge tensors (one clone per array, like DGL)

    def __iter__(self):
        return self

    def __len__(self):
        return max(self.total_count, 0)

    def _launch(self, i):
        """Distribute (host scheduler) and enqueue sample + gather of this epoch's batch i: one native call
        (pinned staging copy of the seeds, H2D, sample graph, gather) - csrc/engine.cu."""
        is_last_iter = i + 1 >= self.total_count  # the scheduler must not start distributing a batch that does not exist
        seeds_cpu = self.scheduler.run(is_last_iter)
        self.engine.launch_host(self._epoch_base + i, seeds_cpu.data_ptr(), seeds_cpu.numel())
        self._launched = i + 1

    def __next__(self):
        if self.counter == self.total_count:
            self.scheduler.drain()
            self.node_distributor.reset()
            if self.engine is not None:
                self._epoch_base = self.engine.launched
                self._launched = 0
            self.counter = 0
            raise StopIteration
        if not self.fused:
            seeds_cpu = self.scheduler.run(self.counter + 1 >= self.total_count)
            self.counter += 1
            batch = self.sampler.sample(self.g, seeds_cpu.to(self.device))
            return self.COALA_GNN_Manager.fetch_feature(batch)
        i = self.counter
        eng = self.engine
        eng.consumer_checkpoint()
        last = min(i + eng.depth, self.total_count - 1)
        while self._launched <= last:
            self._launch(self._launched)  # batch i, then the prefetch: sampled while batch i is gathered and consumed
        self.counter += 1
        # views into the slot's buffers: valid until `engine.depth` more batches have been drawn (the slot is then resampled)
        slot, counts = eng.collect(self._epoch_base + i)
        input_nodes, output_nodes, blocks = slot.sampler.blocks_from_counts(counts, copy=self.copy_blocks)
        return input_nodes, output_nodes, blocks, slot.out[:counts[0]]

    def print_stats(self):
        self.COALA_GNN_Manager.print_stats()
        agg_time = self.COALA_GNN_Manager.get_aggregate_time()
        print(f"Aggregation time: {agg_time}")
        self.COALA_GNN_Manager.aggregation_timer = 0.0
