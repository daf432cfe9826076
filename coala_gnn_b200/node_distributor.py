"""Mirror of S/Training_node_distributor.py (Node_Distributor) over the GPU distributor."""
import torch

from . import binding


class Node_Distributor(object):
    def __init__(self, comm_manager, index_tensor, batch_size, color_file: str, topk_file: str, score_file: str,
                 parsing_method="node_color"):
        self.index_tensor = index_tensor.to("cpu").contiguous()
        self.index_offset = 0
        self.parsing_method = parsing_method
        self.batch_size = batch_size
        self.comm_manager = comm_manager

        self.domain_batch_size = batch_size * comm_manager.local_size
        self.global_batch_size = batch_size * comm_manager.global_size

        # the seed list is uploaded once; every distribute call then reads its slice in HBM
        # (the reference walks the CPU tensor, M/node_distributor_pybind.cuh:164)
        self._dev_index = None
        if torch.cuda.is_available():
            self._dev_index = self.index_tensor.to("cuda:%d" % comm_manager.device_index)
        item_ptr = self._dev_index.data_ptr() if self._dev_index is not None else self.index_tensor.data_ptr()
        self.distribute_manager = binding.mod().Node_distributor_pybind(item_ptr, self.comm_manager.node_id, self.batch_size,
                                                                        comm_manager.local_size, comm_manager.num_master_process,
                                                                        color_file, topk_file, score_file, device=comm_manager.device_index)
        self.num_colors = self.distribute_manager.get_num_colors()
        # one entry per colour id 0..num_colors (the reference sizes these num_colors and reads one past
        # the end for the last colour)
        self.meta_len = self.num_colors + 1
        self.distribute_manager.set_meta_len(self.meta_len)

        self.parsed_training_nodes_buffer = [torch.zeros(self.domain_batch_size, dtype=torch.int64).contiguous() for _ in range(2)]
        self.parsed_training_nodes_buffer_header = 0

        D = comm_manager.num_master_process
        self.cache_color_double_buffer = [[torch.zeros(self.meta_len, dtype=torch.int32) for _ in range(D)] for _ in range(2)]
        self.cache_color_db_header = 0

    def gather_cache_meta(self, gpu_cache_meta):
        self.comm_manager.gather_cache_meta(gpu_cache_meta, self.cache_color_double_buffer[self.cache_color_db_header])

    def parse_domain_training_nodes(self, color_buf_read_header):
        h = self.parsed_training_nodes_buffer_header
        if self.parsing_method == "baseline":  # contiguous slice per domain (S/Training_node_distributor.py:41-47)
            node_id = self.comm_manager.master_process_index
            lo = self.index_offset + node_id * self.domain_batch_size
            self.parsed_training_nodes_buffer[h] = self.index_tensor[lo:lo + self.domain_batch_size]
            self.index_offset += self.global_batch_size
            return self.parsed_training_nodes_buffer[h]
        elif self.parsing_method == "node_color" and self.comm_manager.num_master_process == 1:
            # one domain: every score is >= 0 > -1 until the single bucket is full and the bucket holds exactly
            # the batch, so the greedy loop (M/node_distributor_pybind.cuh:150-222) returns the seeds in arrival order
            lo = self.index_offset
            self.parsed_training_nodes_buffer[h] = self.index_tensor[lo:lo + self.domain_batch_size]
            self.index_offset += self.global_batch_size
            return self.parsed_training_nodes_buffer[h]
        elif self.parsing_method == "node_color":  # (:49-58)
            gather_ptr = [gt.data_ptr() for gt in self.cache_color_double_buffer[color_buf_read_header]]
            self.distribute_manager.distribute_node_with_affinity(self.index_offset, self.parsed_training_nodes_buffer[h].data_ptr(),
                                                                  gather_ptr)
            self.index_offset += self.global_batch_size
            return self.parsed_training_nodes_buffer[h]
        raise ValueError(f"Unsupported parsing method: {self.parsing_method}")

    def reset(self):
        self.index_offset = 0
        self.parsed_training_nodes_buffer_header = 0
        self.cache_color_db_header = 0
