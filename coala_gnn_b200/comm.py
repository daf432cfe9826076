"""Topology + collectives of the data path, re-backed by torch.distributed (no MPI / mpi4py / cupy).

Mirrors S/Shared_Tensor.py: ``MPI_Comm_Manager`` (:24-112) keeps its attribute surface
(global_rank, local_rank, local_size, node_id, is_master, num_master_process,
master_process_index, ...) and its three collectives; ``Shared_UVA_Tensor_Manager`` (:118-179)
keeps get_tensor / write_np_array.  A "node" is a cache DOMAIN: the GPUs that share one
partitioned cache.  On one 8-GPU box several domains can be emulated (D x L = world size), which
is what the reference gets from SLURM_NODEID (examples/sbatch_ssd_gnn_train.py:249-262).
"""
import os

import numpy as np
import torch
import torch.distributed as dist

from . import binding
from .pybind import tensor_from_ptr


class _Comm:
    """Tiny stand-in for an mpi4py communicator over a torch process group."""

    def __init__(self, group, ranks, my_global_rank):
        self.group, self.ranks, self._me = group, list(ranks), my_global_rank

    def rank(self):
        return self.ranks.index(self._me)

    Get_rank = rank

    def Get_size(self):
        return len(self.ranks)

    def barrier(self):
        if len(self.ranks) > 1:
            dist.barrier(group=self.group)

    Barrier = barrier

    def Bcast(self, array, root=0):
        if len(self.ranks) > 1:
            t = torch.from_numpy(array)
            dist.broadcast(t, src=self.ranks[root], group=self.group)

    def allgather(self, obj):
        if len(self.ranks) == 1:
            return [obj]
        out = [None] * len(self.ranks)
        dist.all_gather_object(out, obj, group=self.group)
        return out


class MPI_Comm_Manager(object):
    def __init__(self, node=-1, gpus_per_domain=None):
        self.global_rank = int(os.environ.get("RANK", 0))
        self.global_size = int(os.environ.get("WORLD_SIZE", 1))
        self.device_index = int(os.environ.get("LOCAL_RANK", 0))  # CUDA device of this process on its machine
        if self.global_size > 1 and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            backend = "cpu:gloo,cuda:nccl" if torch.cuda.is_available() else "gloo"
            if torch.cuda.is_available():
                torch.cuda.set_device(self.device_index)
            # no device_id= here: binding one makes torch create NCCL communicators eagerly and turns every
            # later new_group() into an ncclCommSplit that non-members of a *gloo* subgroup wait in forever
            dist.init_process_group(backend, world_size=self.global_size, rank=self.global_rank)
        if gpus_per_domain is None:
            gpus_per_domain = int(os.environ.get("COALA_GPUS_PER_DOMAIN", self.global_size))
        node = int(node)
        self.node_id = node if node >= 0 else self.global_rank // gpus_per_domain
        world = list(range(self.global_size))
        self.global_comm = _Comm(dist.group.WORLD if self.global_size > 1 else None, world, self.global_rank)
        node_ids = self.global_comm.allgather(self.node_id)
        # local communicator = ranks with my node id, ordered by global rank (MPI Split(color=node, key=rank), :30)
        self.dist_local_rank_list = [r for r in world if node_ids[r] == self.node_id]
        self.local_rank = self.dist_local_rank_list.index(self.global_rank)
        self.local_size = len(self.dist_local_rank_list)
        self.is_master = self.local_rank == 0
        seen, self.master_process_list, self.global_dist_local_rank_list = set(), [], []
        for r in world:  # masters in rank order (:40-50)
            if node_ids[r] not in seen:
                seen.add(node_ids[r])
                self.master_process_list.append(r)
                self.global_dist_local_rank_list.append([q for q in world if node_ids[q] == node_ids[r]])
        self.num_master_process = len(self.master_process_list)
        self.master_process_id = self.dist_local_rank_list[0]
        self.master_process_index = self.master_process_list.index(self.master_process_id)
        self.local_gloo_scatter_array, self.local_gloo_gather_array, self.nccl_cache_gather_array = [], [], []
        self.local_gloo_gather = self.local_gloo_scatter = self.nccl_cache_gather = self.master_gloo_gather = None
        self.local_comm = _Comm(None, self.dist_local_rank_list, self.global_rank)
        self._groups_ready = False
        if self.global_size > 1:
            # the local communicator exists from construction on in the reference (MPI Split); here it
            # needs a process group, so the nested groups are created eagerly
            self.initialize_nested_process_group("nccl" if torch.cuda.is_available() else "gloo")

    def initialize_nested_process_group(self, cache_backend="nvshmem"):
        """S/Shared_Tensor.py:62-92 — every rank creates every group, in the same order."""
        if self.global_size == 1 or self._groups_ready:
            return
        for ranks in self.global_dist_local_rank_list:
            self.local_gloo_scatter_array.append(dist.new_group(ranks=ranks, backend="gloo"))
            self.local_gloo_gather_array.append(dist.new_group(ranks=ranks, backend="gloo"))
            if torch.cuda.is_available():
                self.nccl_cache_gather_array.append(dist.new_group(ranks=ranks, backend="nccl"))
        self.local_gloo_gather = self.local_gloo_gather_array[self.master_process_index]
        self.local_gloo_scatter = self.local_gloo_scatter_array[self.master_process_index]
        if self.nccl_cache_gather_array:
            self.nccl_cache_gather = self.nccl_cache_gather_array[self.master_process_index]
        self.master_gloo_gather = dist.new_group(ranks=self.master_process_list, backend="gloo")
        self.local_comm.group = self.local_gloo_scatter
        self._groups_ready = True
        dist.barrier()

    def gather_cache_meta(self, gpu_cache_meta, gathered_data):
        """SUM over the domain, then all-gather across domain masters (S/Shared_Tensor.py:96-100); the gathered
        histograms are then broadcast inside each domain so that every rank can run the (deterministic)
        distributor itself instead of receiving its seeds from the master every batch."""
        if self.global_size == 1:
            gathered_data[0].copy_(gpu_cache_meta)
            return
        if self.local_size > 1:
            dist.all_reduce(gpu_cache_meta, op=dist.ReduceOp.SUM, group=self.local_gloo_gather)
        if self.num_master_process == 1:
            gathered_data[0].copy_(gpu_cache_meta)
            return
        if self.is_master:
            dist.all_gather(gathered_data, gpu_cache_meta, group=self.master_gloo_gather)
        if self.local_size > 1:
            flat = torch.stack(gathered_data) if self.is_master else torch.empty((len(gathered_data),) + tuple(gathered_data[0].shape),
                                                                                  dtype=gathered_data[0].dtype)
            dist.broadcast(flat, src=self.master_process_id, group=self.local_gloo_gather)
            if not self.is_master:
                for dst, src in zip(gathered_data, flat):
                    dst.copy_(src)

    def broadcast_training_nodes(self, parsed_training_node_list):
        """(S/Shared_Tensor.py:102-103)"""
        if self.local_size > 1:
            dist.broadcast(parsed_training_node_list, src=self.master_process_id, group=self.local_gloo_scatter)

    def destroy_process_group(self):
        if self.global_size > 1 and dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()


class Shared_UVA_Tensor_Manager(object):
    """S/Shared_Tensor.py:118-179: one POSIX shm segment per tensor, created by local rank 0,
    mapped + pinned by every rank of the machine, exposed as a torch "cuda" tensor (UVA alias)."""

    def __init__(self, comm_manager, path, tensor_size, zero=True):
        self.comm_manager = comm_manager
        self.tensor_size = int(tensor_size)
        self.device = "cuda:" + str(comm_manager.device_index)
        self.memory_handle = binding.make_shared_uva(path, self.tensor_size, comm_manager.node_id, comm_manager.local_comm,
                                                     comm_manager.device_index, zero)
        self.device_ptr = self.memory_handle.get_device_ptr()
        self.host_ptr = self.memory_handle.get_host_ptr()

    def get_tensor(self, dtype, device, tensor_shape):
        name = np.dtype(dtype).name if not isinstance(dtype, torch.dtype) else str(dtype).replace("torch.", "")
        return tensor_from_ptr(self.device_ptr, tuple(tensor_shape), name, self.device, owner=self.memory_handle)

    def get_host_array(self, dtype, tensor_shape):
        """numpy view of the same memory (host side) — lets rank 0 fill it without a staging copy."""
        import ctypes
        n = int(np.prod(tensor_shape)) * np.dtype(dtype).itemsize
        buf = (ctypes.c_char * n).from_address(self.host_ptr)
        return np.frombuffer(buf, dtype=dtype).reshape(tensor_shape)

    def write_np_array(self, uva_tensor, np_array):
        if self.comm_manager.local_rank == 0:
            if tuple(uva_tensor.shape) != tuple(np_array.shape):
                raise ValueError(f"Tensor shape {uva_tensor.shape} does not match numpy array shape {np_array.shape}")
            self.get_host_array(np_array.dtype, np_array.shape)[...] = np_array
        self.comm_manager.local_comm.Barrier()

    def cleanup(self):
        self.memory_handle.cleanup()
