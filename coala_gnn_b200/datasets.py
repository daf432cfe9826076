"""Synthetic stand-in for IGBDatast_Shared_CSC_UVA / OGBDataset_Shared_CSC_UVA
(examples/ssd_gnn_dataloader.py:365-569, 687-854): feature table and CSC arrays live in POSIX
shared memory, pinned and device-mapped by every rank of the machine; local rank 0 fills them.
Differences: nothing is read from disk (shapes in synth.SHAPES), the graph is generated straight
into the shared segments (no rank-0 np.load + second copy), and the feature table is written by a
GPU kernel through the UVA mapping."""
import os
import tempfile
import time

import numpy as np
import torch

from . import synth
from ._lib import lib, check
from .comm import Shared_UVA_Tensor_Manager
from .sampler import UVAGraph


class SyntheticSharedCSCUVA(object):
    def __init__(self, shape, comm_manager, device, num_classes=19, num_colors=None, feat_cpu=True, tag=None, verbose=False):
        n, e, dim, gseed, fseed = synth.SHAPES[shape] if isinstance(shape, str) else shape
        self.num_nodes, self.num_edges, self.dim, self.feat_seed = n, e, dim, fseed
        self.comm = comm_manager
        self.device = device
        cd = lib().coala_cache_dim_for(dim)
        self.cache_dim = cd
        tag = tag or os.environ.get("MASTER_PORT", str(os.getpid()))
        rank0 = comm_manager.local_rank == 0
        t0 = time.time()

        def seg(name, nbytes, zero=False):
            return Shared_UVA_Tensor_Manager(comm_manager, "/coala_b200_%s_%s" % (tag, name), nbytes, zero=zero)

        self.feat_data = None
        self._segs = []
        if feat_cpu:  # --feat_cpu: features pinned in host memory, row stride = cache_dim floats
            m = seg("feat", n * cd * 4)
            self._segs.append(m)
            self.feat_data = m.get_tensor(np.float32, device, (n, cd))
            if rank0:
                check(lib().coala_synth_fill_features(self.feat_data.data_ptr(), fseed, 0, n, dim, cd, None))
                torch.cuda.synchronize()
            comm_manager.local_comm.Barrier()
        t_feat = time.time() - t0
        m1, m2, m3 = seg("indptr", (n + 1) * 8), seg("indices", e * 8), seg("eids", e * 8)
        self._segs += [m1, m2, m3]
        if rank0:
            out = (m1.get_host_array(np.int64, (n + 1,)), m2.get_host_array(np.int64, (e,)), m3.get_host_array(np.int64, (e,)))
            synth.make_csc(n, e, gseed, out=out, edge_ids="arange" if e > 50_000_000 else "perm")
        comm_manager.local_comm.Barrier()
        self.csc_host = (m1.get_host_array(np.int64, (n + 1,)), m2.get_host_array(np.int64, (e,)), m3.get_host_array(np.int64, (e,)))
        indptr = m1.get_tensor(np.int64, device, (n + 1,))
        indices = m2.get_tensor(np.int64, device, (e,))
        eids = m3.get_tensor(np.int64, device, (e,))
        self.graph = UVAGraph(indptr, indices, eids, n, device)  # dgl.graph(('csc', ...)) (:523)
        labels = torch.from_numpy(synth.make_labels(n, num_classes, gseed)).to(device)
        self.graph.ndata["label"] = self.graph.ndata["labels"] = labels
        n_train, n_val = int(n * 0.6), int(n * 0.2)  # (:546-557)
        for name, lo, hi in (("train_mask", 0, n_train), ("val_mask", n_train, n_train + n_val), ("test_mask", n_train + n_val, n)):
            mask = torch.zeros(n, dtype=torch.bool, device=device)
            mask[lo:hi] = True
            self.graph.ndata[name] = mask
        # colour metadata files in the reference's on-disk layout
        self.meta_dir = os.path.join(tempfile.gettempdir(), "coala_b200_%s_meta" % tag)
        if rank0:
            os.makedirs(self.meta_dir, exist_ok=True)
            # COALA_SYNTH_COLOR = block (default: id-block stand-in) | graph (the reference's colouring, host mode,
            # ~15 s at 10 M nodes) | graph_gpu (same seeds, GPU mode, ~0.3 s) - SURVEY 8(d) "colour metadata for each
            # graph comes from the colouring with the reference call order"
            how = self._color_mode(n)
            if how == "block":
                C = num_colors or max(16, int(0.005 * n))
                color, topk, score = synth.block_color_meta(n, C)
            else:
                color, topk, score = _graph_color_meta(self.csc_host[0], self.csc_host[1], n, how == "graph_gpu", comm_manager.device_index)
            np.save(os.path.join(self.meta_dir, "color.npy"), color)
            np.save(os.path.join(self.meta_dir, "topk.npy"), topk)
            np.save(os.path.join(self.meta_dir, "score.npy"), score)
        comm_manager.local_comm.Barrier()
        self.color_source = {"block": "id-block stand-in (synth.block_color_meta), not derived from the graph",
                             "graph": "Graph_Coloring host mode on this graph = the reference's colouring (M/graph_coloring.cpp), reference call order",
                             "graph_gpu": "Graph_Coloring GPU mode on this graph (level-synchronous variant of the reference's colouring, same seeds)"}[self._color_mode(n)]
        self.color_file, self.topk_file, self.score_file = (os.path.join(self.meta_dir, x) for x in ("color.npy", "topk.npy", "score.npy"))
        if verbose and comm_manager.global_rank == 0:
            import sys
            sys.stderr.write("dataset %s ready in %.1fs (features %.1fs)\n" % (shape, time.time() - t0, t_feat))

    @staticmethod
    def _color_mode(n):
        """COALA_SYNTH_COLOR = graph (the reference's colouring in host mode: its bits; ~15 s at 10 M nodes, one thread) |
        graph_gpu (same seeds, GPU mode, ~0.3 s) | block (id-block stand-in).  Default: graph up to 20 M nodes, graph_gpu
        above ("colour metadata for each graph comes from the colouring with the reference call order")."""
        how = os.environ.get("COALA_SYNTH_COLOR", "auto")
        if how == "auto":
            how = "graph" if n <= 20_000_000 else "graph_gpu"
        if how not in ("block", "graph", "graph_gpu"):
            raise ValueError("COALA_SYNTH_COLOR must be auto, block, graph or graph_gpu")
        return how

    def feat_host_array(self):
        """numpy view of the pinned feature table (host side) — used by the CPU baseline only."""
        return self._segs[0].get_host_array(np.float32, (self.num_nodes, self.cache_dim))

    def __getitem__(self, i):
        return self.graph

    def cleanup(self):
        for m in self._segs:
            m.cleanup()
        self._segs = []

    def unlink_segments(self):
        """Remove the shm names without unmapping (emergency exit while kernels may still read the mappings)."""
        import ctypes
        if self.comm.local_rank == 0:
            libc = ctypes.CDLL(None)
            for m in self._segs:
                path = m.memory_handle._path
                libc.shm_unlink(path if isinstance(path, bytes) else path.encode())


def _copy_npy_into(dst, path, threads=8, chunk_rows=None):
    """Stream a .npy file into `dst` (a numpy view of shared pinned memory) without materialising it:
    memory-map the file and copy row chunks on a few threads (numpy releases the GIL in the copy).
    The reference has rank 0 np.load the whole file and then copy it a second time
    (examples/ssd_gnn_dataloader.py:419-437)."""
    from concurrent.futures import ThreadPoolExecutor
    src = np.load(path, mmap_mode="r")
    if src.shape[0] != dst.shape[0]:
        raise ValueError("%s: leading dimension %d, expected %d" % (path, src.shape[0], dst.shape[0]))
    n = src.shape[0]
    row_bytes = max(1, src[0:1].nbytes)
    chunk_rows = chunk_rows or max(1, (64 << 20) // row_bytes)

    def work(lo):
        hi = min(n, lo + chunk_rows)
        if src.ndim == 2 and dst.shape[1] != src.shape[1]:
            dst[lo:hi, :src.shape[1]] = src[lo:hi]  # rows padded to the cache row stride
        else:
            dst[lo:hi] = src[lo:hi]
    with ThreadPoolExecutor(max_workers=threads) as ex:
        list(ex.map(work, range(0, n, chunk_rows)))


def _graph_color_meta(indptr, indices, n, gpu, device_index):
    """Graph_Coloring in the call order of examples/color_info_gen/generate_color_data.py:20-54 (reference)."""
    import ctypes
    from .pybind import Graph_Coloring
    train = np.arange(int(n * 0.6), dtype=np.int64)
    ctypes.CDLL("libc.so.6").srand(1)  # the reference never seeds rand()
    tool = Graph_Coloring(n)
    tool.set_adj_csc(indptr.ctypes.data, indices.ctypes.data)
    color = np.zeros(n, dtype=np.int64)
    tool.set_color_buffer(color.ctypes.data)
    if gpu:
        tool.gpu_color_graph_optimized(train.ctypes.data, len(train), device_index)
    else:
        tool.cpu_color_graph_optimized(train.ctypes.data, len(train))
    nc = tool.get_num_color()
    topk, score = np.zeros(nc * 10, dtype=np.int64), np.zeros(nc * 10, dtype=np.float64)
    tool.set_topk_color_buffer(topk.ctypes.data)
    tool.set_topk_affinity_buffer(score.ctypes.data)
    if gpu:
        tool.gpu_calculate_color_affinity(device_index)
    else:
        tool.cpu_calculate_color_affinity()
    return color, topk.reshape(nc, 10), score.reshape(nc, 10)


def npy_shape(path):
    a = np.load(path, mmap_mode="r")
    return tuple(a.shape), a.dtype


class NpySharedCSCUVA(object):
    """On-disk counterpart of IGBDatast_Shared_CSC_UVA (examples/ssd_gnn_dataloader.py:365-569): the
    reference's directory layout and file names,
        <path>/<size>/processed/paper/node_feat.npy, node_label_19.npy | node_label_2K.npy
        <path>/<size>/processed/paper__cites__paper/csc_indptr.npy, csc_indices.npy, csc_edge_ids.npy
        <path>/<size>/color.npy, topk.npy, score.npy          (examples/sbatch_ssd_gnn_train.py:52-60)
    loaded straight into shared pinned memory (every rank of the machine maps the same segments)."""

    def __init__(self, path, size, comm_manager, device, num_classes=19, feat_cpu=True, tag=None):
        import os.path as osp
        self.comm, self.device = comm_manager, device
        root = osp.join(path, size)
        paper, cites = osp.join(root, "processed", "paper"), osp.join(root, "processed", "paper__cites__paper")
        tag = tag or os.environ.get("MASTER_PORT", str(os.getpid()))
        rank0 = comm_manager.local_rank == 0
        (n, dim), fdt = npy_shape(osp.join(paper, "node_feat.npy"))
        if fdt != np.float32:
            raise ValueError("node_feat.npy must be float32")
        (np1,), _ = npy_shape(osp.join(cites, "csc_indptr.npy"))
        (e,), _ = npy_shape(osp.join(cites, "csc_indices.npy"))
        if np1 != n + 1:
            raise ValueError("csc_indptr.npy has %d entries for %d nodes" % (np1, n))
        self.num_nodes, self.num_edges, self.dim = n, e, dim
        self.cache_dim = cd = lib().coala_cache_dim_for(dim)
        if cd < 0:
            raise RuntimeError("Only Feature Embedding Size less than 8KB is supported\n")

        def seg(name, nbytes):
            return Shared_UVA_Tensor_Manager(comm_manager, "/coala_b200_%s_%s" % (tag, name), nbytes, zero=False)
        self._segs = []
        self.feat_data = None
        if feat_cpu:
            m = seg("feat", n * cd * 4)
            self._segs.append(m)
            if rank0:
                host = m.get_host_array(np.float32, (n, cd))
                if cd != dim:
                    host[:, dim:] = 0
                _copy_npy_into(host, osp.join(paper, "node_feat.npy"))
            comm_manager.local_comm.Barrier()
            self.feat_data = m.get_tensor(np.float32, device, (n, cd))
        arrays = []
        for name, fname, count in (("indptr", "csc_indptr.npy", n + 1), ("indices", "csc_indices.npy", e), ("eids", "csc_edge_ids.npy", e)):
            m = seg(name, count * 8)
            self._segs.append(m)
            if rank0:
                _copy_npy_into(m.get_host_array(np.int64, (count,)), osp.join(cites, fname))
            arrays.append(m)
        comm_manager.local_comm.Barrier()
        self.csc_host = tuple(m.get_host_array(np.int64, (c,)) for m, c in zip(arrays, (n + 1, e, e)))
        indptr, indices, eids = (m.get_tensor(np.int64, device, (c,)) for m, c in zip(arrays, (n + 1, e, e)))
        self.graph = UVAGraph(indptr, indices, eids, n, device)
        label_file = "node_label_19.npy" if num_classes == 19 else "node_label_2K.npy"
        labels = torch.from_numpy(np.load(osp.join(paper, label_file))).to(torch.long).to(device)  # (:519)
        self.graph.ndata["label"] = self.graph.ndata["labels"] = labels
        n_train, n_val = int(n * 0.6), int(n * 0.2)  # (:546-557)
        for name, lo, hi in (("train_mask", 0, n_train), ("val_mask", n_train, n_train + n_val), ("test_mask", n_train + n_val, n)):
            mask = torch.zeros(n, dtype=torch.bool, device=device)
            mask[lo:hi] = True
            self.graph.ndata[name] = mask
        self.color_file, self.topk_file, self.score_file = (osp.join(root, x) for x in ("color.npy", "topk.npy", "score.npy"))

    def feat_host_array(self):
        return self._segs[0].get_host_array(np.float32, (self.num_nodes, self.cache_dim))

    def __getitem__(self, i):
        return self.graph

    def cleanup(self):
        for m in self._segs:
            m.cleanup()
        self._segs = []
