"""Graph / sampler / block objects that stand in for the DGL pieces the reference's data path uses:
``dgl.graph(('csc', (indptr, indices, edge_ids)))`` on UVA tensors (examples/ssd_gnn_dataloader.py:523),
``dgl.dataloading.MultiLayerNeighborSampler(fanouts)`` (examples/sbatch_ssd_gnn_train.py:70-72) and its
``sample(g, seeds) -> (input_nodes, output_nodes, blocks)`` (S/COALA_GNN_DataLoader.py:162).
The sampling and to_block arithmetic runs in the CUDA kernels of csrc/sampler.cu."""
import ctypes as C

from ._lib import lib, check, check_handle
from .pybind import tensor_from_ptr

MASK = (1 << 64) - 1


def _mix64(z):
    z = (z + 0x9E3779B97F4A7C15) & MASK
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK
    return z ^ (z >> 31)


def batch_key(seed, batch_idx, rank=0):
    """RNG key of one mini-batch (oracle.batch_key is the same arithmetic)."""
    return _mix64((seed & MASK) ^ _mix64((batch_idx * 0x100000001B3 + rank) & MASK))


class UVAGraph:
    """CSC graph whose arrays are device-visible tensors (pinned host through UVA, or HBM).
    ``ndata`` holds per-node device tensors ('label', 'labels', masks) like the DGL graph."""

    def __init__(self, indptr, indices, edge_ids=None, num_nodes=None, device=None):
        self.indptr, self.indices, self.edge_ids = indptr, indices, edge_ids
        self._num_nodes = int(num_nodes if num_nodes is not None else indptr.numel() - 1)
        self.device = device if device is not None else indptr.device
        self.ndata = {}

    def number_of_nodes(self):
        return self._num_nodes

    num_nodes = number_of_nodes

    def formats(self, *_):
        return self


class _Frame(dict):
    """Per-block node data, gathered from the graph's ndata by node id on first access."""

    def __init__(self, graph, nids):
        super().__init__()
        self._g, self._nids = graph, nids

    def __missing__(self, key):
        v = self._g.ndata[key][self._nids]
        self[key] = v
        return v

    def keys(self):
        return set(super().keys()) | set(self._g.ndata.keys())


class _EdgeFrame(dict):
    """edata: '_ID' (dgl.EID) is looked up from the graph's edge-id array on first access."""

    def __init__(self, block):
        super().__init__()
        self._b = block

    def __missing__(self, key):
        if key != "_ID":
            raise KeyError(key)
        self[key] = self._b.edge_id
        return self[key]

    def keys(self):
        return set(super().keys()) | {"_ID"}


class Block:
    """One bipartite layer: edges (edge_src[e] -> edge_dst[e]) in local ids; dst nodes are the first
    ``num_dst_nodes`` src nodes. srcdata['_ID'] / dstdata['_ID'] are the global node ids (dgl.NID).
    ``edge_pos`` are the CSC positions of the sampled edges; ``edge_id`` / edata['_ID'] (the graph's
    edge ids at those positions) are fetched from the UVA edge-id array on first use.
    The id / edge tensors are views of the sampler's device buffers cut on first access (the iterator hands out
    a batch per 0.1 ms: a training loop pays for the views it touches, not for eight slices per batch)."""

    __slots__ = ("_g", "_v", "_n_src", "_n_dst", "_ne", "_t", "_edge_id", "_srcdata", "_dstdata", "_edata")

    def __init__(self, graph, views, n_src, n_dst, n_edges, copy=False):
        self._g, self._v = graph, views
        self._n_src, self._n_dst, self._ne = int(n_src), int(n_dst), int(n_edges)
        self._t = [None, None, None, None]
        self._edge_id = self._srcdata = self._dstdata = self._edata = None
        if copy:  # the batch owns its arrays (a later sample with the same sampler object overwrites the buffers)
            self._t = [views[0][:self._n_src].clone()] + [views[j][:self._ne].clone() for j in (1, 2, 3)]

    def _view(self, j):
        t = self._t[j]
        if t is None:
            t = self._t[j] = self._v[j][:self._n_src if j == 0 else self._ne]
        return t

    src_nodes = property(lambda self: self._view(0))
    edge_src = property(lambda self: self._view(1))
    edge_dst = property(lambda self: self._view(2))
    edge_pos = property(lambda self: self._view(3))

    @property
    def dst_nodes(self):
        return self._view(0)[:self._n_dst]

    @property
    def srcdata(self):
        if self._srcdata is None:
            self._srcdata = _Frame(self._g, self.src_nodes)
            self._srcdata["_ID"] = self.src_nodes
        return self._srcdata

    @property
    def dstdata(self):
        if self._dstdata is None:
            self._dstdata = _Frame(self._g, self.dst_nodes)
            self._dstdata["_ID"] = self.dst_nodes
        return self._dstdata

    @property
    def edata(self):
        if self._edata is None:
            self._edata = _EdgeFrame(self)
        return self._edata

    @property
    def edge_id(self):
        if self._edge_id is None:
            import torch
            eids = getattr(self._g, "edge_ids", None)
            if eids is None:
                self._edge_id = self.edge_pos
            else:
                out = torch.empty_like(self.edge_pos)
                check(lib().coala_lookup_edge_ids(eids.data_ptr(), self.edge_pos.data_ptr(), self.edge_pos.numel(), out.data_ptr(),
                                                  torch.cuda.current_stream().cuda_stream), "edge id lookup")
                self._edge_id = out
        return self._edge_id

    def number_of_src_nodes(self):
        return self._n_src

    def number_of_dst_nodes(self):
        return self._n_dst

    num_src_nodes, num_dst_nodes = number_of_src_nodes, number_of_dst_nodes

    def num_edges(self):
        return self._ne

    def edges(self):
        return self.edge_src, self.edge_dst

    def int(self):  # the training loop calls block.int().to(device) (examples/sbatch_ssd_gnn_train.py:139)
        return self

    def to(self, device):
        return self


class MultiLayerNeighborSampler:
    """Uniform in-neighbour sampling without replacement, one round per fan-out in reverse order,
    each followed by to_block. `seed` fixes the RNG stream; the k-th call uses batch_key(seed, k, rank)."""

    def __init__(self, fanouts, seed=0, rank=0, hbm_indptr=None, indptr_from=None, hbm_graph=None):
        import os
        # indptr (8 (N + 1) bytes) is copied to HBM unless COALA_HBM_INDPTR=0 / hbm_indptr=False: the two dependent host reads of
        # every sampled row become one (+12 % mini-batches/s where the pipeline is not host-link bound).
        # `indptr_from`: another sampler of the same graph and device whose copy this one uses (the slots of one engine).
        self.hbm_indptr = bool(int(os.environ.get("COALA_HBM_INDPTR", "1"))) if hbm_indptr is None else bool(hbm_indptr)
        # COALA_HBM_GRAPH=1 / hbm_graph=True: the neighbour lists (8 bytes per edge) are copied as well - opt-in, because every rank
        # of a node then holds its own copy of what the reference shares once in host memory.
        self.hbm_graph = bool(int(os.environ.get("COALA_HBM_GRAPH", "0"))) if hbm_graph is None else bool(hbm_graph)
        if self.hbm_graph:
            self.hbm_indptr = True
        self._indptr_from = indptr_from
        self.fanouts = [int(f) for f in fanouts]
        self.seed, self.rank, self.calls = int(seed), int(rank), 0
        self._h, self._graph_id, self._max_seeds = None, None, 0

    def _ensure(self, g, n_seeds):
        import torch
        key = (g.indptr.data_ptr(), g.indices.data_ptr())
        if self._h is not None and self._graph_id == key and n_seeds <= self._max_seeds:
            return
        if self._h is not None:
            lib().coala_sampler_destroy(self._h)
            self._h = None
        self._views, self._views_for = None, None  # a new native object may reuse the old one's address
        dev = torch.device(g.device)
        dev_index = dev.index if dev.index is not None else torch.cuda.current_device()
        fan = (C.c_int * len(self.fanouts))(*self.fanouts)
        eptr = g.edge_ids.data_ptr() if g.edge_ids is not None else 0
        self._h = check_handle(lib().coala_sampler_create(dev_index, g.indptr.data_ptr(), g.indices.data_ptr(), eptr,
                                                          g.number_of_nodes(), int(n_seeds), fan, len(self.fanouts)), "sampler")
        if self.hbm_indptr:
            src = self._indptr_from
            if src is not None and src._h is not None and src._graph_id == key and src.hbm_indptr and src.hbm_graph == self.hbm_graph:
                check(lib().coala_sampler_share_indptr(self._h, src._h), "share indptr")
            else:
                self._indptr_from = None
                check(lib().coala_sampler_stage_indptr(self._h), "stage indptr")
                if self.hbm_graph:
                    check(lib().coala_sampler_stage_indices(self._h), "stage indices")
        self._graph_id, self._max_seeds = key, int(n_seeds)
        self._device = "cuda:%d" % dev_index

    def trace(self):
        """Average ms between consecutive enqueued operations (needs COALA_SAMPLER_TRACE=1)."""
        buf = (C.c_double * 64)()
        n = lib().coala_sampler_trace(self._h, buf, 64)
        return [buf[i] for i in range(n)]

    def max_sample(self):
        return int(lib().coala_sampler_max_sample(self._h))

    def graph_nodes(self):
        """Nodes of the CUDA graph a sample is launched as (0 before the first sample)."""
        return int(lib().coala_sampler_graph_nodes(self._h))

    def sample_async(self, g, seeds, rng_key=None, stream=0):
        """Enqueue the whole multi-layer sample; results stay in the sampler's device buffers."""
        self._ensure(g, seeds.numel())
        if rng_key is None:
            rng_key = batch_key(self.seed, self.calls, self.rank)
        self.calls += 1
        check(lib().coala_sampler_sample_async(self._h, seeds.data_ptr(), seeds.numel(), rng_key, stream), "sample")
        self._last_g = g

    def input_nodes_dev_ptrs(self):
        """(device pointer of the input-node id buffer, device pointer of its int64 count)."""
        src = C.c_uint64()
        check(lib().coala_sampler_layer_ptrs(self._h, 0, C.byref(src), None, None, None))
        return src.value, int(lib().coala_sampler_input_count_dev_ptr(self._h))

    def _wrap_buffers(self):
        """Full-capacity torch views of the sampler's device buffers, built once (collect() then only slices)."""
        import torch
        L = len(self.fanouts)
        ns = [self._max_seeds]
        ne = []
        for f in reversed(self.fanouts):  # sampling order
            ne.append(ns[-1] * f)
            ns.append(ns[-1] + ne[-1])
        self._views = []
        for k in range(L):  # reference block order: k = 0 is the input layer = last sampled round
            l = L - 1 - k
            p = [C.c_uint64() for _ in range(4)]
            check(lib().coala_sampler_layer_ptrs(self._h, k, *[C.byref(x) for x in p]))
            caps = (ns[l + 1], ne[l], ne[l], ne[l])
            self._views.append([tensor_from_ptr(ptr.value, (cap,), torch.int64, self._device, owner=self) for ptr, cap in zip(p, caps)])

    def collect(self, copy=True):
        """Wait for the last sample_async and wrap the buffers. copy=False returns views that a later
        sample with this sampler object overwrites."""
        counts = (C.c_int64 * (2 * len(self.fanouts) + 1))()
        check(lib().coala_sampler_counts(self._h, counts), "sampler counts")
        return self.blocks_from_counts(counts, copy)

    def blocks_from_counts(self, counts, copy=False):
        """(input_nodes, output_nodes, blocks) of the finished sample whose sizes are `counts` (coala_sampler_counts order)."""
        L = len(self.fanouts)
        if getattr(self, "_views_for", None) != self._h:
            self._wrap_buffers()
            self._views_for = self._h
        g = self._last_g
        blocks = []
        for k in range(L):
            n_dst = counts[2 * (k + 1)] if k + 1 < L else counts[2 * L]
            blocks.append(Block(g, self._views[k], counts[2 * k], n_dst, counts[2 * k + 1], copy))
        return blocks[0].src_nodes, blocks[-1].dst_nodes, blocks

    def sample(self, g, seeds, rng_key=None):
        import torch
        seeds = seeds.to(g.device) if str(seeds.device) != str(g.device) else seeds
        self.sample_async(g, seeds.contiguous(), rng_key, torch.cuda.current_stream().cuda_stream)
        return self.collect()

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().coala_sampler_destroy(self._h)
                self._h = None
        except Exception:
            pass
