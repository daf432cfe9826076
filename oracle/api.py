"""TEST INFRASTRUCTURE ONLY — numpy-facing wrappers over liboracle.so / libcoala_ref.so."""
import ctypes as C
import os

import numpy as np

from . import lib, ref, _p, _u64, _vp

__all__ = [
    "mix64", "fill_features", "distribute", "ref_distribute", "CacheModel", "sample_layer",
    "to_block", "sample_blocks_detailed", "sample_blocks_fast", "index_select", "ref_color_graph",
    "layer_key", "batch_key", "draw_train_seeds", "levelsync_color_graph", "affinity_from_counts",
]

C1 = 0xA24BAED4963EE407
MASK = (1 << 64) - 1


def mix64(z):
    return int(lib().orc_mix64(C.c_uint64(z & MASK)))


def batch_key(seed, batch_idx, rank=0):
    """RNG key of one mini-batch (same arithmetic as coala_gnn_b200.sampler)."""
    return mix64((seed & MASK) ^ mix64((batch_idx * 0x100000001B3 + rank) & MASK))


def layer_key(bkey, layer):
    """Key of sampling round `layer` (0 = the round nearest the seeds)."""
    return mix64((bkey + (layer + 1) * C1) & MASK)


def fill_features(nrows, dim, seed, row0=0, stride=None):
    stride = stride or dim
    out = np.zeros((nrows, stride), dtype=np.float32)
    lib().orc_fill_features(_p(out), seed, row0, nrows, dim, stride)
    return out


def _i64a(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def distribute(items, offset, node_id, batch, local_size, n_domains, color, topk, score, metas):
    """C restatement of distribute_node_with_affinity. Returns (out[batch*local_size], bucket_len[D])."""
    items, color, topk = _i64a(items), _i64a(color), _i64a(topk)
    score = np.ascontiguousarray(score, dtype=np.float64)
    K = topk.shape[1]
    metas = [np.ascontiguousarray(m, dtype=np.int32) for m in metas]
    ptrs = (C.c_void_p * n_domains)(*[m.ctypes.data for m in metas])
    out = np.zeros(batch * local_size, dtype=np.int64)
    bl = np.zeros(n_domains, dtype=np.int32)
    lib().orc_distribute(_p(items), offset, node_id, batch, local_size, n_domains, _p(color), _p(topk),
                         _p(score), K, C.cast(ptrs, _vp), _p(out), _p(bl))
    return out, bl


def ref_distribute(items, offset, node_id, batch, local_size, n_domains, color_file, topk_file, score_file, metas):
    """The reference's own Node_distributor_pybind compiled from /root/reference (oracle/_ref)."""
    R = ref()
    items = _i64a(items)
    h = R.ref_nd_create(items.ctypes.data, node_id, batch, local_size, n_domains,
                        color_file.encode(), topk_file.encode(), score_file.encode())
    if not h:
        raise RuntimeError("reference distributor failed to load the .npy tables")
    metas = [np.ascontiguousarray(m, dtype=np.int32) for m in metas]
    ptrs = (C.c_uint64 * n_domains)(*[m.ctypes.data for m in metas])
    out = np.zeros(batch * local_size, dtype=np.int64)
    R.ref_nd_distribute(h, offset, out.ctypes.data, C.cast(ptrs, _vp), n_domains)
    nc = R.ref_nd_num_colors(h)
    R.ref_nd_destroy(h)
    return out, nc


class CacheModel:
    """Sequential model of the reference cache; mode 'isolated' | 'partitioned' (= nvshmem / nccl hashing)."""

    def __init__(self, cache_mb, page_size, n_gpus=1, mode="isolated", color=None, num_colors=0, store_lines=True):
        self.color = _i64a(color) if color is not None else None
        self.page_size = page_size
        self.n_gpus = n_gpus
        self.mode = 0 if mode == "isolated" else 1
        self.num_colors = num_colors
        self.h = lib().orc_cache_create(cache_mb, page_size, n_gpus, self.mode, _p(self.color), num_colors, int(store_lines))
        self.num_sets = int(lib().orc_cache_num_sets(self.h))

    def read(self, idx, feat, dim=None):
        idx = _i64a(idx)
        feat = np.ascontiguousarray(feat, dtype=np.float32)
        dim = dim or feat.shape[1]
        assert feat.shape[1] * 4 == self.page_size, "reference cache rows are cache_dim*4 bytes"
        out = np.zeros((len(idx), dim), dtype=np.float32)
        lib().orc_cache_read(self.h, _p(idx), len(idx), _p(feat), _p(out), dim)
        return out

    def stats(self):
        h, m = C.c_uint64(), C.c_uint64()
        lib().orc_cache_stats(self.h, C.byref(h), C.byref(m))
        return h.value, m.value

    def _view(self, fn, gpu, n, dt):
        ptr = getattr(lib(), fn)(self.h, gpu)
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(dt)), shape=(n,)).copy()

    def keys(self, gpu=0):
        return self._view("orc_cache_keys", gpu, self.num_sets * 32, C.c_uint64)

    def color_meta(self, gpu=0):
        return self._view("orc_cache_color_meta", gpu, self.num_sets * 32, C.c_uint64)

    def color_counters(self, gpu=0):
        return self._view("orc_cache_color_counters", gpu, self.num_colors + 1, C.c_int32)

    def set_cnt(self, gpu=0):
        return self._view("orc_cache_set_cnt", gpu, self.num_sets, C.c_uint32)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_cache_destroy(self.h)
            self.h = None


def sample_layer(indptr, indices, edge_ids, seeds, fanout, lkey):
    if fanout > 64:
        raise ValueError("oracle sampler: fan-out above MAX_FANOUT (64, oracle/c/sampler.c)")
    indptr, indices, seeds = _i64a(indptr), _i64a(indices), _i64a(seeds)
    edge_ids = _i64a(edge_ids) if edge_ids is not None else None
    deg = indptr[seeds + 1] - indptr[seeds]
    cap = int(deg.sum()) if fanout < 0 else int(np.minimum(deg, fanout).sum())
    src = np.zeros(cap + 1, dtype=np.int64)
    dst = np.zeros(cap + 1, dtype=np.int64)
    eid = np.zeros(cap + 1, dtype=np.int64)
    ne = lib().orc_sample_layer(_p(indptr), _p(indices), _p(edge_ids), _p(seeds), len(seeds), fanout, lkey,
                                cap + 1, _p(src), _p(dst), _p(eid))
    assert ne == cap
    return src[:ne], dst[:ne], eid[:ne]


def to_block(seeds, e_src):
    seeds, e_src = _i64a(seeds), _i64a(e_src)
    nodes = np.zeros(len(seeds) + len(e_src) + 1, dtype=np.int64)
    loc = np.zeros(len(e_src) + 1, dtype=np.int64)
    n = lib().orc_to_block(_p(seeds), len(seeds), _p(e_src), len(e_src), _p(nodes), _p(loc))
    return nodes[:n], loc[:len(e_src)]


def sample_blocks_detailed(indptr, indices, edge_ids, seeds, fanouts, bkey):
    """Full result in the reference's return shape: (input_nodes, output_nodes, blocks) with
    blocks ordered input layer first; each block is a dict of numpy arrays."""
    cur = _i64a(seeds)
    blocks = []
    for l, f in enumerate(reversed(list(fanouts))):
        src, dst, eid = sample_layer(indptr, indices, edge_ids, cur, f, layer_key(bkey, l))
        nodes, loc = to_block(cur, src)
        blocks.insert(0, dict(src_nodes=nodes, dst_nodes=cur.copy(), edge_src=loc, edge_dst=dst, edge_id=eid))
        cur = nodes
    return cur, _i64a(seeds), blocks


def sample_blocks_fast(indptr, indices, seeds, fanouts, bkey, max_sample, threads=None):
    """OpenMP multi-layer sampler (the timed CPU baseline). Returns (input_nodes, total_edges)."""
    threads = threads or os.cpu_count()
    seeds = _i64a(seeds)
    fan = np.ascontiguousarray(fanouts, dtype=np.int32)
    out = np.zeros(max_sample, dtype=np.int64)
    tot = C.c_int64()
    n = lib().orc_sample_blocks(_p(indptr), _p(indices), None, _p(seeds), len(seeds), _p(fan), len(fan), bkey,
                                max_sample, _p(out), C.byref(tot), threads)
    return out[:n], tot.value


def index_select(feat, idx, threads=None):
    threads = threads or os.cpu_count()
    idx = _i64a(idx)
    out = np.empty((len(idx), feat.shape[1]), dtype=np.float32)
    lib().orc_index_select(_p(feat), feat.shape[1], _p(idx), len(idx), _p(out), threads)
    return out


def ref_color_graph(indptr, indices, train_ids, topk=10):
    """Run the reference Graph_Coloring in the call order of
    examples/color_info_gen/generate_color_data.py:20-67. Returns (color, topk[num_c,K], score[num_c,K])."""
    R = ref()
    indptr, indices, train_ids = _i64a(indptr), _i64a(indices), _i64a(train_ids)
    n = len(indptr) - 1
    R.ref_srand(1)  # glibc rand() is never seeded by the reference -> seed 1
    h = R.ref_gc_create(n)
    R.ref_gc_set_adj_csc(h, indptr.ctypes.data, indices.ctypes.data)
    color = np.zeros(n, dtype=np.int64)
    R.ref_gc_set_color_buffer(h, color.ctypes.data)
    R.ref_gc_color_graph_optimized(h, train_ids.ctypes.data, len(train_ids))
    nc = int(R.ref_gc_get_num_color(h))
    tk = np.zeros(nc * topk, dtype=np.int64)
    sc = np.zeros(nc * topk, dtype=np.float64)
    R.ref_gc_set_topk_color_buffer(h, tk.ctypes.data)
    R.ref_gc_set_topk_affinity_buffer(h, sc.ctypes.data)
    R.ref_gc_calculate_color_affinity(h)
    R.ref_gc_destroy(h)
    return color, tk.reshape(nc, topk), sc.reshape(nc, topk)


# ---------------------------------------------------------------------------------------------
# Specification of the GPU mode of the colour-metadata generator (coala_gnn_b200/csrc/coloring_gpu.cu).
# The reference has no such mode; its own algorithm (M/graph_coloring.cpp) stays pinned by
# ref_color_graph + tests/golden/coloring_*.npz.  What is restated here is the order-independent
# definition the kernels implement, in numpy, so the kernels can be checked bit for bit.
# ---------------------------------------------------------------------------------------------
def draw_train_seeds(num_nodes, train_ids, base_rate=0.005):
    """Seeds of cpu_color_graph_optimized (M/graph_coloring.cpp:77-92): glibc rand() stream from seed 1,
    float(rand()) / RAND_MAX <= base_rate(float) * min(20, N / n_train)(double); train ids are distinct.
    Returns (nodes, colours)."""
    libc = C.CDLL("libc.so.6")
    libc.srand(1)
    rate = float(np.float32(base_rate)) * min(20.0, float(num_nodes) / float(len(train_ids)))
    rmax = np.float32(2147483647)
    nodes = [v for v in _i64a(train_ids).tolist() if float(np.float32(libc.rand()) / rmax) <= rate]
    nodes = np.array(nodes, dtype=np.int64)
    return nodes, np.arange(1, len(nodes) + 1, dtype=np.int64)


def _expand(indptr, indices, frontier, color):
    deg = indptr[frontier + 1] - indptr[frontier]
    if deg.sum() == 0:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    starts = np.repeat(indptr[frontier], deg)
    within = np.arange(deg.sum()) - np.repeat(np.cumsum(deg) - deg, deg)
    return indices[starts + within], np.repeat(color[frontier], deg)


def _claim(n, u, c, keep):
    claim = np.full(n, np.iinfo(np.int64).max, dtype=np.int64)
    np.minimum.at(claim, u[keep], c[keep])
    new = np.nonzero(claim != np.iinfo(np.int64).max)[0]
    return new, claim[new]


def levelsync_color_graph(indptr, indices, train_ids, max_hop=10, base_rate=0.005):
    """Level-synchronous colouring: level 1 = seeds + their uncoloured training in-neighbours, levels
    2..max_hop = uncoloured in-neighbours of the previous level; the smallest claiming colour wins.
    Returns (color int64[N], hop uint16[N], num_colors)."""
    indptr, indices = _i64a(indptr), _i64a(indices)
    n = len(indptr) - 1
    is_train = np.zeros(n, dtype=bool)
    is_train[_i64a(train_ids)] = True
    seeds, cols = draw_train_seeds(n, train_ids, base_rate)
    color = np.zeros(n, dtype=np.int64)
    hop = np.zeros(n, dtype=np.uint16)
    if max_hop < 1 or len(seeds) == 0:
        return color, hop, len(seeds)
    color[seeds] = cols
    hop[seeds] = 1
    u, c = _expand(indptr, indices, seeds, color)
    new, col = _claim(n, u, c, (color[u] == 0) & is_train[u])
    color[new] = col
    hop[new] = 1
    frontier = np.concatenate([seeds, new])
    for level in range(2, max_hop + 1):
        u, c = _expand(indptr, indices, frontier, color)
        new, col = _claim(n, u, c, color[u] == 0)
        color[new] = col
        hop[new] = level
        frontier = new
    return color, hop, len(seeds)


def affinity_from_counts(indptr, indices, color, hop, num_c, topk=10):
    """score(c, nc) = sum_h count(c, nc, h) * exp(-0.5 h) (ascending h, separately rounded multiply and add),
    top-k by (score desc, nc asc), divided by the in-edge count of c; colours 1..num_c-1 get a row
    (loop bound of M/graph_coloring.cpp:266). Returns (topk int64[num_c,K], score float64[num_c,K])."""
    import math
    indptr, indices, color = _i64a(indptr), _i64a(indices), _i64a(color)
    n = len(indptr) - 1
    deg = np.diff(indptr)
    v = np.repeat(np.arange(n), deg)
    c, nc, h = color[v], color[indices], hop[indices].astype(np.int64)
    ok = (c > 0) & (c < num_c) & (nc != 0) & (nc != c)
    in_edges = np.bincount(color, weights=deg.astype(np.float64), minlength=num_c + 1)
    key = (c[ok] << 32) | nc[ok]
    pairs = np.unique(key)
    score = np.zeros(len(pairs), dtype=np.float64)
    for hv in range(16):
        sel = h[ok] == hv
        cnt = np.zeros(len(pairs), dtype=np.float64)
        if sel.any():
            kk, kc = np.unique(key[sel], return_counts=True)
            cnt[np.searchsorted(pairs, kk)] = kc
        score = score + cnt * math.exp(-0.5 * hv)
    pc, pn = pairs >> 32, pairs & 0xFFFFFFFF
    order = np.lexsort((pn, -score, pc))
    tk = np.zeros((num_c, topk), dtype=np.int64)
    sc = np.zeros((num_c, topk), dtype=np.float64)
    pc_o, pn_o, s_o = pc[order], pn[order], score[order]
    first = np.searchsorted(pc_o, np.arange(num_c + 2))
    for col in range(1, num_c):
        a, b = first[col], min(first[col + 1], first[col] + topk)
        tk[col - 1, :b - a] = pn_o[a:b]
        sc[col - 1, :b - a] = s_o[a:b] / in_edges[col]
    return tk, sc
