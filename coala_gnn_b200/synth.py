"""Synthetic IGB/OGB-shaped inputs: CSC graphs, colour metadata, labels.

Stands in for the reference's on-disk datasets (examples/ssd_gnn_dataloader.py:401-563 reads
csc_indptr/csc_indices/csc_edge_ids/node_feat .npy files); nothing is downloaded.  Graph
structure is generated with numpy on the host once at start-up — it is input preparation, not
the hot path.  Feature tables are generated procedurally (a counter hash of (seed,row,col))
by the CUDA kernel behind ``Lib.synth_fill_features`` so any row can be re-derived for
verification without a second copy.
"""
import numpy as np

SHAPES = {
    # name: (nodes, edges, dim, graph seed, feature seed)
    "tiny": (20_000, 240_000, 128, 7, 17),
    "products": (2_449_029, 61_859_140, 100, 1, 11),
    "igb-small": (1_000_000, 12_000_000, 1024, 2, 12),
    "igb-medium": (10_000_000, 120_000_000, 1024, 3, 13),
    "papers100m": (111_059_956, 1_615_685_872, 128, 4, 14),
    "igb-large": (100_000_000, 1_200_000_000, 1024, 5, 15),
}


def _degrees(num_nodes, num_edges, rng):
    mean = num_edges / num_nodes
    sigma = 1.0
    raw = rng.lognormal(mean=np.log(mean) - sigma * sigma / 2, sigma=sigma, size=num_nodes)
    deg = np.floor(raw).astype(np.int64)
    diff = int(num_edges - deg.sum())  # fix the total to exactly num_edges
    if diff > 0:
        np.add.at(deg, rng.integers(0, num_nodes, size=diff), 1)
    while diff < 0:
        cand = np.unique(rng.integers(0, num_nodes, size=-diff * 2))
        cand = cand[deg[cand] > 0][:-diff]
        deg[cand] -= 1
        diff += len(cand)
    return deg


def make_csc(num_nodes, num_edges, seed, local_frac=0.8, window=4096, out=None, chunk=1 << 20, edge_ids="perm"):
    """CSC graph with a skewed in-degree distribution (>=10% rows with deg<=5, heavy tail)
    and community structure (``local_frac`` of in-neighbours fall within +-``window`` ids, the
    rest follow a popularity-skewed global draw) so colour affinity is meaningful.
    Generated in row chunks so the 120M-edge shapes need no E-sized temporaries; `out` may hold
    preallocated (indptr, indices, edge_ids) arrays (e.g. views of shared pinned memory).
    edge_ids: "perm" (random permutation, small graphs), "arange", or None.
    Returns (indptr int64[N+1], indices int64[E], edge_ids int64[E] | None)."""
    rng = np.random.default_rng(seed)
    deg = _degrees(num_nodes, num_edges, rng)
    if out is not None:
        indptr, indices, eids = out
    else:
        indptr = np.empty(num_nodes + 1, dtype=np.int64)
        indices = np.empty(num_edges, dtype=np.int64)
        eids = np.empty(num_edges, dtype=np.int64) if edge_ids else None
    indptr[0] = 0
    np.cumsum(deg, out=indptr[1:])
    for lo in range(0, num_nodes, chunk):
        hi = min(lo + chunk, num_nodes)
        crng = np.random.default_rng([seed, lo])
        e0, e1 = int(indptr[lo]), int(indptr[hi])
        n_e = e1 - e0
        dst = np.repeat(np.arange(lo, hi, dtype=np.int64), deg[lo:hi])
        local = np.clip(dst + crng.integers(-window, window + 1, size=n_e), 0, num_nodes - 1)
        u = crng.random(n_e)
        glob = np.minimum((u * u * u * num_nodes).astype(np.int64), num_nodes - 1)  # skew to low ids
        indices[e0:e1] = np.where(crng.random(n_e) < local_frac, local, glob)
    if eids is not None:
        if edge_ids == "perm" and num_edges <= 50_000_000:
            eids[:] = rng.permutation(num_edges)
        else:
            for lo in range(0, num_edges, 1 << 24):
                hi = min(lo + (1 << 24), num_edges)
                eids[lo:hi] = np.arange(lo, hi, dtype=np.int64)
    return indptr, indices, eids


def block_color_meta(num_nodes, num_colors, K=10):
    """Cheap stand-in colouring for the big shapes (the real colouring is single-threaded host
    code, minutes at 10M nodes): colour = contiguous id block (communities are id-local in
    make_csc), top-k = the nearest blocks with geometrically decaying affinity. Same on-disk
    layout and quirks (1-based ids, last row zero) as the reference's tables."""
    color = (np.arange(num_nodes, dtype=np.int64) * num_colors // num_nodes) + 1
    base = np.arange(1, num_colors + 1, dtype=np.int64)[:, None]
    offs = np.array([(k // 2 + 1) * (1 if k % 2 == 0 else -1) for k in range(K)], dtype=np.int64)[None, :]
    topk = base + offs
    topk[(topk < 1) | (topk > num_colors)] = 0
    score = np.where(topk != 0, 0.25 * 0.7 ** (np.abs(offs) - 1), 0.0).astype(np.float64)
    topk[-1, :] = 0
    score[-1, :] = 0.0
    return color, topk, score


def make_labels(num_nodes, num_classes, seed):
    rng = np.random.default_rng(seed + 1000)
    return rng.integers(0, num_classes, size=num_nodes, dtype=np.int64)


def train_ids(num_nodes, frac=0.6, seed=0):
    """First 60 % of the nodes are training nodes (examples/ssd_gnn_dataloader.py:546-552),
    shuffled like examples/sbatch_ssd_gnn_train.py:62-65 but with a fixed seed."""
    n_train = int(num_nodes * frac)
    rng = np.random.default_rng(seed)
    return rng.permutation(n_train).astype(np.int64)


def random_color_meta(num_nodes, num_colors, K, seed):
    """Random colour tables in the on-disk layout (color int64[N], topk int64[C,K], score f8[C,K]);
    used where running the colouring itself would be too slow (distributor benchmarks)."""
    rng = np.random.default_rng(seed)
    color = rng.integers(0, num_colors + 1, size=num_nodes, dtype=np.int64)
    color[rng.random(num_nodes) < 0.05] = 0
    topk = rng.integers(0, num_colors + 1, size=(num_colors, K), dtype=np.int64)
    topk[rng.random((num_colors, K)) < 0.15] = 0
    score = rng.random((num_colors, K)) * 0.2
    topk[-1, :] = 0  # the reference never fills the last colour's row (graph_coloring.cpp:268)
    score[-1, :] = 0.0
    return color, topk, score


def random_cache_meta(num_colors, n_domains, seed=21):
    """int32 uniform[0,4096) with 30 % zeros."""
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n_domains):
        m = rng.integers(0, 4096, size=num_colors + 1, dtype=np.int32)
        m[rng.random(num_colors + 1) < 0.3] = 0
        out.append(m)
    return out
