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


def make_csc(num_nodes, num_edges, seed, local_frac=0.8, window=4096):
    """CSC graph with a skewed in-degree distribution (>=10% rows with deg<=5, heavy tail)
    and community structure (``local_frac`` of in-neighbours fall within +-``window`` ids, the
    rest follow a popularity-skewed global draw) so colour affinity is meaningful.
    Returns (indptr int64[N+1], indices int64[E], edge_ids int64[E])."""
    rng = np.random.default_rng(seed)
    mean = num_edges / num_nodes
    sigma = 1.0
    raw = rng.lognormal(mean=np.log(mean) - sigma * sigma / 2, sigma=sigma, size=num_nodes)
    deg = np.floor(raw).astype(np.int64)
    # fix the total to exactly num_edges
    diff = int(num_edges - deg.sum())
    if diff > 0:
        add = rng.integers(0, num_nodes, size=diff)
        np.add.at(deg, add, 1)
    elif diff < 0:
        diff = -diff
        while diff > 0:
            cand = rng.integers(0, num_nodes, size=diff * 2)
            cand = cand[deg[cand] > 0][:diff]
            cand = np.unique(cand)
            deg[cand] -= 1
            diff -= len(cand)
    indptr = np.zeros(num_nodes + 1, dtype=np.int64)
    np.cumsum(deg, out=indptr[1:])
    E = int(indptr[-1])
    dst = np.repeat(np.arange(num_nodes, dtype=np.int64), deg)
    is_local = rng.random(E) < local_frac
    off = rng.integers(-window, window + 1, size=E)
    local = np.clip(dst + off, 0, num_nodes - 1)
    u = rng.random(E)
    glob = np.minimum((u * u * u * num_nodes).astype(np.int64), num_nodes - 1)  # skew to low ids
    indices = np.where(is_local, local, glob).astype(np.int64)
    del dst, is_local, off, local, u, glob
    edge_ids = rng.permutation(E).astype(np.int64) if E <= 50_000_000 else np.arange(E, dtype=np.int64)
    return indptr, indices, edge_ids


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
