"""GPU mode of the colour-metadata generator through the C ABI.
* colouring: bit-exact against the numpy specification of the level-synchronous traversal
  (oracle.levelsync_color_graph; same seeds / colour ids as the reference's rand() stream);
* affinity: bit-exact against oracle.affinity_from_counts, and against the REFERENCE's own
  cpu_calculate_color_affinity output (golden fixtures) on the reference's colouring: scores within
  1e-12 relative (summation order differs), top-k lists identical wherever scores are not tied."""
import ctypes

import numpy as np
import pytest

import oracle
from coala_gnn_b200 import synth

pytestmark = pytest.mark.gpu
K = 10


def _golden(name):
    import os
    return np.load(os.path.join(os.path.dirname(__file__), "golden", name + ".npz"))


def _graph(name):
    g = _golden(name)
    n, e, seed = int(g["n"]), int(g["e"]), int(g["seed"])
    indptr, indices, _ = synth.make_csc(n, e, seed)
    return g, n, indptr, indices, np.arange(int(n * 0.6), dtype=np.int64)


def _new_gc(n, indptr, indices):
    from coala_gnn_b200.pybind import Graph_Coloring
    ctypes.CDLL("libc.so.6").srand(1)  # the reference never seeds rand()
    gc = Graph_Coloring(n)
    gc.set_adj_csc(indptr.ctypes.data, indices.ctypes.data)
    color = np.zeros(n, dtype=np.int64)
    gc.set_color_buffer(color.ctypes.data)
    return gc, color


def _affinity(gc, mode):
    nc = gc.get_num_color()
    tk, sc = np.zeros(nc * K, dtype=np.int64), np.zeros(nc * K, dtype=np.float64)
    gc.set_topk_color_buffer(tk.ctypes.data)
    gc.set_topk_affinity_buffer(sc.ctypes.data)
    if mode == "gpu":
        gc.gpu_calculate_color_affinity(0)
    else:
        gc.cpu_calculate_color_affinity()
    return tk.reshape(nc, K), sc.reshape(nc, K)


@pytest.mark.parametrize("name", ["coloring_2k", "coloring_20k"])
def test_gpu_colouring_matches_levelsync_specification(name):
    g, n, indptr, indices, train = _graph(name)
    exp_color, exp_hop, exp_nc = oracle.levelsync_color_graph(indptr, indices, train)
    gc, color = _new_gc(n, indptr, indices)
    gc.gpu_color_graph_optimized(train.ctypes.data, len(train), 0)
    assert gc.get_num_color() == exp_nc == g["topk"].shape[0]  # same seeds and colour ids as the reference
    assert np.array_equal(color, exp_color)
    assert np.array_equal(gc.get_hops(n), exp_hop)
    assert gc.get_num_color_node() == int((exp_color != 0).sum())
    tk, sc = _affinity(gc, "gpu")
    etk, esc = oracle.affinity_from_counts(indptr, indices, exp_color, exp_hop, exp_nc, K)
    assert np.array_equal(tk, etk)
    assert np.array_equal(sc.view(np.uint64), esc.view(np.uint64))
    assert not tk[-1].any()  # the last colour's row stays empty, like the reference


@pytest.mark.parametrize("parts", [None, "3", "1000000"])
@pytest.mark.parametrize("name", ["coloring_2k", "coloring_20k"])
def test_gpu_affinity_on_reference_colouring(name, parts, monkeypatch):
    """parts: the sort runs in that many passes over disjoint colour ranges (graphs above 2^30 edges do that on
    their own; forced here on small graphs, "1000000" = one colour per pass)."""
    if parts:
        monkeypatch.setenv("COALA_GC_PARTS", parts)
    g, n, indptr, indices, train = _graph(name)
    gc, color = _new_gc(n, indptr, indices)
    gc.cpu_color_graph_optimized(train.ctypes.data, len(train))  # host mode == reference (pinned by the golden fixture)
    assert np.array_equal(color, g["color"])
    tk, sc = _affinity(gc, "gpu")
    etk, esc = oracle.affinity_from_counts(indptr, indices, color, gc.get_hops(n), gc.get_num_color(), K)
    assert np.array_equal(tk, etk)
    assert np.array_equal(sc.view(np.uint64), esc.view(np.uint64))
    ref_tk, ref_sc = g["topk"], g["score"]
    assert np.allclose(sc, ref_sc, rtol=1e-12, atol=0)
    differ = tk != ref_tk
    if differ.any():  # only inside groups of tied scores
        r, c = np.nonzero(differ)
        for i, j in zip(r, c):
            tied = np.isclose(ref_sc[i], ref_sc[i, j], rtol=1e-12, atol=0)
            if not tied[-1]:  # a tie group cut by K may keep different members; its scores are compared above
                assert sorted(tk[i, tied]) == sorted(ref_tk[i, tied])
    # the host mode still writes the reference's bits after the GPU mode ran on the same object
    tk2, sc2 = _affinity(gc, "cpu")
    assert np.array_equal(tk2, ref_tk) and np.array_equal(sc2.view(np.uint64), ref_sc.view(np.uint64))


def test_gpu_colour_metadata_feeds_the_distributor():
    """The distributor consumes GPU-mode metadata like the host mode's (same shapes / value ranges)."""
    g, n, indptr, indices, train = _graph("coloring_2k")
    gc, color = _new_gc(n, indptr, indices)
    gc.gpu_color_graph_optimized(train.ctypes.data, len(train), 0)
    tk, sc = _affinity(gc, "gpu")
    nc = gc.get_num_color()
    assert color.max() == nc and color.min() >= 0
    assert tk.max() <= nc and (sc >= 0).all() and (sc <= 1.0 + 1e-12).all()
    D, L, B = 2, 1, 64
    from coala_gnn_b200.pybind import Node_distributor_pybind
    items = np.ascontiguousarray(synth.train_ids(n, seed=4), dtype=np.int64)
    metas = [np.ascontiguousarray(m, dtype=np.int32) for m in synth.random_cache_meta(nc, D)]
    outs = []
    for j in range(D):
        nd = Node_distributor_pybind.from_arrays(items.ctypes.data, j, B, L, D, color, tk, sc, device=0)
        nd.set_meta_len(nc + 1)
        out = np.zeros(B * L, dtype=np.int64)
        nd.distribute_node_with_affinity(0, out.ctypes.data, [m.ctypes.data for m in metas])
        exp, _ = oracle.distribute(items, 0, j, B, L, D, color, tk, sc, metas)
        assert np.array_equal(out, exp)
        outs.append(out)
    assert np.array_equal(np.sort(np.concatenate(outs)), np.sort(items[:D * L * B]))


def test_generate_color_data_script_gpu_mode(tmp_path):
    """examples/generate_color_data.py --mode gpu over the on-disk CSC layout: files equal the specification's arrays."""
    import importlib.util
    import os
    g, n, indptr, indices, train = _graph("coloring_2k")
    cites = os.path.join(str(tmp_path), "tiny", "processed", "paper__cites__paper")
    os.makedirs(cites)
    np.save(os.path.join(cites, "csc_indptr.npy"), indptr)
    np.save(os.path.join(cites, "csc_indices.npy"), indices)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("generate_color_data", os.path.join(root, "examples", "generate_color_data.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    out = os.path.join(str(tmp_path), "meta")
    mod.main(["--path", str(tmp_path), "--dataset_size", "tiny", "--mode", "gpu", "--out_path", out])
    exp_color, exp_hop, exp_nc = oracle.levelsync_color_graph(indptr, indices, train)
    etk, esc = oracle.affinity_from_counts(indptr, indices, exp_color, exp_hop, exp_nc, K)
    assert np.array_equal(np.load(os.path.join(out, "color.npy")), exp_color)
    assert np.array_equal(np.load(os.path.join(out, "topk.npy")), etk)
    assert np.array_equal(np.load(os.path.join(out, "score.npy")).view(np.uint64), esc.view(np.uint64))
