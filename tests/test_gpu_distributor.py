"""GPU parity tests of the seed distributor (stage 1): bit-exact bucket assignment against the
reference's golden vectors and the oracle, for D x L layouts, ties, full buckets, negative counters."""
import os

import numpy as np
import pytest

import oracle
from coala_gnn_b200 import synth
from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _run(items, off, node_id, B, L, D, color, topk, score, metas, device_ptrs=False):
    import torch
    from coala_gnn_b200.pybind import Node_distributor_pybind
    items = np.ascontiguousarray(items, dtype=np.int64)
    nd = Node_distributor_pybind.from_arrays(items.ctypes.data, node_id, B, L, D, color, topk, score, device=0)
    nd.set_meta_len(topk.shape[0] + 1)
    metas = [np.ascontiguousarray(m, dtype=np.int32) for m in metas]
    if device_ptrs:
        dm = [torch.from_numpy(m).cuda() for m in metas]
        out = torch.zeros(B * L, dtype=torch.int64, device="cuda")
        nd.distribute_node_with_affinity(off, out.data_ptr(), [m.data_ptr() for m in dm])
        res = out.cpu().numpy()
    else:
        res = np.zeros(B * L, dtype=np.int64)
        nd.distribute_node_with_affinity(off, res.ctypes.data, [m.ctypes.data for m in metas])
    return res, nd.last_bucket_len(D)


def test_matches_reference_golden_vectors():
    g = np.load(os.path.join(GOLDEN, "distributor_2k.npz"))
    col = np.load(os.path.join(GOLDEN, "coloring_2k.npz"))
    for k in range(int(g["n_cases"])):
        D, L, B, off = (int(x) for x in g["c%d_DLB" % k])
        metas = list(g["c%d_meta" % k])
        for node_id in range(D):
            out, bl = _run(g["items"], off, node_id, B, L, D, col["color"], col["topk"], col["score"], metas,
                           device_ptrs=(k % 2 == 0))
            assert np.array_equal(out, g["c%d_out" % k][node_id]), (k, node_id)
            assert bl == [B * L] * D


@pytest.mark.parametrize("D,L", [(1, 8), (2, 4), (4, 2), (8, 1), (16, 1)])
def test_config_shaped_batches_match_oracle(D, L):
    n, C, K, B = 200_000, 5000, 10, 1024
    color, topk, score = synth.random_color_meta(n, C, K, seed=D)
    items = synth.train_ids(n, seed=2)
    metas = synth.random_cache_meta(C, D, seed=21)
    for node_id in sorted({0, D // 2, D - 1}):
        exp, ebl = oracle.distribute(items, 4096, node_id, B, L, D, color, topk, score, metas)
        got, bl = _run(items, 4096, node_id, B, L, D, color, topk, score, metas)
        assert np.array_equal(got, exp)
        assert bl == list(ebl)


def test_all_ties_and_skewed_scores_fill_buckets_in_order():
    n, C, K, B, D = 50_000, 200, 10, 256, 8
    color, topk, score = synth.random_color_meta(n, C, K, seed=7)
    items = synth.train_ids(n, seed=8)
    zeros = [np.zeros(C + 1, dtype=np.int32) for _ in range(D)]
    skew = [np.full(C + 1, 1000 if j == 5 else 1, dtype=np.int32) for j in range(D)]  # everyone prefers domain 5
    for metas in (zeros, skew):
        for node_id in (0, 5, 7):
            exp, _ = oracle.distribute(items, 0, node_id, B, 1, D, color, topk, score, metas)
            got, _ = _run(items, 0, node_id, B, 1, D, color, topk, score, metas)
            assert np.array_equal(got, exp)


def test_negative_counters_take_the_exact_serial_path():
    n, C, K, B, D = 20_000, 100, 10, 64, 4
    color, topk, score = synth.random_color_meta(n, C, K, seed=11)
    items = synth.train_ids(n, seed=12)
    rng = np.random.default_rng(0)
    metas = [rng.integers(-50, 50, size=C + 1).astype(np.int32) for _ in range(D)]
    for node_id in range(D):
        exp, ebl = oracle.distribute(items, 10, node_id, B, 2, D, color, topk, score, metas)
        got, bl = _run(items, 10, node_id, B, 2, D, color, topk, score, metas)
        n_valid = min(B * 2, int(ebl[node_id]))
        assert np.array_equal(got[:n_valid], exp[:n_valid])
        assert bl == list(ebl)


def test_npy_files_roundtrip(tmp_path):
    import torch
    from coala_gnn_b200.pybind import Node_distributor_pybind
    col = np.load(os.path.join(GOLDEN, "coloring_2k.npz"))
    cf, tf, sf = (str(tmp_path / x) for x in ("color.npy", "topk.npy", "score.npy"))
    np.save(cf, col["color"]), np.save(tf, col["topk"]), np.save(sf, col["score"])
    items = torch.from_numpy(synth.train_ids(2000, seed=5))
    nd = Node_distributor_pybind(items.data_ptr(), 1, 32, 2, 3, cf, tf, sf)
    C = nd.get_num_colors()
    assert C == col["topk"].shape[0]
    metas = [torch.from_numpy(m[:C].copy()) for m in synth.random_cache_meta(C, 3)]  # reference-sized histograms
    out = torch.zeros(64, dtype=torch.int64)
    nd.distribute_node_with_affinity(5, out.data_ptr(), [m.data_ptr() for m in metas])
    padded = [np.concatenate([m.numpy(), [0]]).astype(np.int32) for m in metas]
    exp, _ = oracle.distribute(items.numpy(), 5, 1, 32, 2, 3, col["color"], col["topk"], col["score"], padded)
    assert np.array_equal(out.numpy(), exp)
