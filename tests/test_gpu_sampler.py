"""GPU parity tests of the UVA neighbour sampler + to_block (stage 2): bit-exact against the
oracle (same RNG definition) at sampled and full fan-out, CSC resident in pinned host memory."""
import numpy as np
import pytest

import oracle
from coala_gnn_b200 import synth

pytestmark = pytest.mark.gpu


def _uva_graph(indptr, indices, eids, pinned=True):
    import torch
    from coala_gnn_b200.pybind import tensor_from_ptr
    from coala_gnn_b200.sampler import UVAGraph
    ts = []
    for a in (indptr, indices, eids):
        if a is None:
            ts.append(None)
        elif pinned:
            h = torch.from_numpy(a).pin_memory()
            ts.append(tensor_from_ptr(h.data_ptr(), a.shape, torch.int64, "cuda:0", owner=h))  # UVA alias, like the reference
        else:
            ts.append(torch.from_numpy(a).cuda())
    return UVAGraph(ts[0], ts[1], ts[2], len(indptr) - 1, "cuda:0")


def _compare(blocks, exp_blocks):
    assert len(blocks) == len(exp_blocks)
    for b, e in zip(blocks, exp_blocks):
        assert np.array_equal(b.src_nodes.cpu().numpy(), e["src_nodes"])
        assert np.array_equal(b.dst_nodes.cpu().numpy(), e["dst_nodes"])
        assert np.array_equal(b.edge_src.cpu().numpy(), e["edge_src"])
        assert np.array_equal(b.edge_dst.cpu().numpy(), e["edge_dst"])
        assert np.array_equal(b.edge_id.cpu().numpy(), e["edge_id"])


@pytest.mark.parametrize("fanouts", [[5, 5], [10, 5, 5], [10, 10, 10], [32]])
@pytest.mark.parametrize("pinned", [True, False])
def test_sampled_blocks_bit_exact_vs_oracle(fanouts, pinned):
    import torch
    from coala_gnn_b200.sampler import MultiLayerNeighborSampler, batch_key
    indptr, indices, eids = synth.make_csc(20000, 240000, 7)
    g = _uva_graph(indptr, indices, eids, pinned)
    sampler = MultiLayerNeighborSampler(fanouts, seed=11, rank=2)
    items = synth.train_ids(20000, seed=3)
    for it in range(3):
        seeds = items[it * 256:(it + 1) * 256] if it < 2 else items[600:610]  # ragged last batch
        inp, outn, blocks = sampler.sample(g, torch.from_numpy(seeds).cuda())
        bkey = oracle.batch_key(11, it, 2)
        assert bkey == batch_key(11, it, 2)
        einp, eout, eblocks = oracle.sample_blocks_detailed(indptr, indices, eids, seeds, fanouts, bkey)
        assert np.array_equal(inp.cpu().numpy(), einp) and np.array_equal(outn.cpu().numpy(), eout)
        _compare(blocks, eblocks)


@pytest.mark.parametrize("indptr_at", ["host", "hbm", "hbm_shared", "graph", "graph_shared"])
def test_graph_placement_does_not_change_the_sample(indptr_at):
    """indptr read through UVA, from the sampler's HBM copy (the default), or from another sampler's copy (the slots of one
    engine share slot 0's); with `graph` the neighbour lists are staged as well (COALA_HBM_GRAPH=1): same blocks, bit for bit."""
    import torch
    from coala_gnn_b200.sampler import MultiLayerNeighborSampler
    indptr, indices, eids = synth.make_csc(20000, 240000, 9)
    g = _uva_graph(indptr, indices, eids, True)
    whole = indptr_at.startswith("graph")
    owner = MultiLayerNeighborSampler([10, 5], seed=4, rank=1, hbm_indptr=True, hbm_graph=whole)
    owner._ensure(g, 256)
    sampler = {"host": lambda: MultiLayerNeighborSampler([10, 5], seed=4, rank=1, hbm_indptr=False),
               "hbm": lambda: MultiLayerNeighborSampler([10, 5], seed=4, rank=1),
               "hbm_shared": lambda: MultiLayerNeighborSampler([10, 5], seed=4, rank=1, indptr_from=owner),
               "graph": lambda: MultiLayerNeighborSampler([10, 5], seed=4, rank=1, hbm_graph=True),
               "graph_shared": lambda: MultiLayerNeighborSampler([10, 5], seed=4, rank=1, hbm_graph=True, indptr_from=owner)}[indptr_at]()
    assert sampler.hbm_indptr == (indptr_at != "host") and sampler.hbm_graph == whole
    items = synth.train_ids(20000, seed=5)
    for it in range(2):
        seeds = items[it * 256:(it + 1) * 256]
        inp, outn, blocks = sampler.sample(g, torch.from_numpy(seeds).cuda())
        einp, eout, eblocks = oracle.sample_blocks_detailed(indptr, indices, eids, seeds, [10, 5], oracle.batch_key(4, it, 1))
        assert np.array_equal(inp.cpu().numpy(), einp) and np.array_equal(outn.cpu().numpy(), eout)
        _compare(blocks, eblocks)
    assert (sampler._indptr_from is owner) == indptr_at.endswith("_shared")
    del sampler, owner


def test_full_fanout_is_exact_csc_and_edge_ids_optional():
    """A graph whose maximum in-degree is <= the fan-out: every row keeps all in-edges in CSC order."""
    import torch
    from coala_gnn_b200.sampler import MultiLayerNeighborSampler
    rng = np.random.default_rng(0)
    n = 5000
    deg = rng.integers(0, 9, n)
    indptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(deg, out=indptr[1:])
    indices = rng.integers(0, n, indptr[-1]).astype(np.int64)
    g = _uva_graph(indptr, indices, None)
    seeds = rng.permutation(n)[:512].astype(np.int64)
    inp, outn, blocks = MultiLayerNeighborSampler([8, 8]).sample(g, torch.from_numpy(seeds).cuda(), rng_key=5)
    b = blocks[-1]
    exp_src = np.concatenate([indices[indptr[v]:indptr[v + 1]] for v in seeds])
    assert np.array_equal(b.src_nodes.cpu().numpy()[b.edge_src.cpu().numpy()], exp_src)
    assert np.array_equal(b.edge_id.cpu().numpy(), np.concatenate([np.arange(indptr[v], indptr[v + 1]) for v in seeds]))
    einp, _, eblocks = oracle.sample_blocks_detailed(indptr, indices, None, seeds, [8, 8], 5)
    assert np.array_equal(inp.cpu().numpy(), einp)
    _compare(blocks, eblocks)


def test_config_shaped_batch_properties():
    """SAGE 5,5 batch 1024 on a 1M-node graph: size-independent properties at BASELINE shape."""
    import torch
    from coala_gnn_b200.sampler import MultiLayerNeighborSampler
    indptr, indices, _ = synth.make_csc(1_000_000, 12_000_000, 2)
    g = _uva_graph(indptr, indices, None)
    g.ndata["labels"] = torch.from_numpy(synth.make_labels(1_000_000, 19, 2)).cuda()
    seeds = synth.train_ids(1_000_000, seed=0)[:1024]
    sampler = MultiLayerNeighborSampler([5, 5], seed=1)
    inp, outn, blocks = sampler.sample(g, torch.from_numpy(seeds).cuda())
    assert sampler.max_sample() == 1024 * 36
    inp_h = inp.cpu().numpy()
    assert len(np.unique(inp_h)) == len(inp_h) <= 1024 * 36
    assert np.array_equal(outn.cpu().numpy(), seeds)
    assert np.array_equal(blocks[-1].dstdata["labels"].cpu().numpy(), synth.make_labels(1_000_000, 19, 2)[seeds])
    for b in blocks:
        dst = b.dst_nodes.cpu().numpy()
        deg = indptr[dst + 1] - indptr[dst]
        per_dst = np.bincount(b.edge_dst.cpu().numpy(), minlength=len(dst))
        assert np.array_equal(per_dst, np.minimum(deg, 5))
        pos = b.edge_id.cpu().numpy()
        assert np.array_equal(b.src_nodes.cpu().numpy()[b.edge_src.cpu().numpy()], indices[pos])
        assert np.all(pos >= indptr[dst][b.edge_dst.cpu().numpy()]) and np.all(pos < indptr[dst + 1][b.edge_dst.cpu().numpy()])
    einp, _, _ = oracle.sample_blocks_detailed(indptr, indices, None, seeds, [5, 5], oracle.batch_key(1, 0))
    assert np.array_equal(inp_h, einp)


def test_degree_boundaries_long_rows_and_empty_batch():
    """Rows of in-degree 0, 1, fan-out-1, fan-out, fan-out+1, 63, 64, 65 and 200 (the whole-row read covers
    <= 64 neighbours, longer rows take the per-pick path), the smallest and the largest supported fan-out; a
    batch of zero seeds; a fan-out above the supported 32 is refused."""
    import torch
    from coala_gnn_b200.sampler import MultiLayerNeighborSampler
    rng = np.random.default_rng(4)
    n = 3000
    deg = np.array([0, 1, 4, 5, 6, 63, 64, 65, 200], dtype=np.int64)[np.arange(n) % 9]
    indptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(deg, out=indptr[1:])
    indices = rng.integers(0, n, indptr[-1]).astype(np.int64)
    eids = rng.permutation(indptr[-1]).astype(np.int64)
    g = _uva_graph(indptr, indices, eids)
    seeds = rng.permutation(n)[:300].astype(np.int64)
    for fanouts in ([5, 5], [1, 32], [32, 2]):
        sampler = MultiLayerNeighborSampler(fanouts, seed=3)
        for it in range(2):
            inp, outn, blocks = sampler.sample(g, torch.from_numpy(seeds).cuda())
            einp, eout, eblocks = oracle.sample_blocks_detailed(indptr, indices, eids, seeds, fanouts, oracle.batch_key(3, it))
            assert np.array_equal(inp.cpu().numpy(), einp) and np.array_equal(outn.cpu().numpy(), eout)
            _compare(blocks, eblocks)
        inp, outn, blocks = sampler.sample(g, torch.zeros(0, dtype=torch.int64, device="cuda"))
        assert inp.numel() == 0 and outn.numel() == 0
        assert all(b.num_edges() == 0 and b.number_of_src_nodes() == 0 for b in blocks)
    with pytest.raises(RuntimeError):
        MultiLayerNeighborSampler([33]).sample(g, torch.from_numpy(seeds).cuda())
