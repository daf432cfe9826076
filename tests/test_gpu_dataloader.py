"""End-to-end GPU test of the drop-in boundary: COALA_GNN_DataLoader over the three cache
backends on one GPU — the invariant the reference states itself
(examples/nvshmem_homogeneous_train.py:186-198): feat[i] == feature_table[input_nodes[i]]."""
import numpy as np
import pytest

import oracle
from coala_gnn_b200 import synth

pytestmark = pytest.mark.gpu


def _dataset(tmp_path, n=20000, e=240000, dim=128):
    import torch
    from COALA_GNN import MPI_Comm_Manager, UVAGraph
    from coala_gnn_b200.pybind import tensor_from_ptr
    indptr, indices, eids = synth.make_csc(n, e, 7)
    col = np.load(__import__("os").path.join(__import__("conftest").GOLDEN, "coloring_20k.npz"))
    assert int(col["n"]) == n  # not the same graph seed, but a valid colour table for n nodes
    cf, tf, sf = (str(tmp_path / x) for x in ("color.npy", "topk.npy", "score.npy"))
    np.save(cf, col["color"]), np.save(tf, col["topk"]), np.save(sf, col["score"])
    comm = MPI_Comm_Manager(0)
    keep = []

    def uva(a):
        h = torch.from_numpy(a).pin_memory()
        keep.append(h)
        return tensor_from_ptr(h.data_ptr(), a.shape, str(a.dtype), "cuda:0", owner=h)
    g = UVAGraph(uva(indptr), uva(indices), uva(eids), n, "cuda:0")
    labels = synth.make_labels(n, 19, 1)
    g.ndata["labels"] = g.ndata["label"] = torch.from_numpy(labels).cuda()
    feat = oracle.fill_features(n, dim, 17)
    feat_uva = uva(feat)
    return dict(comm=comm, g=g, feat=feat, feat_uva=feat_uva, files=(cf, tf, sf), csc=(indptr, indices, eids), labels=labels, n=n, dim=dim)


@pytest.fixture(params=["ctypes", "pybind11"])
def binding_kind(request):
    """Every loader test runs over both bindings of the C ABI: the ctypes classes and the compiled pybind11 module
    (COALA_GNN_Pybind/_native*.so, what a maintainer would build in place of M/COALA_GNN_Pybind.cu)."""
    from coala_gnn_b200 import binding
    binding.use_native(request.param == "pybind11")
    yield request.param
    binding.use_native(False)


@pytest.mark.parametrize("backend", ["isolated", "nvshmem", "nccl"])
@pytest.mark.parametrize("distribution", ["node_color", "baseline"])
def test_dataloader_iterates_with_bit_exact_features(tmp_path, backend, distribution, binding_kind):
    import torch
    from COALA_GNN import Node_Distributor, SSD_INFO, COALA_GNN_DataLoader, MultiLayerNeighborSampler
    d = _dataset(tmp_path)
    B, fan = 128, [5, 5]
    train = torch.from_numpy(synth.train_ids(d["n"], seed=1))
    nd = Node_Distributor(d["comm"], train, B, *d["files"], parsing_method=distribution)
    sampler = MultiLayerNeighborSampler(fan, seed=5)
    loader = COALA_GNN_DataLoader(SSD_INFO(1, 512, 1024, 0), nd, d["g"], sampler, B, d["dim"], fan, 2, "cuda:0",
                                  refresh_counter=3, cache_backend=backend, sim_buf=d["feat_uva"])
    indptr, indices, eids = d["csc"]
    for epoch in range(2):
        steps = 0
        for step, (input_nodes, seeds, blocks, feat) in enumerate(loader):
            seeds_h = seeds.cpu().numpy()
            assert np.array_equal(seeds_h, train.numpy()[step * B:(step + 1) * B])  # one domain: in-order striping
            inp = input_nodes.cpu().numpy()
            assert np.array_equal(feat.cpu().numpy().view(np.uint32), d["feat"][inp].view(np.uint32))
            einp, _, eblocks = oracle.sample_blocks_detailed(indptr, indices, eids, seeds_h, fan,
                                                             oracle.batch_key(5, epoch * len(loader) + step))
            assert np.array_equal(inp, einp)
            assert np.array_equal(blocks[0].edge_src.cpu().numpy(), eblocks[0]["edge_src"])
            assert np.array_equal(blocks[-1].dstdata["labels"].cpu().numpy(), d["labels"][seeds_h])
            steps += 1
        assert steps == len(loader) == len(train) // B - 1
    h, m, b = loader.COALA_GNN_Manager.COALA_GNN_Cache.stats()
    assert h > 0 and m > 0 and h + m + b > 0
    loader.print_stats()


def test_nccl_backend_reference_shaped_choreography(tmp_path, monkeypatch):
    """COALA_NCCL_LEGACY=1: the exchange driven from Python through the module's own split_node_list / nccl_get_feature /
    map_feat_data (the call sequence of S/COALA_GNN_Manager.py:143-211) instead of the single native call."""
    import torch
    from COALA_GNN import Node_Distributor, SSD_INFO, COALA_GNN_DataLoader, MultiLayerNeighborSampler
    monkeypatch.setenv("COALA_NCCL_LEGACY", "1")
    d = _dataset(tmp_path)
    B, fan = 128, [5, 5]
    train = torch.from_numpy(synth.train_ids(d["n"], seed=1))
    nd = Node_Distributor(d["comm"], train, B, *d["files"], parsing_method="baseline")
    loader = COALA_GNN_DataLoader(SSD_INFO(1, 512, 1024, 0), nd, d["g"], MultiLayerNeighborSampler(fan, seed=5), B, d["dim"], fan, 2, "cuda:0",
                                  refresh_counter=3, cache_backend="nccl", sim_buf=d["feat_uva"])
    assert loader.engine is None and loader.COALA_GNN_Manager.nccl is None
    for step, (input_nodes, seeds, blocks, feat) in enumerate(loader):
        inp = input_nodes.cpu().numpy()
        assert np.array_equal(feat.cpu().numpy().view(np.uint32), d["feat"][inp].view(np.uint32))
        if step == 5:
            break


@pytest.mark.parametrize("backend", ["isolated", "nvshmem", "nccl"])
@pytest.mark.parametrize("prefetch", [True, False])
def test_device_pipeline_two_stream_prefetch_is_bit_exact(tmp_path, backend, prefetch):
    """DevicePipeline (GPU distributor -> sampler -> gather on two streams, batch t+1 in flight while
    batch t is returned): rows, input nodes and blocks of every step equal the oracle's, across an
    epoch wrap, with slot reuse."""
    import torch
    from COALA_GNN import Node_Distributor, SSD_INFO, COALA_GNN_DataLoader, MultiLayerNeighborSampler
    from coala_gnn_b200.pipeline import DevicePipeline
    d = _dataset(tmp_path)
    B, fan = 256, [5, 5]
    train = torch.from_numpy(synth.train_ids(d["n"], seed=1))
    nd = Node_Distributor(d["comm"], train, B, *d["files"], parsing_method="node_color")
    sampler = MultiLayerNeighborSampler(fan, seed=9)
    loader = COALA_GNN_DataLoader(SSD_INFO(1, 512, 1024, 0), nd, d["g"], sampler, B, d["dim"], fan, 2, "cuda:0",
                                  refresh_counter=4, cache_backend=backend, sim_buf=d["feat_uva"])
    pipe = DevicePipeline(loader, record_gather_events=True, prefetch=prefetch)
    spe = pipe.steps_per_epoch()
    indptr, indices, eids = d["csc"]
    n_steps = spe + 5
    for t in range(n_steps):
        out = pipe.step()
        torch.cuda.current_stream().synchronize()
        seeds = train.numpy()[(t % spe) * B:(t % spe + 1) * B]
        einp, _, eblocks = oracle.sample_blocks_detailed(indptr, indices, eids, seeds, fan, oracle.batch_key(9, t))
        inp, outn, blocks = pipe.current_sampler().collect()
        assert np.array_equal(outn.cpu().numpy(), seeds)
        assert np.array_equal(inp.cpu().numpy(), einp)
        assert np.array_equal(blocks[0].edge_src.cpu().numpy(), eblocks[0]["edge_src"])
        assert np.array_equal(blocks[0].edge_id.cpu().numpy(), eblocks[0]["edge_id"])
        rows = out[:len(einp)].cpu().numpy()
        assert np.array_equal(rows.view(np.uint32), d["feat"][einp].view(np.uint32))
    g_ms = pipe.gather_ms()  # [mean CUDA-event time of the gather launches] from the native engine
    assert len(g_ms) == 1 and 0.0 < g_ms[0] < 50.0


@pytest.mark.parametrize("backend", ["isolated", "nvshmem"])
def test_papers100m_shaped_batch_gat_fanout_full_batch_size(tmp_path, backend):
    """BASELINE config C4's regime at test size: 128-d rows (R = 512 B: one 16-byte vector per lane), three layers of
    fan-out 10,10,10 at the full batch size of 1024 (up to 1.36 M rows per batch): rows bitwise equal to the table,
    input nodes and every block equal to the oracle's, over batches that overflow a small cache."""
    import torch
    from COALA_GNN import Node_Distributor, SSD_INFO, COALA_GNN_DataLoader, MultiLayerNeighborSampler
    d = _dataset(tmp_path)  # 20 000 nodes x 128-d: a 10,10,10 batch of 1024 seeds touches most of the graph
    B, fan = 1024, [10, 10, 10]
    train = torch.from_numpy(synth.train_ids(d["n"], seed=2))
    nd = Node_Distributor(d["comm"], train, B, *d["files"], parsing_method="node_color")
    sampler = MultiLayerNeighborSampler(fan, seed=21)
    loader = COALA_GNN_DataLoader(SSD_INFO(1, 512, 1024, 0), nd, d["g"], sampler, B, d["dim"], fan, 4, "cuda:0",  # 4 MB: 8192 lines < the ~19 K input nodes
                                  cache_backend=backend, sim_buf=d["feat_uva"])
    assert loader.COALA_GNN_Manager.max_sample_size == 1024 * 11 ** 3
    indptr, indices, eids = d["csc"]
    for step, (input_nodes, seeds, blocks, feat) in enumerate(loader):
        if step >= 4:
            break
        torch.cuda.synchronize()
        seeds_h = seeds.cpu().numpy()
        assert np.array_equal(seeds_h, train.numpy()[step * B:(step + 1) * B])
        einp, eout, eblocks = oracle.sample_blocks_detailed(indptr, indices, eids, seeds_h, fan, oracle.batch_key(21, step, 0))
        assert np.array_equal(input_nodes.cpu().numpy(), einp)
        assert len(blocks) == 3
        for b, e in zip(blocks, eblocks):
            assert np.array_equal(b.src_nodes.cpu().numpy(), e["src_nodes"]) and np.array_equal(b.edge_src.cpu().numpy(), e["edge_src"])
            assert np.array_equal(b.edge_dst.cpu().numpy(), e["edge_dst"]) and np.array_equal(b.edge_id.cpu().numpy(), e["edge_id"])
        rows = feat.cpu().numpy()
        assert rows.shape == (len(einp), d["dim"])
        assert np.array_equal(rows.view(np.uint32), d["feat"][einp].view(np.uint32))
    h, m, b = loader.COALA_GNN_Manager.COALA_GNN_Cache.stats()
    assert m > 0 and h + m + b > 0


def test_ids_outside_the_feature_table_are_surfaced_by_print_stats(tmp_path):
    """The manager tells the cache how many rows the host table has (sim_buf.shape[0]); the asynchronous gather
    zero-fills the row of an id beyond it instead of dereferencing it (the reference reads whatever lies there,
    M/isolated_cache.h:335-475 has no bounds check) and print_stats raises on the count."""
    import torch
    from COALA_GNN import Node_Distributor, SSD_INFO, COALA_GNN_DataLoader, MultiLayerNeighborSampler
    d = _dataset(tmp_path)
    B, fan, rows = 128, [5, 5], d["n"] - 5000
    train = torch.from_numpy(synth.train_ids(d["n"], seed=1))
    nd = Node_Distributor(d["comm"], train, B, *d["files"], parsing_method="baseline")
    loader = COALA_GNN_DataLoader(SSD_INFO(1, 512, 1024, 0), nd, d["g"], MultiLayerNeighborSampler(fan, seed=5), B, d["dim"], fan, 2,
                                  "cuda:0", cache_backend="isolated", sim_buf=d["feat_uva"][:rows])  # a table 5000 rows short
    seen_outside = 0
    for step, (input_nodes, seeds, blocks, feat) in enumerate(loader):
        inp = input_nodes.cpu().numpy()
        inside = inp < rows
        seen_outside += int((~inside).sum())
        got = feat.cpu().numpy()
        assert np.array_equal(got[inside].view(np.uint32), d["feat"][inp[inside]].view(np.uint32))
        assert not got[~inside].any()
        if step == 3:
            break
    assert seen_outside > 0
    with pytest.raises(RuntimeError, match="outside the feature table"):
        loader.print_stats()
    loader.COALA_GNN_Manager.check_refused_ids()  # the count was reset by the failing check
