"""SURVEY 8(f) rows next to the path: the on-disk IGB layout loaded straight into shared pinned memory
(NpySharedCSCUVA, counterpart of examples/ssd_gnn_dataloader.py:365-569) and the training loop of
examples/sbatch_ssd_gnn_train.py:129-145 driven by the drop-in loader with a torch SAGE."""
import os

import numpy as np
import pytest

import oracle
from coala_gnn_b200 import synth

pytestmark = pytest.mark.gpu


def _write_igb_layout(root, n=6000, e=60000, dim=128, classes=19):
    size = "tiny"
    paper = os.path.join(root, size, "processed", "paper")
    cites = os.path.join(root, size, "processed", "paper__cites__paper")
    os.makedirs(paper), os.makedirs(cites)
    indptr, indices, eids = synth.make_csc(n, e, 5)
    feat = oracle.fill_features(n, dim, 23)
    labels = synth.make_labels(n, classes, 5)
    np.save(os.path.join(paper, "node_feat.npy"), feat)
    np.save(os.path.join(paper, "node_label_19.npy"), labels.astype(np.float32))  # the reference stores labels as float (:383-401)
    np.save(os.path.join(cites, "csc_indptr.npy"), indptr)
    np.save(os.path.join(cites, "csc_indices.npy"), indices)
    np.save(os.path.join(cites, "csc_edge_ids.npy"), eids)
    color, topk, score = synth.block_color_meta(n, 30)
    for name, a in (("color.npy", color), ("topk.npy", topk), ("score.npy", score)):
        np.save(os.path.join(root, size, name), a)
    return size, indptr, indices, eids, feat, labels


def test_npy_dataset_and_training_loop(tmp_path):
    import torch
    from COALA_GNN import MPI_Comm_Manager, Node_Distributor, SSD_INFO, COALA_GNN_DataLoader, MultiLayerNeighborSampler
    from coala_gnn_b200.datasets import NpySharedCSCUVA
    from coala_gnn_b200.nn import DistSAGE
    size, indptr, indices, eids, feat, labels = _write_igb_layout(str(tmp_path))
    comm = MPI_Comm_Manager(0)
    ds = NpySharedCSCUVA(str(tmp_path), size, comm, "cuda:0", tag="t%d" % os.getpid())
    try:
        assert (ds.num_nodes, ds.dim, ds.cache_dim) == (6000, 128, 128)
        assert np.array_equal(ds.csc_host[0], indptr) and np.array_equal(ds.csc_host[1], indices) and np.array_equal(ds.csc_host[2], eids)
        assert np.array_equal(ds.feat_host_array().view(np.uint32), feat.view(np.uint32))
        g = ds[0]
        assert np.array_equal(g.ndata["label"].cpu().numpy(), labels)
        B, fan = 128, [5, 5]
        train_nid = torch.nonzero(g.ndata["train_mask"], as_tuple=True)[0].cpu()
        nd = Node_Distributor(comm, train_nid, B, ds.color_file, ds.topk_file, ds.score_file)
        loader = COALA_GNN_DataLoader(SSD_INFO(1, ds.cache_dim * 4, 1024, 0), nd, g, MultiLayerNeighborSampler(fan, seed=2), B, ds.dim,
                                      fan, 2, "cuda:0", cache_backend="nvshmem", sim_buf=ds.feat_data)
        torch.manual_seed(0)
        model = DistSAGE(ds.dim, 32, 19, 2).cuda()
        opt = torch.optim.Adam(model.parameters(), lr=0.01)
        losses = []
        for step, (input_nodes, seeds, blocks, x) in enumerate(loader):
            assert np.array_equal(x.cpu().numpy().view(np.uint32), feat[input_nodes.cpu().numpy()].view(np.uint32))
            y = blocks[-1].dstdata["labels"]
            assert np.array_equal(y.cpu().numpy(), labels[seeds.cpu().numpy()])
            blocks = [b.int().to("cuda:0") for b in blocks]
            loss = torch.nn.functional.cross_entropy(model(blocks, x), y.view(-1))
            opt.zero_grad()
            loss.backward()
            opt.step()
            losses.append(float(loss))
        assert len(losses) == len(loader) and all(np.isfinite(losses))
        # SAGE layer arithmetic against a dense numpy restatement on the last batch's first block
        b0 = blocks[0]
        h = x.detach().cpu().numpy().astype(np.float64)
        src, dst = b0.edge_src.cpu().numpy(), b0.edge_dst.cpu().numpy()
        nd_ = b0.number_of_dst_nodes()
        agg = np.zeros((nd_, h.shape[1]))
        np.add.at(agg, dst, h[src])
        deg = np.maximum(np.bincount(dst, minlength=nd_), 1)[:, None]
        lay = model.layers[0]
        exp = h[:nd_] @ lay.fc_self.weight.detach().cpu().numpy().T.astype(np.float64) + (agg / deg) @ lay.fc_neigh.weight.detach().cpu().numpy().T.astype(np.float64) \
            + lay.bias.detach().cpu().numpy()
        got = lay(b0, x).detach().cpu().numpy()
        assert np.allclose(got, exp, rtol=1e-3, atol=1e-3)  # fp32 layer vs fp64 restatement
        loader.print_stats()
    finally:
        ds.cleanup()


def test_loader_copy_blocks_owns_its_tensors(tmp_path):
    """Default: blocks are views of the prefetch slot (valid for `depth` more batches); copy_blocks=True clones."""
    import torch
    from COALA_GNN import MPI_Comm_Manager, Node_Distributor, SSD_INFO, COALA_GNN_DataLoader, MultiLayerNeighborSampler
    from coala_gnn_b200.datasets import NpySharedCSCUVA
    size, indptr, indices, eids, feat, labels = _write_igb_layout(str(tmp_path))
    comm = MPI_Comm_Manager(0)
    ds = NpySharedCSCUVA(str(tmp_path), size, comm, "cuda:0", tag="c%d" % os.getpid())
    try:
        g = ds[0]
        train_nid = torch.nonzero(g.ndata["train_mask"], as_tuple=True)[0].cpu()
        nd = Node_Distributor(comm, train_nid, 64, ds.color_file, ds.topk_file, ds.score_file)
        loader = COALA_GNN_DataLoader(SSD_INFO(1, 512, 1024, 0), nd, g, MultiLayerNeighborSampler([5, 5], seed=4), 64, ds.dim, [5, 5], 2,
                                      "cuda:0", cache_backend="isolated", sim_buf=ds.feat_data)
        loader.copy_blocks = True
        kept = []
        for step, (input_nodes, seeds, blocks, x) in enumerate(loader):
            kept.append((input_nodes, seeds, blocks[0].edge_src))
            if step == 9:
                break
        torch.cuda.synchronize()
        for step, (inp, seeds, esrc) in enumerate(kept):  # still intact after many more batches were drawn
            einp, _, eb = oracle.sample_blocks_detailed(indptr, indices, eids, seeds.cpu().numpy(), [5, 5], oracle.batch_key(4, step))
            assert np.array_equal(inp.cpu().numpy(), einp) and np.array_equal(esrc.cpu().numpy(), eb[0]["edge_src"])
    finally:
        ds.cleanup()
