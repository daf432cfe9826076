"""Minimal torch consumers of `Block` — the step right after the data path.

The reference trains DGL models (examples/models.py:8-60 `DistSAGE` = SAGEConv("mean") stack) on the
blocks its loader yields; DGL is absent here, so this module provides the same layer arithmetic in
plain torch so that the iterator contract `(input_nodes, seeds, blocks, feat)` can be exercised end to
end (`model(blocks, feat)`, `blocks[-1].dstdata['labels']`).  It is a consumer of the path, not part
of it: no custom kernels, no tuning.
"""
import torch
import torch.nn as nn


class SAGEConv(nn.Module):
    """dgl.nn.SAGEConv(in, out, "mean"): out = W_self · h_dst + W_neigh · mean_{u in N(v)} h_u + b."""

    def __init__(self, in_feats, out_feats):
        super().__init__()
        self.fc_self = nn.Linear(in_feats, out_feats, bias=False)
        self.fc_neigh = nn.Linear(in_feats, out_feats, bias=False)
        self.bias = nn.Parameter(torch.zeros(out_feats))

    def forward(self, block, h):
        n_dst = block.number_of_dst_nodes()
        src, dst = block.edges()
        agg = torch.zeros((n_dst, h.shape[1]), dtype=h.dtype, device=h.device)
        agg.index_add_(0, dst, h[src])
        deg = torch.zeros(n_dst, dtype=h.dtype, device=h.device)
        deg.index_add_(0, dst, torch.ones_like(dst, dtype=h.dtype))
        agg = agg / deg.clamp(min=1).unsqueeze(1)
        return self.fc_self(h[:n_dst]) + self.fc_neigh(agg) + self.bias


class DistSAGE(nn.Module):
    """examples/models.py:8-60 with the DGL layer replaced by the torch layer above."""

    def __init__(self, in_feats, n_hidden, n_classes, n_layers, activation=torch.relu, dropout=0.2):
        super().__init__()
        dims = [in_feats] + [n_hidden] * (n_layers - 1) + [n_classes]
        self.layers = nn.ModuleList(SAGEConv(dims[i], dims[i + 1]) for i in range(n_layers))
        self.dropout = nn.Dropout(dropout)
        self.activation = activation

    def forward(self, blocks, x):
        h = x
        for i, (layer, block) in enumerate(zip(self.layers, blocks)):
            h = layer(block, h)
            if i != len(self.layers) - 1:
                h = self.dropout(self.activation(h))
        return h
