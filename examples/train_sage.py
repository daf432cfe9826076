#!/usr/bin/env python
"""Counterpart of the reference's examples/sbatch_ssd_gnn_train.py (:50-195, :200-303) over this repo's
drop-in packages: same flags for the data path (--feat_cpu --distribution --cache_backend --cache_size
--fan_out --batch_size --refresh_counter), same loop `for step, (input_nodes, seeds, blocks, feat) in
enumerate(train_loader)`, SLURM/MPI bootstrap replaced by torchrun's env, DGL model replaced by the
torch SAGE of coala_gnn_b200.nn.  Data: a synthetic shape (--shape) or an on-disk IGB-layout directory
(--path/--dataset_size, the reference's file names).

    python examples/train_sage.py --shape tiny --epochs 1
    torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 examples/train_sage.py --shape igb-medium
"""
import argparse
import os
import sys
import time

import torch
import torch.nn as nn
import torch.optim as optim

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from COALA_GNN import MPI_Comm_Manager, Node_Distributor, SSD_INFO, COALA_GNN_DataLoader, MultiLayerNeighborSampler  # noqa: E402
from coala_gnn_b200.datasets import SyntheticSharedCSCUVA, NpySharedCSCUVA  # noqa: E402
from coala_gnn_b200.nn import DistSAGE  # noqa: E402


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--path", type=str, default=None, help="IGB-layout dataset root (reference flag); omit to use --shape")
    p.add_argument("--dataset_size", type=str, default="tiny")
    p.add_argument("--shape", type=str, default="tiny", help="synthetic shape from coala_gnn_b200.synth.SHAPES")
    p.add_argument("--num_classes", type=int, default=19)
    p.add_argument("--feat_cpu", action="store_true", default=True)
    p.add_argument("--fan_out", type=str, default="5,5")
    p.add_argument("--batch_size", type=int, default=1024)
    p.add_argument("--num_layers", type=int, default=2)
    p.add_argument("--hidden_channels", type=int, default=128)
    p.add_argument("--learning_rate", type=float, default=0.01)
    p.add_argument("--decay", type=float, default=0.001)
    p.add_argument("--epochs", type=int, default=1)
    p.add_argument("--max_steps", type=int, default=0, help="stop each epoch after this many steps (0 = full epoch)")
    p.add_argument("--cache_size", type=int, default=4096, help="MB per GPU")
    p.add_argument("--distribution", type=str, default="node_color", choices=["node_color", "baseline"])
    p.add_argument("--cache_backend", type=str, default="nvshmem", choices=["nvshmem", "p2p", "isolated", "nccl"])
    p.add_argument("--refresh_counter", type=int, default=10)
    p.add_argument("--gpus_per_domain", type=int, default=None, help="cache-domain size (stands in for SLURM_NODEID grouping)")
    return p.parse_args()


def train(g, args, device, comm, ds):
    train_nid = torch.nonzero(g.ndata["train_mask"], as_tuple=True)[0].clone().cpu()
    gen = torch.Generator().manual_seed(0)  # same shuffle on every rank (the reference relies on an unseeded global RNG)
    train_nid = train_nid[torch.randperm(train_nid.size(0), generator=gen)]
    nd = Node_Distributor(comm, train_nid, args.batch_size, ds.color_file, ds.topk_file, ds.score_file,
                          parsing_method=args.distribution)
    fan_out = [int(f) for f in args.fan_out.split(",")]
    sampler = MultiLayerNeighborSampler(fan_out, seed=0, rank=comm.global_rank)
    loader = COALA_GNN_DataLoader(SSD_INFO(1, ds.cache_dim * 4, 1024, 0), nd, g, sampler, args.batch_size, ds.dim, fan_out,
                                  args.cache_size, device, refresh_counter=args.refresh_counter, cache_backend=args.cache_backend,
                                  sim_buf=ds.feat_data, shuffle=False)
    model = DistSAGE(ds.dim, args.hidden_channels, args.num_classes, args.num_layers).to(device)
    if comm.global_size > 1:
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[comm.device_index])
    loss_fcn = nn.CrossEntropyLoss().to(device)
    optimizer = optim.Adam(model.parameters(), lr=args.learning_rate, weight_decay=args.decay)
    model.train()
    count = num_sampled_nodes = 0
    for epoch in range(args.epochs):
        epoch_start = time.time()
        for step, (input_nodes, seeds, blocks, fetch_feature) in enumerate(loader):
            num_sampled_nodes += len(input_nodes)
            batch_labels = blocks[-1].dstdata["labels"]
            blocks = [block.int().to(device) for block in blocks]
            batch_pred = model(blocks, fetch_feature)
            loss = loss_fcn(batch_pred, batch_labels.view(-1).to(device))
            optimizer.zero_grad()
            loss.backward()
            optimizer.step()
            count += 1
            if step % 100 == 0 and comm.global_rank == 0:
                print(f"epoch {epoch} step {step} loss {loss.item():.4f}", flush=True)
            if args.max_steps and step + 1 >= args.max_steps:
                break
        torch.cuda.synchronize()
        if comm.global_rank == 0:
            print(f"Epoch Time: {time.time() - epoch_start:.3f}  iterations: {count}  sampled nodes: {num_sampled_nodes}")
        loader.print_stats()
        if args.max_steps:
            break
    return loss.item()


def main():
    args = parse_args()
    comm = MPI_Comm_Manager(-1, gpus_per_domain=args.gpus_per_domain)
    torch.cuda.set_device(comm.device_index)
    device = "cuda:%d" % comm.device_index
    if args.path:
        ds = NpySharedCSCUVA(args.path, args.dataset_size, comm, device, num_classes=args.num_classes)
    else:
        ds = SyntheticSharedCSCUVA(args.shape, comm, device, num_classes=args.num_classes)
    try:
        loss = train(ds[0], args, device, comm, ds)
        if comm.global_rank == 0:
            print("final loss %.4f" % loss)
    finally:
        comm.global_comm.Barrier()
        ds.cleanup()


if __name__ == "__main__":
    main()
