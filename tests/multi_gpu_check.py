"""TEST INFRASTRUCTURE (launched by tests/test_gpu_multi.py, or by hand). Multi-process parity check of the N>1 path (run under torchrun, one rank per GPU):

    torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/multi_gpu_check.py [--gpus_per_domain L]

Every rank iterates the drop-in loader over a synthetic graph and checks, per batch: gathered rows bitwise
equal to the procedural feature table (re-derived on the GPU), input nodes / blocks equal to the oracle's for
this rank's seeds, seeds consistent with the distribution (1 x N domain: in-order striping; D x L domains:
the oracle's greedy assignment from the same metadata), and that all ranks of a domain together cover the
domain's seed list.  Backends: nvshmem (P2P partitioned, fd-shared shards), isolated, nccl."""
import argparse
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402  (checker only)
from COALA_GNN import MPI_Comm_Manager, Node_Distributor, SSD_INFO, COALA_GNN_DataLoader, MultiLayerNeighborSampler  # noqa: E402
from coala_gnn_b200._lib import lib, check  # noqa: E402
from coala_gnn_b200.datasets import SyntheticSharedCSCUVA  # noqa: E402
from coala_gnn_b200 import synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--gpus_per_domain", type=int, default=None)
ap.add_argument("--backends", default="nvshmem,isolated,nccl")
ap.add_argument("--steps", type=int, default=12)
args = ap.parse_args()

comm = MPI_Comm_Manager(-1, gpus_per_domain=args.gpus_per_domain)
torch.cuda.set_device(comm.device_index)
device = "cuda:%d" % comm.device_index
ds = SyntheticSharedCSCUVA("tiny", comm, device, num_colors=64)
B, fan = 256, [5, 5]
train = torch.from_numpy(synth.train_ids(ds.num_nodes, seed=3))
indptr, indices, eids = ds.csc_host
D, Lsz = comm.num_master_process, comm.local_size
ok = True
for backend in args.backends.split(","):
    for distribution in ("node_color", "baseline"):
        nd = Node_Distributor(comm, train, B, ds.color_file, ds.topk_file, ds.score_file, parsing_method=distribution)
        sampler = MultiLayerNeighborSampler(fan, seed=11, rank=comm.global_rank)
        loader = COALA_GNN_DataLoader(SSD_INFO(1, ds.cache_dim * 4, 1024, 0), nd, ds.graph, sampler, B, ds.dim, fan, 8, device,
                                      refresh_counter=3, cache_backend=backend, sim_buf=ds.feat_data)
        color, topk, score = np.load(ds.color_file), np.load(ds.topk_file), np.load(ds.score_file)
        n_checked = 0
        for step, (input_nodes, seeds, blocks, feat) in enumerate(loader):
            if step >= args.steps:
                break
            bad = C.c_int64()
            check(lib().coala_synth_check_rows(feat.data_ptr(), input_nodes.data_ptr(), input_nodes.numel(), ds.dim, ds.dim, ds.feat_seed, C.byref(bad)))
            assert bad.value == 0, "%s/%s step %d: %d corrupt rows" % (backend, distribution, step, bad.value)
            seeds_h = seeds.cpu().numpy()
            einp, _, eblocks = oracle.sample_blocks_detailed(indptr, indices, eids, seeds_h, fan, oracle.batch_key(11, step, comm.global_rank))
            assert np.array_equal(input_nodes.cpu().numpy(), einp)
            assert np.array_equal(blocks[0].edge_src.cpu().numpy(), eblocks[0]["edge_src"])
            off = step * B * comm.global_size
            if D == 1 or distribution == "baseline":
                lo = off + comm.master_process_index * B * Lsz + comm.local_rank * B
                assert np.array_equal(seeds_h, train.numpy()[lo:lo + B]), "seed striping"
            # the domain's ranks together hold a duplicate-free slice of this step's global batch
            gathered = [torch.zeros(B, dtype=torch.int64, device=device) for _ in range(comm.global_size)]
            dist.all_gather(gathered, seeds.to(device))
            allseeds = torch.cat(gathered).cpu().numpy()
            assert len(np.unique(allseeds)) == len(allseeds), "duplicate seeds across ranks"
            # every bucket ends full (D * cap seeds, cap per bucket), so the ranks partition the global batch
            assert set(allseeds.tolist()) == set(train.numpy()[off:off + B * comm.global_size].tolist()), "ranks do not partition the global batch"
            n_checked += 1
        h, m, b = loader.COALA_GNN_Manager.COALA_GNN_Cache.stats()
        if comm.global_rank == 0:
            print("ok %-8s %-10s %dx%d: %d batches, hit %d miss %d bypass %d" % (backend, distribution, D, Lsz, n_checked, h, m, b), flush=True)
        loader.scheduler.drain()
        if True:
            # the device-resident form of the same path (seeds from the GPU distributor, histograms all-gathered on the
            # device): rows, blocks and this rank's seed list against the oracle's distribution of the same histograms
            from bench import Checker, check_pipeline
            from coala_gnn_b200.pipeline import DevicePipeline
            pipe = DevicePipeline(loader, seed=23)
            chk = Checker(comm, ds, train, fan, nd, device)
            check_pipeline(pipe, chk, 3 * args.steps, oracle_batches=3 * args.steps)
            psum = chk.summary()
            assert psum["ok"] and psum["seed_lists_checked"] == 3 * args.steps * comm.global_size, psum
            if comm.global_rank == 0:
                print("ok %-8s %-10s %dx%d device pipeline: %s" % (backend, distribution, D, Lsz, psum), flush=True)
            del pipe
        del loader, nd, sampler
        torch.cuda.synchronize()
        dist.barrier()
comm.global_comm.Barrier()
ds.cleanup()
if comm.global_rank == 0:
    print("multi_gpu_check: all good")
