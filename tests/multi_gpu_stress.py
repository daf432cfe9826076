"""TEST INFRASTRUCTURE (launched by tests/test_gpu_multi.py, `bench.py --stress`, or by hand).  Cross-GPU eviction storm on
the partitioned cache (the seqlock reader and the remote fill path under real concurrency):

    torchrun --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 tests/multi_gpu_stress.py [--launches 10000]

Every rank of one 1 x G domain hammers the SAME few sets of every shard at once: the id pool is the ids that fall into
sets 0..n_sets-1 of each owner (156 ids per 32-way set on the `tiny` table with a 1 MB shard), so every launch mixes
local hits, peer hits over NVLink, local fills, remote fills under the owner's set lock, lines being evicted while peers
read them, and bypasses.  Launches alternate between the two copy engines and are NOT synchronised across ranks.  Every
output row of every launch is checked bitwise against the procedural feature table (coala_synth_check_rows =
the reference's invariant feat[i] == table[idx[i]], examples/nvshmem_homogeneous_train.py:186-198); at the end the
shards are exported and checked for duplicate lines and the counters for hit + miss + bypass == requests."""
import argparse
import ctypes as C
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from COALA_GNN import MPI_Comm_Manager, Node_Distributor, SSD_INFO, COALA_GNN_DataLoader, MultiLayerNeighborSampler  # noqa: E402
from coala_gnn_b200._lib import lib, check  # noqa: E402
from coala_gnn_b200.datasets import SyntheticSharedCSCUVA  # noqa: E402
from coala_gnn_b200 import synth  # noqa: E402


def eviction_storm(comm, device, launches, rows=4096, sets=2, check_every=1, tag=None):
    """Runs the storm on the ranks of `comm` (one 1 x G domain) and returns the summed counters (same dict on every rank)."""
    G = comm.local_size
    ds = SyntheticSharedCSCUVA("tiny", comm, device, num_colors=64, tag=tag)
    try:
        train = torch.from_numpy(synth.train_ids(ds.num_nodes, seed=3))
        nd = Node_Distributor(comm, train, 256, ds.color_file, ds.topk_file, ds.score_file)
        loader = COALA_GNN_DataLoader(SSD_INFO(1, ds.cache_dim * 4, 1024, 0), nd, ds.graph, MultiLayerNeighborSampler([5, 5], seed=1, rank=comm.global_rank),
                                      256, ds.dim, [5, 5], 1, device, cache_backend="nvshmem", sim_buf=ds.feat_data)  # 1 MB shard: 64 sets x 32 ways
        cache = loader.COALA_GNN_Manager.COALA_GNN_Cache
        cache.set_num_nodes(ds.num_nodes)
        S = cache.num_sets()
        ids = np.arange(ds.num_nodes, dtype=np.int64)
        pool = ids[((ids // G) % S) < sets]  # partitioned hashing: owner id % G, set (id / G) % S (M/cache_kernel.cu:11, M/nvshmem_cache.h:193-196)
        # ids per (owner, set) = N / (G * S): 156 at G = 2, 39 at G = 8 on the `tiny` table - always more than the 32 ways
        assert len(pool) > 32 * sets * G, "pool must overflow the sets it maps to"
        rng = np.random.default_rng(100 + comm.global_rank)
        n_lists = 64
        lists = [torch.from_numpy(pool[rng.integers(0, len(pool), rows)]).to(device) for _ in range(n_lists)]
        out = torch.empty((rows, ds.dim), dtype=torch.float32, device=device)
        st = torch.cuda.current_stream().cuda_stream
        bad = C.c_int64()
        total_bad, total = 0, 0
        comm.global_comm.Barrier()
        t0 = time.time()
        for it in range(launches):
            idx = lists[(it * 7 + comm.global_rank) % n_lists]
            cache.set_copy_engine(it & 1 if (it >> 3) & 1 else (it >> 1) & 1)  # both engines, in changing phase against the peers
            cache.read_feature_async(out.data_ptr(), idx.data_ptr(), idx.numel(), 0, st)
            total += idx.numel()
            if it % check_every == 0:
                check(lib().coala_synth_check_rows(out.data_ptr(), idx.data_ptr(), idx.numel(), ds.dim, ds.dim, ds.feat_seed, C.byref(bad)))
                total_bad += int(bad.value)
                if bad.value:
                    print("rank %d launch %d: %d corrupt rows" % (comm.global_rank, it, bad.value), flush=True)
                    break
        torch.cuda.synchronize()
        dt = time.time() - t0
        h, m, b = cache.stats()
        refused = cache.refused_ids()
        comm.global_comm.Barrier()  # every rank has stopped writing: the shards are quiescent now
        keys, cmeta, scnt = cache.export_state()
        res = keys != np.uint64(0xFFFFFFFFFFFFFFFF)
        line_ids = keys[res].astype(np.int64)
        dup = len(line_ids) - len(np.unique(line_ids))
        wrong_owner = int(np.count_nonzero(line_ids % G != comm.local_rank))
        t = torch.tensor([total_bad, total, h, m, b, refused, dup, wrong_owner], dtype=torch.int64, device=device)
        if comm.global_size > 1:
            dist.all_reduce(t)
        tb, tot, hs, ms, bs, rf, dups, wo = [int(x) for x in t.tolist()]
        ok = tb == 0 and hs + ms + bs == tot and rf == 0 and dups == 0 and wo == 0 and ms > 0 and hs > 0
        del loader, cache, nd
        torch.cuda.synchronize()
        comm.global_comm.Barrier()
        return {"ranks": comm.global_size, "launches_per_rank": launches, "rows_per_launch": rows, "seconds": round(dt, 2), "rows_checked": tot,
                "corrupt_rows": tb, "hit": hs, "miss": ms, "bypass": bs, "refused": rf, "duplicate_lines": dups, "lines_in_wrong_shard": wo, "ok": ok}
    finally:
        ds.cleanup()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--launches", type=int, default=10000)
    ap.add_argument("--rows", type=int, default=4096, help="ids per launch")
    ap.add_argument("--sets", type=int, default=2, help="sets of every shard the id pool falls into")
    ap.add_argument("--check_every", type=int, default=1)
    args = ap.parse_args()
    comm = MPI_Comm_Manager(-1)
    torch.cuda.set_device(comm.device_index)
    r = eviction_storm(comm, "cuda:%d" % comm.device_index, args.launches, args.rows, args.sets, args.check_every)
    if comm.global_rank == 0:
        print("stress " + " ".join("%s=%s" % kv for kv in r.items()), flush=True)
        print("multi_gpu_stress: all good" if r["ok"] else "multi_gpu_stress: FAILED", flush=True)
    sys.exit(0 if r["ok"] else 1)
