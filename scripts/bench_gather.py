"""Micro-benchmark of the gather kernel alone, per regime, in one process:

    python scripts/bench_gather.py [--nodes N] [--cache_mb MB] [--dim D] [--gpus 1|2] [--rows R]

  iso_cold / iso_hit      isolated cache on GPU 0: all-miss (host link) and all-hit (HBM)
  part1_hit               partitioned, 2 shards emulated on GPU 0 (system-scope code path, local HBM)
  p2p_hit_one / _both     --gpus 2: shards on GPU 0 and GPU 1 mapped into each other; GPU 0 gathers
                          alone, then both gather at once (half of the rows cross NVLink)
  p2p_remote_only         ids owned by the peer only: every row crosses NVLink
Every output row is verified bitwise against the procedural feature hash.  L2 is flushed between
timed launches.  Prints one JSON dict."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from coala_gnn_b200._lib import lib, check  # noqa: E402
from coala_gnn_b200.pybind import Isolated_Cache, SSD_GNN_NVSHMEM_Cache, NVSHMEM_Manager, Node_distributor_pybind, SSD_GNN_SSD_Controllers  # noqa: E402
from coala_gnn_b200 import synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--nodes", type=int, default=1_000_000)
ap.add_argument("--cache_mb", type=int, default=4096)
ap.add_argument("--dim", type=int, default=1024)
ap.add_argument("--gpus", type=int, default=1)
ap.add_argument("--rows", type=int, default=25000)
ap.add_argument("--iters", type=int, default=6)
ap.add_argument("--regimes", default="", help="comma list of labels to time (default all); the warm-up launches they need still run")
args = ap.parse_args()
N, DIM, R = args.nodes, args.dim, args.rows
L = lib()
CD = L.coala_cache_dim_for(DIM)
RB = CD * 4
SEED = 12
res = {"config": vars(args)}

torch.cuda.set_device(0)
host = torch.empty((N, CD), dtype=torch.float32).pin_memory()
check(L.coala_synth_fill_features(host.data_ptr(), SEED, 0, N, DIM, CD, None))
torch.cuda.synchronize()
color, topk, score = synth.random_color_meta(N, 5000, 10, seed=1)
items = np.arange(64, dtype=np.int64)
gb = C.c_double()
check(L.coala_probe_read_bandwidth(host.data_ptr(), min(N * RB, 2 << 30), 3, C.byref(gb)))
res["host_link_read_GBs"] = gb.value


def per_device(d):
    torch.cuda.set_device(d)
    nd = Node_distributor_pybind.from_arrays(items.ctypes.data, 0, 8, 1, 1, color, topk, score, device=d)
    ctrl = SSD_GNN_SSD_Controllers(1, RB, 1024, 0, d, DIM, True)
    out = torch.empty((R, DIM), dtype=torch.float32, device="cuda:%d" % d)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:%d" % d)
    return dict(nd=nd, ctrl=ctrl, out=out, flush=flush, dev=d)


def verify(D, idx_t, n):
    torch.cuda.set_device(D["dev"])
    bad = C.c_int64()
    check(L.coala_synth_check_rows(D["out"].data_ptr(), idx_t.data_ptr(), n, DIM, DIM, SEED, C.byref(bad)))
    assert bad.value == 0, "%d corrupt rows" % bad.value


WANT = set(x for x in args.regimes.split(",") if x)
# warm-up launches a timed regime depends on
NEEDS = {"iso_cold": ("iso_hit", "iso_miss01", "iso_miss03", "iso_miss10"), "iso_seq_warm": ("iso_seq_hit",), "iso_clustered_warm": ("iso_clustered_hit",),
         "part1_cold": ("part1_hit", "part1_miss01", "part1_miss03", "part1_miss10"),
         "p2p_cold_one": ("p2p_hit_one", "p2p_hit_both", "p2p_miss01_both", "p2p_miss03_both", "p2p_miss10_both"),
         "p2p_seq_warm": ("p2p_seq_hit_both",), "p2p_clustered_warm": ("p2p_clustered_hit_both", "p2p_clustered_hit_one"),
         "p2p_remote_only_warm": ("p2p_remote_only_one", "p2p_remote_only_both"), "p2p_local_only_warm": ("p2p_local_only_one",)}


def wanted(label):
    return not WANT or label in WANT or any(x in WANT for x in NEEDS.get(label, ()))


def timed(label, jobs, batches_per_job, iters=None):
    """jobs: list of (D, cache); batches_per_job: list (per job) of lists of id arrays. Launches all jobs
    back to back per iteration (they run concurrently on different GPUs)."""
    if not wanted(label):
        return
    iters = iters or len(batches_per_job[0])
    ts = [[] for _ in jobs]
    for it in range(iters):
        tens, evs = [], []
        for (D, cache), bl in zip(jobs, batches_per_job):
            torch.cuda.set_device(D["dev"])
            tens.append(torch.from_numpy(bl[it % len(bl)]).to("cuda:%d" % D["dev"]))
            # evict L2 with a READ sweep (a write sweep would leave 126 MB of dirty lines whose write-back
            # then competes with the timed kernel)
            check(L.coala_probe_read_bandwidth(D["flush"].data_ptr(), D["flush"].numel(), 1, C.byref(gb)))
        for j, (D, cache) in enumerate(jobs):
            torch.cuda.synchronize(D["dev"])
        for j, (D, cache) in enumerate(jobs):
            torch.cuda.set_device(D["dev"])
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            st = torch.cuda.current_stream().cuda_stream
            e0.record()
            cache.read_feature_async(D["out"].data_ptr(), tens[j].data_ptr(), tens[j].numel(), 0, st)
            e1.record()
            evs.append((e0, e1))
        for j, (D, cache) in enumerate(jobs):
            torch.cuda.synchronize(D["dev"])
            ts[j].append(evs[j][0].elapsed_time(evs[j][1]))
            verify(D, tens[j], tens[j].numel())
    out = {}
    for j, (D, cache) in enumerate(jobs):
        ms = float(np.median(ts[j]))
        n = len(batches_per_job[j][0])
        out["gpu%d" % D["dev"]] = dict(ms=ms, min_ms=float(np.min(ts[j])), rows=n, out_GBs=n * RB / ms / 1e6, hbm_alg_GBs=2 * n * RB / ms / 1e6,
                                       stats=cache.stats(reset=True))
    res[label] = out
    print(label, json.dumps(out), flush=True)


rng = np.random.default_rng(0)
perm = rng.permutation(N)
nb = min(8, (N * 6 // 10) // R)
batches = [np.ascontiguousarray(perm[i * R:(i + 1) * R]).astype(np.int64) for i in range(nb)]
cold_pool = perm[N * 6 // 10:]  # never warmed: every use of one of these ids is a miss
cold_used = [0]


def mixed(frac, parity=None):
    """one batch per timed iteration: warmed ids with `frac` of them replaced by never-seen ids"""
    out = []
    k = int(R * frac)
    for i in range(args.iters):
        b = batches[i % nb].copy()
        pool = cold_pool if parity is None else cold_pool[cold_pool % 2 == parity]
        c = pool[cold_used[0]:cold_used[0] + k]
        cold_used[0] += k
        pos = rng.choice(R, size=len(c), replace=False)
        b[pos] = c
        out.append(b)
    return out



def clustered(n_lists, seed=5):
    """id lists shaped like a sampled mini-batch on a graph with id-local communities: ~1000 centres,
    each with ~25 ids within +-4096 (unique inside a list)"""
    r = np.random.default_rng(seed)
    out = []
    for _ in range(n_lists):
        c = r.integers(0, N, size=R // 20)
        ids = (c[:, None] + r.integers(-4096, 4097, size=(len(c), 40))).ravel()
        ids = np.unique(np.clip(ids, 0, N - 1))
        r.shuffle(ids)
        out.append(np.ascontiguousarray(ids[:R]).astype(np.int64))
    return out


D0 = per_device(0)
# ---- isolated
iso = Isolated_Cache(D0["ctrl"], D0["nd"], 0, 1, args.cache_mb, host.data_ptr())
timed("iso_cold", [(D0, iso)], [batches], iters=nb)
timed("iso_hit", [(D0, iso)], [batches], iters=args.iters)
seq = [np.arange(i * R, (i + 1) * R, dtype=np.int64) for i in range(4)]  # contiguous ids: adjacent sets, same way
timed("iso_seq_warm", [(D0, iso)], [seq], iters=len(seq))
timed("iso_seq_hit", [(D0, iso)], [seq], iters=8)
cl = clustered(6)
timed("iso_clustered_warm", [(D0, iso)], [cl], iters=len(cl))
timed("iso_clustered_hit", [(D0, iso)], [cl], iters=len(cl))
for f in (0.01, 0.03, 0.10):
    timed("iso_miss%02d" % int(f * 100), [(D0, iso)], [mixed(f)], iters=args.iters)
del iso
torch.cuda.empty_cache()

# ---- partitioned, 2 shards emulated on GPU 0 (SYS code path on local memory)
emu = []
for g in range(2):
    NVSHMEM_Manager.current_rank = g
    emu.append(SSD_GNN_NVSHMEM_Cache(D0["ctrl"], D0["nd"], g, 2, args.cache_mb // 2, host.data_ptr()))
NVSHMEM_Manager.current_rank = None
emu[0].attach_local(1, emu[1])
emu[1].attach_local(0, emu[0])
timed("part1_cold", [(D0, emu[0])], [batches], iters=nb)
timed("part1_hit", [(D0, emu[0])], [batches], iters=args.iters)
for f in (0.01, 0.03, 0.10):
    timed("part1_miss%02d" % int(f * 100), [(D0, emu[0])], [mixed(f)], iters=args.iters)
del emu
torch.cuda.empty_cache()

if args.gpus >= 2:
    D1 = per_device(1)
    caches = []
    for g, D in enumerate((D0, D1)):
        torch.cuda.set_device(g)
        NVSHMEM_Manager.current_rank = g
        caches.append(SSD_GNN_NVSHMEM_Cache(D["ctrl"], D["nd"], g, 2, args.cache_mb, host.data_ptr()))
    NVSHMEM_Manager.current_rank = None
    caches[0].attach_local(1, caches[1])
    caches[1].attach_local(0, caches[0])
    b1 = batches[::-1]
    timed("p2p_cold_one", [(D0, caches[0])], [batches], iters=nb)
    timed("p2p_hit_one", [(D0, caches[0])], [batches], iters=args.iters)
    timed("p2p_hit_both", [(D0, caches[0]), (D1, caches[1])], [batches, b1], iters=args.iters)
    timed("p2p_seq_warm", [(D0, caches[0]), (D1, caches[1])], [seq, seq[::-1]], iters=len(seq))
    timed("p2p_seq_hit_both", [(D0, caches[0]), (D1, caches[1])], [seq, seq[::-1]], iters=8)
    cl0, cl1 = clustered(6, 7), clustered(6, 8)
    timed("p2p_clustered_warm", [(D0, caches[0]), (D1, caches[1])], [cl0, cl1], iters=6)
    timed("p2p_clustered_hit_both", [(D0, caches[0]), (D1, caches[1])], [cl0, cl1], iters=6)
    timed("p2p_clustered_hit_one", [(D0, caches[0])], [cl0], iters=6)
    for f in (0.01, 0.03, 0.10):
        timed("p2p_miss%02d_both" % int(f * 100), [(D0, caches[0]), (D1, caches[1])], [mixed(f), mixed(f)], iters=args.iters)
    odd = [np.ascontiguousarray(perm[perm % 2 == 1][i * R:(i + 1) * R]).astype(np.int64) for i in range(min(4, (N // 2) // R))]
    even = [np.ascontiguousarray(perm[perm % 2 == 0][i * R:(i + 1) * R]).astype(np.int64) for i in range(min(4, (N // 2) // R))]
    timed("p2p_remote_only_warm", [(D0, caches[0])], [odd], iters=len(odd))
    timed("p2p_remote_only_one", [(D0, caches[0])], [odd], iters=args.iters)
    timed("p2p_local_only_warm", [(D0, caches[0])], [even], iters=len(even))
    timed("p2p_local_only_one", [(D0, caches[0])], [even], iters=args.iters)
    timed("p2p_remote_only_both", [(D0, caches[0]), (D1, caches[1])], [odd, even], iters=args.iters)
    # peer-link streaming-read probe: GPU 0 reads GPU 1's memory (a plain cudaMalloc buffer needs classic peer access)
    torch.cuda.set_device(1)
    peer_buf = torch.empty(1 << 30, dtype=torch.uint8, device="cuda:1")
    torch.cuda.set_device(0)
    rc = L.coala_enable_peer_access(0, 1)
    if int(rc) == 0:
        torch.cuda.synchronize()
        check(L.coala_probe_read_bandwidth(peer_buf.data_ptr(), 1 << 30, 5, C.byref(gb)))
        res["nvlink_read_GBs"] = gb.value
    else:
        res["nvlink_read_GBs"] = "peer access unavailable (cudaDeviceEnablePeerAccess rc %s)" % rc
torch.cuda.set_device(0)
devbuf = torch.empty(1 << 30, dtype=torch.uint8, device="cuda:0")
check(L.coala_probe_read_bandwidth(devbuf.data_ptr(), 1 << 30, 5, C.byref(gb)))
res["hbm_read_GBs"] = gb.value
print(json.dumps(res))
