"""Practical ceiling of the N-GPU all-hit regime (coala_probe_gather_all): a 4 KB-row table hash-partitioned over G GPUs like
the cache (owner id % G), every GPU gathering 25 000 random rows AT ONCE with no cache protocol - (G-1)/G of each GPU's rows
come over NVLink while it serves its peers.  One process drives all G GPUs.  Prints one JSON dict.

    python scripts/probe_gather_all.py --gpus 8 [--rows 25000] [--part_gb 4]"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from coala_gnn_b200._lib import lib, check  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--gpus", type=int, default=2)
ap.add_argument("--rows", type=int, default=25000)
ap.add_argument("--part_gb", type=float, default=4.0, help="size of every GPU's part of the table")
ap.add_argument("--iters", type=int, default=16)
args = ap.parse_args()
L = lib()
R, n, iters = 4096, args.rows, args.iters
G = min(args.gpus, torch.cuda.device_count())
part_rows = int(args.part_gb * (1 << 30)) // R
res = {"config": vars(args), "gpus": G, "part_rows": part_rows}
tables, idxs, outs = [], [], []
for g in range(G):
    torch.cuda.set_device(g)
    tables.append(torch.empty(part_rows * R, dtype=torch.uint8, device="cuda:%d" % g))
    rng = np.random.default_rng(10 + g)
    idxs.append(torch.from_numpy(rng.integers(0, 1 << 62, size=n * (iters + 1)).astype(np.int64)).to("cuda:%d" % g))
    outs.append(torch.empty(n * R, dtype=torch.uint8, device="cuda:%d" % g))
for a in range(G):
    for b in range(G):
        if a != b:
            check(L.coala_enable_peer_access(a, b), "peer access %d -> %d" % (a, b))
torch.cuda.set_device(0)
U = C.c_uint64 * G
t_arr, i_arr, o_arr = U(*[t.data_ptr() for t in tables]), U(*[t.data_ptr() for t in idxs]), U(*[t.data_ptr() for t in outs])
us = (C.c_double * G)()
for groups in sorted({1, 2, 4, G}):  # 1 = one GPU alone against its own part only; G = everybody at once
    if groups > G:
        continue
    lists = [ix.remainder(part_rows * groups) for ix in idxs[:groups]]  # ids of a table spread over `groups` parts
    for g in range(groups):
        torch.cuda.synchronize(g)
    i_arr = U(*([t.data_ptr() for t in lists] + [0] * (G - groups)))
    r = {}
    for rif in (1, 2):
        for ctas in (2, 4, 8):
            check(L.coala_probe_gather_all(groups, t_arr, i_arr, o_arr, n, rif, ctas, iters, us), "probe")
            v = [us[g] for g in range(groups)]
            r["rif%d_ctas%d" % (rif, ctas)] = {"us_max": round(max(v), 2), "us_mean": round(sum(v) / len(v), 2)}
    best = min(r.items(), key=lambda kv: kv[1]["us_max"])
    inbound = n * R * (groups - 1) / groups
    res["G%d" % groups] = dict(r, best=best[0], best_us_max=best[1]["us_max"], rows_GBs=round(n * R / best[1]["us_max"] / 1e3, 1),
                               nvlink_in_GBs=round(inbound / best[1]["us_max"] / 1e3, 1))
    print("G%d" % groups, json.dumps(res["G%d" % groups]), flush=True)
print(json.dumps(res))
