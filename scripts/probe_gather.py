"""Practical ceiling of the gather's access pattern, without the cache protocol (coala_probe_gather):
out[i] = table[idx[i]] for 25 000 random 4 KB rows out of a table much larger than L2, table in local HBM / a peer's HBM /
pinned host memory, by rows in flight per warp and resident CTAs per SM.  Prints one JSON dict.

    python scripts/probe_gather.py [--gpus 2] [--rows 25000] [--row_bytes 4096]"""
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
ap.add_argument("--gpus", type=int, default=1)
ap.add_argument("--rows", type=int, default=25000)
ap.add_argument("--row_bytes", type=int, default=4096)
ap.add_argument("--table_gb", type=float, default=8.0)
ap.add_argument("--iters", type=int, default=8)
args = ap.parse_args()
L = lib()
R, n, iters = args.row_bytes, args.rows, args.iters
n_rows = int(args.table_gb * (1 << 30)) // R
rng = np.random.default_rng(0)
torch.cuda.set_device(0)
idx = torch.from_numpy(rng.integers(0, n_rows, size=n * (iters + 1)).astype(np.int64)).cuda()  # list 0 = warm-up, 1..iters timed
out = torch.empty(n * R, dtype=torch.uint8, device="cuda:0")
res = {"config": vars(args), "table_rows": n_rows}
us = C.c_double()


def sweep(label, table_ptr):
    r = {}
    for rif in (1, 2):
        for ctas in (1, 2, 3, 4, 6, 8):
            check(L.coala_probe_gather(table_ptr, idx.data_ptr(), out.data_ptr(), n, R, rif, ctas, iters, C.byref(us)))
            r["rif%d_ctas%d" % (rif, ctas)] = {"us": round(us.value, 2), "row_GBs": round(n * R / us.value / 1e3, 1)}
    best = min(r.items(), key=lambda kv: kv[1]["us"])
    res[label] = dict(r, best=best[0], best_us=best[1]["us"], best_row_GBs=best[1]["row_GBs"])
    print(label, json.dumps(res[label]), flush=True)


local = torch.empty(n_rows * R, dtype=torch.uint8, device="cuda:0")
sweep("local_hbm", local.data_ptr())
del local
torch.cuda.empty_cache()
if args.gpus >= 2:
    torch.cuda.set_device(1)
    peer = torch.empty(n_rows * R, dtype=torch.uint8, device="cuda:1")
    torch.cuda.set_device(0)
    rc = int(L.coala_enable_peer_access(0, 1))
    if rc == 0:
        sweep("peer_hbm_nvlink", peer.data_ptr())
    else:
        res["peer_hbm_nvlink"] = "peer access unavailable (rc %d)" % rc
    del peer
host = torch.empty(min(n_rows, (4 << 30) // R) * R, dtype=torch.uint8).pin_memory()
idx.remainder_(host.numel() // R)
sweep("pinned_host", host.data_ptr())
print(json.dumps(res))
