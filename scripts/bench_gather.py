"""Micro-benchmark of the gather kernel alone (all-miss, all-hit, mixed) + link probes."""
import ctypes as C
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from coala_gnn_b200._lib import lib, check
from coala_gnn_b200.pybind import Isolated_Cache, Node_distributor_pybind, SSD_GNN_SSD_Controllers
from coala_gnn_b200 import synth

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
DIM = 1024
CACHE_MB = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
n_rows = 36864
L = lib()
t0 = time.time()
host = torch.empty((N, DIM), dtype=torch.float32).pin_memory()
print("pinned alloc %.1fs" % (time.time() - t0))
t0 = time.time()
check(L.coala_synth_fill_features(host.data_ptr(), 12, 0, N, DIM, DIM, None))
torch.cuda.synchronize()
print("gpu fill %.2fs" % (time.time() - t0))
color, topk, score = synth.random_color_meta(N, 5000, 10, seed=1)
items = np.arange(64, dtype=np.int64)
nd = Node_distributor_pybind.from_arrays(items.ctypes.data, 0, 8, 1, 1, color, topk, score, device=0)
ctrl = SSD_GNN_SSD_Controllers(1, 4096, 1024, 0, 0, DIM, True)
cache = Isolated_Cache(ctrl, nd, 0, 1, CACHE_MB, host.data_ptr())
res = {}
gb = C.c_double()
check(L.coala_probe_read_bandwidth(host.data_ptr(), min(N * DIM * 4, 2 << 30), 3, C.byref(gb)))
res["host_link_read_GBs"] = gb.value
dev = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
check(L.coala_probe_read_bandwidth(dev.data_ptr(), 1 << 30, 5, C.byref(gb)))
res["hbm_read_GBs"] = gb.value
rng = np.random.default_rng(0)
out = torch.empty((n_rows, DIM), dtype=torch.float32, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(idx_list, label):
    ts = []
    for idx in idx_list:
        t = torch.from_numpy(idx).cuda()
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        cache.read_feature_async(out.data_ptr(), t.data_ptr(), len(idx), 0, torch.cuda.current_stream().cuda_stream)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
        bad = C.c_int64()
        check(L.coala_synth_check_rows(out.data_ptr(), t.data_ptr(), len(idx), DIM, DIM, 12, C.byref(bad)))
        assert bad.value == 0, bad.value
    ms = float(np.median(ts))
    res[label] = dict(ms=ms, out_GBs=n_rows * 4096 / ms / 1e6, stats=cache.stats(reset=True))
    print(label, res[label])


perm = rng.permutation(N)
batches = [perm[i * n_rows:(i + 1) * n_rows].astype(np.int64) for i in range(8)]
timed(batches, "cold_all_miss")
timed(batches, "warm_all_hit")
timed(batches[::-1], "warm_all_hit_2")
mixed = [np.concatenate([b[:n_rows // 2], perm[(10 + i) * n_rows:(10 + i) * n_rows + n_rows // 2]]).astype(np.int64) for i, b in enumerate(batches)]
timed(mixed, "half_miss")
print(json.dumps(res))
