"""Offline colour-metadata generation, host mode (the reference's algorithm, single thread) against the GPU mode:

    python scripts/bench_coloring.py [--nodes N] [--deg D] [--train_frac F]

Times colouring and affinity separately (wall clock around the synchronous C-ABI calls; the GPU numbers include
the host<->device copies of the graph, the colour / hop arrays and the result). Prints one JSON dict."""
import argparse
import ctypes
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from coala_gnn_b200 import synth  # noqa: E402
from coala_gnn_b200.pybind import Graph_Coloring  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--nodes", type=int, default=2_000_000)
ap.add_argument("--deg", type=int, default=12)
ap.add_argument("--train_frac", type=float, default=0.6)
args = ap.parse_args()
N = args.nodes
indptr, indices, _ = synth.make_csc(N, N * args.deg, 7)
train = np.arange(int(N * args.train_frac), dtype=np.int64)
K = 10
res = {"config": vars(args), "edges": int(indptr[-1])}


def run(mode):
    ctypes.CDLL("libc.so.6").srand(1)
    gc = Graph_Coloring(N)
    gc.set_adj_csc(indptr.ctypes.data, indices.ctypes.data)
    color = np.zeros(N, dtype=np.int64)
    gc.set_color_buffer(color.ctypes.data)
    t0 = time.perf_counter()
    if mode == "gpu":
        gc.gpu_color_graph_optimized(train.ctypes.data, len(train), 0)
    else:
        gc.cpu_color_graph_optimized(train.ctypes.data, len(train))
    t1 = time.perf_counter()
    nc = gc.get_num_color()
    tk, sc = np.zeros(nc * K, dtype=np.int64), np.zeros(nc * K, dtype=np.float64)
    gc.set_topk_color_buffer(tk.ctypes.data)
    gc.set_topk_affinity_buffer(sc.ctypes.data)
    t2 = time.perf_counter()
    if mode == "gpu":
        gc.gpu_calculate_color_affinity(0)
    else:
        gc.cpu_calculate_color_affinity()
    t3 = time.perf_counter()
    return dict(color_s=t1 - t0, affinity_s=t3 - t2, num_colors=nc, colored=gc.get_num_color_node(),
                mean_top1_score=float(sc.reshape(nc, K)[:, 0].mean()))


import torch  # noqa: E402  (device memory / context only)
torch.cuda.init()
run("gpu")  # context creation, module load
res["gpu"] = run("gpu")
res["host"] = run("host")
res["speedup_color"] = res["host"]["color_s"] / res["gpu"]["color_s"]
res["speedup_affinity"] = res["host"]["affinity_s"] / res["gpu"]["affinity_s"]
print(json.dumps(res))
