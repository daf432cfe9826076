"""Offline colour metadata for a dataset in the reference's on-disk layout — the counterpart of
examples/color_info_gen/generate_color_data.py (reference), same call order on Graph_Coloring (:20-54) and
the same three output files (:57-67):

    python examples/generate_color_data.py --path <root> --dataset_size <size> [--mode host|gpu] [--out_path DIR]

reads   <root>/<size>/processed/paper__cites__paper/csc_indptr.npy, csc_indices.npy
writes  <out_path or root/size>/color.npy  int64[N], topk.npy int64[num_colors, K], score.npy float64[num_colors, K]

Training nodes are the first 60 % of the ids, as the reference's IGB loaders define them
(examples/ssd_gnn_dataloader.py:546-557).  --mode host is the reference's algorithm bit for bit (single
thread); --mode gpu draws the same seeds and runs the traversal and the affinity pass on the device
(include/coala_b200.h, "GPU mode")."""
import argparse
import ctypes
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from COALA_GNN_Pybind import Graph_Coloring  # noqa: E402


def color_graph(indptr, indices, train_nid, topk=10, mode="host", device=0):
    """Returns (color int64[N], topk int64[C, K], score float64[C, K], num_colored_nodes)."""
    indptr = np.ascontiguousarray(indptr, dtype=np.int64)
    indices = np.ascontiguousarray(indices, dtype=np.int64)
    train_nid = np.ascontiguousarray(train_nid, dtype=np.int64)
    num_nodes = len(indptr) - 1
    tool = Graph_Coloring(num_nodes)
    tool.set_adj_csc(indptr.ctypes.data, indices.ctypes.data)
    color = np.zeros(num_nodes, dtype=np.int64)
    tool.set_color_buffer(color.ctypes.data)
    if mode == "gpu":
        tool.gpu_color_graph_optimized(train_nid.ctypes.data, len(train_nid), device)
    else:
        tool.cpu_color_graph_optimized(train_nid.ctypes.data, len(train_nid))
    num_colors = tool.get_num_color()
    topk_color = np.zeros(num_colors * topk, dtype=np.int64)
    topk_affinity = np.zeros(num_colors * topk, dtype=np.float64)
    tool.set_topk_color_buffer(topk_color.ctypes.data)
    tool.set_topk_affinity_buffer(topk_affinity.ctypes.data)
    if mode == "gpu":
        tool.gpu_calculate_color_affinity(device)
    else:
        tool.cpu_calculate_color_affinity()
    colored = tool.get_num_color_node()
    del tool
    return color, topk_color.reshape(num_colors, topk), topk_affinity.reshape(num_colors, topk), colored


def main(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--path", type=str, required=True, help="path containing the datasets")
    p.add_argument("--dataset_size", type=str, default="experimental")
    p.add_argument("--out_path", type=str, default=None, help="directory for color.npy / topk.npy / score.npy")
    p.add_argument("--topk", type=int, default=10)
    p.add_argument("--device", type=int, default=0)
    p.add_argument("--mode", choices=["host", "gpu"], default="host")
    p.add_argument("--srand", type=int, default=1, help="glibc rand() seed; the reference never seeds it (= 1)")
    args = p.parse_args(argv)
    if args.topk != 10:
        raise SystemExit("topk is fixed at 10 in Graph_Coloring (M/graph_coloring.h:22)")
    root = os.path.join(args.path, args.dataset_size)
    cites = os.path.join(root, "processed", "paper__cites__paper")
    indptr = np.load(os.path.join(cites, "csc_indptr.npy"))
    indices = np.load(os.path.join(cites, "csc_indices.npy"))
    n = len(indptr) - 1
    train_nid = np.arange(int(n * 0.6), dtype=np.int64)
    print("number of nodes: ", n)
    ctypes.CDLL("libc.so.6").srand(args.srand)
    t0 = time.perf_counter()
    color, topk, score, colored = color_graph(indptr, indices, train_nid, args.topk, args.mode, args.device)
    print(f"num_colors: {topk.shape[0]}\nnum colored node: {colored}\n{args.mode} mode: {time.perf_counter() - t0:.2f} s")
    out = args.out_path or root
    os.makedirs(out, exist_ok=True)
    np.save(os.path.join(out, "color.npy"), color)
    np.save(os.path.join(out, "topk.npy"), topk)
    np.save(os.path.join(out, "score.npy"), score)
    print("Saving is done")


if __name__ == "__main__":
    main()
