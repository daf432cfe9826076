"""TEST INFRASTRUCTURE (run by hand; uses the CPU oracle only).  Host-side simulation of the experiment the
reference's examples/Distribution_compare_script.sh:26-35 runs: D cache domains with private caches, seeds
distributed `node_color` (M/node_distributor_pybind.cuh:150-222 on the domain-summed cache colour histograms,
refreshed every `refresh` batches, S/COALA_GNN_DataLoader.py:46-59) against `baseline` striping
(S/Training_node_distributor.py:41-47).  Prints the steady-state hit ratio of both.  It answers, without a GPU,
whether colour affinity can raise the hit ratio on the synthetic graphs of coala_gnn_b200.synth.

    python tests/sim_distribution.py [--nodes 1000000] [--domains 4] [--cache_frac 0.1] [--steps 300]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402
from coala_gnn_b200 import synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=1_000_000)
    ap.add_argument("--deg", type=int, default=12)
    ap.add_argument("--domains", type=int, default=4)
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--cache_frac", type=float, default=0.1, help="lines per domain / nodes")
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--refresh", type=int, default=10)
    ap.add_argument("--fan_out", default="5,5")
    ap.add_argument("--color", default="graph", choices=["graph", "block"])
    args = ap.parse_args()
    n, D, B = args.nodes, args.domains, args.batch
    fan = [int(f) for f in args.fan_out.split(",")]
    indptr, indices, _ = synth.make_csc(n, n * args.deg, 2, edge_ids=None)
    train = synth.train_ids(n, seed=0)
    if args.color == "graph":
        color, topk, score = oracle.ref_color_graph(indptr, indices, np.arange(int(n * 0.6), dtype=np.int64))
    else:
        color, topk, score = synth.block_color_meta(n, max(16, int(0.005 * n)))
    C = topk.shape[0]
    print("colours: %d, coloured nodes %.3f, mean top-1 score %.4f" % (C, float((color != 0).mean()), float(score[:, 0].mean())), flush=True)
    page = 512
    feat = np.zeros((n, page // 4), dtype=np.float32)
    cache_mb = max(1, int(args.cache_frac * n * page / (1 << 20)))
    max_sample = B
    for f in fan:
        max_sample *= f + 1
    res = {}
    for method in ("baseline", "node_color"):
        caches = [oracle.CacheModel(cache_mb, page, 1, "isolated", color=color, num_colors=C, store_lines=False) for _ in range(D)]
        metas = [np.zeros(C + 1, dtype=np.int32) for _ in range(D)]
        hist = []
        spe = len(train) // (B * D) - 1
        for t in range(args.steps):
            off = (t % spe) * B * D
            if t % args.refresh == 0 and t > 0:
                metas = [c.color_counters(0).astype(np.int32) for c in caches]
            h0 = [c.stats() for c in caches]
            for d in range(D):
                if method == "baseline":
                    seeds = train[off + d * B: off + (d + 1) * B]
                else:
                    seeds, _ = oracle.distribute(train, off, d, B, 1, D, color, topk, score, metas)
                inp, _ = oracle.sample_blocks_fast(indptr, indices, seeds, fan, oracle.batch_key(1, t, d), max_sample)
                caches[d].read(inp, feat, page // 4)
            h1 = [c.stats() for c in caches]
            hits = sum(b[0] - a[0] for a, b in zip(h0, h1))
            tot = sum(b[0] + b[1] - a[0] - a[1] for a, b in zip(h0, h1))
            hist.append(hits / tot)
            if t % 50 == 49:
                print("%-10s step %4d hit ratio (last 50) %.4f" % (method, t + 1, float(np.mean(hist[-50:]))), flush=True)
        res[method] = float(np.mean(hist[-50:]))
    print("steady state: baseline %.4f  node_color %.4f" % (res["baseline"], res["node_color"]))


if __name__ == "__main__":
    main()
