"""Generates tests/golden/*.npz by running the REFERENCE's own compiled host code
(oracle/_ref/libcoala_ref.so, built from /root/reference by oracle/Makefile) on seeded inputs.
Run in the build container (where /root/reference exists):  python tests/golden/make_golden.py
The fixtures are small and committed; the GPU box only reads them.
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from coala_gnn_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def coloring_case(name, n, e, seed):
    indptr, indices, _ = synth.make_csc(n, e, seed)
    train = np.arange(int(n * 0.6), dtype=np.int64)  # nonzero(train_mask): first 60 %, ascending
    color, topk, score = oracle.ref_color_graph(indptr, indices, train)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), n=n, e=e, seed=seed, color=color, topk=topk, score=score)
    return indptr, indices, color, topk, score


def distributor_cases(name, color, topk, score, n):
    d = tempfile.mkdtemp()
    cf, tf, sf = (os.path.join(d, x) for x in ("color.npy", "topk.npy", "score.npy"))
    np.save(cf, color), np.save(tf, topk), np.save(sf, score)
    items = synth.train_ids(n, seed=5)
    rng = np.random.default_rng(99)
    cases = {}
    k = 0
    for (D, L, B) in [(1, 8, 16), (2, 4, 32), (4, 2, 32), (8, 1, 64), (3, 1, 50)]:
        for variant in ("random", "zeros", "ties"):
            C = topk.shape[0]
            if variant == "random":
                metas = synth.random_cache_meta(C, D, seed=21 + k)
            elif variant == "zeros":
                metas = [np.zeros(C + 1, dtype=np.int32) for _ in range(D)]
            else:  # identical histograms in every domain: every score ties, capacity closure decides
                m = rng.integers(0, 50, size=C + 1, dtype=np.int32)
                metas = [m.copy() for _ in range(D)]
            offset = int(rng.integers(0, len(items) - D * L * B))
            outs = []
            for node_id in range(D):
                out, nc = oracle.ref_distribute(items, offset, node_id, B, L, D, cf, tf, sf, metas)
                assert nc == C
                outs.append(out)
            cases["c%d_DLB" % k] = np.array([D, L, B, offset], dtype=np.int64)
            cases["c%d_meta" % k] = np.stack(metas)
            cases["c%d_out" % k] = np.stack(outs)
            k += 1
    np.savez_compressed(os.path.join(OUT, name + ".npz"), items=items, n_cases=k, **cases)


if __name__ == "__main__":
    if not oracle.have_ref():
        raise SystemExit("oracle/_ref missing: run `make -C oracle` where /root/reference exists")
    indptr, indices, color, topk, score = coloring_case("coloring_2k", 2000, 24000, 7)
    coloring_case("coloring_20k", 20000, 240000, 8)
    distributor_cases("distributor_2k", color, topk, score, 2000)
    print("golden fixtures written to", OUT)
