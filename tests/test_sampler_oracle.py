"""CPU tests of the sampler oracle. DGL is absent, so these pin the oracle to DGL's documented
semantics: exact CSC slices at full fan-out, min(deg,f) distinct in-neighbours
with uniform marginals otherwise, to_block numbering = dst first then first occurrence."""
import numpy as np

import oracle
from coala_gnn_b200 import synth


def _graph():
    return synth.make_csc(20000, 240000, 7)


def test_full_fanout_is_exact_csc_slices():
    indptr, indices, eids = _graph()
    seeds = np.random.default_rng(0).permutation(20000)[:500].astype(np.int64)
    src, dst, eid = oracle.sample_layer(indptr, indices, eids, seeds, -1, 123)
    exp_src = np.concatenate([indices[indptr[v]:indptr[v + 1]] for v in seeds])
    exp_eid = np.concatenate([eids[indptr[v]:indptr[v + 1]] for v in seeds])
    exp_dst = np.repeat(np.arange(len(seeds)), indptr[seeds + 1] - indptr[seeds])
    assert np.array_equal(src, exp_src) and np.array_equal(eid, exp_eid) and np.array_equal(dst, exp_dst)


def test_sampled_fanout_counts_subset_no_duplicates():
    indptr, indices, eids = _graph()
    seeds = np.random.default_rng(1).permutation(20000)[:2000].astype(np.int64)
    f = 5
    src, dst, eid = oracle.sample_layer(indptr, indices, None, seeds, f, 999)
    deg = indptr[seeds + 1] - indptr[seeds]
    assert np.array_equal(np.bincount(dst, minlength=len(seeds)), np.minimum(deg, f))
    # eid == CSC position when edge_ids is None: positions inside the row, distinct, ascending
    for i in np.unique(dst)[:400]:
        p = eid[dst == i]
        assert np.all(p >= indptr[seeds[i]]) and np.all(p < indptr[seeds[i] + 1])
        assert np.all(np.diff(p) > 0)
        assert np.array_equal(src[dst == i], indices[p])


def test_sampled_fanout_is_uniform_chi_square():
    """Each in-edge of a degree-d row must be kept with probability f/d."""
    d, f, trials = 12, 5, 6000
    indptr = np.array([0, d], dtype=np.int64)
    indices = np.arange(d, dtype=np.int64)
    counts = np.zeros(d)
    pair = np.zeros((d, d))
    for t in range(trials):
        _, _, eid = oracle.sample_layer(indptr, indices, None, np.array([0]), f, oracle.layer_key(t, 0))
        counts[eid] += 1
        pair[np.ix_(eid, eid)] += 1
    exp = trials * f / d
    chi2 = ((counts - exp) ** 2 / exp).sum()
    assert chi2 < 40.0, chi2  # 11 dof, p ~ 1e-5 cut-off
    off = pair[~np.eye(d, dtype=bool)]
    exp2 = trials * f * (f - 1) / (d * (d - 1))
    assert abs(off.mean() - exp2) / exp2 < 0.02 and off.std() / exp2 < 0.1


def test_to_block_numbering():
    seeds = np.array([7, 3, 9], dtype=np.int64)
    e_src = np.array([5, 3, 8, 5, 7, 11, 8], dtype=np.int64)
    nodes, loc = oracle.to_block(seeds, e_src)
    assert list(nodes) == [7, 3, 9, 5, 8, 11]
    assert list(loc) == [3, 1, 4, 3, 0, 5, 4]


def test_multilayer_matches_fast_path_and_shapes():
    indptr, indices, eids = _graph()
    seeds = synth.train_ids(20000, seed=3)[:256]
    bkey = oracle.batch_key(1, 5)
    inp, outn, blocks = oracle.sample_blocks_detailed(indptr, indices, eids, seeds, [10, 5, 5], bkey)
    assert np.array_equal(outn, seeds)
    assert np.array_equal(blocks[-1]["dst_nodes"], seeds)
    for a, b in zip(blocks[:-1], blocks[1:]):
        assert np.array_equal(a["dst_nodes"], b["src_nodes"])
    assert np.array_equal(inp, blocks[0]["src_nodes"]) and len(np.unique(inp)) == len(inp)
    for b in blocks:
        assert np.array_equal(b["src_nodes"][:len(b["dst_nodes"])], b["dst_nodes"])
        pos = np.argsort(eids)[b["edge_id"]]  # edge id -> CSC position
        assert np.array_equal(b["src_nodes"][b["edge_src"]], indices[pos])
    inp2, tot = oracle.sample_blocks_fast(indptr, indices, seeds, [10, 5, 5], bkey, 256 * 11 * 6 * 6)
    assert np.array_equal(inp, inp2) and tot == sum(len(b["edge_src"]) for b in blocks)


def test_native_batch_key_matches_oracle_and_python():
    """The native batch engine derives a batch's RNG key itself (csrc/engine.cu): same arithmetic as oracle.batch_key."""
    from coala_gnn_b200._lib import lib
    from coala_gnn_b200.sampler import batch_key
    for seed, t, rank in [(1, 0, 0), (11, 5, 3), (2 ** 63 + 5, 123456789, 7), (77, 2 ** 40, 15), (1001, 999999, 1)]:
        k = lib().coala_batch_key(seed & (2 ** 64 - 1), t, rank)
        assert k == batch_key(seed, t, rank) == oracle.batch_key(seed, t, rank)
