"""coala_split_node_list_static: the allocation-free form of split_node_list.

Some users of libcoala_b200.so cannot let it allocate GPU memory: a plugin inside a process whose device memory is owned
by one pool allocator (torch's caching allocator, an RMM pool, a MIG slice sized to the byte), or a serving loop in
which cudaMalloc / cudaMallocAsync are off limits because they may lock or synchronise the device.  For them the
partition's block-count scratch sits in a static device array inside the library (include/coala_b200.h documents the
2 097 152-id limit, the per-device set-up in the first call, and that overlapping calls race on the shared array).

This file is named so that pytest collects it last: with `-x` a failure here must not hide any other test.  It passes
both alone in a fresh process (the first call below then does the per-device set-up) and after the rest of the suite.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _split(fn, idx_dev, n, G, max_sample):
    import torch
    from coala_gnn_b200._lib import check
    # poison the outputs so that a word the kernels fail to write cannot pass by accident
    node = torch.full((G * max_sample,), -7, dtype=torch.int64, device="cuda")
    mp = torch.full((G * max_sample,), -7, dtype=torch.int64, device="cuda")
    cnt = torch.full((G,), -7, dtype=torch.int64, device="cuda")
    check(fn(idx_dev.data_ptr(), n, node.data_ptr(), mp.data_ptr(), cnt.data_ptr(), G, max_sample, None), "split_node_list")
    torch.cuda.synchronize()
    return node.cpu().numpy(), mp.cpu().numpy(), cnt.cpu().numpy()


def test_static_scratch_split_equals_allocating_split_twice_in_a_row():
    """Two back-to-back calls of the static-scratch entry point, with different sizes and GPU counts so the second one
    meets the first one's left-overs in the shared array; each must equal coala_split_node_list exactly (integer work:
    the tolerance is zero, as in test_gpu_cache.py) and the numpy definition of a stable partition by id % G."""
    import torch
    from coala_gnn_b200._lib import lib
    L = lib()
    rng = np.random.default_rng(23)
    # (n, G): a many-block list first, then a shorter ragged one with another G over the same scratch words, then empty,
    # then exactly the maximum size
    cases = []
    for n, G in ((300_001, 8), (70_003, 3), (0, 5), (L.coala_split_node_list_static_max_ids(), 2)):
        idx = rng.integers(0, 10 ** 7, max(n, 1)).astype(np.int64)[:n]
        t = torch.from_numpy(idx).cuda() if n else torch.zeros(1, dtype=torch.int64, device="cuda")
        cases.append((n, G, n + 5, idx, t))
    ref = [_split(L.coala_split_node_list, t, n, G, ms) for n, G, ms, _, t in cases]
    # the static-scratch calls run one straight after the other, nothing else of the library in between
    got = [_split(L.coala_split_node_list_static, t, n, G, ms) for n, G, ms, _, t in cases]
    for (n, G, max_sample, idx, _), r, g_ in zip(cases, ref, got):
        for a, b, what in zip(r, g_, ("node", "map", "counter")):
            assert np.array_equal(a, b), "%s differs at n=%d G=%d" % (what, n, G)
        node, mp, cnt = g_
        for g in range(G):
            pos = np.nonzero(idx % G == g)[0]
            assert cnt[g] == len(pos)
            assert np.array_equal(node[g * max_sample: g * max_sample + len(pos)], idx[pos])
            assert np.array_equal(mp[g * max_sample: g * max_sample + len(pos)], pos)


def test_static_scratch_split_makes_no_allocation_and_refuses_oversize():
    """The call must not touch the device allocator (free memory before == after, torch's pool untouched), and a list
    beyond the built-in scratch is refused with a message instead of being truncated or launched."""
    import torch
    from coala_gnn_b200._lib import lib, last_error
    L = lib()
    limit = L.coala_split_node_list_static_max_ids()
    assert limit == 2_097_152
    G, n = 4, 50_000
    t = torch.arange(n, dtype=torch.int64, device="cuda")
    node = torch.zeros(G * n, dtype=torch.int64, device="cuda")
    mp = torch.zeros(G * n, dtype=torch.int64, device="cuda")
    cnt = torch.zeros(G, dtype=torch.int64, device="cuda")
    args = (t.data_ptr(), n, node.data_ptr(), mp.data_ptr(), cnt.data_ptr(), G, n, None)
    assert L.coala_split_node_list_static(*args) == 0  # set-up call (first on this device when the file runs alone)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    for _ in range(3):
        assert L.coala_split_node_list_static(*args) == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0
    assert torch.cuda.memory_reserved() == reserved0
    assert cnt.cpu().tolist() == [n // G] * G
    # one id too many: rejected before anything is launched (pointers are never dereferenced)
    assert L.coala_split_node_list_static(t.data_ptr(), limit + 1, node.data_ptr(), mp.data_ptr(), cnt.data_ptr(), G, n, None) != 0
    msg = last_error()
    assert "exceeds the built-in scratch" in msg and str(limit) in msg
    torch.cuda.synchronize()
    assert cnt.cpu().tolist() == [n // G] * G
