#!/usr/bin/env python
"""bench.py — mini-batches/s of the COALA-GNN data path (seed distribution -> UVA neighbour sampling
-> cached feature gather) on synthetic IGB-shaped data.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One JSON line on rank 0.  A "step" is one mini-batch (batch 1024, SAGE fan-out 5,5) per GPU.
  value    : device-resident pipeline (coala_gnn_b200.pipeline.DevicePipeline), CUDA-event timed in steady state:
             K-step windows recorded back to back (no drain in between) for >= --min_seconds, bracketed by
             barrier + synchronize, max over ranks per window, the MEDIAN window is reported;
  e2e      : the drop-in COALA_GNN_DataLoader iterator with host seed lists / host metadata (same windows, host clock);
  parity   : after each timed leg the next K batches of EVERY rank are checked bit for bit: gathered rows against
             the procedural feature table, sampled blocks (and the seed distribution) against the CPU oracle;
  roofline : the gather kernel, timed live with CUDA events inside the timed region;
  cpu_baseline : oracle sampler (OpenMP) + torch.index_select on the host cores (bounded sample);
  extras   : the same path in the other regimes of the reference's experiment matrix (whole table resident,
             node_color vs baseline over D x L cache domains, cache backends), each warmed and timed the same way.
`--impl reference` times the CPU path as its own arm.
"""
import argparse
import ctypes as C
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

def metric_string(args):
    """BASELINE.json's metric, worded for the shape / fan-out / batch actually run (a papers100m 10,10,10 run is not labelled IGB-medium 5,5)."""
    return ("mini-batches/s (%s-shaped synthetic graph, %d-layer neighbour sampling fan-out %s, batch %d per GPU; seed distribution + UVA sampling + cached feature gather)"
            % (args.shape, len(args.fan_out.split(",")), args.fan_out, args.batch_size))


NVL_PEAK = 770.0  # GB/s per direction: the measured peer-copy figure of B200_PROFILING.md (900 nominal); scripts/bench_gather.py reads 777 with a kernel


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=100)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--shape", default=os.environ.get("COALA_BENCH_SHAPE", "igb-medium"))
    p.add_argument("--cache_size", type=int, default=16384, help="MB per GPU (reference flag name)")
    p.add_argument("--cache_backend", default="nvshmem", choices=["nvshmem", "p2p", "isolated", "nccl"])
    p.add_argument("--distribution", default="node_color", choices=["node_color", "baseline"])
    p.add_argument("--gpus_per_domain", type=int, default=0, help="L: GPUs sharing one partitioned cache (0 = all: one 1xN domain); "
                                                                  "D = N / L cache domains stand in for the reference's machines (SLURM_NODEID)")
    p.add_argument("--batch_size", type=int, default=1024)
    p.add_argument("--fan_out", default="5,5")
    p.add_argument("--refresh_counter", type=int, default=10)
    p.add_argument("--prewarm", type=int, default=-1, help="untimed batches that warm the cache (-1: budgeted by --prewarm_seconds)")
    p.add_argument("--prewarm_seconds", type=float, default=25.0)
    p.add_argument("--min_seconds", type=float, default=2.0, help="timed windows of --steps are repeated until they cover this long")
    p.add_argument("--cpu_baseline_seconds", type=float, default=12.0)
    p.add_argument("--no_cpu_baseline", action="store_true")
    p.add_argument("--extras", default=os.environ.get("COALA_BENCH_EXTRAS", "auto"),
                   help="auto | none | comma list of all_resident,distribution,backends,stress")
    p.add_argument("--extras_budget", type=float, default=300.0, help="seconds after which the extras are abandoned and the main line printed")
    p.add_argument("--hbm_indptr", type=int, default=None, choices=[0, 1], help="sampler keeps a copy of indptr (80 MB at igb-medium) in HBM "
                   "(default: COALA_HBM_INDPTR, which defaults to 1)")
    p.add_argument("--hbm_graph", action="store_true", help="B200 option: the neighbour lists (indices, 0.96 GB at igb-medium) in HBM as well")
    p.add_argument("--no_prefetch", action="store_true", help="diagnostic: sample and gather of a batch back to back, no batch in flight")
    p.add_argument("--profile_e2e", action="store_true", help="cProfile the host side of the e2e loop (rank 0) to stderr")
    p.add_argument("--cuda_profiler_range", action="store_true",
                   help="bracket the timed region with cudaProfilerStart/Stop (for `ncu --profile-from-start off`)")
    return p.parse_args()


class ClockSampler(object):
    """nvidia-smi clocks + throttle reasons (B200_PROFILING.md recipe), sampled from process start; stop(t0, t1)
    keeps the samples that fall inside the timed region."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, regions):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        rows = [r for ts, r in self.rows if len(r) >= 7 and any(a <= ts <= b + 0.1 for a, b in regions)]
        sm = [float(r[0]) for r in rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm), "samples_whole_run": len(self.rows)}


def workload_string(args, ds, D, L):
    """One description of the measured workload, identical in both arms (the driver compares the two lines' config)."""
    return ("%s (%d nodes, %d edges, %d-d fp32 features pinned in host memory), %d-layer fan-out %s, batch %d/GPU, %d MB cache/GPU, backend %s (%dx%d domain), distribution %s"
            % (args.shape, ds.num_nodes, ds.num_edges, ds.dim, len(args.fan_out.split(",")), args.fan_out, args.batch_size, args.cache_size, args.cache_backend, D, L, args.distribution))


def domains_for(args):
    """(D, L) of a run with --gpus N and --gpus_per_domain (0 = one domain spanning the box), as MPI_Comm_Manager lays them out."""
    L = args.gpus_per_domain if args.gpus_per_domain else args.gpus
    return max(1, args.gpus // max(1, L)), L


def cpu_path(ds_host, train, B, fan, seconds=None, steps=None, warmup=2, threads=None, first=0):
    """The reference's CPU formulation of the path (BASELINE.json: "DGL CPU sampler + torch index_select gather"):
    the oracle's OpenMP restatement of the DGL CPU sampler + to_block (DGL is absent from the image), then
    torch.index_select on the host feature table with all host threads."""
    import oracle
    import torch
    indptr, indices, feat = ds_host
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    feat_t = torch.from_numpy(feat)
    max_sample = B
    for f in fan:
        max_sample *= f + 1
    out = torch.empty((max_sample, feat.shape[1]), dtype=torch.float32)
    times, rows = [], 0
    i = 0
    t_end = time.time() + seconds if seconds else None
    while True:
        seeds = train[((first + i) % (len(train) // B - 1)) * B:][:B]
        t0 = time.perf_counter()
        inp, _ = oracle.sample_blocks_fast(indptr, indices, seeds, fan, oracle.batch_key(1, first + i), max_sample, threads)
        torch.index_select(feat_t, 0, torch.from_numpy(inp), out=out[:len(inp)])
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
            rows += len(inp)
        i += 1
        if steps is not None and len(times) >= steps:
            break
        if t_end is not None and time.time() > t_end and len(times) >= 3:
            break
    tot = float(np.sum(times))
    return {"batches_per_s": len(times) / tot, "ms_per_step": 1e3 * tot / len(times), "rows_per_step": rows / len(times),
            "gather_GBs": rows * feat.shape[1] * 4 / tot / 1e9, "steps": len(times), "threads": threads}


def cpu_path_windows(ds_host, train, B, fan, steps, warmup, min_seconds, settle_seconds=10.0, threads=None):
    """The CPU arm under the GPU arm's protocol: warm up until the rate has settled (two consecutive windows within 5 %, at
    least `warmup` batches, at most `settle_seconds`: the first seconds run 25-30 % slow while the OpenMP team, torch's
    thread pool and the page tables of the 41 GB table warm up), then windows of `steps` batches until they cover
    `min_seconds`; value = median window."""
    w = cpu_path(ds_host, train, B, fan, steps=max(steps, warmup), warmup=0, threads=threads)
    t_end, n_warm, prev = time.time() + settle_seconds, w["steps"], w["batches_per_s"]
    while time.time() < t_end:
        w = cpu_path(ds_host, train, B, fan, steps=steps, warmup=0, threads=threads, first=n_warm)
        n_warm += w["steps"]
        settled = abs(w["batches_per_s"] - prev) <= 0.05 * prev
        prev = w["batches_per_s"]
        if settled:
            break
    wins, covered = [], 0.0
    while covered < min_seconds or len(wins) < 3:
        w = cpu_path(ds_host, train, B, fan, steps=steps, warmup=0, threads=threads, first=n_warm + len(wins) * steps)
        wins.append(w)
        covered += w["steps"] / w["batches_per_s"]
    wins.sort(key=lambda x: x["batches_per_s"])
    med = dict(wins[len(wins) // 2])
    med.update(windows=len(wins), warm_batches=n_warm, window_rate_min_med_max=[wins[0]["batches_per_s"], med["batches_per_s"], wins[-1]["batches_per_s"]])
    return med


# ------------------------------------------------------------------------------------------------------------
# helpers shared by every leg
# ------------------------------------------------------------------------------------------------------------
def barrier_sync(comm):
    import torch
    import torch.distributed as dist
    torch.cuda.synchronize()
    if comm.global_size > 1:
        dist.barrier()
    torch.cuda.synchronize()


def reduce_ranks(comm, xs, device, op="sum"):
    import torch
    import torch.distributed as dist
    xs = [float(x) for x in xs]
    if comm.global_size == 1:
        return xs
    t = torch.tensor(xs, dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
    return [float(v) for v in t.tolist()]


def n_windows_for(comm, rate_per_rank, K, min_seconds, device):
    """How many K-step windows cover min_seconds at the observed rate; the same number on every rank."""
    n = int(math.ceil(min_seconds * max(rate_per_rank, 1.0) / K))
    n = max(3, min(n, 4000))
    return int(reduce_ranks(comm, [n], device, "max")[0])


def timed_windows(comm, step_fn, K, n_windows, refill, device, profiler_range=False):
    """Steady-state timing.  barrier + synchronize; `refill` untimed steps (the synchronize completed every prefetched
    batch: without them the first `depth` timed steps would be free); then n_windows windows of exactly K steps each,
    one CUDA event between windows on the consumer stream and NO synchronisation in between; barrier + synchronize.
    Returns (per-window ms as the max over ranks, host time stamps of the region)."""
    import torch
    barrier_sync(comm)
    for _ in range(refill):
        step_fn()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(n_windows + 1)]
    if profiler_range:
        torch.cuda.profiler.start()
    t_a = time.time()
    evs[0].record()
    for w in range(n_windows):
        for _ in range(K):
            step_fn()
        evs[w + 1].record()
    torch.cuda.synchronize()
    if profiler_range:
        torch.cuda.profiler.stop()
    t_b = time.time()
    barrier_sync(comm)
    ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(n_windows)]
    return reduce_ranks(comm, ms, device, "max"), (t_a, t_b)


def build_world(args, comm=None, ds=None, cache_size=None, backend=None, distribution=None, gpus_per_domain=None):
    """Comm manager (D x L domains), synthetic dataset, distributor, sampler, drop-in loader."""
    import torch
    from COALA_GNN import MPI_Comm_Manager, Node_Distributor, SSD_INFO, COALA_GNN_DataLoader, MultiLayerNeighborSampler
    from coala_gnn_b200.datasets import SyntheticSharedCSCUVA
    from coala_gnn_b200 import synth
    if comm is None:
        L = gpus_per_domain if gpus_per_domain is not None else args.gpus_per_domain
        comm = MPI_Comm_Manager(-1, gpus_per_domain=L if L else None)  # default: one cache domain spanning all ranks of the box
    torch.cuda.set_device(comm.device_index)
    device = "cuda:%d" % comm.device_index
    if ds is None:
        ds = SyntheticSharedCSCUVA(args.shape, comm, device, verbose=True)
    fan = [int(f) for f in args.fan_out.split(",")]
    train = torch.from_numpy(synth.train_ids(ds.num_nodes, seed=0))
    nd = Node_Distributor(comm, train, args.batch_size, ds.color_file, ds.topk_file, ds.score_file,
                          parsing_method=distribution or args.distribution)
    sampler = MultiLayerNeighborSampler(fan, seed=1, rank=comm.global_rank, hbm_indptr=None if args.hbm_indptr is None else bool(args.hbm_indptr),
                                        hbm_graph=True if args.hbm_graph else None)
    loader = COALA_GNN_DataLoader(SSD_INFO(1, ds.cache_dim * 4, 1024, 0), nd, ds.graph, sampler, args.batch_size, ds.dim, fan,
                                  cache_size or args.cache_size, device, refresh_counter=args.refresh_counter,
                                  cache_backend=backend or args.cache_backend, sim_buf=ds.feat_data)
    loader.COALA_GNN_Manager.COALA_GNN_Cache.set_num_nodes(ds.num_nodes)
    return comm, ds, train, fan, nd, sampler, loader, device


class Checker(object):
    """Bitwise parity of what a leg produced, on this rank: gathered rows against the procedural feature table
    (re-derived on the GPU, coala_synth_check_rows — the reference's own invariant feat[i] == table[input_nodes[i]],
    examples/nvshmem_homogeneous_train.py:186-198), sampled blocks and the seed distribution against the CPU oracle."""

    def __init__(self, comm, ds, train, fan, nd, device):
        import oracle
        self.oracle = oracle
        self.comm, self.ds, self.fan, self.nd, self.device = comm, ds, fan, nd, device
        self.train = train.numpy()
        self.tot = dict(batches=0, rows_checked=0, mismatches=0, blocks_checked=0, block_mismatches=0, seed_lists_checked=0,
                        seed_mismatches=0, refused_ids=0)
        self._tables = None

    def rows(self, feat_ptr, idx_ptr, n):
        from coala_gnn_b200._lib import lib, check
        bad = C.c_int64()
        check(lib().coala_synth_check_rows(int(feat_ptr), int(idx_ptr), int(n), self.ds.dim, self.ds.dim, self.ds.feat_seed, C.byref(bad)))
        self.tot["batches"] += 1
        self.tot["rows_checked"] += int(n)
        self.tot["mismatches"] += int(bad.value)

    def blocks(self, seeds_h, key, input_nodes_h, edge_src_h):
        indptr, indices, _ = self.ds.csc_host
        einp, _, eblocks = self.oracle.sample_blocks_detailed(indptr, indices, None, seeds_h, self.fan, key)
        ok = np.array_equal(input_nodes_h, einp) and np.array_equal(edge_src_h, eblocks[0]["edge_src"])
        self.tot["blocks_checked"] += 1
        self.tot["block_mismatches"] += 0 if ok else 1

    def seeds(self, seeds_h, offset, metas_h, method):
        """this rank's B seeds of the step whose global slice starts at `offset`"""
        comm, B = self.comm, self.nd.batch_size
        L, D = comm.local_size, comm.num_master_process
        if method == "baseline" or D == 1:
            lo = offset + (comm.master_process_index * B * L if method == "baseline" else 0) + comm.local_rank * B
            exp = self.train[lo:lo + B]
        else:
            if self._tables is None:
                self._tables = tuple(np.load(f) for f in (self.ds.color_file, self.ds.topk_file, self.ds.score_file))
            color, topk, score = self._tables
            dom, _ = self.oracle.distribute(self.train, offset, comm.master_process_index, B, L, D, color, topk, score, list(metas_h))
            exp = dom[comm.local_rank * B:(comm.local_rank + 1) * B]
        self.tot["seed_lists_checked"] += 1
        self.tot["seed_mismatches"] += 0 if np.array_equal(seeds_h, exp) else 1

    def summary(self):
        keys = sorted(self.tot)
        vals = reduce_ranks(self.comm, [self.tot[k] for k in keys], self.device)
        out = {k: int(v) for k, v in zip(keys, vals)}
        out["ranks"] = self.comm.global_size
        out["ok"] = out["mismatches"] == 0 and out["block_mismatches"] == 0 and out["seed_mismatches"] == 0 and out["refused_ids"] == 0
        return out


def check_pipeline(pipe, chk, K, oracle_batches=4):
    """The next K batches of the device-resident pipeline, checked one by one (untimed)."""
    import torch
    from coala_gnn_b200._lib import lib, check
    from coala_gnn_b200.pybind import tensor_from_ptr
    from coala_gnn_b200.sampler import batch_key
    L = len(chk.fan)
    for i in range(K):
        out = pipe.step()
        torch.cuda.synchronize()
        t = pipe.step_idx - 1
        sl = pipe.engine.slot_of(t)
        cnts = (C.c_int64 * (2 * L + 1))()
        check(lib().coala_sampler_counts(sl.sampler._h, cnts))
        n = int(cnts[0])
        idx_ptr, _ = sl.sampler.input_nodes_dev_ptrs()
        chk.rows(out.data_ptr(), idx_ptr, n)
        if i < oracle_batches:
            seeds_h = sl.last_seeds.cpu().numpy()
            inp = tensor_from_ptr(idx_ptr, (n,), torch.int64, chk.device).cpu().numpy()
            es = C.c_uint64()
            check(lib().coala_sampler_layer_ptrs(sl.sampler._h, 0, None, C.byref(es), None, None))
            edge_src = tensor_from_ptr(es.value, (int(cnts[1]),), torch.int64, chk.device).cpu().numpy()
            chk.blocks(seeds_h, batch_key(pipe.engine.seed, t, pipe.engine.rank), inp, edge_src)
            chk.seeds(seeds_h, pipe.seed_offset(t), sl.meta_used.cpu().numpy(), pipe.nd.parsing_method)
    chk.tot["refused_ids"] += pipe.mgr.COALA_GNN_Cache.refused_ids()


def measure_pipeline(args, comm, loader, device, K, min_seconds, rate_hint, warmup, seed=1001, profiler_range=False):
    """value-style measurement of one configured loader through DevicePipeline.  Returns (pipe, result dict)."""
    import torch
    from coala_gnn_b200.pipeline import DevicePipeline
    cache = loader.COALA_GNN_Manager.COALA_GNN_Cache
    pipe = DevicePipeline(loader, record_gather_events=True, seed=seed, prefetch=not args.no_prefetch)
    for _ in range(warmup):
        pipe.step()
    nw = n_windows_for(comm, rate_hint, K, min_seconds, device)
    torch.cuda.synchronize()
    cache.stats(reset=True)
    pipe.engine.reset_times()
    ms, region = timed_windows(comm, pipe.step, K, nw, pipe.engine.depth, device, profiler_range)
    h, m, b = cache.stats(reset=True)
    g_ms = pipe.gather_ms()
    s_ms = pipe.engine.sample_ms()
    steps_done = K * nw + pipe.engine.depth
    med = float(np.median(ms))
    hs, ms_, bs = reduce_ranks(comm, [h, m, b], device)
    res = {"value": comm.global_size * K / (med * 1e-3), "ms_per_step": med / K, "windows": nw, "window_steps": K,
           "window_ms_min_med_max": [float(np.min(ms)), med, float(np.max(ms))], "mean_value_all_windows": comm.global_size * K * nw / (float(np.sum(ms)) * 1e-3),
           "hit_ratio": hs / max(1.0, hs + ms_ + bs), "rows_per_step": (h + m + b) / steps_done, "miss_rows_per_step": (m + b) / steps_done,
           "gather_kernel_ms": float(np.mean(g_ms)) if g_ms else None, "sample_stream_ms": float(np.mean(s_ms)) if s_ms else None,
           "region": region}
    return pipe, res


def gather_links(res, R, G, hbm_peak, host_peak, peak_src):
    """Per-link algorithmic bytes of one gather launch against each link's peak: a row is 2R of HBM traffic
    (line read + output write on a hit, line write + output write on a miss), R over the host link on a miss, and R over
    NVLink into this GPU when its owner is a peer (fraction (G-1)/G of the rows under id % G routing)."""
    rows, miss_rows, t = res["rows_per_step"], res["miss_rows_per_step"], res["gather_kernel_ms"] * 1e-3
    remote = (G - 1) / G
    alg = {"hbm": 2.0 * R * rows, "host_link": R * miss_rows, "nvlink_in": R * (rows - miss_rows) * remote}
    peaks = {"hbm": hbm_peak, "host_link": host_peak, "nvlink_in": NVL_PEAK}
    srcs = {"hbm": peak_src, "host_link": "coala_probe_read_bandwidth (UVA streaming read, this run, all ranks probing at once)",
            "nvlink_in": "B200_PROFILING.md measured peer copy"}
    links = {k: {"achieved": alg[k] / t / 1e9, "peak": peaks[k], "unit": "GB/s", "frac": alg[k] / t / 1e9 / peaks[k],
                 "alg_bytes_per_launch": alg[k], "peak_source": srcs[k]} for k in alg}
    binding = max(alg, key=lambda k: alg[k] / peaks[k])  # the link with the largest lower bound on the kernel's time
    return links, binding


def limiter_of(res, depth):
    """What bounds a step: the gather stream is serial (one launch per step), samples overlap `depth`-fold."""
    g, s, step = res["gather_kernel_ms"] or 0.0, res["sample_stream_ms"] or 0.0, res["ms_per_step"]
    if g >= 0.85 * step:
        return "gather kernel (%.0f %% of the step)" % (100 * g / step)
    if s / max(1, depth) >= 0.7 * step:
        return "sampling chain (%.3f ms per sample, %d in flight)" % (s, depth)
    return "mixed: gather %.3f ms, sample %.3f ms / %d in flight, step %.3f ms" % (g, s, depth, step)


def replay_gather(comm, pipe, cache, fan, device, engines):
    """The gather kernel alone on real id lists whose rows are resident: a few more batches are sampled now and their lists
    gathered afterwards one by one with nothing else running on this GPU and the ranks aligned by a barrier (each list
    is used once per engine and the other lists, ~0.7 GB, pass through L2 in between)."""
    import torch
    from coala_gnn_b200._lib import lib, check
    from coala_gnn_b200.pybind import tensor_from_ptr
    torch.cuda.synchronize()
    lists = []
    for _ in range(8):
        pipe.step()
        torch.cuda.synchronize()
        sl = pipe.engine.slot_of(pipe.step_idx - 1)
        cnts = (C.c_int64 * (2 * len(fan) + 1))()
        check(lib().coala_sampler_counts(sl.sampler._h, cnts))
        idx_ptr, _ = sl.sampler.input_nodes_dev_ptrs()
        lists.append(tensor_from_ptr(idx_ptr, (int(cnts[0]),), torch.int64, device).clone())
    for _ in range(pipe.engine.depth):
        pipe.step()  # drain the prefetch: the rows of the lists above are resident now
    torch.cuda.synchronize()
    cache.stats(reset=True)
    out_buf = pipe.engine.slots[0].out
    res = {"rows": int(np.mean([x.numel() for x in lists]))}
    try:
        for eng in engines:
            cache.set_copy_engine(eng)
            r = []
            for ids in lists:
                barrier_sync(comm)
                r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                r0.record()
                cache.read_feature_async(out_buf.data_ptr(), ids.data_ptr(), ids.numel(), 0, torch.cuda.current_stream().cuda_stream)
                r1.record()
                torch.cuda.synchronize()
                r.append(r0.elapsed_time(r1))
            h, m, b = cache.stats(reset=True)
            res["engine_%d" % eng] = {"ms": float(np.median(r)), "hit_ratio": h / max(1, h + m + b)}
    finally:
        cache.set_copy_engine(int(os.environ.get("COALA_GATHER_ENGINE", "0")))
    return res


def preload_cache(comm, loader, ds, device):
    """Fill a cache through the real gather path with every id this rank's shard owns (partitioned: id % L == rank)."""
    import torch
    mgr = loader.COALA_GNN_Manager
    cache, ms = mgr.COALA_GNN_Cache, mgr.max_sample_size
    out = mgr.out_buffers(1)[0]
    part = mgr.cache_backend in ("nvshmem", "nccl") and comm.local_size > 1
    ids = torch.arange(comm.local_rank if part else 0, ds.num_nodes, comm.local_size if part else 1, dtype=torch.int64, device=device)
    st = torch.cuda.current_stream().cuda_stream
    for lo in range(0, ids.numel(), ms):
        chunk = ids[lo:lo + ms]
        cache.read_feature_async(out.data_ptr(), chunk.data_ptr(), chunk.numel(), 0, st)
    torch.cuda.synchronize()
    cache.stats(reset=True)



class LoaderDriver(object):
    """Draws batches from the drop-in iterator across epochs; every step copies the result header (the seeds) back to
    pinned host memory, the D2H read a training loop does for its loss."""

    def __init__(self, loader, B):
        import torch
        self.loader, self.it, self.last = loader, iter(loader), None
        self.pin = torch.empty(B, dtype=torch.int64).pin_memory()

    def step(self):
        try:
            self.last = next(self.it)
        except StopIteration:
            self.it = iter(self.loader)
            self.last = next(self.it)
        self.pin.copy_(self.last[1], non_blocking=True)


def measure_loader(comm, drv, K, n_windows, device, warmup):
    """e2e-style measurement on the host clock (the iterator blocks on the sampler's counts every step, so the host clock
    sees the device): barrier+synchronize, refill, n_windows windows of K steps, final synchronize in the last window."""
    import torch
    loader = drv.loader
    cache = loader.COALA_GNN_Manager.COALA_GNN_Cache
    for _ in range(max(3, warmup)):
        drv.step()
    barrier_sync(comm)
    cache.stats(reset=True)
    depth = loader.engine.depth if loader.engine is not None else 0
    for _ in range(depth):
        drv.step()
    t_a = time.time()
    stamps = [time.perf_counter()]
    for w in range(n_windows):
        for _ in range(K):
            drv.step()
        stamps.append(time.perf_counter())
    torch.cuda.synchronize()
    t_end = time.perf_counter()
    region = (t_a, time.time())
    e_ms = reduce_ranks(comm, [1e3 * (stamps[i + 1] - stamps[i]) for i in range(n_windows)], device, "max")
    e_total = reduce_ranks(comm, [t_end - stamps[0]], device, "max")[0]
    barrier_sync(comm)
    h, m, b = cache.stats(reset=True)
    hs, ms_, bs = reduce_ranks(comm, [h, m, b], device)
    med = float(np.median(e_ms))
    N = comm.global_size
    return {"value": N * K / (med * 1e-3), "ms_per_step": med / K, "windows": n_windows, "mean_value_all_windows": N * K * n_windows / e_total,
            "hit_ratio": hs / max(1.0, hs + ms_ + bs), "miss_rows_per_step": (m + b) / (K * n_windows + depth),
            "rows_per_step": (h + m + b) / (K * n_windows + depth), "region": region}


def check_loader(drv, chk, sampler, K, oracle_batches=4):
    """The next K batches of the iterator, checked one by one (untimed)."""
    import torch
    from coala_gnn_b200.sampler import batch_key
    loader = drv.loader
    for i in range(K):
        drv.step()
        input_nodes, seeds, blocks, feat = drv.last
        torch.cuda.synchronize()
        chk.rows(feat.data_ptr(), input_nodes.data_ptr(), input_nodes.numel())
        if i < oracle_batches:
            t_eng = (loader._epoch_base + loader.counter - 1) if loader.engine is not None else sampler.calls - 1
            chk.blocks(seeds.cpu().numpy(), batch_key(sampler.seed, t_eng, sampler.rank), input_nodes.cpu().numpy(), blocks[0].edge_src.cpu().numpy())
    chk.tot["refused_ids"] += loader.COALA_GNN_Manager.COALA_GNN_Cache.refused_ids()


# ------------------------------------------------------------------------------------------------------------
# main measurement
# ------------------------------------------------------------------------------------------------------------
def run_ours(args, state):
    import torch
    from coala_gnn_b200._lib import lib, check
    comm, ds, train, fan, nd, sampler, loader, device = build_world(args)
    state.update(comm=comm, ds=ds)
    N, B, R, K = comm.global_size, args.batch_size, ds.cache_dim * 4, args.steps
    clocks = ClockSampler(comm.device_index) if comm.global_rank == 0 else None
    cache = loader.COALA_GNN_Manager.COALA_GNN_Cache
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm_peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    if os.path.exists(peaks_path):
        hbm_peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json"
    gb = C.c_double()
    barrier_sync(comm)
    check(lib().coala_probe_read_bandwidth(ds.feat_data.data_ptr(), min(ds.num_nodes * R, 4 << 30), 2, C.byref(gb)))
    host_peak = gb.value

    # ---- untimed cache pre-warm through the real path (steady state is what an epoch sees)
    from coala_gnn_b200.pipeline import DevicePipeline
    warm = DevicePipeline(loader, seed=77, prefetch=not args.no_prefetch)
    t0 = time.time()
    n_pre = 0
    budget = args.prewarm if args.prewarm >= 0 else 10 ** 9
    while n_pre < budget:
        for _ in range(50):
            warm.step()
        torch.cuda.synchronize()
        n_pre += 50
        if args.prewarm < 0:  # every rank must leave after the same batch (the path holds collectives)
            if reduce_ranks(comm, [float(time.time() - t0 > args.prewarm_seconds)], device, "max")[0] > 0:
                break
    rate = n_pre / max(1e-3, time.time() - t0)
    del warm
    cache.stats(reset=True)

    # ---- value: device-resident pipeline, steady state
    chk = Checker(comm, ds, train, fan, nd, device)
    pipe, res = measure_pipeline(args, comm, loader, device, K, args.min_seconds, rate, args.warmup, profiler_range=args.cuda_profiler_range)
    regions = [res.pop("region")]
    if os.environ.get("COALA_SAMPLER_TRACE") and comm.global_rank == 0:
        sys.stderr.write("sampler trace (ms between ops): %s\n" % ["%.4f" % x for x in pipe.engine.slots[0].sampler.trace()])
    check_pipeline(pipe, chk, K)
    parity_value = chk.summary()

    # ---- the gather kernel alone on resident rows, both copy engines
    use_e1 = N == 1 or os.environ.get("COALA_BENCH_REPLAY_ENGINE1", "1") == "1"
    replay = replay_gather(comm, pipe, cache, fan, device, [0, 1] if use_e1 else [0])
    n_last = replay["rows"]

    # ---- e2e: the drop-in iterator (host seed list, host colour metadata, per-step D2H of the result header)
    loader.scheduler.refresh_counter = args.refresh_counter
    drv = LoaderDriver(loader, B)
    nw_e = n_windows_for(comm, res["value"] / N, K, args.min_seconds, device)
    prof = None
    if args.profile_e2e and comm.global_rank == 0:
        import cProfile
        prof = cProfile.Profile()
        prof.enable()
    e2e = measure_loader(comm, drv, K, nw_e, device, args.warmup)
    if prof is not None:
        import pstats
        prof.disable()
        pstats.Stats(prof, stream=sys.stderr).sort_stats("cumulative").print_stats(35)
    regions.append(e2e.pop("region"))
    chk2 = Checker(comm, ds, train, fan, nd, device)
    check_loader(drv, chk2, sampler, K)
    parity_e2e = chk2.summary()
    clk = clocks.stop(regions) if clocks else None

    L = comm.local_size
    # per step and rank: seeds H2D through the pinned staging buffer; D2H of the block counts, of the step header the loop
    # reads back, and (every refresh_counter steps) of the cache colour histogram
    h2d = B * 8
    d2h = B * 8 + 8 * (2 * len(fan) + 1) + nd.meta_len * 4 / args.refresh_counter
    out = None
    G = L if args.cache_backend in ("nvshmem", "p2p") else 1
    remote = (G - 1) / G
    links, binding = gather_links(res, R, G, hbm_peak, host_peak, peak_src)
    state.update(train=train, fan=fan, device=device, hbm_peak=hbm_peak, host_peak=host_peak, peak_src=peak_src, rate=res["value"] / N,
                 main_res=res, main_loader=loader, main_pipe=pipe)
    if comm.global_rank == 0:
        # dram__bytes_read.sum + dram__bytes_write.sum per launch from the ncu --set full capture of this same workload
        # (profiles/r2_gather_n1_bench_full_raw.csv: 78.8 MB + 89.0 MB; pcie__read_bytes 39.7 MB for 36.0 MB of missed rows);
        # only quoted for the configuration it was captured on
        traffic = 167.8e6 if (N == 1 and args.shape == "igb-medium" and args.cache_size == 16384 and args.fan_out == "5,5" and B == 1024) else None
        t_meas = res["gather_kernel_ms"] * 1e-3

        def replay_block(r):
            return {"ms": r["ms"], "hit_ratio": r["hit_ratio"],
                    "hbm": {"achieved": 2.0 * R * n_last / (r["ms"] * 1e-3) / 1e9, "peak": hbm_peak, "frac": 2.0 * R * n_last / (r["ms"] * 1e-3) / 1e9 / hbm_peak},
                    "nvlink_in": {"achieved": R * n_last * remote / (r["ms"] * 1e-3) / 1e9, "peak": NVL_PEAK, "frac": R * n_last * remote / (r["ms"] * 1e-3) / 1e9 / NVL_PEAK}}
        out = {
            "metric": metric_string(args), "value": res["value"], "unit": "mini-batches/s", "n_gpus": N, "steps": K, "warmup": args.warmup,
            "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "impl": "ours",
            "config": {"workload": workload_string(args, ds, comm.num_master_process, L),
                       "l2": "each step gathers ~%.0f MB of rows chosen by fresh sampling (a different row set every step; the table is %.0f GB, L2 is 126 MB)" % (res["rows_per_step"] * R / 1e6, ds.num_nodes * R / 1e9),
                       "cache_prewarm_batches": n_pre, "refresh_counter": args.refresh_counter,
                       "colour_metadata": ds.color_source,
                       "cache_protocol": "read-back ordered writers (COALA_CACHE_FENCED=0, validated on B200 only)" if os.environ.get("COALA_CACHE_FENCED", "1") == "0"
                                         else "fenced writers (PTX acquire / fence.acq_rel, default)",
                       "graph": "CSC indptr/indices/edge_ids in pinned host memory (UVA)" + (", indptr and indices copies in HBM" if sampler.hbm_graph else
                                                                                            ", indptr copy in HBM" if sampler.hbm_indptr else "")},
            "timing": {"protocol": "barrier+synchronize, %d untimed refill steps, then %d back-to-back windows of %d steps (one CUDA event per window on the consumer stream, "
                                   "no drain in between), barrier+synchronize; max over ranks per window; value = median window" % (pipe.engine.depth, res["windows"], K),
                       "windows": res["windows"], "window_ms_min_med_max": res["window_ms_min_med_max"], "mean_value_all_windows": res["mean_value_all_windows"],
                       "scaling_refers_to": "this line: %d MB cache per GPU, so the per-GPU hit ratio rises with N (capacity effect); extras.all_resident is the same-regime line" % args.cache_size},
            "hit_ratio": res["hit_ratio"], "rows_per_step": res["rows_per_step"],
            "gather_GBs": res["rows_per_step"] * R / t_meas / 1e9, "sample_stream_ms": res["sample_stream_ms"],
            "limiter": limiter_of(res, pipe.engine.depth),
            "parity": {"value_leg": parity_value, "e2e_leg": parity_e2e,
                       "rows_checked": parity_value["rows_checked"] + parity_e2e["rows_checked"],
                       "mismatches": parity_value["mismatches"] + parity_e2e["mismatches"] + parity_value["block_mismatches"] + parity_e2e["block_mismatches"]
                       + parity_value["seed_mismatches"] + parity_e2e["seed_mismatches"],
                       "ranks": N, "sampler_note": "blocks are compared with this repo's CPU restatement of DGL's documented semantics (DGL is absent: parity with DGL itself is unpinned)"},
            # the dominant kernel against the link that bounds it (T_bound = max over links of bytes/peak);
            # `hbm` is always listed too, and `all_hit_replay` times the same kernel on resident rows (the HBM / NVLink regime)
            "roofline": {"bound": binding, "achieved": links[binding]["achieved"], "peak": links[binding]["peak"], "unit": "GB/s",
                         "frac": links[binding]["frac"], "traffic": traffic,
                         "traffic_source": ("profiles/r2_gather_n1_bench_full_raw.csv (ncu --set full of this workload, dram__bytes_read.sum + dram__bytes_write.sum per launch; "
                                            "not re-measured in this run)") if traffic else None,
                         "peak_source": links[binding]["peak_source"],
                         "kernel": "coala_gather_kernel", "kernel_ms": res["gather_kernel_ms"], "kernel_share_of_step": res["gather_kernel_ms"] / res["ms_per_step"],
                         "links": links,
                         "all_hit_replay": dict(rows=n_last, **{k: replay_block(v) for k, v in replay.items() if k.startswith("engine_")})},
            "e2e": {"value": e2e["value"], "unit": "mini-batches/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "windows": nw_e, "mean_value_all_windows": e2e["mean_value_all_windows"],
                    "host_link_feature_bytes_per_step": int(R * e2e["miss_rows_per_step"]), "hit_ratio": e2e["hit_ratio"],
                    "api": "COALA_GNN_DataLoader.__next__"},
            "gpu_launches": pipe.kernels_per_step * K * res["windows"],
            "gpu_launches_per_step": pipe.kernels_per_step,
            "clocks": clk,
        }
    return out


# ------------------------------------------------------------------------------------------------------------
# extras: the other regimes of the reference's experiment matrix, measured the same way in the same process
# ------------------------------------------------------------------------------------------------------------
def summarize(res, pipe, R, G, st):
    links, binding = gather_links(res, R, G, st["hbm_peak"], st["host_peak"], st["peak_src"])
    return {"value": res["value"], "ms_per_step": res["ms_per_step"], "hit_ratio": res["hit_ratio"], "windows": res["windows"],
            "gather_kernel_ms": res["gather_kernel_ms"], "sample_stream_ms": res["sample_stream_ms"], "rows_per_step": res["rows_per_step"],
            "binding_link": binding, "binding_link_frac": links[binding]["frac"], "binding_link_GBs": links[binding]["achieved"],
            "limiter": limiter_of(res, pipe.engine.depth)}


def extra_all_resident(args, st):
    """Same-regime scaling line: the cache of every N holds the whole table (hit ratio 1.0), so 1 -> 8 compares like with like."""
    import torch
    comm, ds, device = st["comm"], st["ds"], st["device"]
    R, K, L = ds.cache_dim * 4, args.steps, comm.local_size
    if st["main_res"]["hit_ratio"] >= 0.9995:
        return dict(summarize(st["main_res"], st["main_pipe"], R, L, st), cache_mb_per_gpu=args.cache_size, same_run_as_main_line=True)
    lines_needed = ds.num_nodes / L
    mb = int(math.ceil(lines_needed * 1.2 * R / (1 << 20) / 1024.0)) * 1024
    _, _, _, fan, nd, sampler, loader, _ = build_world(args, comm=comm, ds=ds, cache_size=mb)
    preload_cache(comm, loader, ds, device)
    pipe, res = measure_pipeline(args, comm, loader, device, K, 1.0, 4000.0, args.warmup, seed=2002)
    res.pop("region")
    chk = Checker(comm, ds, st["train"], fan, nd, device)
    check_pipeline(pipe, chk, min(K, 8), oracle_batches=2)
    replay = replay_gather(comm, pipe, loader.COALA_GNN_Manager.COALA_GNN_Cache, fan, device, [0, 1])
    out = dict(summarize(res, pipe, R, L, st), cache_mb_per_gpu=mb, same_run_as_main_line=False, parity=chk.summary())
    n_last = replay["rows"]
    for k, v in replay.items():
        if k.startswith("engine_"):
            out["gather_alone_" + k] = {"ms": v["ms"], "hbm_frac": 2.0 * R * n_last / (v["ms"] * 1e-3) / 1e9 / st["hbm_peak"],
                                       "nvlink_in_frac": R * n_last * (L - 1) / L / (v["ms"] * 1e-3) / 1e9 / NVL_PEAK, "rows": n_last}
    del pipe, loader
    torch.cuda.empty_cache()
    return out


def extra_distribution(args, st):
    """The reference's headline comparison (examples/Distribution_compare_script.sh:26-35): node_color vs baseline over D x L
    cache domains at 4096 MB per GPU, colour metadata from the colouring of the graph.  Stage 1 does real work here:
    nd_score_kernel / nd_assign_kernel run every step on the domain histograms refreshed every `refresh_counter` batches."""
    import torch
    from COALA_GNN import MPI_Comm_Manager
    comm0, ds, device = st["comm"], st["ds"], st["device"]
    N, K, R = comm0.global_size, args.steps, ds.cache_dim * 4
    layouts = [L for L in (1, 2, 4) if N % L == 0 and N // L >= 2]
    out = {"cache_mb_per_gpu": 4096, "colour_metadata": ds.color_source, "layouts": {}}
    for L in layouts:
        comm = MPI_Comm_Manager(-1, gpus_per_domain=L)
        row = {}
        for method in ("baseline", "node_color"):
            _, _, _, fan, nd, sampler, loader, _ = build_world(args, comm=comm, ds=ds, cache_size=4096, backend="nvshmem", distribution=method)
            from coala_gnn_b200.pipeline import DevicePipeline
            warm = DevicePipeline(loader, seed=55)
            for _ in range(1200):  # several cache turnovers: 1 M lines per GPU, ~2e4 fills per batch
                warm.step()
            torch.cuda.synchronize()
            del warm
            pipe, res = measure_pipeline(args, comm, loader, device, K, 1.0, 1000.0, args.warmup, seed=3003)
            res.pop("region")
            chk = Checker(comm, ds, st["train"], fan, nd, device)
            check_pipeline(pipe, chk, min(K, 8), oracle_batches=3)
            row[method] = dict(summarize(res, pipe, R, L, st), parity=chk.summary(), gpu_launches_per_step=pipe.kernels_per_step)
            if method == "node_color":  # stage 1 alone: score + assign of one global batch, on this rank
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                dom = pipe.domain_seeds[0]
                ptrs = [pipe.meta[d].data_ptr() for d in range(pipe.D)]
                torch.cuda.synchronize()
                e0.record()
                for _ in range(20):
                    nd.distribute_manager.distribute_async(0, dom.data_ptr(), ptrs, torch.cuda.current_stream().cuda_stream, slot=1)
                e1.record()
                torch.cuda.synchronize()
                row[method]["distribute_kernels_ms"] = e0.elapsed_time(e1) / 20
            del pipe, loader, nd
            torch.cuda.empty_cache()
        row["hit_ratio_gain"] = row["node_color"]["hit_ratio"] - row["baseline"]["hit_ratio"]
        row["speedup_node_color_over_baseline"] = row["node_color"]["value"] / row["baseline"]["value"]
        out["layouts"]["%dx%d" % (N // L, L)] = row
    return out


def extra_backends(args, st):
    """examples/Cache_compare_script.sh:28-35: partitioned P2P cache vs isolated caches vs the NCCL all-to-all variant, 4096 MB per GPU,
    one domain spanning the box."""
    import torch
    comm, ds, device = st["comm"], st["ds"], st["device"]
    K, R, L = args.steps, ds.cache_dim * 4, comm.local_size
    out = {"cache_mb_per_gpu": 4096}
    for backend in ("nvshmem", "isolated", "nccl"):
        _, _, _, fan, nd, sampler, loader, _ = build_world(args, comm=comm, ds=ds, cache_size=4096, backend=backend)
        chk = Checker(comm, ds, st["train"], fan, nd, device)
        if loader.engine is None:  # a backend outside the prefetch engine: through the iterator, host clock
            drv = LoaderDriver(loader, args.batch_size)
            for _ in range(600):
                drv.step()
            res = measure_loader(comm, drv, K, n_windows_for(comm, 300.0, K, 1.0, device), device, args.warmup)
            res.pop("region")
            check_loader(drv, chk, sampler, min(K, 8), oracle_batches=2)
            out[backend] = {"value": res["value"], "ms_per_step": res["ms_per_step"], "hit_ratio": res["hit_ratio"], "windows": res["windows"],
                            "through": "COALA_GNN_DataLoader iterator (host clock)", "parity": chk.summary()}
            del drv, loader, nd
            torch.cuda.empty_cache()
            continue
        from coala_gnn_b200.pipeline import DevicePipeline
        warm = DevicePipeline(loader, seed=66)
        for _ in range(1200):
            warm.step()
        torch.cuda.synchronize()
        del warm
        pipe, res = measure_pipeline(args, comm, loader, device, K, 1.0, 1000.0, args.warmup, seed=4004)
        res.pop("region")
        check_pipeline(pipe, chk, min(K, 8), oracle_batches=2)
        out["p2p" if backend == "nvshmem" else backend] = dict(summarize(res, pipe, R, L if backend == "nvshmem" else 1, st), parity=chk.summary(),
                                                               through="DevicePipeline (CUDA events)")
        del pipe, loader, nd
        torch.cuda.empty_cache()
    return out


def extra_stress(args, st):
    """Cross-GPU eviction storm (tests/multi_gpu_stress.py): every rank hammers the same two sets of every shard of a 1 MB
    partitioned cache at once, launches alternating between the copy engines, every row of every launch checked bitwise."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from multi_gpu_stress import eviction_storm
    return eviction_storm(st["comm"], st["device"], 3000, tag="%s_storm" % os.environ.get("MASTER_PORT", str(os.getpid())))


def run_extras(args, st, out):
    comm = st["comm"]
    N = comm.global_size
    want = args.extras
    if want == "none":
        return
    names = ["all_resident"] + (["distribution", "backends", "stress"] if N >= 2 else [])
    if want != "auto":
        names = [n for n in want.split(",") if n in ("all_resident", "distribution", "backends", "stress") and (N >= 2 or n == "all_resident")]
    if comm.num_master_process != 1:
        names = [n for n in names if n == "all_resident"]
    extras = {}
    fns = {"all_resident": extra_all_resident, "distribution": extra_distribution, "backends": extra_backends, "stress": extra_stress}
    for n in names:
        t0 = time.time()
        try:
            extras[n] = fns[n](args, st)
        except Exception as e:  # an extra must never cost the main line (every rank takes the same path: same inputs)
            import traceback
            extras[n] = {"error": "%s: %s" % (type(e).__name__, e), "where": traceback.format_exc().strip().splitlines()[-3:]}
        if isinstance(extras[n], dict):
            extras[n]["seconds"] = round(time.time() - t0, 1)
        if out is not None:
            out["extras"] = extras  # visible to the watchdog as soon as a section is complete


def finalize_parity(out):
    """Make the line self-verifying.  The driver keeps `e2e`, `roofline` and `config` whole but only the NAMES of other keys,
    so a compact parity summary is copied into the first two and the headline extras into `config`; a main-line parity
    failure voids the number (`value`: null) and the process exits non-zero.  Returns the exit code."""
    par = out.get("parity", {})
    legs = [par.get("value_leg", {}), par.get("e2e_leg", {})]
    ok = bool(par) and par.get("mismatches", 1) == 0 and all(l.get("ok", False) for l in legs)
    extras = out.get("extras", {}) or {}

    def walk(d, path, bad):  # every nested block that carries an `ok` flag next to parity-style counters
        if isinstance(d, dict):
            if "ok" in d and ("mismatches" in d or "corrupt_rows" in d) and not d["ok"]:
                bad.append(path)
            for k, v in d.items():
                walk(v, path + "." + str(k), bad)
    bad = []
    walk(extras, "extras", bad)
    stress = extras.get("stress", {}) if isinstance(extras.get("stress"), dict) else {}
    summary = {"parity_ok": ok, "parity_rows_checked": par.get("rows_checked"), "parity_ranks": par.get("ranks"),
               "seed_lists_checked": sum(l.get("seed_lists_checked", 0) for l in legs),
               "blocks_checked": sum(l.get("blocks_checked", 0) for l in legs),
               "stress_corrupt_rows": stress.get("corrupt_rows"), "extras_parity_ok": not bad}
    out["e2e"].update(summary)
    out["roofline"].update(summary)
    out["extras_parity_ok"] = not bad
    if bad:
        out["extras_parity_failed"] = bad

    def pick(d, *keys):
        for k in keys:
            d = d.get(k) if isinstance(d, dict) else None
        return round(d, 4) if isinstance(d, float) else d
    head = {"all_resident.value": pick(extras, "all_resident", "value"),
            "all_resident.binding_link_frac": pick(extras, "all_resident", "binding_link_frac")}
    dist = extras.get("distribution")
    if isinstance(dist, dict):
        for shape, row in (dist.get("layouts") or {}).items():
            if isinstance(row, dict):
                for m in ("baseline", "node_color"):
                    if isinstance(row.get(m), dict):
                        head["distribution.%s.%s.hit_ratio" % (shape, m)] = pick(row, m, "hit_ratio")
                        head["distribution.%s.%s.value" % (shape, m)] = pick(row, m, "value")
    back = extras.get("backends")
    if isinstance(back, dict):
        for b in ("p2p", "isolated", "nccl"):
            if isinstance(back.get(b), dict):
                head["backends.%s.value" % b] = pick(back, b, "value")
    out["config"]["extras_summary"] = {k: v for k, v in head.items() if v is not None}
    if not ok:
        out["value_voided_by_parity"] = out.get("value")
        out["value"] = None
        out["e2e"]["value_voided_by_parity"] = out["e2e"].get("value")
        out["e2e"]["value"] = None
        return 3
    return 0


def main():
    args = parse_args()
    if os.environ.get("COALA_HANG_DUMP"):  # debugging aid: dump every thread's stack if the run stalls
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["COALA_HANG_DUMP"]), repeat=True, exit=False)
    world = int(os.environ.get("WORLD_SIZE", 1))
    rank = int(os.environ.get("RANK", 0))
    if args.impl == "reference":
        if rank != 0:
            return 0
        os.environ["WORLD_SIZE"], os.environ["RANK"] = "1", "0"
        from COALA_GNN import MPI_Comm_Manager
        from coala_gnn_b200.datasets import SyntheticSharedCSCUVA
        from coala_gnn_b200 import synth
        comm = MPI_Comm_Manager(0)
        os.environ["COALA_SYNTH_COLOR"] = "block"  # the CPU arm never reads colour metadata
        ds = SyntheticSharedCSCUVA(args.shape, comm, "cuda:0", verbose=True)
        try:
            fan = [int(f) for f in args.fan_out.split(",")]
            train = synth.train_ids(ds.num_nodes, seed=0)
            r = cpu_path_windows((ds.csc_host[0], ds.csc_host[1], ds.feat_host_array()), train, args.batch_size, fan, args.steps, args.warmup,
                                 args.min_seconds)
            line = {"metric": metric_string(args), "value": r["batches_per_s"], "unit": "mini-batches/s", "n_gpus": args.gpus, "steps": args.steps,
                    "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                    "dtype": "f32", "data": "synthetic", "impl": "reference",
                    "config": {"workload": workload_string(args, ds, *domains_for(args)),
                               "arm": "the reference's CPU formulation of the same workload: CPU sampler (OpenMP restatement of the DGL CPU sampler, DGL absent) + "
                                      "torch.index_select gather on %d host threads; the cache / backend / distribution fields of `workload` name the GPU arm's setting" % r["threads"]},
                    "cpu_baseline": {"value": r["batches_per_s"], "unit": "mini-batches/s", "cores": r["threads"], "kind": "port",
                                     "sample": "median of %d windows of %d mini-batches of %d seeds (%.0f gathered rows each) after %d warm-up batches; window rates min/med/max %.0f/%.0f/%.0f"
                                               % (r["windows"], r["steps"], args.batch_size, r["rows_per_step"], r["warm_batches"], *r["window_rate_min_med_max"])},
                    "timing": {"protocol": "warm-up until two consecutive windows agree within 5 % (<= 10 s), then windows of --steps batches for >= --min_seconds; value = median window",
                               "windows": r["windows"], "warm_batches": r["warm_batches"]},
                    "gather_GBs": r["gather_GBs"],
                    "e2e": {"value": r["batches_per_s"], "unit": "mini-batches/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
            print(json.dumps(line), flush=True)
        finally:
            ds.cleanup()
        return 0

    if args.gpus > 1 and world == 1:  # convenience: plain `python bench.py --gpus N` re-launches itself under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus), "--master-addr", "127.0.0.1",
               "--master-port", "29511", os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    state = {}
    printed = threading.Event()
    out_box = [None]

    exit_code = [0]

    def emit(note=None):
        if printed.is_set():
            return
        printed.set()
        if out_box[0] is not None:
            if note:
                out_box[0].setdefault("extras", {})["abandoned"] = note
            if "roofline" in out_box[0]:  # rank 0's line
                exit_code[0] = finalize_parity(out_box[0])
            print(json.dumps(out_box[0]), flush=True)

    def bail():  # the extras must never cost the main line: print what is complete and leave (every rank has the same timer)
        emit("extras exceeded --extras_budget of %.0f s; sections listed are complete" % args.extras_budget)
        ds = state.get("ds")
        try:
            if ds is not None:
                ds.unlink_segments()  # kernels may still be reading the mappings: drop the names only
        finally:
            os._exit(exit_code[0])

    comm = None
    try:
        out_box[0] = run_ours(args, state)
        comm = state["comm"]
        timer = threading.Timer(args.extras_budget, bail)
        timer.daemon = True
        timer.start()
        try:
            run_extras(args, state, out_box[0])
        except Exception as e:
            if out_box[0] is not None:
                out_box[0].setdefault("extras", {})["error"] = "%s: %s" % (type(e).__name__, e)
        finally:
            timer.cancel()
        if comm.global_rank == 0:
            ds, train, fan = state["ds"], state["train"], state["fan"]
            if world == 1 and not args.no_cpu_baseline:
                # the same protocol as `--impl reference` (settle, then windows of --steps, median), so the two CPU numbers of one box agree
                r = cpu_path_windows((ds.csc_host[0], ds.csc_host[1], ds.feat_host_array()), train.numpy(), args.batch_size, fan, args.steps, args.warmup,
                                     args.cpu_baseline_seconds)
                out_box[0]["cpu_baseline"] = {"value": r["batches_per_s"], "unit": "mini-batches/s", "cores": r["threads"], "kind": "port",
                                              "sample": "median of %d windows of %d mini-batches of %d seeds after %d warm-up batches (oracle OpenMP sampler + torch.index_select, "
                                                        "%.1f GB/s gathered); window rates min/med/max %.0f/%.0f/%.0f"
                                                        % (r["windows"], r["steps"], args.batch_size, r["warm_batches"], r["gather_GBs"], *r["window_rate_min_med_max"])}
            emit()
    finally:
        comm = state.get("comm")
        if comm is not None and comm.global_size > 1:
            import torch.distributed as dist
            try:
                dist.barrier()
            except Exception:
                pass
        ds = state.get("ds")
        if ds is not None:
            ds.cleanup()
        if comm is not None and comm.global_size > 1:
            try:
                import torch.distributed as dist
                dist.destroy_process_group()
            except Exception:
                pass
    return exit_code[0]


if __name__ == "__main__":
    sys.exit(main())
