#!/usr/bin/env python
"""bench.py — mini-batches/s of the COALA-GNN data path (seed distribution -> UVA neighbour sampling
-> cached feature gather) on synthetic IGB-shaped data.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One JSON line on rank 0.  A "step" is one mini-batch (batch 1024, SAGE fan-out 5,5) per GPU.
  value : device-resident pipeline (coala_gnn_b200.pipeline.DevicePipeline), CUDA-event timed;
  e2e   : the drop-in COALA_GNN_DataLoader iterator with host seed lists / host metadata;
  roofline : the gather kernel, timed live with CUDA events inside the timed region;
  cpu_baseline : oracle sampler (OpenMP) + index_select on the host cores (bounded sample).
`--impl reference` times that CPU path as its own arm.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "mini-batches/s (IGB-medium-shaped synthetic graph, SAGE fan-out 5,5, batch 1024 per GPU; seed distribution + UVA sampling + cached feature gather)"


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=500)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--shape", default=os.environ.get("COALA_BENCH_SHAPE", "igb-medium"))
    p.add_argument("--cache_size", type=int, default=16384, help="MB per GPU (reference flag name)")
    p.add_argument("--cache_backend", default="nvshmem", choices=["nvshmem", "p2p", "isolated", "nccl"])
    p.add_argument("--distribution", default="node_color", choices=["node_color", "baseline"])
    p.add_argument("--batch_size", type=int, default=1024)
    p.add_argument("--fan_out", default="5,5")
    p.add_argument("--refresh_counter", type=int, default=10)
    p.add_argument("--prewarm", type=int, default=-1, help="untimed batches that warm the cache (-1: one pass budgeted by --prewarm_seconds)")
    p.add_argument("--prewarm_seconds", type=float, default=25.0)
    p.add_argument("--cpu_baseline_seconds", type=float, default=12.0)
    p.add_argument("--no_cpu_baseline", action="store_true")
    p.add_argument("--hbm_indptr", action="store_true", help="B200 option: sampler keeps a copy of indptr (80 MB) in HBM")
    p.add_argument("--no_prefetch", action="store_true", help="diagnostic: sample and gather of a batch back to back, no batch in flight")
    p.add_argument("--profile_e2e", action="store_true", help="cProfile the host side of the e2e loop (rank 0) to stderr")
    p.add_argument("--cuda_profiler_range", action="store_true",
                   help="bracket the timed region with cudaProfilerStart/Stop (for `ncu --profile-from-start off`)")
    return p.parse_args()


class ClockSampler(object):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
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
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def cpu_path(ds_host, train, B, fan, seconds=None, steps=None, warmup=2, threads=None):
    """The reference's CPU formulation of the path (BASELINE.json: "DGL CPU sampler + torch index_select
    gather"): oracle OpenMP sampler + to_block, then a row gather from the host feature table."""
    import oracle
    indptr, indices, feat = ds_host
    threads = threads or os.cpu_count()
    max_sample = B
    for f in fan:
        max_sample *= f + 1
    times, rows = [], 0
    i = 0
    t_end = time.time() + seconds if seconds else None
    while True:
        seeds = train[(i % (len(train) // B - 1)) * B:][:B]
        t0 = time.perf_counter()
        inp, _ = oracle.sample_blocks_fast(indptr, indices, seeds, fan, oracle.batch_key(1, i), max_sample, threads)
        out = oracle.index_select(feat, inp, threads)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
            rows += len(inp)
        i += 1
        if steps is not None and len(times) >= steps:
            break
        if t_end is not None and time.time() > t_end and len(times) >= 3:
            break
    del out
    tot = float(np.sum(times))
    return {"batches_per_s": len(times) / tot, "ms_per_step": 1e3 * tot / len(times), "rows_per_step": rows / len(times),
            "gather_GBs": rows * feat.shape[1] * 4 / tot / 1e9, "steps": len(times), "threads": threads}


def build_world(args):
    import torch
    from COALA_GNN import MPI_Comm_Manager, Node_Distributor, SSD_INFO, COALA_GNN_DataLoader, MultiLayerNeighborSampler
    from coala_gnn_b200.datasets import SyntheticSharedCSCUVA
    from coala_gnn_b200 import synth
    comm = MPI_Comm_Manager(-1)  # one cache domain spanning all ranks of the box
    torch.cuda.set_device(comm.device_index)
    device = "cuda:%d" % comm.device_index
    ds = SyntheticSharedCSCUVA(args.shape, comm, device, verbose=True)
    fan = [int(f) for f in args.fan_out.split(",")]
    train = torch.from_numpy(synth.train_ids(ds.num_nodes, seed=0))
    nd = Node_Distributor(comm, train, args.batch_size, ds.color_file, ds.topk_file, ds.score_file, parsing_method=args.distribution)
    sampler = MultiLayerNeighborSampler(fan, seed=1, rank=comm.global_rank, hbm_indptr=True if args.hbm_indptr else None)
    backend = args.cache_backend
    loader = COALA_GNN_DataLoader(SSD_INFO(1, ds.cache_dim * 4, 1024, 0), nd, ds.graph, sampler, args.batch_size, ds.dim, fan,
                                  args.cache_size, device, refresh_counter=args.refresh_counter, cache_backend=backend,
                                  sim_buf=ds.feat_data)
    return comm, ds, train, fan, nd, sampler, loader, device


def barrier_sync(comm):
    import torch
    import torch.distributed as dist
    torch.cuda.synchronize()
    if comm.global_size > 1:
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(comm, x, device):
    import torch
    import torch.distributed as dist
    if comm.global_size == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(comm, xs, device):
    import torch
    import torch.distributed as dist
    if comm.global_size == 1:
        return list(xs)
    t = torch.tensor(list(xs), dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(v) for v in t.tolist()]


def run_ours(args):
    import ctypes as C
    import torch
    from coala_gnn_b200._lib import lib, check
    from coala_gnn_b200.pipeline import DevicePipeline
    comm, ds, train, fan, nd, sampler, loader, device = build_world(args)
    N, B, R = comm.global_size, args.batch_size, ds.cache_dim * 4
    cache = loader.COALA_GNN_Manager.COALA_GNN_Cache
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm_peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    if os.path.exists(peaks_path):
        hbm_peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json"
    gb = C.c_double()
    check(lib().coala_probe_read_bandwidth(ds.feat_data.data_ptr(), min(ds.num_nodes * R, 4 << 30), 2, C.byref(gb)))
    host_peak = gb.value

    pipe = DevicePipeline(loader, record_gather_events=True, seed=1001, prefetch=not args.no_prefetch)  # its own sample stream: the e2e leg must not replay these batches
    spe = pipe.steps_per_epoch()

    def pipe_steps(n):
        for _ in range(n):
            pipe.step()  # epochs wrap inside the pipeline

    # ---- untimed cache pre-warm through the real path (steady state is what an epoch sees)
    t0 = time.time()
    n_pre = 0
    budget = args.prewarm if args.prewarm >= 0 else 10 ** 9
    while n_pre < budget:
        pipe_steps(50)
        torch.cuda.synchronize()
        n_pre += 50
        if args.prewarm < 0:  # every rank must leave after the same batch (the path holds collectives)
            if max_over_ranks(comm, float(time.time() - t0 > args.prewarm_seconds), device) > 0:
                break
    pre_stats = cache.stats(reset=True)
    pipe.engine.gather_times.clear()
    pipe.engine.sample_times.clear()

    # ---- value: device-resident pipeline
    pipe_steps(args.warmup)
    barrier_sync(comm)
    cache.stats(reset=True)
    pipe.engine.gather_times.clear()
    pipe.engine.sample_times.clear()
    clocks = ClockSampler(comm.device_index) if comm.global_rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if args.cuda_profiler_range:
        torch.cuda.profiler.start()
    e0.record()
    pipe_steps(args.steps)
    e1.record()
    if args.cuda_profiler_range:
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    barrier_sync(comm)
    dev_ms = max_over_ranks(comm, e0.elapsed_time(e1), device)
    clk = clocks.stop() if clocks else None
    h, m, b = cache.stats(reset=True)
    gather_ms = float(np.mean(pipe.gather_ms()))
    sample_ms = float(np.mean(pipe.engine.sample_ms()))
    if os.environ.get("COALA_SAMPLER_TRACE") and comm.global_rank == 0:
        sys.stderr.write("sampler trace (ms between ops): %s\n" % ["%.4f" % x for x in pipe.engine.slots[0].sampler.trace()])
    rows = (h + m + b) / args.steps
    value = N * args.steps / (dev_ms * 1e-3)

    # ---- replay: id lists of a few more batches, sampled now and gathered afterwards one by one with nothing
    # else running on this GPU and the ranks aligned by a barrier: the gather kernel alone, in the
    # multi-process mapping, on real id lists (each list used once, so L2 holds none of its rows)
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
    for _ in range(3):
        pipe.step()  # the sampled rows of the lists above are resident now; drain the prefetch
    torch.cuda.synchronize()
    cache.stats(reset=True)
    replay, out_buf = [], pipe.engine.slots[0].out
    for ids in lists:
        barrier_sync(comm)
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        r0.record()
        cache.read_feature_async(out_buf.data_ptr(), ids.data_ptr(), ids.numel(), 0, torch.cuda.current_stream().cuda_stream)
        r1.record()
        torch.cuda.synchronize()
        replay.append(r0.elapsed_time(r1))
    replay_ms = float(np.median(replay))
    n_last = int(np.mean([x.numel() for x in lists]))
    rh, rm, rb = cache.stats(reset=True)
    # the same lists once more with copy engine 1 (hit rows as bulk asynchronous copies through shared memory); the 7 other
    # lists (~0.7 GB) have passed through L2 since a list was last touched
    replay_e1_ms, replay_e1_err = None, "not run (multi-process runs only with COALA_BENCH_REPLAY_ENGINE1=1)"
    try:
        # engine 1 over peer mappings has been measured in-process (scripts/bench_gather.py --gpus 2), not yet across
        # processes: a device fault there could not be caught, so multi-rank runs keep to engine 0 unless asked
        r = []
        use_e1 = N == 1 or os.environ.get("COALA_BENCH_REPLAY_ENGINE1") == "1"
        if use_e1:
            cache.set_copy_engine(1)
        for ids in (lists if use_e1 else []):
            barrier_sync(comm)
            r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            r0.record()
            cache.read_feature_async(out_buf.data_ptr(), ids.data_ptr(), ids.numel(), 0, torch.cuda.current_stream().cuda_stream)
            r1.record()
            torch.cuda.synchronize()
            r.append(r0.elapsed_time(r1))
        if r:
            replay_e1_ms = float(np.median(r))
    except RuntimeError as e:  # symmetric on all ranks (same launch, same arguments): the barriers stay matched
        replay_e1_err = str(e)
    finally:
        cache.set_copy_engine(0)
        cache.stats(reset=True)

    # ---- e2e: the drop-in iterator (host seed list, host colour metadata, per-step D2H of the result ids)
    it = iter(loader)
    loader.scheduler.refresh_counter = args.refresh_counter
    pin = torch.empty(B, dtype=torch.int64).pin_memory()

    def loader_steps(n):
        nonlocal it
        last = None
        for _ in range(n):
            try:
                last = next(it)
            except StopIteration:
                it = iter(loader)
                last = next(it)
            pin.copy_(last[1], non_blocking=True)  # D2H read of the step's result header (seeds)
        return last

    loader_steps(max(3, args.warmup // 2))
    barrier_sync(comm)
    cache.stats(reset=True)
    prof = None
    if args.profile_e2e and comm.global_rank == 0:
        import cProfile
        prof = cProfile.Profile()
        prof.enable()
    t0 = time.perf_counter()
    loader_steps(args.steps)
    torch.cuda.synchronize()
    e2e_local = time.perf_counter() - t0
    if prof is not None:
        import pstats
        prof.disable()
        pstats.Stats(prof, stream=sys.stderr).sort_stats("cumulative").print_stats(35)
    e2e_s = max_over_ranks(comm, e2e_local, device)
    barrier_sync(comm)
    eh, em, eb = cache.stats(reset=True)
    e2e_value = N * args.steps / e2e_s
    L = comm.local_size
    # per step and rank: seeds H2D through the pinned staging buffer; D2H of the block counts, of the step header the loop
    # reads back, and (every refresh_counter steps) of the cache colour histogram
    h2d = B * 8
    d2h = B * 8 + 8 * (2 * len(fan) + 1) + nd.meta_len * 4 / args.refresh_counter

    hs, ms_, bs = sum_over_ranks(comm, [h, m, b], device)
    out = None
    if comm.global_rank == 0:
        # per-link algorithmic bytes of one gather launch: a row is 2R of HBM traffic (line read +
        # output write on a hit, line write + output write on a miss), R over the host link on a miss, and R over
        # NVLink into this GPU when its owner is a peer (fraction (G-1)/G of the rows under id % G routing)
        miss_rows = (m + b) / args.steps
        G = comm.local_size if args.cache_backend in ("nvshmem", "p2p") else 1
        remote = (G - 1) / G
        alg = {"hbm": 2.0 * R * rows, "host_link": R * miss_rows, "nvlink_in": R * (rows - miss_rows) * remote}
        nvl_peak = 770.0  # measured peer-copy figure of B200_PROFILING.md (900 nominal); scripts/bench_gather.py reads 777 with a kernel
        peaks = {"hbm": hbm_peak, "host_link": host_peak, "nvlink_in": nvl_peak}
        srcs = {"hbm": peak_src, "host_link": "coala_probe_read_bandwidth (UVA streaming read, this run)",
                "nvlink_in": "B200_PROFILING.md measured peer copy"}
        t_meas = gather_ms * 1e-3
        links = {k: {"achieved": alg[k] / t_meas / 1e9, "peak": peaks[k], "unit": "GB/s", "frac": alg[k] / t_meas / 1e9 / peaks[k],
                     "alg_bytes_per_launch": alg[k], "peak_source": srcs[k]} for k in alg}
        binding = max(alg, key=lambda k: alg[k] / peaks[k])  # the link with the largest lower bound on the kernel's time
        hbm_replay = 2.0 * R * n_last / (replay_ms * 1e-3) / 1e9
        # dram__bytes_read.sum + dram__bytes_write.sum per launch from the ncu --set full capture of this same workload
        # (profiles/r1b_gather_full_raw.csv: 77.1 MB + 88.9 MB); only quoted for the configuration it was captured on
        traffic = 166.0e6 if (N == 1 and args.shape == "igb-medium" and args.cache_size == 16384 and args.fan_out == "5,5" and B == 1024) else None
        out = {
            "metric": METRIC, "value": value, "unit": "mini-batches/s", "n_gpus": N, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "impl": "ours",
            "config": {"workload": "%s (%d nodes, %d edges, %d-d fp32 features pinned in host memory), SAGE fan-out %s, batch %d/GPU, %d MB cache/GPU, backend %s (%dx%d domain), distribution %s"
                       % (args.shape, ds.num_nodes, ds.num_edges, ds.dim, args.fan_out, B, args.cache_size, args.cache_backend, comm.num_master_process, L, args.distribution),
                       "l2": "each step gathers ~%.0f MB of rows chosen by fresh sampling (a different row set every step; the table is %.0f GB, L2 is 126 MB)" % (rows * R / 1e6, ds.num_nodes * R / 1e9),
                       "cache_prewarm_batches": n_pre, "refresh_counter": args.refresh_counter,
                       "graph": "CSC indptr/indices/edge_ids in pinned host memory (UVA)" + (", indptr copy in HBM" if sampler.hbm_indptr else "")},
            "hit_ratio": hs / max(1.0, hs + ms_ + bs), "rows_per_step": rows,
            "gather_GBs": rows * R / t_meas / 1e9, "sample_stream_ms": sample_ms,
            # the dominant kernel against the link that bounds it (T_bound = max over links of bytes/peak);
            # `hbm` is always listed too, and `all_hit_replay` times the same kernel on resident rows (the HBM / NVLink regime)
            "roofline": {"bound": binding, "achieved": links[binding]["achieved"], "peak": links[binding]["peak"], "unit": "GB/s",
                         "frac": links[binding]["frac"], "traffic": traffic, "peak_source": links[binding]["peak_source"],
                         "kernel": "coala_gather_kernel", "kernel_ms": gather_ms, "kernel_share_of_step": gather_ms / (dev_ms / args.steps),
                         "links": links,
                         "all_hit_replay": {"ms": replay_ms, "rows": n_last, "hit_ratio": rh / max(1, rh + rm + rb),
                                            "copy_engine_1": ({"ms": replay_e1_ms, "hbm_frac": 2.0 * R * n_last / (replay_e1_ms * 1e-3) / 1e9 / hbm_peak,
                                                               "nvlink_in_frac": R * n_last * remote / (replay_e1_ms * 1e-3) / 1e9 / nvl_peak}
                                                              if replay_e1_ms else {"error": replay_e1_err}),
                                            "hbm": {"achieved": hbm_replay, "peak": hbm_peak, "frac": hbm_replay / hbm_peak},
                                            "nvlink_in": {"achieved": R * n_last * remote / (replay_ms * 1e-3) / 1e9, "peak": nvl_peak,
                                                          "frac": R * n_last * remote / (replay_ms * 1e-3) / 1e9 / nvl_peak}}},
            "e2e": {"value": e2e_value, "unit": "mini-batches/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "host_link_feature_bytes_per_step": int(R * (em + eb) / args.steps), "hit_ratio": eh / max(1, eh + em + eb),
                    "api": "COALA_GNN_DataLoader.__next__"},
            "gpu_launches": pipe.kernels_per_step * args.steps,
            "clocks": clk,
        }
    return comm, ds, train, fan, out


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
        ds = SyntheticSharedCSCUVA(args.shape, comm, "cuda:0", verbose=True)
        try:
            fan = [int(f) for f in args.fan_out.split(",")]
            train = synth.train_ids(ds.num_nodes, seed=0)
            r = cpu_path((ds.csc_host[0], ds.csc_host[1], ds.feat_host_array()), train, args.batch_size, fan, steps=args.steps, warmup=args.warmup)
            line = {"metric": METRIC, "value": r["batches_per_s"], "unit": "mini-batches/s", "n_gpus": args.gpus, "steps": args.steps,
                    "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                    "dtype": "f32", "data": "synthetic", "impl": "reference",
                    "config": {"workload": "%s, SAGE fan-out %s, batch %d: CPU sampler (OpenMP restatement of the DGL CPU sampler, DGL absent) + index_select gather on %d host threads"
                               % (args.shape, args.fan_out, args.batch_size, r["threads"])},
                    "cpu_baseline": {"value": r["batches_per_s"], "unit": "mini-batches/s", "cores": r["threads"], "kind": "port",
                                     "sample": "%d mini-batches of %d seeds (%.0f gathered rows each)" % (r["steps"], args.batch_size, r["rows_per_step"])},
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
    comm = ds = None
    try:
        comm, ds, train, fan, out = run_ours(args)
        if comm.global_rank == 0:
            if world == 1 and not args.no_cpu_baseline:
                r = cpu_path((ds.csc_host[0], ds.csc_host[1], ds.feat_host_array()), train.numpy(), args.batch_size, fan,
                             seconds=args.cpu_baseline_seconds)
                out["cpu_baseline"] = {"value": r["batches_per_s"], "unit": "mini-batches/s", "cores": r["threads"], "kind": "port",
                                       "sample": "%d mini-batches of %d seeds in %.1f s (oracle OpenMP sampler + index_select, %.1f GB/s gathered)"
                                                 % (r["steps"], args.batch_size, r["steps"] / r["batches_per_s"], r["gather_GBs"])}
            print(json.dumps(out), flush=True)
    finally:
        if comm is not None and comm.global_size > 1:
            import torch.distributed as dist
            try:
                dist.barrier()
            except Exception:
                pass
        if ds is not None:
            ds.cleanup()
    return 0


if __name__ == "__main__":
    sys.exit(main())
