#!/bin/bash
# round-2 GPU session 2 (2 GPUs): multi-GPU parity + eviction storm, bench N=2 (extras: D x L distribution, backends incl. native NCCL),
# gather microbench (baseline for the kernel work), ncu link metric names
mkdir -p gpurun_out
(ncu --query-metrics 2>/dev/null | grep -i -E "nvl|pcie" | head -80) > gpurun_out/r2_2_ncu_link_metrics.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/r2_2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_2_pytest.log
tail -15 gpurun_out/r2_2_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 500 $TR bench.py --gpus 2 --shape igb-small --cache_size 1024 --steps 20 --warmup 5 --prewarm_seconds 3 > gpurun_out/r2_2_small.json 2> gpurun_out/r2_2_small.err; echo "small rc=$?"
tail -c 1200 gpurun_out/r2_2_small.err
timeout 700 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_2_n2.json 2> gpurun_out/r2_2_n2.err; echo "n2 rc=$?"
tail -c 1200 gpurun_out/r2_2_n2.err
timeout 400 python scripts/bench_gather.py --gpus 2 --nodes 2000000 --cache_mb 8192 > gpurun_out/r2_2_gather_micro.log 2>&1; echo "micro rc=$?"
grep -E "^(iso_hit|iso_miss|p2p_hit|p2p_miss|p2p_remote_only_one|p2p_local_only_one)" gpurun_out/r2_2_gather_micro.log | cut -c1-260
