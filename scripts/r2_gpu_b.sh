#!/bin/bash
# round-2 GPU session B (2 GPUs): multi-GPU parity (both bindings, D x L device pipeline, native NCCL exchange), bench N=2 with extras
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/r2b_env.txt; nproc >> gpurun_out/r2b_env.txt; free -g >> gpurun_out/r2b_env.txt
(ncu --query-metrics 2>/dev/null | grep -i -E "nvl|pcie" | head -60) > gpurun_out/r2b_ncu_link_metrics.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -5 gpurun_out/r2b_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR bench.py --gpus 2 --shape igb-small --cache_size 1024 --steps 20 --warmup 5 --prewarm_seconds 3 > gpurun_out/r2b_small.json 2> gpurun_out/r2b_small.err; echo "small rc=$?"
tail -c 1500 gpurun_out/r2b_small.err
timeout 900 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2b_n2.json 2> gpurun_out/r2b_n2.err; echo "n2 rc=$?"
tail -c 1500 gpurun_out/r2b_n2.err
