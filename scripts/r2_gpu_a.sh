#!/bin/bash
# round-2 GPU session A (1 GPU): parity tests, bench on igb-small (fast) then the driver's command on the default config
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/r2a_env.txt; nproc >> gpurun_out/r2a_env.txt; free -g >> gpurun_out/r2a_env.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
timeout 600 python bench.py --shape igb-small --cache_size 4096 --steps 20 --warmup 5 --prewarm_seconds 3 --cpu_baseline_seconds 3 > gpurun_out/r2a_small.json 2> gpurun_out/r2a_small.err; echo "small rc=$?"
tail -c 600 gpurun_out/r2a_small.err
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2a_n1.json 2> gpurun_out/r2a_n1.err; echo "n1 rc=$?"
tail -c 600 gpurun_out/r2a_n1.err
