#!/bin/bash
# round-2 GPU session 5 (2 GPUs): where does a gathered row's time go?  protocol-free row-gather probe (the practical ceiling per link),
# then ncu --set full (+ NVLink / PCIe byte counters) of single gather launches per regime
mkdir -p gpurun_out
timeout 300 python scripts/probe_gather.py --gpus 2 > gpurun_out/r2_5_probe_gather.log 2>&1; echo "probe rc=$?"
grep -E "^(local|peer|pinned)" gpurun_out/r2_5_probe_gather.log | cut -c1-1200
NCU="ncu --set full --import-source on --clock-control none --metrics nvlrx__bytes.sum,nvltx__bytes.sum,pcie__read_bytes.sum,pcie__write_bytes.sum -k regex:coala_gather -f"
A="--nodes 2000000 --cache_mb 8192 --iters 2"
timeout 300 $NCU -s 9 -c 1 -o gpurun_out/r2_5_iso_hit_e0 python scripts/bench_gather.py --regimes iso_hit $A > gpurun_out/r2_5_ncu_iso_hit_e0.log 2>&1; echo "ncu1 rc=$?"
COALA_GATHER_ENGINE=1 timeout 300 $NCU -s 9 -c 1 -o gpurun_out/r2_5_iso_hit_e1 python scripts/bench_gather.py --regimes iso_hit $A > gpurun_out/r2_5_ncu_iso_hit_e1.log 2>&1; echo "ncu2 rc=$?"
timeout 300 $NCU -s 5 -c 1 -o gpurun_out/r2_5_remote_only_e0 python scripts/bench_gather.py --gpus 2 --regimes p2p_remote_only_one $A > gpurun_out/r2_5_ncu_remote_only_e0.log 2>&1; echo "ncu3 rc=$?"
timeout 300 $NCU -s 9 -c 1 -o gpurun_out/r2_5_hit_one_e0 python scripts/bench_gather.py --gpus 2 --regimes p2p_hit_one $A > gpurun_out/r2_5_ncu_hit_one_e0.log 2>&1; echo "ncu4 rc=$?"
timeout 300 $NCU -s 10 -c 1 -o gpurun_out/r2_5_miss03_e0 python scripts/bench_gather.py --gpus 2 --regimes p2p_miss03_both $A > gpurun_out/r2_5_ncu_miss03_e0.log 2>&1; echo "ncu5 rc=$?"
ls -la gpurun_out/*.ncu-rep
