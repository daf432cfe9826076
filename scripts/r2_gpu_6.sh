#!/bin/bash
# round-2 GPU session 6 (2 GPUs): gather with deferred re-check + interleaved front end — parity (cache, multi-GPU, storm),
# microbench both engines, protocol-free probe on all three links, ncu of the NVLink / mixed regimes
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cache.py tests/test_gpu_multi.py tests/test_gpu_dataloader.py -x -q > gpurun_out/r2_6_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_6_pytest.log
tail -6 gpurun_out/r2_6_pytest.log
for e in 0 1; do
COALA_GATHER_ENGINE=$e timeout 400 python scripts/bench_gather.py --gpus 2 --nodes 2000000 --cache_mb 8192 --regimes iso_hit,iso_miss03,iso_miss10,p2p_hit_one,p2p_hit_both,p2p_miss03_both,p2p_miss10_both,p2p_remote_only_one,p2p_local_only_one > gpurun_out/r2_6_gather_micro_e$e.log 2>&1; echo "micro e$e rc=$?"
grep -E "^(iso_hit|iso_miss|p2p_hit|p2p_miss|p2p_remote_only_one|p2p_local_only_one)" gpurun_out/r2_6_gather_micro_e$e.log | cut -c1-120
done
timeout 300 python scripts/probe_gather.py --gpus 2 > gpurun_out/r2_6_probe_gather.log 2>&1; echo "probe rc=$?"
grep -E "^(local|peer|pinned)" gpurun_out/r2_6_probe_gather.log | sed 's/.*"best"/"best"/'
NCU="ncu --set full --import-source on --clock-control none --metrics nvlrx__bytes.sum,nvltx__bytes.sum,pcie__read_bytes.sum,pcie__write_bytes.sum -k regex:coala_gather -f"
A="--nodes 2000000 --cache_mb 8192 --iters 2"
timeout 300 $NCU -s 9 -c 1 -o gpurun_out/r2_6_iso_hit_e0 python scripts/bench_gather.py --regimes iso_hit $A > gpurun_out/r2_6_ncu_iso_hit_e0.log 2>&1; echo "ncu1 rc=$?"
timeout 300 $NCU -s 5 -c 1 -o gpurun_out/r2_6_remote_only_e0 python scripts/bench_gather.py --gpus 2 --regimes p2p_remote_only_one $A > gpurun_out/r2_6_ncu_remote_only_e0.log 2>&1; echo "ncu3 rc=$?"
timeout 300 $NCU -s 10 -c 1 -o gpurun_out/r2_6_miss03_e0 python scripts/bench_gather.py --gpus 2 --regimes p2p_miss03_both $A > gpurun_out/r2_6_ncu_miss03_e0.log 2>&1; echo "ncu5 rc=$?"
ls -la gpurun_out/r2_6*.ncu-rep
