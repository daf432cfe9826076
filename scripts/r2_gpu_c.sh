#!/bin/bash
# round-2 GPU session C (1 GPU): new sampler parity + effect; hbm_indptr / copy engine 1 in the pipeline
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2c_pytest.log
tail -4 gpurun_out/r2c_pytest.log
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_sampler.py -x -q > gpurun_out/r2c_memcheck_sampler.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/r2c_memcheck_sampler.log
tail -4 gpurun_out/r2c_memcheck_sampler.log
B="python bench.py --shape igb-small --cache_size 4096 --steps 20 --warmup 5 --prewarm_seconds 3 --no_cpu_baseline --extras none"
timeout 300 $B > gpurun_out/r2c_small_base.json 2> gpurun_out/r2c_small_base.err; echo "base rc=$?"
timeout 300 $B --hbm_indptr > gpurun_out/r2c_small_hbm.json 2> gpurun_out/r2c_small_hbm.err; echo "hbm rc=$?"
COALA_GATHER_ENGINE=1 timeout 300 $B > gpurun_out/r2c_small_e1.json 2> gpurun_out/r2c_small_e1.err; echo "e1 rc=$?"
COALA_GATHER_ENGINE=1 timeout 300 $B --hbm_indptr > gpurun_out/r2c_small_e1hbm.json 2> gpurun_out/r2c_small_e1hbm.err; echo "e1hbm rc=$?"
COALA_SAMPLER_TRACE=1 timeout 300 $B > gpurun_out/r2c_small_trace.json 2> gpurun_out/r2c_small_trace.err; echo "trace rc=$?"
grep "sampler trace" gpurun_out/r2c_small_trace.err
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 --hbm_indptr > gpurun_out/r2c_n1_hbm.json 2> gpurun_out/r2c_n1_hbm.err; echo "n1 rc=$?"
tail -c 500 gpurun_out/r2c_n1_hbm.err
