#!/bin/bash
# round-2 GPU session 1 (1 GPU): parity tests, sampler memcheck, igb-small variants, the driver's default command
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/r2_1_env.txt; nproc >> gpurun_out/r2_1_env.txt; free -g >> gpurun_out/r2_1_env.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_1_pytest.log
tail -5 gpurun_out/r2_1_pytest.log
timeout 400 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_sampler.py -x -q > gpurun_out/r2_1_memcheck_sampler.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/r2_1_memcheck_sampler.log
tail -4 gpurun_out/r2_1_memcheck_sampler.log
B="python bench.py --shape igb-small --cache_size 4096 --steps 20 --warmup 5 --prewarm_seconds 3 --no_cpu_baseline --extras none"
timeout 300 $B > gpurun_out/r2_1_small_base.json 2> gpurun_out/r2_1_small_base.err; echo "base rc=$?"
timeout 300 $B --hbm_indptr > gpurun_out/r2_1_small_hbm.json 2> gpurun_out/r2_1_small_hbm.err; echo "hbm rc=$?"
COALA_GATHER_ENGINE=1 timeout 300 $B > gpurun_out/r2_1_small_e1.json 2> gpurun_out/r2_1_small_e1.err; echo "e1 rc=$?"
COALA_SAMPLER_TRACE=1 timeout 300 $B > gpurun_out/r2_1_small_trace.json 2> gpurun_out/r2_1_small_trace.err; echo "trace rc=$?"
grep "sampler trace" gpurun_out/r2_1_small_trace.err | head
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_1_n1.json 2> gpurun_out/r2_1_n1.err; echo "n1 rc=$?"
tail -c 800 gpurun_out/r2_1_n1.err
for f in gpurun_out/r2_1_small_*.json gpurun_out/r2_1_n1.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ("value","ms_per_step","parity")}, d.get("e2e",{}).get("value"), d.get("roofline",{}).get("frac"))
except Exception as e: print("ERR",e)
PY
done
