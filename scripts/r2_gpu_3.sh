#!/bin/bash
# round-2 GPU session 3 (1 GPU): native batch engine — parity tests, e2e vs value on igb-small (host-bound regime), host profile
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_3_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_3_pytest.log
tail -15 gpurun_out/r2_3_pytest.log
B="python bench.py --shape igb-small --cache_size 4096 --steps 20 --warmup 5 --prewarm_seconds 3 --no_cpu_baseline --extras none"
timeout 300 $B > gpurun_out/r2_3_small.json 2> gpurun_out/r2_3_small.err; echo "small rc=$?"
timeout 300 $B --hbm_indptr --profile_e2e > gpurun_out/r2_3_small_hbm.json 2> gpurun_out/r2_3_small_hbm.err; echo "hbm rc=$?"
grep -A45 "cumulative" gpurun_out/r2_3_small_hbm.err | head -60
for f in gpurun_out/r2_3_small*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], 'value',d['value'],'e2e',d['e2e']['value'],'ok',d['parity']['value_leg']['ok'],d['parity']['e2e_leg']['ok'],'gather_ms',d['roofline']['kernel_ms'],'sample_ms',d['sample_stream_ms'])
except Exception as e: print("ERR",e)
PY
done
