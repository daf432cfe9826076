#!/bin/bash
# round-2 GPU session 13 (1 GPU): config C4 — papers100M-shaped graph (111 M nodes, 1.6 B edges, 128-d rows = 512 B), three layers of
# fan-out 10,10,10, batch 1024, 16 GB cache; the reference arm line format check on igb-small
mkdir -p gpurun_out
free -g | head -2 > gpurun_out/r2_13_env.txt
timeout 1500 python bench.py --gpus 1 --shape papers100m --fan_out 10,10,10 --cache_size 16384 --steps 10 --warmup 3 --prewarm_seconds 15 --cpu_baseline_seconds 8 --extras none > gpurun_out/r2_13_c4.json 2> gpurun_out/r2_13_c4.err; echo "c4 rc=$?"
tail -c 800 gpurun_out/r2_13_c4.err
timeout 300 python bench.py --impl reference --shape igb-small --gpus 1 --steps 5 --warmup 2 > gpurun_out/r2_13_ref_small.json 2> gpurun_out/r2_13_ref_small.err; echo "ref rc=$?"; head -c 700 gpurun_out/r2_13_ref_small.json
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r2_13_c4.json').read().strip().splitlines()[-1])
    r=d['roofline']
    print('value',d['value'],'e2e',d['e2e']['value'],'ms',d['ms_per_step'],'hit',d['hit_ratio'],'rows',d['rows_per_step'],'kernel_ms',r['kernel_ms'],r['bound'],r['frac'],'sample_ms',d['sample_stream_ms'],'limiter',d['limiter'],'cpu',d.get('cpu_baseline'),'parity',d['parity']['mismatches'])
    print(r['links']); print(r['all_hit_replay'])
except Exception as e: print('ERR',e)
PY
