#!/bin/bash
# round-2 GPU session 20 (2 GPUs): neighbour lists in HBM (--hbm_graph) at N=2, no extras
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29561 bench.py --gpus 2 --steps 20 --warmup 5 --hbm_graph --extras none --no_cpu_baseline > gpurun_out/r2_20_n2_graph.json 2> gpurun_out/r2_20_n2_graph.err; echo "rc=$?"
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r2_20_n2_graph.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('value',d['value'],'e2e',d['e2e']['value'],'ms',d['ms_per_step'],'hit',d['hit_ratio'],'kernel_ms',r['kernel_ms'],'bound',r['bound'],r['frac'],'sample_ms',d['sample_stream_ms'],'parity',d['parity']['mismatches'])
except Exception as e: print('ERR',e)
PY
