#!/bin/bash
# round-2 GPU session 9 (8 GPUs): the driver's N=8 command (extras: all-resident, D x L distribution, backends, storm), then
# parity of every batch of every rank at 1x8 and 4x2, and the 8-rank eviction storm
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv | head -3 > gpurun_out/r2_9_env.txt; nproc >> gpurun_out/r2_9_env.txt; free -g >> gpurun_out/r2_9_env.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29541 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_9_n8.json 2> gpurun_out/r2_9_n8.err; echo "n8 rc=$?"
tail -c 600 gpurun_out/r2_9_n8.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r2_9_n8.json').read().strip().splitlines()[-1])
    r=d['roofline']
    print('value',d['value'],'e2e',d['e2e']['value'],'ms',d['ms_per_step'],'hit',d['hit_ratio'],'kernel_ms',r['kernel_ms'],'bound',r['bound'],r['frac'],'sample_ms',d['sample_stream_ms'],'parity',d['parity']['mismatches'])
    print({k:v['ms'] for k,v in r['all_hit_replay'].items() if k.startswith('engine')})
    x=d.get('extras',{})
    print('extras', list(x.keys()), x.get('abandoned'))
    if 'distribution' in x:
        for k,v in x['distribution']['layouts'].items(): print(k, {m:(round(v[m]['value']),round(v[m]['hit_ratio'],3)) for m in ('baseline','node_color')}, v['node_color'].get('distribute_kernels_ms'))
    if 'backends' in x: print({k:(round(v['value']),round(v['hit_ratio'],3)) for k,v in x['backends'].items() if isinstance(v,dict)})
    if 'stress' in x: print(x['stress'])
except Exception as e: print('ERR',e)
PY
timeout 400 $TR --master-port 29542 tests/multi_gpu_check.py --steps 6 > gpurun_out/r2_9_check_1x8.log 2>&1; echo "check 1x8 rc=$?"; grep -E "^ok|all good" gpurun_out/r2_9_check_1x8.log | cut -c1-200 | tail -14
timeout 400 $TR --master-port 29543 tests/multi_gpu_check.py --steps 6 --gpus_per_domain 2 --backends nvshmem > gpurun_out/r2_9_check_4x2.log 2>&1; echo "check 4x2 rc=$?"; grep -E "^ok|all good" gpurun_out/r2_9_check_4x2.log | cut -c1-200 | tail -6
timeout 300 $TR --master-port 29544 tests/multi_gpu_stress.py --launches 10000 > gpurun_out/r2_9_storm_8.log 2>&1; echo "storm rc=$?"; grep -E "^stress|multi_gpu_stress" gpurun_out/r2_9_storm_8.log | cut -c1-400
