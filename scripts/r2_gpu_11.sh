#!/bin/bash
# round-2 GPU session 11 (8 GPUs): the driver's N=8 command (extras: all-resident, D x L distribution, backends, storm) + the 8-rank storm
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 700 $TR --master-port 29541 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_11_n8.json 2> gpurun_out/r2_11_n8.err; echo "n8 rc=$?"
tail -c 300 gpurun_out/r2_11_n8.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r2_11_n8.json').read().strip().splitlines()[-1])
    r=d['roofline']
    print('value',d['value'],'e2e',d['e2e']['value'],'ms',d['ms_per_step'],'hit',d['hit_ratio'],'kernel_ms',r['kernel_ms'],'bound',r['bound'],r['frac'],'sample_ms',d['sample_stream_ms'],'parity',d['parity']['mismatches'])
    print({k:v['ms'] for k,v in r['all_hit_replay'].items() if k.startswith('engine')})
    x=d.get('extras',{})
    print('extras', list(x.keys()), x.get('abandoned'), x.get('error'))
    if 'distribution' in x and 'layouts' in x['distribution']:
        for k,v in x['distribution']['layouts'].items(): print(k, {m:(round(v[m]['value']),round(v[m]['hit_ratio'],3)) for m in ('baseline','node_color')}, v['node_color'].get('distribute_kernels_ms'))
    if 'backends' in x: print({k:(round(v['value']),round(v['hit_ratio'],3)) for k,v in x['backends'].items() if isinstance(v,dict) and 'value' in v})
    print(x.get('stress'))
except Exception as e: print('ERR',e)
PY
timeout 200 $TR --master-port 29544 tests/multi_gpu_stress.py --launches 10000 > gpurun_out/r2_11_storm_8.log 2>&1; echo "storm rc=$?"; grep -E "^stress|multi_gpu_stress" gpurun_out/r2_11_storm_8.log | cut -c1-400
