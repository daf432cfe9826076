#!/bin/bash
# round-2 GPU session 17 (2 GPUs, head with indptr in HBM): the driver's N=4 commands (ours, then the reference arm under torchrun: rank 0 alone works)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 700 $TR --master-port 29551 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_17_n2.json 2> gpurun_out/r2_17_n2.err; echo "n4 rc=$?"
tail -c 300 gpurun_out/r2_17_n2.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r2_17_n2.json').read().strip().splitlines()[-1])
    r=d['roofline']
    print('value',d['value'],'e2e',d['e2e']['value'],'ms',d['ms_per_step'],'hit',d['hit_ratio'],'kernel_ms',r['kernel_ms'],'bound',r['bound'],r['frac'],'sample_ms',d['sample_stream_ms'],'parity',d['parity']['mismatches'])
    print({k:v['ms'] for k,v in r['all_hit_replay'].items() if k.startswith('engine')})
    x=d.get('extras',{}); print('extras', list(x.keys()), x.get('abandoned'), x.get('error'))
    a=x.get('all_resident'); print(a and (a['value'],a['ms_per_step'],a['gather_kernel_ms'],a['binding_link_frac'],a.get('gather_alone_engine_0'),a.get('gather_alone_engine_1')))
    if 'distribution' in x and 'layouts' in x['distribution']:
        for k,v in x['distribution']['layouts'].items(): print(k, {m:(round(v[m]['value']),round(v[m]['hit_ratio'],3)) for m in ('baseline','node_color')})
    if 'backends' in x: print({k:(round(v['value']),round(v['hit_ratio'],3)) for k,v in x['backends'].items() if isinstance(v,dict) and 'value' in v})
    print(x.get('stress'))
except Exception as e: print('ERR',e)
PY
