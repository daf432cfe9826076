#!/bin/bash
# round-3 GPU session 3 (2 GPUs): fenced writers as the default - whole GPU suite, the 2-domain check with its output kept,
# bench N=1 (full line + reference arm under the new window protocol), bench N=2 in both protocols, probe ceilings restated
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r3_3_pytest_gpu_n2.log 2>&1; echo "suite rc=$?"; tail -2 gpurun_out/r3_3_pytest_gpu_n2.log
timeout 300 $TR --master-port 29611 tests/multi_gpu_check.py --gpus_per_domain 1 --backends isolated,nvshmem > gpurun_out/r3_3_check_2x1.log 2>&1; echo "check 2x1 rc=$?"; grep -E "^ok|all good|Error|assert" gpurun_out/r3_3_check_2x1.log | head -12
timeout 300 python scripts/probe_gather.py --gpus 2 > gpurun_out/r3_3_probe_gather.log 2>&1; echo "probe rc=$?"; grep -E "^(local_hbm|peer_hbm_nvlink|pinned_host) " gpurun_out/r3_3_probe_gather.log | sed -E 's/.*"best": "([^"]*)", "best_us": ([0-9.]*), "best_row_GBs": ([0-9.]*).*/best \1 \2 us \3 GB\/s/'
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r3_3_ref_n1.json 2> gpurun_out/r3_3_ref_n1.err; echo "ref rc=$?"
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r3_3_n1.json 2> gpurun_out/r3_3_n1.err; echo "n1 rc=$?"
for f in 1 0; do
  COALA_CACHE_FENCED=$f timeout 500 $TR --master-port 2963$f bench.py --gpus 2 --steps 20 --warmup 5 --extras none --no_cpu_baseline > gpurun_out/r3_3_n2_f$f.json 2> gpurun_out/r3_3_n2_f$f.err; echo "n2 f$f rc=$?"
done
python - <<'PY'
import json
def last(p):
    return json.loads(open(p).read().strip().splitlines()[-1])
try:
    d=last('gpurun_out/r3_3_ref_n1.json'); print('REF value',d['value'],d['cpu_baseline']['sample'][:160])
except Exception as e: print('ERR ref',e)
for name in ('n1','n2_f1','n2_f0'):
    try:
        d=last('gpurun_out/r3_3_%s.json'%name); r=d['roofline']
        print(name,'value',d['value'],'e2e',d['e2e']['value'],'ms',d['ms_per_step'],'hit',d.get('hit_ratio'),'kernel_ms',r.get('kernel_ms'),r['bound'],round(r['frac'],3),'parity_ok',d['e2e'].get('parity_ok'),'cpu',d.get('cpu_baseline',{}).get('value'),d['config'].get('cache_protocol'),d['config'].get('extras_summary'))
    except Exception as e: print('ERR',name,e)
PY
