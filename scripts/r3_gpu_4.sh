#!/bin/bash
# round-3 GPU session 4 (2 GPUs): NCCL variant with the size exchange first + exact-length id lists + growing serve buffers:
# GPU suite, 2-rank check of the nccl backend, bench N=2 on the nccl backend; reference arm and N=1 line once more (CPU numbers under one protocol)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r3_4_pytest_gpu_n2.log 2>&1; echo "suite rc=$?"; tail -2 gpurun_out/r3_4_pytest_gpu_n2.log
timeout 300 $TR --master-port 29711 tests/multi_gpu_check.py --backends nccl > gpurun_out/r3_4_check_nccl.log 2>&1; echo "check nccl rc=$?"; grep -E "^ok|all good|Error|assert" gpurun_out/r3_4_check_nccl.log | cut -c1-260 | head -8
timeout 500 $TR --master-port 29721 bench.py --gpus 2 --steps 20 --warmup 5 --cache_backend nccl --extras none --no_cpu_baseline > gpurun_out/r3_4_n2_nccl.json 2> gpurun_out/r3_4_n2_nccl.err; echo "n2 nccl rc=$?"
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r3_4_ref_n1.json 2> gpurun_out/r3_4_ref_n1.err; echo "ref rc=$?"
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 --extras none > gpurun_out/r3_4_n1.json 2> gpurun_out/r3_4_n1.err; echo "n1 rc=$?"
python - <<'PY'
import json
def last(p):
    return json.loads(open(p).read().strip().splitlines()[-1])
for name in ('ref_n1','n1','n2_nccl'):
    try:
        d=last('gpurun_out/r3_4_%s.json'%name)
        print(name,'value',d['value'],'e2e',d['e2e']['value'],'ms',d.get('ms_per_step'),'hit',d.get('hit_ratio'),'parity_ok',d['e2e'].get('parity_ok'),'cpu',d.get('cpu_baseline',{}).get('value'))
    except Exception as e: print('ERR',name,e)
PY
