#!/bin/bash
# round-2 GPU session 4 (2 GPUs): dynamic-queue gather kernels — full parity suite, eviction storm, microbench per regime with both
# copy engines, bench N=2
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_4_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_4_pytest.log
tail -12 gpurun_out/r2_4_pytest.log
for e in 0 1; do
COALA_GATHER_ENGINE=$e timeout 400 python scripts/bench_gather.py --gpus 2 --nodes 2000000 --cache_mb 8192 > gpurun_out/r2_4_gather_micro_e$e.log 2>&1; echo "micro e$e rc=$?"
grep -E "^(iso_hit|iso_miss|p2p_hit|p2p_miss|p2p_remote_only_one|p2p_local_only_one)" gpurun_out/r2_4_gather_micro_e$e.log | cut -c1-150
done
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 700 $TR bench.py --gpus 2 --steps 20 --warmup 5 --extras all_resident > gpurun_out/r2_4_n2.json 2> gpurun_out/r2_4_n2.err; echo "n2 rc=$?"
tail -c 600 gpurun_out/r2_4_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_4_n2.json').read().strip().splitlines()[-1])
r=d['roofline']
print('value',d['value'],'e2e',d['e2e']['value'],'hit',d['hit_ratio'],'kernel_ms',r['kernel_ms'],'bound',r['bound'],r['frac'],'parity',d['parity']['mismatches'])
print({k:v['ms'] for k,v in r['all_hit_replay'].items() if k.startswith('engine')})
print(d['extras']['all_resident']['value'], d['extras']['all_resident']['gather_kernel_ms'], d['extras']['all_resident'].get('gather_alone_engine_0'), d['extras']['all_resident'].get('gather_alone_engine_1'))
PY
