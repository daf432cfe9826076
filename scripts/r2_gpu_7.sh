#!/bin/bash
# round-2 GPU session 7 (2 GPUs): miss-prefetch pass + busy-line wait — parity, microbench (with / without the pass), bench N=2
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cache.py tests/test_gpu_multi.py tests/test_gpu_dataloader.py -x -q > gpurun_out/r2_7_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_7_pytest.log
tail -6 gpurun_out/r2_7_pytest.log
R="iso_hit,iso_miss01,iso_miss03,iso_miss10,p2p_hit_one,p2p_hit_both,p2p_miss01_both,p2p_miss03_both,p2p_miss10_both,p2p_remote_only_one,p2p_local_only_one"
for pf in 1 0; do
COALA_GATHER_PREFETCH=$pf timeout 400 python scripts/bench_gather.py --gpus 2 --nodes 2000000 --cache_mb 8192 --regimes $R > gpurun_out/r2_7_gather_micro_pf$pf.log 2>&1; echo "micro prefetch=$pf rc=$?"
grep -E "^(iso_hit|iso_miss|p2p_hit|p2p_miss|p2p_remote_only_one|p2p_local_only_one)" gpurun_out/r2_7_gather_micro_pf$pf.log | cut -c1-100
done
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 700 $TR bench.py --gpus 2 --steps 20 --warmup 5 --extras all_resident > gpurun_out/r2_7_n2.json 2> gpurun_out/r2_7_n2.err; echo "n2 rc=$?"
tail -c 400 gpurun_out/r2_7_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_7_n2.json').read().strip().splitlines()[-1])
r=d['roofline']
print('value',d['value'],'e2e',d['e2e']['value'],'hit',d['hit_ratio'],'kernel_ms',r['kernel_ms'],'bound',r['bound'],r['frac'],'parity',d['parity']['mismatches'])
print({k:v['ms'] for k,v in r['all_hit_replay'].items() if k.startswith('engine')})
a=d['extras']['all_resident']; print(a['value'], a['gather_kernel_ms'], a.get('gather_alone_engine_0'), a.get('gather_alone_engine_1'))
PY
