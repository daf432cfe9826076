#!/bin/bash
# round-2 GPU session 8 (2 GPUs): fence-free (read-back ordered) fill path vs fenced — parity, storms in both modes, microbench, bench N=2
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cache.py tests/test_gpu_multi.py tests/test_gpu_dataloader.py -x -q > gpurun_out/r2_8_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_8_pytest.log
tail -6 gpurun_out/r2_8_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
for fz in 0 1; do
COALA_CACHE_FENCED=$fz timeout 300 $TR tests/multi_gpu_stress.py --launches 6000 --sets 1 > gpurun_out/r2_8_storm_fenced$fz.log 2>&1; echo "storm fenced=$fz rc=$?"; grep -E "^stress|multi_gpu_stress" gpurun_out/r2_8_storm_fenced$fz.log | cut -c1-400
done
R="iso_hit,iso_miss01,iso_miss03,iso_miss10,p2p_hit_both,p2p_miss01_both,p2p_miss03_both,p2p_miss10_both"
for fz in 0 1; do
COALA_CACHE_FENCED=$fz timeout 400 python scripts/bench_gather.py --gpus 2 --nodes 2000000 --cache_mb 8192 --regimes $R > gpurun_out/r2_8_gather_micro_fenced$fz.log 2>&1; echo "micro fenced=$fz rc=$?"
grep -E "^(iso_hit|iso_miss|p2p_hit|p2p_miss)" gpurun_out/r2_8_gather_micro_fenced$fz.log | cut -c1-100
done
timeout 700 $TR bench.py --gpus 2 --steps 20 --warmup 5 --extras all_resident,backends > gpurun_out/r2_8_n2.json 2> gpurun_out/r2_8_n2.err; echo "n2 rc=$?"
tail -c 400 gpurun_out/r2_8_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_8_n2.json').read().strip().splitlines()[-1])
r=d['roofline']
print('value',d['value'],'e2e',d['e2e']['value'],'hit',d['hit_ratio'],'kernel_ms',r['kernel_ms'],'bound',r['bound'],r['frac'],'parity',d['parity']['mismatches'])
print({k:v['ms'] for k,v in r['all_hit_replay'].items() if k.startswith('engine')})
a=d['extras']['all_resident']; print(a['value'], a['gather_kernel_ms'], a.get('gather_alone_engine_0'), a.get('gather_alone_engine_1'))
b=d['extras']['backends']; print({k:(round(v['value']),round(v['hit_ratio'],3),round(v.get('binding_link_frac',0),3)) for k,v in b.items() if isinstance(v,dict)})
PY
