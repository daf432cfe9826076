#!/bin/bash
# round-3 GPU session 2 (2 GPUs): price of the fenced (PTX-model) writer protocol after its fill path went from 3 to 2 fences:
# gather microbench, 2-GPU storm, GPU suite, bench N=1 and N=2 - each in both modes (COALA_CACHE_FENCED=0|1)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
COALA_CACHE_FENCED=1 timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/r3_2_pytest_fenced.log 2>&1; echo "suite fenced rc=$?"; tail -2 gpurun_out/r3_2_pytest_fenced.log
for f in 0 1; do
  COALA_CACHE_FENCED=$f timeout 400 python scripts/bench_gather.py --gpus 2 --nodes 2000000 --cache_mb 8192 > gpurun_out/r3_2_micro_f$f.log 2>&1; echo "micro f$f rc=$?"
  grep -E "^(iso_cold|iso_hit|iso_miss03|iso_miss10|part1_miss03|p2p_hit_both|p2p_miss03_both|p2p_miss10_both) " gpurun_out/r3_2_micro_f$f.log | cut -c1-220
  COALA_CACHE_FENCED=$f timeout 300 $TR --master-port 2951$f tests/multi_gpu_stress.py --launches 5000 > gpurun_out/r3_2_storm_f$f.log 2>&1; echo "storm f$f rc=$?"; grep -E "^stress" gpurun_out/r3_2_storm_f$f.log | cut -c1-300
done
for f in 0 1; do
  COALA_CACHE_FENCED=$f timeout 500 python bench.py --gpus 1 --steps 20 --warmup 5 --extras none --no_cpu_baseline > gpurun_out/r3_2_n1_f$f.json 2> gpurun_out/r3_2_n1_f$f.err; echo "n1 f$f rc=$?"
  COALA_CACHE_FENCED=$f timeout 500 $TR --master-port 2953$f bench.py --gpus 2 --steps 20 --warmup 5 --extras none --no_cpu_baseline > gpurun_out/r3_2_n2_f$f.json 2> gpurun_out/r3_2_n2_f$f.err; echo "n2 f$f rc=$?"
done
python - <<'PY'
import json
for n in (1,2):
    for f in (0,1):
        try:
            d=json.loads(open('gpurun_out/r3_2_n%d_f%d.json'%(n,f)).read().strip().splitlines()[-1]); r=d['roofline']
            print('N',n,'fenced',f,'value',d['value'],'e2e',d['e2e']['value'],'ms',d['ms_per_step'],'hit',d.get('hit_ratio'),'kernel_ms',r.get('kernel_ms'),r['bound'],r['frac'],'parity',d['parity'].get('mismatches'))
        except Exception as e: print('ERR',n,f,e)
PY
