#!/bin/bash
# round-2 GPU session 12 (1 GPU): the driver's N=1 commands (ours + reference arm), C2 (igb-small, 4 GB), full GPU test suite,
# ncu launch list of a timed region and ncu --set full of one gather launch inside the bench
mkdir -p gpurun_out
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_12_n1.json 2> gpurun_out/r2_12_n1.err; echo "n1 rc=$?"
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_12_ref.json 2> gpurun_out/r2_12_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --shape igb-small --cache_size 4096 --steps 20 --warmup 5 --prewarm_seconds 5 > gpurun_out/r2_12_c2.json 2> gpurun_out/r2_12_c2.err; echo "c2 rc=$?"
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_12_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_12_pytest.log; tail -3 gpurun_out/r2_12_pytest.log
B="python bench.py --gpus 1 --steps 2 --warmup 3 --prewarm 600 --min_seconds 0.001 --extras none --no_cpu_baseline --cuda_profiler_range"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 400 --csv --log-file gpurun_out/r2_12_launches_n1.csv $B > gpurun_out/r2_12_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --import-source on --clock-control none --profile-from-start off --metrics pcie__read_bytes.sum,pcie__write_bytes.sum -k regex:coala_gather -s 2 -c 1 -f -o gpurun_out/r2_12_gather_n1 $B > gpurun_out/r2_12_ncu_full.log 2>&1; echo "ncu full rc=$?"
for f in n1 ref c2; do python - gpurun_out/r2_12_$f.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d.get('roofline',{})
    print(sys.argv[1],'value',d['value'],'e2e',d['e2e']['value'],'hit',d.get('hit_ratio'),'kernel_ms',r.get('kernel_ms'),r.get('bound'),r.get('frac'),'cpu',d.get('cpu_baseline',{}).get('value'),'parity',d.get('parity',{}).get('mismatches'))
    if 'all_hit_replay' in r: print({k:(v['ms'],round(v['hbm']['frac'],3)) for k,v in r['all_hit_replay'].items() if k.startswith('engine')})
    a=d.get('extras',{}).get('all_resident'); 
    if a: print('all_resident',a['value'],a['gather_kernel_ms'],a.get('binding_link_frac'),a.get('gather_alone_engine_0'),a.get('gather_alone_engine_1'))
except Exception as e: print('ERR',sys.argv[1],e)
PY
done
ls -la gpurun_out/r2_12*
