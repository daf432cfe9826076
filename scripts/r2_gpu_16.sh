#!/bin/bash
# round-2 GPU session 16 (1 GPU): indptr in HBM by default, indices in HBM as an option: tests, smoke, igb-small (all resident) and igb-medium
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_16_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_16_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_16_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2_16_smoke.log
show() { python - $1 $2 <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d['roofline']
    print('%-22s value %7.0f e2e %7.0f ms/step %.4f gather %.4f frac %.3f sample %.4f parity %s' % (sys.argv[2], d['value'], d['e2e']['value'], d['ms_per_step'], r['kernel_ms'], r['frac'], d['sample_stream_ms'], d['parity']['mismatches']))
except Exception as e: print(sys.argv[2],'ERR',e)
PY
}
S="python bench.py --shape igb-small --cache_size 4096 --steps 20 --warmup 5 --prewarm_seconds 3 --no_cpu_baseline --extras none"
timeout 200 $S > gpurun_out/r2_16_small_default.json 2> gpurun_out/r2_16_small_default.err; show gpurun_out/r2_16_small_default.json small_default
timeout 200 $S --hbm_indptr 0 > gpurun_out/r2_16_small_uva.json 2> gpurun_out/r2_16_small_uva.err; show gpurun_out/r2_16_small_uva.json small_uva
timeout 200 $S --hbm_graph > gpurun_out/r2_16_small_graph.json 2> gpurun_out/r2_16_small_graph.err; show gpurun_out/r2_16_small_graph.json small_graph
COALA_ENGINE_PRIORITY=gather timeout 200 $S > gpurun_out/r2_16_small_prio_gather.json 2> gpurun_out/r2_16_small_prio_gather.err; show gpurun_out/r2_16_small_prio_gather.json small_prio_gather
COALA_ENGINE_PRIORITY=none timeout 200 $S > gpurun_out/r2_16_small_prio_none.json 2> gpurun_out/r2_16_small_prio_none.err; show gpurun_out/r2_16_small_prio_none.json small_prio_none
M="python bench.py --gpus 1 --steps 20 --warmup 5 --no_cpu_baseline --extras none"
timeout 600 $M > gpurun_out/r2_16_n1_default.json 2> gpurun_out/r2_16_n1_default.err; show gpurun_out/r2_16_n1_default.json n1_default
timeout 600 $M --hbm_graph > gpurun_out/r2_16_n1_graph.json 2> gpurun_out/r2_16_n1_graph.err; show gpurun_out/r2_16_n1_graph.json n1_graph
