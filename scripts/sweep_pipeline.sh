#!/bin/bash
# Knob sweep of the whole pipeline (run on a GPU box): one bench.py line per setting, appended to gpurun_out/sweep_<tag>.log.
#   scripts/sweep_pipeline.sh <tag> <n_gpus> [extra bench.py args...]
# Knobs (all read at object creation): COALA_GATHER_ENGINE (0 registers | 1 bulk copies), COALA_GATHER_WAVES,
# COALA_GATHER_CTAS_PER_SM, COALA_PREFETCH_DEPTH, COALA_HBM_INDPTR.  Multi-rank runs are wrapped in `timeout`.
set -u
tag=$1; n=$2; shift 2
out=gpurun_out/sweep_${tag}.log
mkdir -p gpurun_out
: > "$out"
run() {
    echo "### $*" >> "$out"
    if [ "$n" -gt 1 ]; then
        env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port 29533 \
            bench.py --gpus "$n" --no_cpu_baseline "${EXTRA[@]}" 2>&1 | grep '^{' >> "$out"
    else
        env "$@" timeout 600 python bench.py --no_cpu_baseline "${EXTRA[@]}" 2>&1 | grep '^{' >> "$out"
    fi
}
EXTRA=("$@")
run COALA_GATHER_ENGINE=0
run COALA_GATHER_ENGINE=1
run COALA_GATHER_ENGINE=1 COALA_GATHER_BULK_WAVES=1
run COALA_HBM_INDPTR=1
run COALA_HBM_INDPTR=1 COALA_GATHER_ENGINE=1
run COALA_PREFETCH_DEPTH=2
run COALA_PREFETCH_DEPTH=4
run COALA_GATHER_CTAS_PER_SM=1
run COALA_SYNTH_COLOR=graph_gpu
python - "$out" <<'PY'
import json, sys
label = None
for line in open(sys.argv[1]):
    if line.startswith("###"):
        label = line[4:].strip()
    elif line.startswith("{"):
        d = json.loads(line)
        r = d["roofline"]
        print("%-55s value %9.1f  e2e %9.1f  gather %.3f ms  sample %.3f ms  %s %.2f" % (
            label, d["value"], d["e2e"]["value"], r["kernel_ms"], d.get("sample_stream_ms", 0), r["bound"], r["frac"]))
PY
