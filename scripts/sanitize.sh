#!/bin/bash
# compute-sanitizer recipe (run on a GPU box: gpurun -- 'bash scripts/sanitize.sh').  memcheck over the cache, sampler and
# distributor parity tests; racecheck (shared-memory hazards) over the cache tests, which cover both copy engines
# (engine 1 stages rows in shared memory behind mbarriers).  Logs land in gpurun_out/sanitize_*.log.
# NOTE (round 2): compute-sanitizer is closed on this GPU pool (profiles/r2_compute_sanitizer_closed.log: the wrapper refuses to
# start it), so this recipe could not be run here; the device-side id bounds checks (coala_cache_set_num_nodes -> stats[3]),
# the eviction-storm stress (tests/multi_gpu_stress.py) and the bitwise row checks of every bench leg stand in for it.
mkdir -p gpurun_out
SAN=${SAN:-compute-sanitizer}
for t in cache sampler distributor; do
  timeout ${SANITIZE_TIMEOUT:-900} $SAN --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_$t.py -x -q > gpurun_out/sanitize_memcheck_$t.log 2>&1
  echo "memcheck $t rc=$?" | tee -a gpurun_out/sanitize_memcheck_$t.log
done
timeout ${SANITIZE_TIMEOUT:-900} $SAN --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_cache.py -x -q > gpurun_out/sanitize_racecheck_cache.log 2>&1
echo "racecheck cache rc=$?" | tee -a gpurun_out/sanitize_racecheck_cache.log
grep -h -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" gpurun_out/sanitize_*.log
