#!/bin/bash
# round-3 GPU session 1 (1 GPU): the static-scratch split alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r3_1_zz_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/r3_1_zz_alone.log
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/r3_1_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r3_1_pytest_gpu.log
