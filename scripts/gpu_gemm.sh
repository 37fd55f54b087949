#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu > gpurun_out/gemm_tests.log 2>&1; echo "tests exit $?"
tail -25 gpurun_out/gemm_tests.log
timeout 120 python scripts/gemm_time.py > gpurun_out/gemm_time.txt 2>&1; echo "time exit $?"; tail -8 gpurun_out/gemm_time.txt
