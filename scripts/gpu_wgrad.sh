#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_convgrad.py -x -q -m gpu > gpurun_out/wgrad_tests.log 2>&1; echo "tests exit $?"
tail -8 gpurun_out/wgrad_tests.log
timeout 300 python scripts/wgrad_time.py > gpurun_out/wgrad_time.txt 2>&1; echo "time exit $?"; cat gpurun_out/wgrad_time.txt | tail -20
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_wgrad.jsonl 2> gpurun_out/bench_wgrad.log; echo "bench exit $?"
