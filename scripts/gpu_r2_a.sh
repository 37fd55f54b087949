#!/bin/bash
# round 2, call A: new V-trace kernel correctness + variant sweep
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_vtrace.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/a_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/a_tests.log
tail -5 gpurun_out/a_tests.log
timeout 600 python scripts/sweep_vtrace_variants.py gpurun_out/a_vtrace_variants.jsonl > gpurun_out/a_sweep.log 2>&1
echo "sweep rc=$?"
tail -3 gpurun_out/a_sweep.log
