#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 1200 python -m pytest tests -q -m gpu -p no:cacheprovider -x > gpurun_out/gpu_all.log 2>&1; echo "gpu tests exit $?" >> gpurun_out/summary.txt
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --steps 10 --warmup 3 --out gpurun_out/bench.jsonl > gpurun_out/bench.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -n 5 gpurun_out/gpu_all.log; tail -n 2 gpurun_out/smoke.log
