#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
for f in tests/test_gpu_vtrace.py tests/test_gpu_head.py tests/test_gpu_learner.py; do
  timeout 900 python -m pytest $f -q -m gpu -p no:cacheprovider --timeout 300 > gpurun_out/$(basename $f .py).log 2>&1
  echo "$f exit $?" >> gpurun_out/summary.txt
done
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --sweep variants --out gpurun_out/sweep_variants.jsonl > gpurun_out/sweep_variants.log 2>&1; echo "variants exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --steps 10 --warmup 3 --out gpurun_out/bench.jsonl > gpurun_out/bench.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --sweep vtrace --out gpurun_out/sweep_vtrace.jsonl > gpurun_out/sweep_vtrace.log 2>&1; echo "sweep vtrace exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
for f in gpurun_out/test_gpu_vtrace.log gpurun_out/test_gpu_head.log gpurun_out/test_gpu_learner.log gpurun_out/smoke.log; do tail -n 3 $f; done
