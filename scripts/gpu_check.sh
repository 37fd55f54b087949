#!/bin/bash
# One gpurun call: GPU parity tests (one process per file so a faulting kernel cannot poison the
# rest), smoke, a short bench and the kernel sweeps.  Logs land in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
for f in tests/test_gpu_vtrace.py tests/test_gpu_head.py tests/test_gpu_learner.py; do
  timeout 900 python -m pytest $f -q -m gpu -p no:cacheprovider --timeout 300 -x > gpurun_out/$(basename $f .py).log 2>&1
  echo "$f exit $?" >> gpurun_out/summary.txt
done
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --steps 5 --warmup 3 --out gpurun_out/bench.jsonl > gpurun_out/bench.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --sweep head --out gpurun_out/sweep_head.jsonl > gpurun_out/sweep_head.log 2>&1; echo "sweep head exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --sweep vtrace --out gpurun_out/sweep_vtrace.jsonl > gpurun_out/sweep_vtrace.log 2>&1; echo "sweep vtrace exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
tail -5 gpurun_out/test_gpu_vtrace.log gpurun_out/test_gpu_head.log gpurun_out/test_gpu_learner.log gpurun_out/smoke.log
tail -c 1500 gpurun_out/bench.log
