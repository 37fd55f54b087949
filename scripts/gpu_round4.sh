#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
for f in tests/test_gpu_vtrace.py tests/test_gpu_head.py tests/test_gpu_learner.py; do
  timeout 900 python -m pytest $f -q -m gpu -p no:cacheprovider --timeout 300 > gpurun_out/$(basename $f .py).log 2>&1
  echo "$f exit $?" >> gpurun_out/summary.txt
done
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --sweep variants --out gpurun_out/sweep_variants.jsonl > gpurun_out/sweep_variants.log 2>&1; echo "variants exit $?" >> gpurun_out/summary.txt
timeout 300 python scripts/profile_kernels.py 3 > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'head_kernel|vtrace_tma' -s 4 -c 4 \
    -o gpurun_out/prof_kernels -f python scripts/profile_kernels.py 3 > gpurun_out/prof_ncu.log 2>&1
echo "ncu kernels exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
for f in gpurun_out/test_gpu_vtrace.log gpurun_out/test_gpu_head.log gpurun_out/test_gpu_learner.log gpurun_out/smoke.log; do tail -n 3 $f; done
