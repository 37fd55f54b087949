#!/bin/bash
# graphed learner: its tests, then the bench at c2 and c1 (info.graphed next to the eager headline)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_learner.py -x -q -m gpu -k "graphed or sparse" > gpurun_out/gpu_graph_tests.log 2>&1; echo "tests exit $?"
tail -5 gpurun_out/gpu_graph_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_graph_c2.jsonl 2> gpurun_out/bench_graph_c2.log; echo "c2 exit $?"
timeout 600 python bench.py --workload c1 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_graph_c1.jsonl 2> gpurun_out/bench_graph_c1.log; echo "c1 exit $?"
tail -3 gpurun_out/bench_graph_c1.log
