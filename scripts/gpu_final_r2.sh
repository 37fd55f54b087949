#!/bin/bash
# end-of-round refresh: full GPU suite (with the parity numbers written out), smoke, bench line, ncu passes
mkdir -p gpurun_out
rm -f gpurun_out/parity.jsonl
MB_PARITY_OUT=gpurun_out/parity.jsonl python -m pytest tests -x -q -m gpu 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | cut -c1-100
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"
tail -c 400 gpurun_out/bench_n1.err
bash scripts/gpu_profile_r2.sh 2>&1 | tail -3
