#!/bin/bash
# One gpurun call: ncu captures (each only after the same command exited 0 plain) + sweeps + bench.
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 300 python scripts/profile_kernels.py 3 > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'head_kernel|vtrace_tma' -s 4 -c 4 \
    -o gpurun_out/prof_kernels -f python scripts/profile_kernels.py 3 > gpurun_out/prof_ncu.log 2>&1
echo "ncu kernels exit $?" >> gpurun_out/summary.txt
timeout 300 python scripts/profile_encoder_kernels.py 2 > gpurun_out/prof_enc_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'conv3x3|linear_tf32x3|maxpool3s2|active_cells' -s 13 -c 13 \
    -o gpurun_out/prof_encoder -f python scripts/profile_encoder_kernels.py 2 > gpurun_out/prof_enc_ncu.log 2>&1
echo "ncu encoder kernels exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/step_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv \
    --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/step_ncu.log 2>&1
echo "ncu launches exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --sweep head --out gpurun_out/sweep_head.jsonl > gpurun_out/sweep_head.log 2>&1; echo "sweep head exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --sweep vtrace --out gpurun_out/sweep_vtrace.jsonl > gpurun_out/sweep_vtrace.log 2>&1; echo "sweep vtrace exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --steps 20 --warmup 5 --out gpurun_out/bench.jsonl > gpurun_out/bench.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 --out gpurun_out/bench_ref.jsonl > gpurun_out/bench_ref.log 2>&1; echo "bench ref exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
