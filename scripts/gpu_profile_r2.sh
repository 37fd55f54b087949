#!/bin/bash
# round 2: ncu captures (each only after the same command exited 0 plain) + launch list of the step
mkdir -p gpurun_out
timeout 300 python scripts/profile_r2.py 2 > gpurun_out/prof_r2_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:'pack_cells_kernel|actor_head_fwd_kernel|actor_head_list_fwd_kernel|cell_grad_kernel|head_kernel|vtrace_tma|conv3x3_tc_kernel|conv1_bits' -s 10 -c 10 \
    -o gpurun_out/prof_r2 -f python scripts/profile_r2.py 2 > gpurun_out/prof_r2_ncu.log 2>&1
echo "ncu r2 kernels exit $?"
timeout 600 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-graph-info --no-tf32-info --no-parity > gpurun_out/step_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv \
    --log-file gpurun_out/launches_r2.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-graph-info --no-tf32-info --no-parity > gpurun_out/step_ncu.log 2>&1
echo "ncu launches exit $?"
ls -la gpurun_out/prof_r2.ncu-rep gpurun_out/launches_r2.csv
