#!/bin/bash
# encoder-side kernels only: ncu --set full (after the same command exited 0 plain), kernel-vs-library timings, warm step profile
mkdir -p gpurun_out
timeout 300 python scripts/profile_encoder_kernels.py 2 > gpurun_out/prof_enc_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'conv3x3|linear_tf32x3|maxpool3s2|active_cells' -s 13 -c 13 \
    -o gpurun_out/prof_encoder -f python scripts/profile_encoder_kernels.py 2 > gpurun_out/prof_enc_ncu.log 2>&1
echo "ncu encoder kernels exit $?"
timeout 300 python scripts/wgrad_time.py > gpurun_out/wgrad_time.txt 2>&1; echo "conv timing exit $?"
timeout 300 python scripts/gemm_time.py > gpurun_out/gemm_time.txt 2>&1; echo "gemm timing exit $?"
timeout 300 python scripts/step_profile.py > gpurun_out/step_dense.txt 2>&1; echo "step profile exit $?"
