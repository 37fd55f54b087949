#!/bin/bash
# usage: gpu_multi.sh N [c3]   (run under gpurun --gpus N); "c3" adds the B=1024 strong-scaling workload
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/gpus_$N.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  scripts/dp_check.py > gpurun_out/dp_check_$N.log 2>&1; echo "dp_check $N exit $?" | tee -a gpurun_out/summary_multi.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 \
  bench.py --gpus $N --steps 10 --warmup 3 --no-tf32-info --out gpurun_out/bench_n$N.jsonl > gpurun_out/bench_n$N.log 2>&1; echo "bench $N exit $?" | tee -a gpurun_out/summary_multi.txt
if [ "$2" = "c3" ]; then
# configs[2]: B=1024 rows in total over the N ranks (strong scaling)
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 \
  bench.py --gpus $N --workload c3 --steps 10 --warmup 3 --no-tf32-info --no-graph-info --out gpurun_out/bench_c3_n$N.jsonl > gpurun_out/bench_c3_n$N.log 2>&1; echo "bench c3 $N exit $?" | tee -a gpurun_out/summary_multi.txt
fi
tail -n 3 gpurun_out/dp_check_$N.log
