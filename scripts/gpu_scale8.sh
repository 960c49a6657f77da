#!/bin/bash
# 8-GPU box: bench at N = 8, 4, 2 (replica scaling) and the sharded MSM / distributed NTT checks at world 8 and 4.
mkdir -p gpurun_out
for n in 8 4 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --skip-cpu-baseline --skip-msm20 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "rc=$?" >> gpurun_out/bench_n$n.err
done
for n in 8 4; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n scripts/dist_check.py 22 24 > gpurun_out/dist_check_world$n.jsonl 2> gpurun_out/dist_check_world$n.err; echo "rc=$?" >> gpurun_out/dist_check_world$n.err
done
echo done
