#!/bin/bash
# Strong scaling of config 5 on this box: sharded batch-32-shaped proof at N = 1, 2, 4, ... visible GPUs, then the
# distributed NTT / sharded MSM checks (NCCL all-to-all, libfqfft as the checker) at N = all.
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
: > gpurun_out/sharded_scale_n$N.jsonl
for k in 1 2 4 8; do
  [ $k -gt $N ] && break
  timeout 600 python scripts/sharded_only.py $k >> gpurun_out/sharded_scale_n$N.jsonl 2> gpurun_out/sharded_$k.err
  tail -1 gpurun_out/sharded_scale_n$N.jsonl | cut -c1-80; tail -c 300 gpurun_out/sharded_scale_n$N.jsonl | grep -o '"n_gpus.*'
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/dist_check.py 22 24 > gpurun_out/dist_check_world$N.jsonl 2> gpurun_out/dist_check_world$N.err
cat gpurun_out/dist_check_world$N.jsonl; tail -3 gpurun_out/dist_check_world$N.err
echo done
