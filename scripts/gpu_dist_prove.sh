#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/dist_prove.jsonl
for n in 8 4 2; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n scripts/dist_prove.py >> gpurun_out/dist_prove.jsonl 2> gpurun_out/dist_prove_n$n.err; echo "n=$n rc=$?" >> gpurun_out/dist_prove.err
done
