#!/bin/bash
# Multi-GPU pass: multi-device context parity, then bench at N = visible GPUs (replicas + the sharded config-5 block).
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_dist_gpu.py -m gpu -x -q > gpurun_out/pytest_multi_n$N.log 2>&1; rc=$?; echo "pytest rc=$rc" >> gpurun_out/pytest_multi_n$N.log
tail -15 gpurun_out/pytest_multi_n$N.log
[ $rc -ne 0 ] && exit 1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?" >> gpurun_out/bench_n$N.err
tail -c 1500 gpurun_out/bench_n$N.json
tail -5 gpurun_out/bench_n$N.err
echo done
