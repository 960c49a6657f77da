#!/bin/bash
# Multi-GPU evidence on this box (N = all visible GPUs): parity tests of the multi-device context and the distributed
# transforms, the hooked reference prover on a multi-device context, strong scaling of config 5 at N = 1, 2, 4, 8,
# the NCCL all-to-all transforms checked against libfqfft, and the bench line under torchrun.
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_dist_gpu.py -m gpu -x -q 2>&1 | tail -4 > gpurun_out/pytest_multi_n$N.log; cat gpurun_out/pytest_multi_n$N.log
DEVS=$(seq -s, 0 $((N-1)))
ZECALE_GPU_DEVICES=$DEVS timeout 900 python -m pytest tests/test_hooks_gpu.py -m gpu -x -q -s -k "batch8 or synthetic" 2>&1 | grep -v "^\*\|^Swap\|^doing\|^Skipping\|^[0-9]*$" | tail -8 > gpurun_out/pytest_hooks_multi_n$N.log; cat gpurun_out/pytest_hooks_multi_n$N.log | cut -c1-400
: > gpurun_out/sharded_scale_n$N.jsonl
for k in 1 2 4 8; do
  [ $k -gt $N ] && break
  timeout 600 python scripts/sharded_only.py $k >> gpurun_out/sharded_scale_n$N.jsonl 2> gpurun_out/sharded_$k.err
  tail -c 300 gpurun_out/sharded_scale_n$N.jsonl | grep -o '"n_gpus.*'
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/dist_check.py 22 24 > gpurun_out/dist_check_world$N.jsonl 2> gpurun_out/dist_check_world$N.err
cat gpurun_out/dist_check_world$N.jsonl | cut -c1-300; tail -3 gpurun_out/dist_check_world$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_n$N.json').read().strip().splitlines()[-1]); print('bench N=$N value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value']); print(d.get('sharded'))"
echo done
