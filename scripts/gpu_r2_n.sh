#!/bin/bash
# bench.py under torchrun at N = visible GPUs: replicas, the NCCL exchange paths (point-sharded MSM, 4-step NTT), the sharded proof.
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_n$N.json').read().strip().splitlines()[-1]); print('N=$N value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value']); print(d.get('nccl_paths')); print(d.get('sharded'))"
tail -3 gpurun_out/bench_n$N.err | cut -c1-300
echo done
