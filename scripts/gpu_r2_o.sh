#!/bin/bash
# Config 5 at N = all GPUs with the phase trace of every context (last proof).
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
ZECALE_GPU_TRACE=1 timeout 600 python scripts/sharded_only.py $N > gpurun_out/sharded_trace_n$N.json 2> gpurun_out/sharded_trace_n$N.err; echo "rc=$?"
cut -c1-200 gpurun_out/sharded_trace_n$N.json | grep -o '"ms_per_proof[^,]*'
grep -n "zecale_gpu trace" gpurun_out/sharded_trace_n$N.err | tail -12
tail -60 gpurun_out/sharded_trace_n$N.err > gpurun_out/sharded_trace_tail_n$N.txt
echo done
