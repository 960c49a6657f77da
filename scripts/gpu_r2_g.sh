#!/bin/bash
# Tensor-core Montgomery reduction prototype: exactness against python integers, timing per variant (rows of the a*b
# product in flight x CTAs per SM), one ncu capture of the best.
mkdir -p gpurun_out
: > gpurun_out/imma_mul.jsonl
for d in 2; do
  timeout 300 ./scripts/probes/imma_mul_d$d /tmp/imma_dump_$d.bin 1000 >> gpurun_out/imma_mul.jsonl 2>> gpurun_out/imma_mul.err; echo "imma d$d rc=$?"
  timeout 300 python scripts/probes/imma_model.py check /tmp/imma_dump_$d.bin >> gpurun_out/imma_mul.jsonl 2>> gpurun_out/imma_mul.err; echo "check rc=$?"
done
cut -c1-230 gpurun_out/imma_mul.jsonl; tail -3 gpurun_out/imma_mul.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_chain --launch-skip 17 -c 1 -o /tmp/prof_imma ./scripts/probes/imma_mul_d2 /tmp/d2.bin 300 > gpurun_out/ncu_imma.log 2>&1
ncu -i /tmp/prof_imma.ncu-rep --page raw --csv > gpurun_out/prof_imma_raw.csv 2>/dev/null
ls -la gpurun_out/prof_imma_raw.csv
echo done
