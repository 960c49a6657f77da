#!/bin/bash
# Tensor-core Montgomery reduction prototype: exactness against python integers, timing, one ncu capture.
mkdir -p gpurun_out
timeout 300 ./scripts/probes/imma_mul /tmp/imma_dump.bin 1000 > gpurun_out/imma_mul.jsonl 2> gpurun_out/imma_mul.err; echo "imma rc=$?"
timeout 300 python scripts/probes/imma_model.py check /tmp/imma_dump.bin >> gpurun_out/imma_mul.jsonl 2>> gpurun_out/imma_mul.err; echo "imma check rc=$?"
cat gpurun_out/imma_mul.jsonl | cut -c1-260; tail -3 gpurun_out/imma_mul.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_chain --launch-skip 5 -c 1 -o /tmp/prof_imma ./scripts/probes/imma_mul /tmp/d2.bin 300 > gpurun_out/ncu_imma.log 2>&1
ncu -i /tmp/prof_imma.ncu-rep --page raw --csv > gpurun_out/prof_imma_raw.csv 2>/dev/null
ls -la gpurun_out/prof_imma_raw.csv
echo done
