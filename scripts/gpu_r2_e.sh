#!/bin/bash
# DFMA prototype: exactness against python integers, timing, and one ncu capture of the chain kernel.
mkdir -p gpurun_out
timeout 300 ./scripts/probes/dfma_mul /tmp/dfma_dump.bin 2000 > gpurun_out/dfma_mul.jsonl 2> gpurun_out/dfma_mul.err; echo "dfma rc=$?"
timeout 300 python scripts/probes/dfma_model.py check /tmp/dfma_dump.bin >> gpurun_out/dfma_mul.jsonl 2>> gpurun_out/dfma_mul.err; echo "dfma check rc=$?"
tail -3 gpurun_out/dfma_mul.jsonl
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_chain --launch-skip 7 -c 1 -o /tmp/prof_dfma ./scripts/probes/dfma_mul /tmp/d2.bin 500 > gpurun_out/ncu_dfma.log 2>&1
ncu -i /tmp/prof_dfma.ncu-rep --page raw --csv > gpurun_out/prof_dfma_raw.csv 2>/dev/null
ls -la gpurun_out/prof_dfma_raw.csv
echo done
