#!/bin/bash
# Evidence pass for the Karatsuba product: launch list and ncu captures (k_accumulate, k_tail_level, k_ntt_pass).
# The .ncu-rep files (20 MB each with sources) are reduced to their raw-page CSV on the box: gpurun_out/ is capped at 64 MiB.
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches.csv python scripts/profile_step.py resident > gpurun_out/ncu_launches.log 2>&1; echo "launches rc=$?"
cap() { # name regex count
  timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:$2 -c $3 \
      -o /tmp/prof_$1 python scripts/profile_step.py resident > gpurun_out/ncu_$1.log 2>&1; echo "$1 rc=$?"
  ncu -i /tmp/prof_$1.ncu-rep --page raw --csv > gpurun_out/prof_$1_raw.csv 2>/dev/null
}
cap accumulate k_accumulate 2
cap tail k_tail_level 3
cap ntt k_ntt_pass 2
ls -la gpurun_out | tail -8
echo done
