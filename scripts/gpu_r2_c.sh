#!/bin/bash
# Evidence pass 2: sanitizer logs, ncu launch list and full captures (exported to CSV on the box; the .ncu-rep files stay there).
mkdir -p gpurun_out
for tool in memcheck racecheck initcheck; do
  timeout 900 compute-sanitizer --tool $tool --error-exitcode 7 python scripts/sanitize_step.py > gpurun_out/sanitizer_$tool.log 2>&1; echo "$tool rc=$?" | tee -a gpurun_out/sanitizer_$tool.log
  tail -4 gpurun_out/sanitizer_$tool.log
done
timeout 300 python scripts/profile_step.py resident > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches.csv python scripts/profile_step.py resident > gpurun_out/ncu_launches.log 2>&1
cap() { # name regex count
  timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:$2 -c $3 \
      -o /tmp/prof_$1 python scripts/profile_step.py resident > gpurun_out/ncu_$1.log 2>&1
  ncu -i /tmp/prof_$1.ncu-rep --page raw --csv > gpurun_out/prof_$1_raw.csv 2>/dev/null
  ncu -i /tmp/prof_$1.ncu-rep --page source --csv --print-source sass > gpurun_out/prof_$1_source.csv 2>/dev/null
  head -c 3000000 gpurun_out/prof_$1_source.csv > gpurun_out/prof_$1_source_head.csv; rm -f gpurun_out/prof_$1_source.csv
  ls -la /tmp/prof_$1.ncu-rep gpurun_out/prof_$1_raw.csv
}
cap accumulate k_accumulate 2
cap tail k_tail_level 3
cap ntt k_ntt_pass 2
echo done1
# --- BASELINE configs 2 and 3 as tables, and the CPU-baseline variants
timeout 1500 python scripts/sweep.py --cpu 2> gpurun_out/sweep.err | grep '^{' > gpurun_out/msm_ntt_sweep.jsonl; echo "sweep rc=${PIPESTATUS[0]}"; tail -3 gpurun_out/sweep.err
timeout 900 python scripts/cpu_variants.py 2> gpurun_out/cpu_variants.err | grep '^{' > gpurun_out/cpu_baseline_variants.jsonl; echo "variants rc=${PIPESTATUS[0]}"
cat gpurun_out/cpu_baseline_variants.jsonl | cut -c1-200
echo done2
