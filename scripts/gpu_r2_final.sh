#!/bin/bash
# Final single-GPU pass of round 2: whole GPU suite, range-checked build, smoke(), both bench arms, launch list, ncu captures
# (reduced to raw-page CSV on the box: gpurun_out/ is capped at 64 MiB), MSM / NTT sweep.
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q -s 2>&1 | grep -v "^\*\|^Swap\|^doing\|^Skipping\|^[0-9]*$" | tail -25 > gpurun_out/pytest_gpu_final.log; echo "pytest rc=${PIPESTATUS[0]}" | tee -a gpurun_out/pytest_gpu_final.log
tail -3 gpurun_out/pytest_gpu_final.log | cut -c1-200
ZECALE_GPU_LIB=$PWD/zecale_b200/libzecale_gpu_checked.so timeout 1500 python -m pytest tests/test_msm_gpu.py tests/test_prove_gpu.py \
    tests/test_keyfile_gpu.py tests/test_batch_exp_gpu.py tests/test_golden_gpu.py -m gpu -x -q -k "not full_size" 2>&1 | tail -15 > gpurun_out/checked_build_tests.txt
echo "checked rc=${PIPESTATUS[0]}" | tee -a gpurun_out/checked_build_tests.txt
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke_final.log
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_final.json').read().strip().splitlines()[-1]); r=d['roofline']
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'd1', d['arm']['ms_per_step_depth1'], 'frac', r['frac'], 'in_step', r['in_step_frac'])
print({k:(v['c'], round(v['accumulate_ms'],2)) for k,v in r['msm_shapes'].items()}); print(d['roofline_ntt']['transforms_ms'], d['roofline_ntt']['frac_imad']); print(d['cpu_baseline']['value'], d['clocks']); print(d.get('sharded')); print(d['msm_g1_2e20'])"
timeout 1200 python bench.py --impl reference --gpus 1 --steps 4 --warmup 1 > gpurun_out/bench_reference_final.json 2> gpurun_out/bench_reference_final.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/bench_reference_final.json
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
timeout 1500 python scripts/sweep.py > gpurun_out/msm_ntt_sweep.jsonl 2> gpurun_out/sweep.err; echo "sweep rc=$?"; wc -l gpurun_out/msm_ntt_sweep.jsonl
echo done
