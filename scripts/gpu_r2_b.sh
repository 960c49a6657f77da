#!/bin/bash
# Single-GPU evidence pass: parity of the batched transforms, bench, sanitizer logs, ncu captures.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ntt_gpu.py tests/test_batch_exp_gpu.py tests/test_prove_gpu.py tests/test_golden_gpu.py tests/test_hooks_gpu.py -m gpu -x -q -k "not batch32 and not large" > gpurun_out/pytest_b.log 2>&1; rc=$?; echo "pytest rc=$rc" >> gpurun_out/pytest_b.log
tail -6 gpurun_out/pytest_b.log
[ $rc -ne 0 ] && exit 1
timeout 600 python bench.py --steps 10 --warmup 3 --skip-cpu-baseline > gpurun_out/bench_b.json 2> gpurun_out/bench_b.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/bench_b.json').read().strip().splitlines()[-1])
print('value %.2f ms %.2f e2e %.2f depth1 %.2f ntt_ms %.3f frac_imad %.3f' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['arm']['ms_per_step_depth1'], d['roofline_ntt']['transforms_ms'], d['roofline_ntt']['frac_imad']))
print(d.get('sharded')); print(d.get('msm_g1_2e20'))
"
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python scripts/sanitize_step.py > gpurun_out/sanitizer_memcheck.log 2>&1; echo "memcheck rc=$?" | tee -a gpurun_out/sanitizer_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 7 python scripts/sanitize_step.py > gpurun_out/sanitizer_racecheck.log 2>&1; echo "racecheck rc=$?" | tee -a gpurun_out/sanitizer_racecheck.log
timeout 900 compute-sanitizer --tool initcheck --error-exitcode 7 python scripts/sanitize_step.py > gpurun_out/sanitizer_initcheck.log 2>&1; echo "initcheck rc=$?" | tee -a gpurun_out/sanitizer_initcheck.log
timeout 300 python scripts/profile_step.py resident > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches.csv python scripts/profile_step.py resident > gpurun_out/ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_accumulate -c 2 \
    -o gpurun_out/prof_accumulate python scripts/profile_step.py resident > gpurun_out/ncu_acc.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_tail_level -c 3 \
    -o gpurun_out/prof_tail python scripts/profile_step.py resident > gpurun_out/ncu_tail.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_ntt_pass -c 2 \
    -o gpurun_out/prof_ntt python scripts/profile_step.py resident > gpurun_out/ncu_ntt.log 2>&1
echo done
