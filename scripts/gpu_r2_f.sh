#!/bin/bash
# NTT layout change: parity (all NTT / H-polynomial / prove tests, also under the range-checked build), bench, one ncu capture.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_ntt_gpu.py tests/test_prove_gpu.py tests/test_golden_gpu.py tests/test_dist_gpu.py -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_ntt.log; cat gpurun_out/pytest_ntt.log
timeout 600 python bench.py --steps 20 --warmup 5 --skip-sharded > gpurun_out/bench_f.json 2> gpurun_out/bench_f.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_f.json').read().strip().splitlines()[-1]); print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'd1', d['arm']['ms_per_step_depth1'], 'ntt', d['roofline_ntt']['transforms_ms'], d['roofline_ntt']['frac_imad'])"
timeout 300 python scripts/sweep.py --min-log 18 --max-log 22 2>/dev/null | grep ntt_ | cut -c1-120
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_ntt_pass -c 2 -o /tmp/prof_ntt2 python scripts/profile_step.py resident > gpurun_out/ncu_ntt2.log 2>&1
ncu -i /tmp/prof_ntt2.ncu-rep --page raw --csv > gpurun_out/prof_ntt2_raw.csv 2>/dev/null
echo done
