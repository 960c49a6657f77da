#!/bin/bash
# k_accumulate with the MULW/MULFIN stash in shared memory (variant library): parity subset, step time, ncu capture.
mkdir -p gpurun_out
export ZECALE_GPU_LIB=$PWD/zecale_b200/libzecale_gpu_ss.so
timeout 900 python -m pytest tests/test_golden_gpu.py tests/test_msm_gpu.py -m gpu -x -q -k "not full_size" 2>&1 | tail -2
timeout 600 python bench.py --steps 20 --warmup 5 --skip-cpu-baseline --skip-sharded --skip-msm20 > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_q.json').read().strip().splitlines()[-1]); r=d['roofline']
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'd1', d['arm']['ms_per_step_depth1'], 'frac', r['frac'], 'in_step', r['in_step_frac'])
print({k:(v['c'], round(v['accumulate_ms'],2)) for k,v in r['msm_shapes'].items()})"
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_accumulate -c 2 \
      -o /tmp/prof_acc python scripts/profile_step.py resident > gpurun_out/ncu_acc_q.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/prof_acc.ncu-rep --page raw --csv > gpurun_out/prof_accumulate_q_raw.csv 2>/dev/null
echo done
