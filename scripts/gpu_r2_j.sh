#!/bin/bash
# In-tree library: MSM / golden / prove parity subset and step time.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_group_gpu.py tests/test_golden_gpu.py tests/test_msm_gpu.py tests/test_prove_gpu.py tests/test_ntt_gpu.py -m gpu -x -q -k "not 23 and not 24" 2>&1 | tail -2
timeout 600 python bench.py --steps 20 --warmup 5 --skip-cpu-baseline --skip-sharded > gpurun_out/bench_j.json 2> gpurun_out/bench_j.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_j.json').read().strip().splitlines()[-1]); r=d['roofline']
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'd1', d['arm']['ms_per_step_depth1'], 'frac', r['frac'], 'in_step', r['in_step_frac'], 'peak', r['peak'], 'fq/s', r.get('fq_products_per_s'))
print({k:(v['c'], round(v['accumulate_ms'],2)) for k,v in r['msm_shapes'].items()}); print(d['roofline_ntt']['transforms_ms'], d['roofline_ntt']['frac_imad'], d['clocks'], d['msm_g1_2e20'])"
echo done
