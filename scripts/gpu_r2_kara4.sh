#!/bin/bash
# Variant libraries (zecale_b200/libzecale_gpu_<v>.so): parity subset and step time.
mkdir -p gpurun_out
for v in ${VARIANTS}; do
  export ZECALE_GPU_LIB=$PWD/zecale_b200/libzecale_gpu_$v.so
  [ ! -f "$ZECALE_GPU_LIB" ] && continue
  echo "== $v"
  timeout 900 python -m pytest tests/test_field_gpu.py tests/test_group_gpu.py tests/test_golden_gpu.py tests/test_ntt_gpu.py tests/test_msm_gpu.py -m gpu -x -q -k "not 2e20 and not 2e18 and not slow and not 22 and not 23 and not 24" 2>&1 | tail -2
  timeout 600 python bench.py --steps 10 --warmup 3 --skip-cpu-baseline --skip-sharded --skip-msm20 > gpurun_out/kara_bench_$v.json 2> gpurun_out/kara_bench_$v.err; echo "bench rc=$?"
  python -c "
import json; d=json.loads(open('gpurun_out/kara_bench_$v.json').read().strip().splitlines()[-1]); r=d['roofline']
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'd1', d['arm']['ms_per_step_depth1'], 'frac', r['frac'], 'in_step', r['in_step_frac'], 'peak', r['peak'], 'fq/s', r.get('fq_products_per_s'))
print({k:(v['c'], round(v['accumulate_ms'],2)) for k,v in r['msm_shapes'].items()}); print(d['roofline_ntt']['transforms_ms'], d['roofline_ntt']['frac_imad'], d['clocks'])"
done
echo done
