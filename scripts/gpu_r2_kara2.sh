#!/bin/bash
# Karatsuba variants at 3 / 4 accumulate CTAs per SM: step time; ncu capture of k_accumulate under the K1 product.
mkdir -p gpurun_out
for v in k1 k1c4 k2c4; do
  export ZECALE_GPU_LIB=$PWD/zecale_b200/libzecale_gpu_$v.so
  [ ! -f "$ZECALE_GPU_LIB" ] && continue
  echo "== $v"
  timeout 600 python bench.py --steps 10 --warmup 3 --skip-cpu-baseline --skip-sharded --skip-msm20 > gpurun_out/kara_bench_$v.json 2> gpurun_out/kara_bench_$v.err; echo "bench rc=$?"
  python -c "
import json; d=json.loads(open('gpurun_out/kara_bench_$v.json').read().strip().splitlines()[-1]); r=d['roofline']
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'd1', d['arm']['ms_per_step_depth1'], 'frac', r['frac'], 'in_step', r['in_step_frac'])
print({k:(v['c'], round(v['accumulate_ms'],2)) for k,v in r['msm_shapes'].items()}); print(d['roofline_ntt']['transforms_ms'], d['clocks'])"
done
for v in k1 k1c4; do
export ZECALE_GPU_LIB=$PWD/zecale_b200/libzecale_gpu_$v.so
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_accumulate -c 1 \
    -o gpurun_out/prof_accumulate_$v python scripts/profile_step.py resident > gpurun_out/ncu_acc_$v.log 2>&1
echo "ncu rc=$?"
done
echo done
