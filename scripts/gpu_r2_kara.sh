#!/bin/bash
# Karatsuba product variants against the interleaved CIOS product: product throughput, field/group/golden parity, step time.
mkdir -p gpurun_out
for v in base k1 k2; do
  if [ $v = base ]; then unset ZECALE_GPU_LIB; else export ZECALE_GPU_LIB=$PWD/zecale_b200/libzecale_gpu_$v.so; fi
  [ $v != base ] && [ ! -f "$ZECALE_GPU_LIB" ] && continue
  echo "== $v"
  timeout 300 python scripts/microbench.py --quick > gpurun_out/kara_microbench_$v.jsonl 2>&1; tail -3 gpurun_out/kara_microbench_$v.jsonl | cut -c1-200
  timeout 600 python -m pytest tests/test_field_gpu.py tests/test_group_gpu.py tests/test_golden_gpu.py -m gpu -x -q 2>&1 | tail -3
  timeout 600 python bench.py --steps 10 --warmup 3 --skip-cpu-baseline --skip-sharded --skip-msm20 > gpurun_out/kara_bench_$v.json 2> gpurun_out/kara_bench_$v.err; echo "bench rc=$?"
  python -c "
import json; d=json.loads(open('gpurun_out/kara_bench_$v.json').read().strip().splitlines()[-1]); r=d['roofline']
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'd1', d['arm']['ms_per_step_depth1'], 'frac', r['frac'], 'in_step', r['in_step_frac'])
print({k:(v['c'], round(v['accumulate_ms'],2)) for k,v in r['msm_shapes'].items()}); print(d['roofline_ntt']['transforms_ms'], d['clocks'])"
done
echo done
