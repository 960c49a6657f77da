#!/bin/bash
# Hybrid batched-affine accumulation: parity under both modes, then step time per (rounds, K).
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_msm_gpu.py tests/test_golden_gpu.py -m gpu -x -q > gpurun_out/pytest_tree.log 2>&1; rc=$?; echo "pytest rc=$rc" >> gpurun_out/pytest_tree.log
tail -5 gpurun_out/pytest_tree.log
[ $rc -ne 0 ] && exit 1
B="python bench.py --steps 10 --warmup 3 --skip-cpu-baseline --skip-msm20 --skip-sharded"
ZECALE_GPU_ACC=xyzz timeout 300 $B > gpurun_out/bench_t_xyzz.json 2> gpurun_out/bench_t_xyzz.err
for R in 1 2 3 4; do
  ZECALE_GPU_ACC=tree ZECALE_GPU_TREE_ROUNDS=$R timeout 300 $B > gpurun_out/bench_t_r$R.json 2> gpurun_out/bench_t_r$R.err
  ZECALE_GPU_ACC=tree ZECALE_GPU_TREE_ROUNDS=$R ZECALE_GPU_TREE_K=32 timeout 300 $B > gpurun_out/bench_t_r${R}_k32.json 2> gpurun_out/bench_t_r${R}_k32.err
done
for f in gpurun_out/bench_t_*.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1])
print(' value %.2f ms %.2f depth1 %.2f acc_ms %s' % (d['value'], d['ms_per_step'], d['arm']['ms_per_step_depth1'], [round(v['accumulate_ms'],2) for v in d['roofline']['msm_shapes'].values()]))
"; done
echo done
