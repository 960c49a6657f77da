#!/bin/bash
# Parity tests (MSM cases run under both accumulation kernels), then the bench step for each mode.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_msm_gpu.py -m gpu -x -q > gpurun_out/pytest_msm.log 2>&1; rc=$?; echo "pytest rc=$rc" >> gpurun_out/pytest_msm.log
[ $rc -ne 0 ] && exit 1
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for mode in xyzz tree; do
  ZECALE_GPU_ACC=$mode timeout 300 python bench.py --skip-cpu-baseline > gpurun_out/bench_$mode.json 2> gpurun_out/bench_$mode.err; echo "rc=$?" >> gpurun_out/bench_$mode.err
done
timeout 120 python scripts/inv_probe.py > gpurun_out/inv_probe.log 2>&1
echo done
