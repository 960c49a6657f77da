#!/bin/bash
mkdir -p gpurun_out
for d in 1 2; do
timeout 600 python bench.py --depth $d --skip-msm20 --skip-cpu-baseline > gpurun_out/bench_d$d.json 2> gpurun_out/bench_d$d.err; echo "rc=$?" >> gpurun_out/bench_d$d.err
done
timeout 900 python -m pytest tests -m gpu -x -q -k "msm or prove or ntt or golden" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
