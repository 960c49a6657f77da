#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --skip-msm20 --skip-cpu-baseline > gpurun_out/bench_d2.json 2> gpurun_out/bench_d2.err; echo "rc=$?" >> gpurun_out/bench_d2.err
timeout 600 python bench.py --depth 1 --skip-msm20 --skip-cpu-baseline > gpurun_out/bench_d1.json 2> gpurun_out/bench_d1.err; echo "rc=$?" >> gpurun_out/bench_d1.err
timeout 600 python bench.py --depth 3 --skip-msm20 --skip-cpu-baseline > gpurun_out/bench_d3.json 2> gpurun_out/bench_d3.err; echo "rc=$?" >> gpurun_out/bench_d3.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
echo done
