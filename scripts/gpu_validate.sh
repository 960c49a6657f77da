#!/bin/bash
# Session check: parity tests (quick MSM subset first so that a hang cannot eat the budget), smoke, one bench line.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_field_gpu.py tests/test_msm_gpu.py -m gpu -x -q > gpurun_out/pytest_msm.log 2>&1; rc=$?; echo "pytest rc=$rc" >> gpurun_out/pytest_msm.log
[ $rc -ne 0 ] && exit 1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
echo done
