#!/bin/bash
# Round 2, first GPU pass: fast parity tests first (fail early), then the whole suite, smoke, bench, launch list.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
lscpu | head -20 > gpurun_out/lscpu.txt
timeout 600 python -m pytest tests/test_field_gpu.py tests/test_group_gpu.py tests/test_msm_gpu.py -m gpu -x -q -k "not full_size" > gpurun_out/pytest_msm.log 2>&1; rc=$?; echo "pytest rc=$rc" >> gpurun_out/pytest_msm.log
tail -5 gpurun_out/pytest_msm.log
[ $rc -ne 0 ] && exit 1
timeout 300 python bench.py --steps 5 --warmup 3 --skip-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?" >> gpurun_out/bench_quick.err
tail -c 600 gpurun_out/bench_quick.err
timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
ZECALE_GPU_TRACE=1 timeout 200 python scripts/trace_step.py > gpurun_out/trace.log 2>&1
timeout 300 python scripts/profile_step.py resident > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches.csv python scripts/profile_step.py resident > gpurun_out/ncu_launches.log 2>&1
echo done
