#!/bin/bash
# Evidence pass of a session: parity tests, smoke, bench (both arms), microbench, size sweep, ncu launch list + full captures.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
lscpu | head -20 > gpurun_out/lscpu.txt
timeout 400 python -m pytest tests/test_field_gpu.py tests/test_msm_gpu.py -m gpu -x -q > gpurun_out/pytest_msm.log 2>&1; rc=$?; echo "pytest rc=$rc" >> gpurun_out/pytest_msm.log
[ $rc -ne 0 ] && exit 1
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 300 python bench.py --depth 3 --skip-msm20 --skip-cpu-baseline > gpurun_out/bench_d3.json 2> gpurun_out/bench_d3.err
timeout 300 python scripts/microbench.py > gpurun_out/microbench.log 2>&1
timeout 300 python scripts/perf_probe.py 16 18 20 22 > gpurun_out/perf_probe.log 2>&1
timeout 300 python scripts/profile_step.py resident > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches.csv python scripts/profile_step.py resident > gpurun_out/ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_accumulate -c 2 \
    -o gpurun_out/prof_accumulate python scripts/profile_step.py resident > gpurun_out/ncu_acc.log 2>&1
timeout 300 python scripts/profile_step.py ntt 18 > gpurun_out/plain3.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_ntt_pass -c 2 \
    -o gpurun_out/prof_ntt python scripts/profile_step.py ntt 18 > gpurun_out/ncu_ntt.log 2>&1
timeout 400 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo done
