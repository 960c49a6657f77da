#!/bin/bash
# Last confirmation pass on the final library: whole GPU suite, smoke(), both bench arms.
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q -s 2>&1 | grep -v "^\*\|^Swap\|^doing\|^Skipping\|^[0-9]*$" | tail -25 > gpurun_out/pytest_gpu_final.log; echo "pytest rc=${PIPESTATUS[0]}" | tee -a gpurun_out/pytest_gpu_final.log
tail -3 gpurun_out/pytest_gpu_final.log | cut -c1-200
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke_final.log
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_final.json').read().strip().splitlines()[-1]); r=d['roofline']
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'd1', d['arm']['ms_per_step_depth1'], 'frac', r['frac'], 'in_step', r['in_step_frac'], 'traffic', r['traffic'])
print({k:(v['c'], round(v['accumulate_ms'],2)) for k,v in r['msm_shapes'].items()}); print(d['roofline_ntt']['transforms_ms'], d['roofline_ntt']['frac_imad']); print(d['cpu_baseline']['value'], d['clocks']); print(d.get('sharded')); print(d['msm_g1_2e20'])"
timeout 1200 python bench.py --impl reference --gpus 1 --steps 4 --warmup 1 > gpurun_out/bench_reference_final.json 2> gpurun_out/bench_reference_final.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/bench_reference_final.json
echo done
