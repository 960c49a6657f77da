#!/bin/bash
# Hooked reference prover after the host-side assembly change (blinding terms during device work, s*A / r*B on two threads),
# and the first-pair special cases of the accumulate chains.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_msm_gpu.py -m gpu -x -q -k "first_pair or edge or sizes" 2>&1 | tail -2
timeout 1500 python -m pytest tests/test_hooks_gpu.py -m gpu -x -q -s -k "not batch32" 2>&1 | grep -v "^\*\|^Swap\|^doing\|^Skipping\|^[0-9]*$" | tail -12 | cut -c1-420 > gpurun_out/pytest_hooks.log; cat gpurun_out/pytest_hooks.log
echo done
