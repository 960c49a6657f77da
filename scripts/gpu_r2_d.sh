#!/bin/bash
# Evidence pass 3: range-checked build under the GPU tests (stands in for compute-sanitizer, profiles/r02_sanitizer.md),
# the DFMA multiplier prototype against the integer one, un-pinned small-MSM latency.
mkdir -p gpurun_out
ZECALE_GPU_LIB=$PWD/zecale_b200/libzecale_gpu_checked.so timeout 1500 python -m pytest tests/test_msm_gpu.py tests/test_prove_gpu.py \
    tests/test_keyfile_gpu.py tests/test_batch_exp_gpu.py tests/test_golden_gpu.py -m gpu -x -q -k "not full_size" 2>&1 | tail -15 > gpurun_out/checked_build_tests.txt
echo "checked rc=${PIPESTATUS[0]}" | tee -a gpurun_out/checked_build_tests.txt
tail -5 gpurun_out/checked_build_tests.txt
# DFMA prototype
timeout 300 ./scripts/probes/dfma_mul /tmp/dfma_dump.bin 2000 > gpurun_out/dfma_mul.jsonl 2> gpurun_out/dfma_mul.err; echo "dfma rc=$?"
timeout 300 python scripts/probes/dfma_model.py check /tmp/dfma_dump.bin >> gpurun_out/dfma_mul.jsonl 2>> gpurun_out/dfma_mul.err; echo "dfma check rc=$?"
timeout 300 python - >> gpurun_out/dfma_mul.jsonl 2>> gpurun_out/dfma_mul.err <<'PY'
import json, sys
sys.path.insert(0, ".")
import zecale_b200 as z
ctx = z.Context(0)
for bps, tpb in ((2, 128), (3, 128), (4, 64)):
    blocks, iters = 148 * bps, 2000
    ms = min(ctx.bench_fq_mul(blocks, tpb, iters) for _ in range(2))
    print(json.dumps({"probe": "integer fq_mul chain (k_bench_fq_mul, the product path's multiplier)", "blocks": blocks, "threads": tpb, "iters": iters,
                      "ms": ms, "fq_products_per_s": blocks * tpb * iters / (ms * 1e-3), "equiv_32x32_products_per_s": 1176 * blocks * tpb * iters / (ms * 1e-3)}))
ctx.close()
PY
cat gpurun_out/dfma_mul.jsonl | cut -c1-250
# un-pinned MSM latency at small sizes (host -> host)
timeout 300 python - > gpurun_out/unpinned_small.jsonl 2> gpurun_out/unpinned_small.err <<'PY'
import json, sys, time
sys.path.insert(0, ".")
import numpy as np
import bench
import zecale_b200 as z
ctx = z.Context(0)
for lg in (8, 10, 12, 14, 16):
    n = 1 << lg
    pts = ctx.synth_bases(n, z.G1, seed=lg)
    sc = bench.random_fr(np.random.default_rng(lg), n)
    ctx.msm_g1(pts, sc)
    ts = []
    for _ in range(5):
        t0 = time.perf_counter(); ctx.msm_g1(pts, sc); ts.append((time.perf_counter() - t0) * 1e3)
    print(json.dumps({"op": "msm_g1 un-pinned host->host", "log_n": lg, "c": z.msm_window_bits(n), "ms_min": min(ts), "ms": ts}), flush=True)
ctx.close()
PY
cat gpurun_out/unpinned_small.jsonl | cut -c1-200
echo done
