"""Fq Montgomery-product throughput (IMAD pipe) on the current GPU: the measured
denominator of the MSM roofline. Prints JSON lines."""
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import zecale_b200 as z

ctx = z.Context(0)
res = []
quick = "--quick" in sys.argv
for blocks_per_sm in ((2, 3) if quick else (1, 2, 3, 4)):
    for tpb in ((128,) if quick else (32, 64, 128)):
        blocks = 148 * blocks_per_sm
        iters = 2000
        ms = min(ctx.bench_fq_mul(blocks, tpb, iters) for _ in range(3))
        muls = blocks * tpb * iters
        r = {"blocks_per_sm": blocks_per_sm, "tpb": tpb, "warps_per_sm": blocks_per_sm * tpb // 32, "ms": ms,
             "fq_mul_per_s": muls / ms * 1e3, "wide_mad_per_s": muls * 1152 / ms * 1e3,
             "us_per_mul_per_thread": ms * 1e3 / iters}
        print(json.dumps(r), flush=True)
ctx.close()
