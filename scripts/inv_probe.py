"""Cost of a batch of Fq inversions (binary Euclid, ALU pipe) relative to Montgomery products."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench, zecale_b200 as z
ctx = z.Context(0)
n = 1 << 19
rng = np.random.default_rng(1)
a = np.zeros((n, 12), dtype=np.uint64)
a[:, :11] = rng.integers(0, 1 << 63, size=(n, 11), dtype=np.uint64)
a[:, 11] = rng.integers(0, 0x0122e824fb83ce0, size=n, dtype=np.uint64)
for op, name in ((2, "mul"), (0, "add"), (4, "inv_euclid"), (7, "inv_fermat")):
    ctx.test_fq(op, a[:1024], a[:1024])
    ts = []
    for _ in range(3):
        t0 = time.perf_counter(); ctx.test_fq(op, a, a); ts.append(time.perf_counter() - t0)
    print(name, "n=%d" % n, "best %.2f ms" % (min(ts) * 1e3), flush=True)
ctx.close()
