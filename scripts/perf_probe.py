"""Quick device-timed MSM / NTT sweep (inputs resident in HBM). Prints JSON lines."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import zecale_b200 as z
from oracle import ref as R

ref = R.Ref()
ctx = z.Context(0)
sizes = [int(a) for a in sys.argv[1:]] or [16, 18, 20]
for lg in sizes:
    n = 1 << lg
    nb = min(n, 1 << 16)
    base = ref.bases_seq(nb, 1)
    bases = np.tile(base, (n // nb, 1))
    scalars = ref.fr_rng(7, nb)
    scalars = np.tile(scalars, (n // nb, 1))
    scalars[:, 0] ^= np.arange(n, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)
    scalars[:, 5] &= np.uint64(0x00FFFFFFFFFFFFFF)
    pinned = ctx.pin_bases(bases)
    d_sc = torch.from_numpy(scalars.view(np.int64)).cuda()
    torch.cuda.synchronize()
    ts = []
    for it in range(3):
        _, ms = ctx.msm_pinned_dev(pinned, d_sc.data_ptr(), n)
        ts.append(ms)
    print(json.dumps({"op": "msm_g1", "log_n": lg, "ms": ts, "c": z.msm_window_bits(n)}), flush=True)
    pinned.release()
    a = torch.from_numpy(scalars.view(np.int64)).cuda()
    ts = []
    for it in range(4):
        ts.append(ctx.ntt_dev(a.data_ptr(), lg, 0))
    print(json.dumps({"op": "ntt", "log_m": lg, "ms": ts}), flush=True)
ctx.close()
