"""Host time to ENQUEUE one proof (the six async C-ABI calls + flush) against the time until its results are back:
how much of a multi-device proof's critical path is the single host thread issuing launches for every device."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench

wl = bench.Workload(0, seed=1, depth=1)
for _ in range(3):
    wl.enqueue_e2e(0)
    wl.ctxs[0].sync()
enq, tot = [], []
for _ in range(10):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    wl.enqueue_e2e(0)
    t1 = time.perf_counter()
    wl.ctxs[0].sync()
    t2 = time.perf_counter()
    enq.append((t1 - t0) * 1e3)
    tot.append((t2 - t0) * 1e3)
print(json.dumps({"enqueue_ms_min": min(enq), "enqueue_ms_median": sorted(enq)[5], "proof_ms_min": min(tot),
                  "launches_per_proof": wl.ctxs[0].launch_count() / 13}))
wl.close()
