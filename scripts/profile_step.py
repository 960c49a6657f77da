"""One batch-8 proof step (bench.py's workload) bracketed by cudaProfilerStart/Stop, for
`ncu --profile-from-start off`. Usage: python scripts/profile_step.py [resident|e2e|msm20|ntt LOGM]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench

mode = sys.argv[1] if len(sys.argv) > 1 else "resident"
cudart = torch.cuda.cudart()
if mode in ("resident", "e2e"):
    wl = bench.Workload(0, seed=1)
    step = wl.step_resident if mode == "resident" else wl.step_e2e
    step()
    step()
    torch.cuda.synchronize()
    cudart.cudaProfilerStart()
    step()
    torch.cuda.synchronize()
    cudart.cudaProfilerStop()
    wl.close()
else:
    import zecale_b200 as z
    ctx = z.Context(0)
    lg = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    n = 1 << lg
    sc = torch.from_numpy(bench.random_fr(np.random.default_rng(5), n).view(np.int64)).cuda()
    if mode == "msm20":
        pts = ctx.synth_bases(n, z.G1, seed=77)
        pin = ctx.pin_bases(pts)
        ctx.msm_pinned_dev(pin, sc.data_ptr(), n)
        torch.cuda.synchronize()
        cudart.cudaProfilerStart()
        ctx.msm_pinned_dev(pin, sc.data_ptr(), n)
        torch.cuda.synchronize()
        cudart.cudaProfilerStop()
        pin.release()
    else:
        ctx.ntt_dev(sc.data_ptr(), lg, 0)
        torch.cuda.synchronize()
        cudart.cudaProfilerStart()
        ctx.ntt_dev(sc.data_ptr(), lg, 0)
        torch.cuda.synchronize()
        cudart.cudaProfilerStop()
    ctx.close()
print("profile_step", mode, "ok")
