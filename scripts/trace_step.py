"""Phase timeline of pipelined proof steps (ZECALE_GPU_TRACE=1 makes the library print marks)."""
import os
import sys
os.environ["ZECALE_GPU_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import time
import torch
import bench
depth = int(sys.argv[1]) if len(sys.argv) > 1 else 1
mode = sys.argv[2] if len(sys.argv) > 2 else "resident"
wl = bench.Workload(0, seed=1, depth=depth)
wl.run(3, mode)
torch.cuda.synchronize()
sys.stderr.write("==== timed\n")
t0 = time.perf_counter()
wl.run(4, mode)
torch.cuda.synchronize()
sys.stderr.write("==== 4 steps in %.2f ms\n" % ((time.perf_counter() - t0) * 1e3))
wl.close()
