"""Multi-GPU check + timing under torchrun (one rank per GPU, NCCL): distributed 4-step NTT
(all-to-all over NVLink) vs the single-GPU transform, and point-sharded MSM vs the single-GPU MSM.
Prints one JSON line per experiment on rank 0.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 scripts/dist_check.py [log_m ...]
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import numpy as np
import torch
import torch.distributed as dist

import bench
import zecale_b200 as z
from zecale_b200 import dist as D

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = z.Context(local)
g = bench.fr_generator()
sizes = [int(a) for a in sys.argv[1:]] or [22, 24]


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


drv = D.DistNtt(D.gpu_phase_fn(ctx), D.nccl_exchange_fn() if world > 1 else None, world, rank)
from oracle import ref as R   # checker only (scripts/ is measurement + verification, not the product path)
REF = R.Ref() if R.available() else None
if REF is not None:
    REF.set_threads(max(1, (os.cpu_count() or 1) // max(world, 1)))
for log_m in sizes:
    m = 1 << log_m
    x = bench.random_fr(np.random.default_rng(log_m), m)            # same on every rank (seeded)
    ia, ib = D.layout_a_indices(log_m, world, rank), D.layout_b_indices(log_m, world, rank)
    res = {"op": "ntt_dist", "log_m": log_m, "world": world}
    for flags, name in ((0, "fft"), (3, "icoset")):
        src_idx, dst_idx = (ib, ia) if flags & 1 else (ia, ib)
        buf = torch.from_numpy(x[src_idx].view(np.int64).copy()).cuda()
        scratch = torch.empty_like(buf)
        keep = buf.clone()
        drv.transform(buf, scratch, log_m, flags, g if flags & 2 else None)        # warm-up (tables)
        ts = []
        out = None
        for _ in range(3):
            buf.copy_(keep)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out, _ = drv.transform(buf, scratch, log_m, flags, g if flags & 2 else None)   # phases + all-to-all chained on one stream
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ts.append(float(t.item()))
        res[name + "_ms"] = min(ts)
        # correctness: every rank checks its shard against libfqfft (oracle/_ref) when it is on the box,
        # else against the single-GPU transform (itself pinned to libfqfft by tests/test_ntt_gpu.py)
        if REF is not None:
            exp = REF.ntt(x, flags, g)[0][dst_idx]
            res[name + "_checked_against"] = "libfqfft basic_radix2_domain (oracle/_ref)"
        else:
            full = torch.from_numpy(x.view(np.int64).copy()).cuda()
            ctx.ntt_dev(full.data_ptr(), log_m, flags, g if flags & 2 else None)
            exp = full.cpu().numpy().view(np.uint64)[dst_idx]
            del full
            res[name + "_checked_against"] = "single-GPU transform"
        ok = torch.tensor([int(np.array_equal(out.cpu().numpy().view(np.uint64), exp))], device="cuda")
        if world > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        res[name + "_ok"] = bool(ok.item())
    if rank == 0:
        print(json.dumps(res), flush=True)

# point-sharded MSM, n = 2^20 and 2^22 G1
for log_n in (20, 22):
    n = 1 << log_n
    lo, hi = D.msm_chunk_range(n, world, rank)
    pts = ctx.synth_bases(n, z.G1, seed=77)[lo:hi]                    # every rank generates, keeps its shard
    pin = ctx.pin_bases(np.ascontiguousarray(pts))
    sc_all = bench.random_fr(np.random.default_rng(5), n)
    sc = torch.from_numpy(sc_all[lo:hi].view(np.int64).copy()).cuda()
    ts = []
    total = None
    for _ in range(3):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        total = D.msm_sharded(lambda: ctx.msm_pinned_dev(pin, sc.data_ptr(), hi - lo)[0], ctx.point_sum, world,
                              device=torch.device("cuda", local))
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ts.append(float(t.item()))
    pin.release()
    res = {"op": "msm_sharded_g1", "log_n": log_n, "world": world, "ms": min(ts)}
    if rank == 0:
        # the whole MSM on this GPU alone: the sharded sum must be the same 288 bytes
        allp = ctx.pin_bases(ctx.synth_bases(n, z.G1, seed=77))
        whole = ctx.msm_pinned(allp, sc_all)
        allp.release()
        res["equals_single_gpu_point"] = bool(np.array_equal(whole, total))
        print(json.dumps(res), flush=True)
ctx.close()
if world > 1:
    dist.destroy_process_group()
