"""Multi-GPU layer of the prove hot path: one process per GPU, torch.distributed for the plumbing
(NCCL over NVLink on the box, gloo in the CPU tests). SURVEY.md §8(e).

* Whole proofs are replicas (bench.py --gpus N): no code here.
* One MSM is sharded by points exactly like libff::multi_exp shards it over OpenMP threads
  (multiexp.tcc:663-685): rank g owns bases [g*one, (g+1)*one), the last rank takes the
  remainder; every rank runs a complete Pippenger on its shard, the `world` partial points
  (288 B each) are all-gathered and summed in rank order (multiexp.tcc:680-685).
* A transform of 2^22 elements and more is a 4-step NTT whose transpose is ONE all-to-all
  (zecale_gpu_ntt_dist_phase in include/zecale_gpu.h defines the two layouts). The
  iFFT -> cosetFFT -> pointwise -> icosetFFT chain of r1cs_to_qap_witness_map
  (r1cs_to_qap.tcc:252-330) alternates layout B -> A -> B -> A, so it needs no other exchange.

The arithmetic is behind two callables so that the same driver runs on the GPU (the C ABI) and,
in tests/, on a CPU stand-in that restates the index maps independently:
    phase_fn(src, dst, log_m, flags, coset_g, world, rank, phase)   -> fills dst (torch tensors (m/world, 6) int64)
    exchange_fn(dst, src)                                           -> all-to-all of `world` equal chunks
"""
import numpy as np


# ------------------------------------------------------------------ MSM sharding
def msm_chunk_range(n, world, rank):
    """[lo, hi) of rank's bases: the chunk arithmetic of libff::multi_exp (multiexp.tcc:663-676)."""
    if n < world or world == 1:
        return (0, n) if rank == 0 else (n, n)   # too few points: rank 0 does them all
    one = n // world
    lo = rank * one
    hi = n if rank == world - 1 else (rank + 1) * one
    return lo, hi


def all_gather_points(partial, world, group=None, device=None):
    """All-gather one 288-byte point per rank; returns a (world, 36) uint64 numpy array in rank order."""
    import torch
    import torch.distributed as dist
    t = torch.from_numpy(np.ascontiguousarray(partial).view(np.int64).copy())
    if device is not None:
        t = t.to(device)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t, group=group)
    return np.stack([o.cpu().numpy().view(np.uint64) for o in out])


def msm_sharded(local_msm_fn, point_sum_fn, world, group=None, device=None):
    """sum over ranks of local_msm_fn() (a (36,) uint64 special-form point), every rank gets the total."""
    part = local_msm_fn()
    if world == 1:
        return part
    return point_sum_fn(all_gather_points(part, world, group, device))


# ------------------------------------------------------------------ 4-step layouts
def four_step_shape(log_m, world):
    lg = world.bit_length() - 1
    if world < 1 or (1 << lg) != world:
        raise ValueError("world size must be a power of two")
    l1 = (log_m + 1) // 2
    l2 = log_m - l1
    if lg >= l2:
        raise ValueError("2^%d elements cannot be split 4-step-wise over %d ranks" % (log_m, world))
    m1, m2 = 1 << l1, 1 << l2
    return {"l1": l1, "l2": l2, "m1": m1, "m2": m2, "c1": m1 // world, "c2": m2 // world}


def layout_a_indices(log_m, world, rank):
    """Global index of each locally stored element, layout A: [j1][j2 - rank*c2] -> j1*m2 + j2."""
    s = four_step_shape(log_m, world)
    j1 = np.arange(s["m1"], dtype=np.int64)[:, None]
    j2 = rank * s["c2"] + np.arange(s["c2"], dtype=np.int64)[None, :]
    return (j1 * s["m2"] + j2).reshape(-1)


def layout_b_indices(log_m, world, rank):
    """Layout B: [k1 - rank*c1][k2] -> k1 + m1*k2."""
    s = four_step_shape(log_m, world)
    k1 = rank * s["c1"] + np.arange(s["c1"], dtype=np.int64)[:, None]
    k2 = np.arange(s["m2"], dtype=np.int64)[None, :]
    return (k1 + s["m1"] * k2).reshape(-1)


class DistNtt:
    """Driver of the distributed transforms. buf / scratch: torch tensors of shape (m/world, 6),
    int64 (the 48-byte Fr rows), on the device the phase function works on. The two buffers ping-pong:
    every method returns (result, other) - which of the two tensors holds the result depends on the
    world size - and nothing is copied back. With the GPU backends below every phase, the all-to-all
    and the pointwise step are enqueued on torch's current stream, so a whole H-polynomial pipeline
    runs without a host synchronisation."""

    def __init__(self, phase_fn, exchange_fn, world, rank):
        self.phase_fn, self.exchange_fn, self.world, self.rank = phase_fn, exchange_fn, world, rank

    def transform(self, buf, scratch, log_m, flags, coset_g=None):
        """flags bit0 inverse, bit1 coset. Forward: layout A in, B out; inverse: B in, A out.
        Input in `buf`; returns (tensor holding the result, the other tensor)."""
        self.phase_fn(buf, scratch, log_m, flags, coset_g, self.world, self.rank, 0)
        if self.world > 1:
            self.exchange_fn(buf, scratch)        # buf <- all-to-all(scratch)
            self.phase_fn(buf, scratch, log_m, flags, coset_g, self.world, self.rank, 1)
            return scratch, buf
        self.phase_fn(scratch, buf, log_m, flags, coset_g, self.world, self.rank, 1)
        return buf, scratch

    def qap_h(self, aA, aB, aC, scratch, log_m, coset_g, pointwise_fn):
        """H polynomial from the three evaluation vectors held in layout B; returns the tensor that holds
        the coefficients (layout A). pointwise_fn(aA, aB, aC) computes (aA*aB - aC) / Z on the coset, in aA.
        All four tensors are clobbered."""
        vec = []
        for v in (aA, aB, aC):
            v, scratch = self.transform(v, scratch, log_m, 1)             # iFFT       B -> A
            v, scratch = self.transform(v, scratch, log_m, 2, coset_g)    # cosetFFT   A -> B
            vec.append(v)
        pointwise_fn(vec[0], vec[1], vec[2])
        h, _ = self.transform(vec[0], scratch, log_m, 3, coset_g)         # icosetFFT  B -> A
        return h


# ------------------------------------------------------------------ GPU backends (C ABI + NCCL)
def gpu_phase_fn(ctx):
    """Phases on torch's current stream (zecale_gpu_ntt_dist_phase_on): ordered with the NCCL collective
    that torch.distributed enqueues behind the same stream, no host synchronisation."""
    import torch

    def phase(src, dst, log_m, flags, coset_g, world, rank, ph):
        ctx.ntt_dist_phase_on(torch.cuda.current_stream().cuda_stream, src.data_ptr(), dst.data_ptr(), log_m, flags, coset_g,
                              world, rank, ph)
    return phase


def gpu_pointwise_fn(ctx, log_m, coset_g):
    import torch

    def pointwise(aA, aB, aC):
        ctx.qap_pointwise_dev_on(torch.cuda.current_stream().cuda_stream, aA.data_ptr(), aB.data_ptr(), aC.data_ptr(), log_m,
                                 coset_g, aA.shape[0])
    return pointwise


def nccl_exchange_fn(group=None):
    import torch.distributed as dist

    def exchange(dst, src):
        dist.all_to_all_single(dst.view(-1), src.view(-1), group=group)   # ordered on the current stream by ProcessGroupNCCL
    return exchange
