#!/usr/bin/env python3
"""bench.py — BW6-761 Groth16 aggregator prove hot path at batch 8 (BASELINE.json metric).

One "step" = the device work of ONE batch-8 aggregator proof as the hooked reference
prover would issue it (libsnark r1cs_gg_ppzksnark_prover, r1cs_gg_ppzksnark.tcc:464-603):
    witness map   : sparse mat-vecs A.w, B.w, C.w (237 702 rows, 1.8 M terms), then
                    3 x (iFFT, cosetFFT) + pointwise + icosetFFT over Fr, m = 2^18
    MSM A  (G1)   : 237 801 bases x witness
    MSM B  (G2,G1): 87 966 sparse pairs x witness[indices]
    MSM H  (G1)   : 262 143 bases x coefficients_for_H
    MSM L  (G1)   : 237 790 bases x witness[11:]
with the sizes of the real batch-8 circuit (SURVEY.md §8 [probe]); bases are synthetic
distinct curve points (k_i * generator), the witness an aggregator-like mix, all seeded.

    value : proofs/s, inputs resident in HBM, timed with CUDA events (max over ranks)
    e2e   : proofs/s through the host-pointer C ABI (pinned host buffers, H2D/D2H inside)
    --impl reference : the SAME statement (same constraint matrices, witness and B index list) through the
            reference's own code (oracle/_ref): libsnark::r1cs_to_qap_witness_map (serial mat-vecs + 7 libfqfft
            transforms), multi_exp_filter_one_zero (A, L), kc_multi_exp_with_mixed_addition (B), multi_exp (H)
            - the five calls of r1cs_gg_ppzksnark_prover - on all host cores, every step full size.

N > 1: one process per GPU, each proving its own stream of proofs (replicas; no data-path
collective) -> weak scaling, value = N * proofs / max-over-ranks time.
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

# one hardware work queue per CUDA stream: the library runs up to 12 streams per proof in flight
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

# batch-8 aggregator circuit (SURVEY.md §0 / §8 [probe])
LOG_M = 18
N_VARS = 237800          # assignment has N_VARS + 1 entries (constant one first)
N_INPUTS = 10
N_A = N_VARS + 1
N_B = 87966              # non-zero entries of pk.B_query
N_H = (1 << LOG_M) - 1
N_L = N_VARS - N_INPUTS
N_CONSTRAINTS = 237702
NNZ = (444246, 496421, 877410)   # non-zero terms of the A / B / C constraint matrices (SURVEY.md §0 [probe])
METRIC = "bw6_761_groth16_batch8_proofs_per_s"
UNIT = "proofs/s"
FR_TOP = 0x01AE3A4617C510EA  # top 64-bit limb of r: a random value with a smaller top limb is < r
# 32x32 products the device executes: a*b by one Karatsuba level = 3 * 12*12, the word-serial reduction
# 24*24 + 24. SURVEY.md 8(d)'s schoolbook figure (24*24 + 24*24 + 24 = 1176 per Fq product, 10 per mixed addition) is kept
# as PRODUCTS_PER_MADD_SCHOOLBOOK for comparison with round 1.
PRODUCTS_MULW = 432
PRODUCTS_REDC = 600
FQ_MUL_PER_MADD = 10         # madd-2008-s: 8M + 2S, every S done as M
# a mixed addition = 10 products but 9 reductions: Y3 = R(Q - X3) - Y1 PPP is reduced once (MULW + MULFIN)
PRODUCTS_PER_MADD = 10 * PRODUCTS_MULW + 9 * PRODUCTS_REDC
# the second point of every chain meets an affine accumulator (VM program madd_aff): 6 products, 5 reductions
PRODUCTS_PER_MADD_AFF = 6 * PRODUCTS_MULW + 5 * PRODUCTS_REDC
PRODUCTS_PER_MADD_SCHOOLBOOK = 10 * 1176


def random_fr(rng, n):
    a = rng.integers(0, 1 << 63, size=(n, 6), dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=(n, 6), dtype=np.uint64)
    a[:, 5] = rng.integers(0, FR_TOP, size=n, dtype=np.uint64)
    return a


def witness_like(rng, n):
    """Aggregator-like assignment in Montgomery form: 1 % zeros, 0.02 % ones, 20 % values
    below 2^16, the rest uniform (SURVEY.md §8(d) filter statistics)."""
    import bw6_constants as C
    w = random_fr(rng, n)
    kind = rng.random(n)
    Rm = (1 << C.FR_RBITS) % C.R
    small = np.nonzero(kind < 0.20)[0]
    vals = rng.integers(2, 1 << 16, size=small.shape[0])
    for i, v in zip(small.tolist(), vals.tolist()):
        x = v * Rm % C.R
        for k in range(6):
            w[i, k] = (x >> (64 * k)) & 0xFFFFFFFFFFFFFFFF
    w[(kind >= 0.20) & (kind < 0.21)] = 0
    one = [(Rm >> (64 * k)) & 0xFFFFFFFFFFFFFFFF for k in range(6)]
    w[(kind >= 0.21) & (kind < 0.2102)] = np.array(one, dtype=np.uint64)
    w[0] = np.array(one, dtype=np.uint64)
    return w


def synthetic_csr(rng, nnz, one):
    """CSR matrix with N_CONSTRAINTS rows and `nnz` terms spread evenly over the rows, columns uniform
    over the assignment, coefficients 90 % one / 10 % random (the aggregator's are mostly +-1 and
    small constants; the cost of a term does not depend on the value)."""
    per = np.full(N_CONSTRAINTS, nnz // N_CONSTRAINTS, dtype=np.uint64)
    per[: nnz - int(per.sum())] += 1
    row_ptr = np.zeros(N_CONSTRAINTS + 1, dtype=np.uint64)
    row_ptr[1:] = np.cumsum(per)
    col = rng.integers(0, N_VARS + 1, size=nnz, dtype=np.uint32)
    val = np.tile(one, (nnz, 1))
    k = nnz // 10
    val[rng.choice(nnz, k, replace=False)] = random_fr(rng, k)
    return row_ptr, col, val


def fr_generator():
    import bw6_constants as C
    x = C.FR_GENERATOR * ((1 << C.FR_RBITS) % C.R) % C.R
    return np.array([(x >> (64 * k)) & 0xFFFFFFFFFFFFFFFF for k in range(6)], dtype=np.uint64)


class Statement:
    """The synthetic batch-8-shaped statement both arms prove: constraint matrices, assignment, the index
    list of the sparse B query and the coset generator. Seeded, so `--impl reference` sees the same bytes."""

    def __init__(self, seed=1):
        rng = np.random.default_rng(seed)
        self.b_index = np.sort(rng.choice(N_A, N_B, replace=False)).astype(np.uint64)
        self.w = witness_like(rng, N_A)
        one = self.w[0].copy()
        self.csr = [synthetic_csr(rng, nz, one) for nz in NNZ]
        self.g = fr_generator()


WORKLOAD = ("batch-8 aggregator Groth16 prove hot path (r1cs_gg_ppzksnark_prover): r1cs_to_qap_witness_map "
            "(A.w, B.w, C.w mat-vecs + 7 NTT, m=2^18) + MSM A / B (G2,G1 pair) / H / L")


def bench_config():
    """Identical in both arms (the driver compares it)."""
    return {"workload": WORKLOAD,
            "msm_sizes": {"A": N_A, "B_g2": N_B, "B_g1": N_B, "H": N_H, "L": N_L}, "ntt_log_m": LOG_M,
            "r1cs": {"constraints": N_CONSTRAINTS, "variables": N_VARS, "inputs": N_INPUTS, "nnz_abc": list(NNZ)},
            "l2": "no flush needed: every step gathers ~3 GB of distinct base points from a 4 GB key image (L2 is 126 MB)"}


def host_cpu():
    """CPU model, logical and physical core counts of this box."""
    model, phys = None, set()
    try:
        pid = cid = None
        for ln in open("/proc/cpuinfo"):
            k, _, v = ln.partition(":")
            k, v = k.strip(), v.strip()
            if k == "model name" and model is None:
                model = v
            elif k == "physical id":
                pid = v
            elif k == "core id":
                cid = v
            elif k == "" and pid is not None:
                phys.add((pid, cid))
                pid = cid = None
    except OSError:
        pass
    return {"model": model, "logical": os.cpu_count(), "physical": len(phys) or os.cpu_count()}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._pump, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


class Workload:
    """Synthetic batch-8 proving key image + witness, on one GPU."""

    def __init__(self, device, seed=1, depth=3):
        import torch
        import zecale_b200 as z
        self.torch, self.z = torch, z
        self.dev = device
        torch.cuda.set_device(device)
        # `depth` proofs are kept in flight, each on its own context (= its own lanes, streams and
        # scratch); the pinned proving key is shared by all of them
        self.ctxs = [z.Context(device) for _ in range(max(1, depth))]
        self.ctx = self.ctxs[0]
        st = Statement(seed)
        rng = np.random.default_rng(seed + 1000)
        ctx = self.ctx
        # proving-key queries: distinct curve points, pinned (device-resident) once, like a loaded pk
        self.pk = {}
        for name, n, grp, sd in (("A", N_A, z.G1, 11), ("B1", N_B, z.G1, 12), ("B2", N_B, z.G2, 13),
                                 ("H", N_H, z.G1, 14), ("L", N_L, z.G1, 15)):
            pts = ctx.synth_bases(n, grp, seed=sd + 100 * seed)
            self.pk[name] = ctx.pin_bases(pts, grp)
            del pts
        self.b_index = st.b_index
        self.g = st.g
        # host inputs (pinned) of one proof
        def pinned(a):
            t = torch.from_numpy(a.view(np.int64)).pin_memory()
            return t, t.numpy().view(np.uint64)
        m = 1 << LOG_M
        self.t_w, self.h_w = pinned(st.w)
        # constraint system image: the evaluations <A_i,w>, <B_i,w>, <C_i,w> are computed
        # on the device from the assignment, as the specialised prover of the hooks does
        self.r1cs = ctx.r1cs_load(N_CONSTRAINTS, N_INPUTS, N_VARS, st.csr)
        assert ctx.r1cs_log_domain(self.r1cs) == LOG_M
        self.t_abc, self.h_abc = [], []
        for _ in range(3):
            t, h = pinned(random_fr(rng, m))
            self.t_abc.append(t)
            self.h_abc.append(h)
        # device-resident copies
        self.d_w = self.t_w.cuda()
        self.d_abc0 = [t.cuda() for t in self.t_abc]
        self.d_abc = [torch.empty_like(t) for t in self.d_abc0]
        self.d_idx = torch.from_numpy(self.b_index.view(np.int64)).cuda()
        self.t_hout = [torch.empty(((1 << LOG_M) + 1, 6), dtype=torch.int64).pin_memory() for _ in self.ctxs]
        self.h_out = [t.numpy().view(np.uint64) for t in self.t_hout]
        torch.cuda.synchronize()
        self.last = None
        self.results = [None] * len(self.ctxs)

    def run(self, steps, mode, depth=None):
        """`steps` proofs, `depth` (default: all contexts) of them in flight: proof i is enqueued on
        context i mod depth as soon as proof i - depth (same context) has delivered its results.
        Returns the results of the last proof."""
        enqueue = self.enqueue_resident if mode == "resident" else self.enqueue_e2e
        d = min(depth or len(self.ctxs), len(self.ctxs))
        for i in range(steps):
            k = i % d
            if self.results[k] is not None:
                self.ctxs[k].sync()
            self.results[k] = enqueue(k)
        last = None
        for j in range(steps, steps + d):     # drain in issue order
            k = j % d
            if self.results[k] is not None:
                self.ctxs[k].sync()
                last = self.results[k]
                self.results[k] = None
        self.last = last
        return last

    def step_resident(self):
        return self.run(1, "resident")

    def step_e2e(self):
        return self.run(1, "e2e")

    def enqueue_resident(self, slot=0):
        """One proof's device work with every input already in HBM: the H polynomial and the H MSM
        on lane 0, the four witness MSMs on lanes 1-4; results are delivered by the context's sync."""
        ctx = self.ctxs[slot]
        out = [np.zeros(36, dtype=np.uint64) for _ in range(5)]
        w, idx = self.d_w.data_ptr(), self.d_idx.data_ptr()
        ctx.qap_h_from_witness_async(0, self.r1cs, w, self.g, device=True)
        ctx.msm_pinned_async(1, self.pk["A"], w, out[0], n=N_A, device=True)
        ctx.msm_pinned_async(2, self.pk["L"], w + 48 * (N_INPUTS + 1), out[4], n=N_L, device=True)
        ctx.msm_pinned_async(3, self.pk["B2"], w, out[1], n=N_B, index=idx, device=True)
        ctx.msm_pinned_async(4, self.pk["B1"], w, out[2], n=N_B, index=idx, device=True)
        ctx.msm_pinned_async(0, self.pk["H"], None, out[3], n=N_H, from_qap=True)
        ctx.flush()
        return out

    def enqueue_e2e(self, slot=0):
        """The same proof from HOST buffers (pinned), exactly the calls the specialised prover of the
        hooks makes (zecale_gpu_hooks::evaluate_fused): the assignment crosses PCIe once (shared by the
        mat-vecs and by A, B, L), five points come back."""
        ctx = self.ctxs[slot]
        out = [np.zeros(36, dtype=np.uint64) for _ in range(5)]
        w = self.h_w
        ctx.qap_h_from_witness_async(0, self.r1cs, w, self.g)
        ctx.msm_pinned_async(1, self.pk["A"], w, out[0], n=N_A)
        ctx.msm_pinned_async(2, self.pk["L"], w[N_INPUTS + 1:], out[4], n=N_L)
        ctx.msm_pinned_async(3, self.pk["B2"], w, out[1], n=N_B, index=self.b_index)
        ctx.msm_pinned_async(4, self.pk["B1"], w, out[2], n=N_B, index=self.b_index)
        ctx.msm_pinned_async(0, self.pk["H"], None, out[3], n=N_H, from_qap=True)
        ctx.flush()
        return out

    def step_blocking(self):
        """Strict drop-in sequence: one blocking C-ABI call per reference call site, host buffers."""
        ctx = self.ctx
        h = ctx.qap_h(self.h_abc[0], self.h_abc[1], self.h_abc[2], self.g)
        w = self.h_w
        return [ctx.msm_pinned(self.pk["A"], w, n=N_A),
                ctx.msm_pinned(self.pk["B2"], w, index=self.b_index, n=N_B),
                ctx.msm_pinned(self.pk["B1"], w, index=self.b_index, n=N_B),
                ctx.msm_pinned(self.pk["H"], h, n=N_H),
                ctx.msm_pinned(self.pk["L"], w[N_INPUTS + 1:], n=N_L)]

    E2E_H2D = N_A * 48 + N_B * 8      # the assignment [1, primary, aux] and the B-query index list
    E2E_D2H = 5 * 288                 # the five MSM results (coefficients_for_H never leave the device)

    def isolated_kernels(self):
        """The dominant kernels timed ALONE (nothing else on the GPU), on the step's own inputs:
        each of the five MSMs and the H polynomial are run once through the blocking device-pointer
        entry points and the library's CUDA-event brackets around k_accumulate (class 0) and the NTT
        transforms (class 1) are read back. Inside the timed region these kernels overlap with the
        kernels of sibling lanes and of the other proof in flight, so event brackets taken there
        measure co-scheduled time, not the kernel. Also returns the mixed additions executed: bucket
        entries (non-zero signed digits of points that are not at infinity) minus accumulate tasks
        (the first entry of a task is a copy)."""
        ctx = self.ctx
        w = self.d_w.data_ptr()
        for dst, src in zip(self.d_abc, self.d_abc0):
            dst.copy_(src)
        self.torch.cuda.current_stream().synchronize()
        ctx.kernel_ms(0)
        ctx.kernel_ms(1)
        ctx.qap_h_dev(self.d_abc[0].data_ptr(), self.d_abc[1].data_ptr(), self.d_abc[2].data_ptr(), LOG_M, self.g)
        ntt_ms, ntt_n = ctx.kernel_ms(1)
        tot, shapes = 0, {}
        calls = (("A", w, N_A, None), ("B2", w, N_B, self.d_idx.data_ptr()), ("B1", w, N_B, self.d_idx.data_ptr()),
                 ("H", self.d_abc[0].data_ptr(), N_H, None), ("L", w + 48 * (N_INPUTS + 1), N_L, None))
        for name, sc, n, idx in calls:
            ctx.msm_pinned_dev(self.pk[name], sc, n, d_index_ptr=idx)
            st = ctx.msm_last_stats()
            ms, _ = ctx.kernel_ms(0)
            st["accumulate_ms"] = ms
            tot += st["entries"] - st["tasks"]
            shapes[name] = st
        acc_ms = sum(v["accumulate_ms"] for v in shapes.values())
        return tot, shapes, acc_ms, ntt_ms, ntt_n

    def close(self):
        for p in self.pk.values():
            p.release()
        self.ctx.r1cs_free(self.r1cs)
        for c in self.ctxs:
            c.close()

    def kernel_ms(self, kernel_class):
        ms, n = 0.0, 0
        for c in self.ctxs:
            a, b = c.kernel_ms(kernel_class)
            ms += a
            n += b
        return ms, n

    def launch_count(self):
        return sum(c.launch_count() for c in self.ctxs)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """DRAM bytes per k_accumulate launch from the committed ncu --set full capture (profiles/r02_traffic.json,
    written from the .ncu-rep by tools/ncu_summary.py); None if no capture of this kernel version exists."""
    p = os.path.join(ROOT, "profiles", "r02_traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return None


PAPER_PEAK_PRODUCTS = 148 * 32 * 1.965e9   # 148 SM x 32 IMAD.WIDE / clk x 1.965 GHz = 9.3e12


def imad_peak(ctx):
    """Measured integer-multiply ceiling of this GPU in 32x32 products per second: independent chains of IMAD.WIDE.U32
    (k_bench_imad, 8 chains per thread) over a few occupancies. Also returns the Fq Montgomery products/s of the
    dependent-chain product kernel (k_bench_fq_mul), the figure rounds 1-2 used as the ceiling."""
    best = 0.0
    for blocks_per_sm, tpb in ((2, 256), (4, 256), (8, 256)):
        blocks, iters = 148 * blocks_per_sm, 4096
        ms = min(ctx.bench_imad(blocks, tpb, iters) for _ in range(3))
        best = max(best, blocks * tpb * iters * 8 / (ms * 1e-3))
    fq = 0.0
    for blocks_per_sm, tpb in ((2, 128), (4, 64)):
        blocks, iters = 148 * blocks_per_sm, 1000
        ms = min(ctx.bench_fq_mul(blocks, tpb, iters) for _ in range(2))
        fq = max(fq, blocks * tpb * iters / (ms * 1e-3))
    return best, fq


def run_b200(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    wl = Workload(local_rank, seed=1 + rank, depth=args.depth)
    ctx = wl.ctx

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(mode, steps, depth=None):
        """(seconds, launches) of `steps` proofs; CUDA events around host-synchronising calls, wall clock kept if larger."""
        l0 = wl.launch_count()
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        t0 = time.perf_counter()
        wl.run(steps, mode, depth)       # returns after every context has synchronised its lanes
        torch.cuda.synchronize()
        ev1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        t = ev0.elapsed_time(ev1) * 1e-3
        barrier()
        # the library runs on its own streams, bracketed by host-synchronising calls: the events enclose it
        return max(t, wall) if t < 0.5 * wall else t, wl.launch_count() - l0

    # ---- device-resident arm (value)
    wl.run(args.warmup, "resident")
    sampler = ClockSampler(local_rank)
    sampler.start()
    t_res, launches = timed("resident", args.steps)
    clocks = sampler.stop()

    # ---- end-to-end arm (host pointers)
    wl.run(max(args.depth, min(args.warmup, 2)), "e2e")   # every context in flight takes one host-pointer step (staging buffers) first
    t_e2e, _ = timed("e2e", args.steps)

    # ---- one proof at a time (what a caller without a pipeline reaches)
    wl.run(2, "e2e", depth=1)
    t_d1, _ = timed("e2e", max(3, min(args.steps, 10)), depth=1)
    t_d1 /= max(3, min(args.steps, 10))

    # ---- secondary headline: G1 MSM 2^20 latency (rank 0, N = 1 only)
    msm20 = None
    if world == 1 and not args.skip_msm20:
        n = 1 << 20
        pts = ctx.synth_bases(n, wl.z.G1, seed=77)
        pin = ctx.pin_bases(pts)
        sc_h = random_fr(np.random.default_rng(5), n)
        t_sc = torch.from_numpy(sc_h.view(np.int64)).pin_memory()
        sc = t_sc.cuda()
        torch.cuda.synchronize()
        ts = [ctx.msm_pinned_dev(pin, sc.data_ptr(), n)[1] for _ in range(4)]
        h_sc = t_sc.numpy().view(np.uint64)
        th = []
        for _ in range(3):
            t0 = time.perf_counter()
            ctx.msm_pinned(pin, h_sc)
            th.append((time.perf_counter() - t0) * 1e3)
        import hashlib
        msm20 = {"device_resident_ms": min(ts[1:]), "host_to_host_pinned_key_ms": min(th), "c": ctx.msm_last_stats()["c"],
                 "point_digest": hashlib.sha256(ctx.msm_pinned(pin, h_sc).tobytes()).hexdigest()[:16]}
        pin.release()
        del sc, pts

    # ---- the two exchange paths north_star names, one process per GPU over NCCL (N > 1 only): the same 2^20 MSM with its
    # points sharded over the ranks (288-byte partial sums all-gathered) and a 4-step NTT of 2^22 with one all-to-all
    nccl = None
    if world > 1 and not args.skip_msm20:
        try:
            nccl = nccl_block(ctx, wl.z, rank, world, local_rank, barrier)
        except Exception as e:  # noqa: BLE001 - the headline line must still be printed
            nccl = {"error": repr(e)}

    if world > 1:
        tt = torch.tensor([t_res, t_e2e, t_d1], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_res, t_e2e, t_d1 = tt.tolist()
    out = None
    if rank == 0:
        peaks, peaks_src = measured_peaks()
        peak_prod, fq_mul_rate = imad_peak(ctx)
        madds, shapes, iso_acc_ms, iso_ntt_ms, iso_ntt_n = wl.isolated_kernels()
        acc_s = iso_acc_ms * 1e-3                            # the step's 5 accumulate launches, each timed alone
        n_tasks = sum(v["tasks"] for v in shapes.values())    # one chain per task; (almost) every chain has a second point
        prod_step = (madds - n_tasks) * PRODUCTS_PER_MADD + n_tasks * PRODUCTS_PER_MADD_AFF
        achieved = prod_step / acc_s if acc_s > 0 else 0.0
        step_s = t_res / args.steps
        m = 1 << LOG_M
        ntt_bytes = 7 * 192 * m                              # 2 passes x (48 B read + 48 B write) per element
        ntt_products = 7 * 300 * (m // 2 * LOG_M + m)        # butterflies + the one scaling pass of each transform
        ntt_s = iso_ntt_ms * 1e-3                            # the step's 7 transforms, timed alone
        traffic = ncu_traffic()
        out = {
            "metric": METRIC, "value": world * args.steps / t_res, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * step_s,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32x24 (761-bit Fq) / u32x12 (377-bit Fr)",
            "data": "synthetic (seeded curve points k_i*G, aggregator-like witness and constraint matrices)",
            "config": bench_config(),
            "arm": {"parallelism": "replicas (one proof stream per GPU, no data-path collective)" if world > 1 else "single GPU",
                    "proofs_in_flight": args.depth,
                    "value_depth1": world / t_d1, "ms_per_step_depth1": 1e3 * t_d1,
                    "value_depth1_note": "host-pointer C ABI, ONE proof in flight: what a caller that proves strictly one at a time reaches"},
            "e2e": {"value": world * args.steps / t_e2e, "unit": UNIT, "h2d_bytes_per_step": Workload.E2E_H2D,
                    "d2h_bytes_per_step": Workload.E2E_D2H},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "imad", "bound_note": "integer-multiply pipe (sm__pipe_fmaheavy): neither HBM- nor tensor-bound, SURVEY.md 8(d)",
                         "kernel": "k_accumulate (MSM bucket accumulation)", "achieved": achieved / 1e12,
                         "peak": peak_prod / 1e12, "unit": "T(32x32 products)/s", "frac": achieved / peak_prod if peak_prod else None,
                         "peak_source": "measured in this run: independent IMAD.WIDE.U32 chains (k_bench_imad, zecale_gpu_bench_imad)",
                         "fq_products_per_s": fq_mul_rate,
                         "products_per_madd": PRODUCTS_PER_MADD,
                         "products_per_madd_affine_accumulator": PRODUCTS_PER_MADD_AFF,
                         "products_note": "32x32 products executed per mixed addition: 10 x 432 (one Karatsuba level on a*b) + 9 x 600 "
                                          "(word-serial reductions; Y3 is reduced once); the second addition of each chain (one per task) "
                                          "has an affine accumulator and costs 6 x 432 + 5 x 600. Round 1 executed 10 x 1176 = 11760 for every "
                                          "addition: madds_per_step x 11760 / step time is the figure comparable with r01's in_step_achieved",
                         "in_step_achieved_schoolbook_equiv": madds * PRODUCTS_PER_MADD_SCHOOLBOOK / step_s / 1e12,
                         "peak_paper": PAPER_PEAK_PRODUCTS / 1e12, "frac_of_paper_peak": achieved / PAPER_PEAK_PRODUCTS,
                         "launches_per_step": 5, "launch_ms_avg": 1e3 * acc_s / 5,
                         "timing": "CUDA events on the launching stream, each of the step's five launches alone on the GPU",
                         "algorithmic_products_per_launch_avg": prod_step / 5,
                         "hbm_algorithmic_bytes_per_launch_avg": (madds + sum(v["tasks"] for v in shapes.values())) * 192 / 5,
                         "in_step_achieved": prod_step / step_s / 1e12,
                         "in_step_frac": prod_step / step_s / peak_prod if peak_prod else None,
                         "in_step_note": "the accumulate products of one step / the driver-timed ms_per_step: a lower bound on the pipe's "
                                         "useful load inside the timed region (tails, transforms and sorting excluded from the numerator)",
                         "madds_per_step": madds, "msm_shapes": shapes,
                         "traffic": traffic["dram_bytes_per_launch"] if traffic else None,
                         "traffic_source": traffic["source"] if traffic else "no ncu capture of this kernel version committed yet"},
            "roofline_ntt": {"bound": "hbm", "kernel": "k_ntt_pass (Fr radix-2 NTT, 7 transforms per step)",
                             "achieved": ntt_bytes / ntt_s / 1e9 if ntt_s > 0 else None, "peak": peaks["hbm_gbs"],
                             "unit": "GB/s", "frac": (ntt_bytes / ntt_s / 1e9) / peaks["hbm_gbs"] if ntt_s > 0 else None,
                             "peak_source": peaks_src,
                             "achieved_imad": ntt_products / ntt_s / 1e12 if ntt_s > 0 else None,
                             "frac_imad": ntt_products / ntt_s / peak_prod if ntt_s > 0 and peak_prod else None,
                             "imad_note": "300 products per Fr product x (m/2 log2 m butterflies + m scalings) per transform; the 377-bit "
                                          "NTT is integer-multiply bound (SURVEY.md 8(d)), the HBM fraction is what north_star asks to see",
                             "transforms_ms": iso_ntt_ms, "share_of_step": ntt_s / step_s, "traffic": None},
            "msm_g1_2e20": msm20,
            "nccl_paths": nccl,
        }
    wl.close()
    del wl
    torch.cuda.empty_cache()
    # ---- config 5 (strong scaling of one proof): rank 0 drives all N GPUs through one multi-device
    # context (the drop-in is a single-process prover); the other ranks have released their GPUs and wait
    if not args.skip_sharded:
        if world > 1:
            cpu_group = dist.new_group(backend="gloo")
            dist.barrier(group=cpu_group)
        if rank == 0:
            try:
                out["sharded"] = sharded_block(args, world, list(range(world)))
            except Exception as e:  # noqa: BLE001 - the headline line must still be printed
                out["sharded"] = {"error": repr(e)}
        if world > 1:
            dist.barrier(group=cpu_group)
    if world > 1:
        dist.destroy_process_group()
    return out


def nccl_block(ctx, z, rank, world, local_rank, barrier):
    """One process per GPU, torch.distributed over NCCL (zecale_b200/dist.py): (1) G1 MSM of 2^20 points sharded by points
    like libff::multi_exp shards over threads (multiexp.tcc:663-685), partial sums all-gathered and added in rank order -
    the digest equals msm_g1_2e20.point_digest of the N = 1 line; (2) forward and inverse-coset 4-step NTT of 2^22 elements,
    layouts A <-> B with ONE all-to-all each (zecale_gpu_ntt_dist_phase), checked by a forward / inverse round trip.
    Times are CUDA events around the calls, max over ranks, best of three. All ranks must call this."""
    import hashlib
    import torch
    import torch.distributed as dist
    from zecale_b200 import dist as D
    dev = torch.device("cuda", local_rank)

    def timed_max(fn, reps=3):
        best, res = None, None
        for _ in range(reps):
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            res = fn()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            best = float(t.item()) if best is None else min(best, float(t.item()))
        return best, res

    out = {"world": world}
    n = 1 << 20
    lo, hi = D.msm_chunk_range(n, world, rank)
    pts = ctx.synth_bases(n, z.G1, seed=77)[lo:hi]
    pin = ctx.pin_bases(np.ascontiguousarray(pts))
    del pts
    sc = torch.from_numpy(random_fr(np.random.default_rng(5), n)[lo:hi].view(np.int64).copy()).to(dev)
    ms, total = timed_max(lambda: D.msm_sharded(lambda: ctx.msm_pinned_dev(pin, sc.data_ptr(), hi - lo)[0], ctx.point_sum, world,
                                                device=dev))
    pin.release()
    out["msm_g1_2e20_point_sharded"] = {"ms": ms, "points_per_rank": hi - lo, "exchange": "all_gather of one 288-byte point per rank",
                                        "point_digest": hashlib.sha256(total.tobytes()).hexdigest()[:16]}
    log_m = 22
    g = fr_generator()
    drv = D.DistNtt(D.gpu_phase_fn(ctx), D.nccl_exchange_fn(), world, rank)
    x = random_fr(np.random.default_rng(log_m), 1 << log_m)[D.layout_a_indices(log_m, world, rank)]
    keep = torch.from_numpy(x.view(np.int64).copy()).to(dev)
    buf, scratch = keep.clone(), torch.empty_like(keep)
    drv.transform(buf, scratch, log_m, 0)                                  # warm-up: twiddle tables
    res = {"log_m": log_m, "exchange": "all_to_all_single of m/world elements per rank, once per transform"}

    def fwd():
        buf.copy_(keep)
        return drv.transform(buf, scratch, log_m, 0)

    # the copy is inside the bracket on purpose only for simplicity of the closure; it is 1/20 of the transform
    res["fft_ms"], (y, other) = timed_max(fwd)
    back, _ = drv.transform(y, other, log_m, 1)                            # inverse: layout B -> A, must restore the input
    ok = torch.tensor([int(torch.equal(back, keep))], device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    res["round_trip_ok"] = bool(ok.item())
    xb = torch.from_numpy(random_fr(np.random.default_rng(log_m + 1), 1 << log_m)[D.layout_b_indices(log_m, world, rank)]
                          .view(np.int64).copy()).to(dev)
    bufb = xb.clone()

    def icoset():
        bufb.copy_(xb)
        return drv.transform(bufb, scratch, log_m, 3, g)

    drv.transform(bufb, scratch, log_m, 3, g)
    res["icoset_fft_ms"], _ = timed_max(icoset)
    out["ntt_4step"] = res
    return out


# ---- BASELINE config 5: ONE batch-32-shaped proof, MSM points sharded over the GPUs (strong scaling) ----
B32 = {"log_m": 20, "n_vars": 920960, "n_inputs": 34, "n_b": 329934, "n_constraints": 920718,
       "nnz": (1720900, 1923000, 3398900)}   # sizes of the real batch-32 aggregator circuit (tests/test_hooks_gpu.py prints them)


def sharded_block(args, world, devices):
    """One proof of the batch-32 shape at a time on a multi-device context spanning `devices` (zecale_gpu_init
    with n_devices = N): every pinned query is sharded by points, each device runs a Pippenger per shard, the
    384-byte partial sums are peer-copied to devices[0] and combined there; the witness map runs on devices[0]
    and the H coefficients are sliced out to the other devices over NVLink. Host-pointer calls (pinned host
    assignment), exactly what the hooked prover issues. Returns ms per proof and a digest of the five points
    (identical at every N: same seeded statement)."""
    import hashlib
    import torch
    import zecale_b200 as z
    global LOG_M, N_VARS, N_INPUTS, N_A, N_B, N_H, N_L, N_CONSTRAINTS, NNZ
    saved = (LOG_M, N_VARS, N_INPUTS, N_A, N_B, N_H, N_L, N_CONSTRAINTS, NNZ)
    LOG_M, N_VARS, N_INPUTS, N_B, N_CONSTRAINTS, NNZ = (B32["log_m"], B32["n_vars"], B32["n_inputs"], B32["n_b"],
                                                       B32["n_constraints"], B32["nnz"])
    N_A, N_H, N_L = N_VARS + 1, (1 << LOG_M) - 1, N_VARS - N_INPUTS
    try:
        torch.cuda.set_device(devices[0])
        st = Statement(32)
        gen = z.Context(devices[0])                 # bases are generated once on one device ...
        ctx = z.Context(devices if len(devices) > 1 else devices[0])   # ... and pinned through the (multi-device) context
        pk = {}
        t0 = time.perf_counter()
        for name, n, grp, sd in (("A", N_A, z.G1, 311), ("B1", N_B, z.G1, 312), ("B2", N_B, z.G2, 313),
                                 ("H", N_H, z.G1, 314), ("L", N_L, z.G1, 315)):
            pts = gen.synth_bases(n, grp, seed=sd)
            pk[name] = ctx.pin_bases(pts, grp)
            del pts
        gen.close()
        t_pin = time.perf_counter() - t0
        t_w = torch.from_numpy(st.w.view(np.int64)).pin_memory()
        h_w = t_w.numpy().view(np.uint64)
        r1cs = ctx.r1cs_load(N_CONSTRAINTS, N_INPUTS, N_VARS, st.csr)
        assert ctx.r1cs_log_domain(r1cs) == LOG_M

        def prove():
            out = [np.zeros(36, dtype=np.uint64) for _ in range(5)]
            ctx.qap_h_from_witness_async(0, r1cs, h_w, st.g)
            ctx.msm_pinned_async(1, pk["A"], h_w, out[0], n=N_A)
            ctx.msm_pinned_async(2, pk["L"], h_w[N_INPUTS + 1:], out[4], n=N_L)
            ctx.msm_pinned_async(3, pk["B2"], h_w, out[1], n=N_B, index=st.b_index)
            ctx.msm_pinned_async(4, pk["B1"], h_w, out[2], n=N_B, index=st.b_index)
            ctx.msm_pinned_async(0, pk["H"], None, out[3], n=N_H, from_qap=True)
            ctx.sync()
            return out

        for _ in range(3):
            out = prove()
        for d in devices:
            torch.cuda.synchronize(d)
        steps = max(3, min(args.steps, 8))
        t0 = time.perf_counter()
        for _ in range(steps):
            out = prove()
        for d in devices:
            torch.cuda.synchronize(d)
        ms = (time.perf_counter() - t0) * 1e3 / steps
        digest = hashlib.sha256(b"".join(o.tobytes() for o in out)).hexdigest()[:16]
        for p_ in pk.values():
            p_.release()
        ctx.r1cs_free(r1cs)
        ctx.close()
        return {"workload": "ONE batch-32-shaped aggregator proof (m=2^20; MSM A %d / B %d x2 / H %d / L %d), MSM points sharded over the "
                            "GPUs of a multi-device context, partial sums combined by peer copies; host-pointer C ABI, one proof in flight"
                            % (N_A, N_B, N_H, N_L),
                "n_gpus": len(devices), "ms_per_proof": ms, "proofs_per_s": 1e3 / ms, "scaling": "strong", "steps": steps,
                "key_pin_s": t_pin, "digest": digest}
    finally:
        LOG_M, N_VARS, N_INPUTS, N_A, N_B, N_H, N_L, N_CONSTRAINTS, NNZ = saved


class ReferenceWorkload:
    """The same statement on the host cores with the UNMODIFIED reference (oracle/_ref), through the
    calls r1cs_gg_ppzksnark_prover makes (r1cs_gg_ppzksnark.tcc:477-580): r1cs_to_qap_witness_map
    (serial mat-vecs + 7 libfqfft transforms), multi_exp_filter_one_zero for A and L,
    kc_multi_exp_with_mixed_addition for the sparse (G2, G1) B query, multi_exp for H; BDLO12_signed,
    special bases, chunks = threads. Always full size. Bases are (i+1)*G (the cost of a Pippenger pass
    does not depend on which distinct points it sums)."""

    def __init__(self, threads=None):
        from oracle import ref as R
        self.ref = R.Ref()
        self.threads = self.ref.set_threads(threads or os.cpu_count())   # torchrun exports OMP_NUM_THREADS=1
        st = Statement(1)
        self.st = st
        self.g1 = self.ref.bases_seq(max(N_A, N_H), 1)
        g2 = self.ref.bases_seq(N_B, 2)
        self.kc = np.ascontiguousarray(np.hstack([g2, self.g1[:N_B]]))
        assert self.ref.session_from_csr(N_CONSTRAINTS, N_INPUTS, N_VARS, st.csr, st.w) == 0

    def step(self):
        """Seconds of one proof's hot path (sum of the reference's own timed calls) and the split."""
        ref, st = self.ref, self.st
        h, t_map = ref.session_witness_map(1 << LOG_M)
        _, tA = ref.msm(self.g1[:N_A], st.w, 1, method=1)
        _, tB = ref.msm_kc(self.kc, st.b_index, N_A, st.w)
        _, tH = ref.msm(self.g1[:N_H], h[:N_H], 1, method=0)
        _, tL = ref.msm(self.g1[:N_L], st.w[N_INPUTS + 1:], 1, method=1)
        return t_map + tA + tB + tH + tL, {"witness_map_s": t_map, "msm_a_s": tA, "msm_b_s": tB, "msm_h_s": tH, "msm_l_s": tL}


def cpu_baseline_sample():
    """cpu_baseline leg of the default run: one full-size step on all host cores, and one on 8 threads
    (the count BASELINE.md's survey-container numbers were taken with)."""
    from oracle import ref as R
    cpu = host_cpu()
    if not R.available():
        return {"value": None, "unit": UNIT, "cores": cpu["logical"], "kind": "reference", "cpu": cpu,
                "sample": "oracle/_ref not built on this box"}
    rw = ReferenceWorkload()
    t, split = rw.step()
    out = {"value": 1.0 / t, "unit": UNIT, "cores": rw.threads, "kind": "reference", "cpu": cpu,
           "sample": "one full-size step, once: libsnark r1cs_to_qap_witness_map (mat-vecs + 7 transforms, m=2^18) + multi_exp_filter_one_zero "
                     "(A, L) + kc_multi_exp_with_mixed_addition (B) + multi_exp (H) at the batch-8 sizes",
           "split_s": split}
    if rw.threads > 8:
        rw.threads = rw.ref.set_threads(8)
        t8, _ = rw.step()
        out["value_8_threads"] = 1.0 / t8
    return out


def run_reference(args, rank, world):
    if rank != 0:
        return None
    from oracle import ref as R
    if not R.available():
        return {"impl": "reference", "unavailable": "oracle/_ref/libzecale_ref.so is not built (needs /root/reference at build time)"}
    rw = ReferenceWorkload()
    cpu = host_cpu()
    n_warm = 1 if args.warmup > 0 else 0
    for _ in range(n_warm):
        rw.step()
    ts, splits = [], []
    for _ in range(max(args.steps, 1)):
        t, sp = rw.step()
        ts.append(t)
        splits.append(sp)
    t = sum(ts) / len(ts)
    split = {k: sum(sp[k] for sp in splits) / len(splits) for k in splits[0]}
    sample = ("every step full size (no sub-sampling, no extrapolation): the unmodified reference's r1cs_to_qap_witness_map + the four "
              "multi-exponentiation calls of r1cs_gg_ppzksnark_prover on the same statement as the b200 arm; %d timed steps after %d "
              "full-size warm-up step (a CPU step is ~5 s; more warm-up would only burn minutes)" % (len(ts), n_warm))
    return {"impl": "reference", "metric": METRIC, "value": 1.0 / t, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64x12 (761-bit Fq) / u64x6 (377-bit Fr), GMP mpn", "data": "synthetic (seeded)",
            "config": bench_config(),
            "arm": {"parallelism": "OpenMP, %d threads" % rw.threads, "cpu": cpu},
            "cpu_baseline": {"value": 1.0 / t, "unit": UNIT, "cores": rw.threads, "kind": "reference", "sample": sample, "cpu": cpu},
            "e2e": {"value": 1.0 / t, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "split": split}


def main():
    # the contract is ONE JSON line on stdout: native libraries (NCCL prints its version banner) write to
    # fd 1 too, so everything but the final line is sent to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        os.write(real_stdout, (json.dumps(obj) + "\n").encode())

    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--depth", type=int, default=3, help="proofs kept in flight per GPU (1 = strictly one at a time)")
    ap.add_argument("--skip-msm20", action="store_true")
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--skip-sharded", action="store_true", help="skip the config-5 block (one batch-32-shaped proof sharded over the GPUs)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        out = run_reference(args, rank, world)
        if out is not None:
            emit(out)
        return
    if args.warmup < 3:
        args.warmup = 3
    out = run_b200(args, rank, world, local_rank)
    if rank == 0:
        if world == 1 and not args.skip_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline_sample()
        emit(out)


if __name__ == "__main__":
    main()
