"""BASELINE configs 2 and 3 as measurement tables (one JSON line per row):

  MSM  G1 and G2, n = 2^16 .. 2^24 random points / scalars:
       key expansion (pin) seconds and bytes, the window the pinned plan uses,
       M2 in its three variants - device-resident pinned, host->host pinned, host->host un-pinned
  NTT  FFT / iFFT / cosetFFT / icosetFFT, m = 2^16 .. 2^24: device-resident and host->host
  CPU  the reference (oracle/_ref: libff::multi_exp / libfqfft) at 2^16 / 2^18 / 2^20 on all cores and on 8 threads

Usage: python scripts/sweep.py [--max-log 24] [--cpu] > profiles/r02_msm_ntt_sweep.jsonl
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
import torch

import bench
import zecale_b200 as z

ap = argparse.ArgumentParser()
ap.add_argument("--max-log", type=int, default=24)
ap.add_argument("--min-log", type=int, default=16)
ap.add_argument("--cpu", action="store_true")
ap.add_argument("--skip-gpu", action="store_true")
args = ap.parse_args()


def emit(row):
    print(json.dumps(row), flush=True)


def best(f, reps=3):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        f()
        ts.append((time.perf_counter() - t0) * 1e3)
    return min(ts)


if not args.skip_gpu:
    ctx = z.Context(0)
    g = bench.fr_generator()
    free0 = torch.cuda.mem_get_info()[0]
    for lg in range(args.min_log, args.max_log + 1, 2):
        n = 1 << lg
        sc_h = bench.random_fr(np.random.default_rng(lg), n)
        t_sc = torch.from_numpy(sc_h.view(np.int64)).pin_memory()
        h_sc = t_sc.numpy().view(np.uint64)
        d_sc = t_sc.cuda()
        for grp, name in ((z.G1, "g1"), (z.G2, "g2")):
            pts = ctx.synth_bases(n, grp, seed=1000 + lg)
            torch.cuda.synchronize()
            f0 = torch.cuda.mem_get_info()[0]
            t0 = time.perf_counter()
            pin = ctx.pin_bases(pts, grp)
            t_pin = time.perf_counter() - t0
            key_bytes = f0 - torch.cuda.mem_get_info()[0]
            dev = [ctx.msm_pinned_dev(pin, d_sc.data_ptr(), n)[1] for _ in range(3)]
            st = ctx.msm_last_stats()
            host_pinned = best(lambda: ctx.msm_pinned(pin, h_sc), 3)
            pin.release()
            f = ctx.msm_g1 if grp == z.G1 else ctx.msm_g2
            host_unpinned = best(lambda: f(pts, h_sc), 2 if lg >= 22 else 3)
            emit({"op": "msm_" + name, "log_n": lg, "c_pinned": st["c"], "base_copies": st["base_copies"], "key_image_bytes": int(key_bytes),
                  "pin_s": t_pin, "device_resident_pinned_ms": min(dev), "host_to_host_pinned_ms": host_pinned,
                  "host_to_host_unpinned_ms": host_unpinned, "c_unpinned": z.msm_window_bits(n),
                  "points_per_s_device_resident": n / (min(dev) * 1e-3)})
            del pts
        for flags, name in ((0, "fft"), (1, "ifft"), (2, "coset_fft"), (3, "icoset_fft")):
            a = d_sc.clone()
            cg = g if flags & 2 else None
            ctx.ntt_dev(a.data_ptr(), lg, flags, cg)
            dev = min(ctx.ntt_dev(a.data_ptr(), lg, flags, cg) for _ in range(4))
            buf = h_sc.copy()
            host = best(lambda: ctx.ntt(buf, flags, cg), 2)
            m = n
            emit({"op": "ntt_" + name, "log_m": lg, "device_resident_ms": dev, "host_to_host_ms": host,
                  "hbm_gbs_algorithmic": (192 if lg <= 22 else 288) * m / (dev * 1e-3) / 1e9,
                  "products_per_s": 300 * (m // 2 * lg + (m if flags else 0)) / (dev * 1e-3)})
            del a
        del d_sc, t_sc
    ctx.close()

if args.cpu:
    from oracle import ref as R
    ref = R.Ref()
    cpu = bench.host_cpu()
    gq = ref.fr_from_ints([15])[0]
    for threads in sorted({os.cpu_count(), min(8, os.cpu_count())}, reverse=True):
        ref.set_threads(threads)
        for lg in (16, 18, 20):
            n = 1 << lg
            sc = ref.fr_rng(lg, n)
            for grp, name in ((1, "g1"), (2, "g2")):
                b = ref.bases_seq(n, grp)
                _, t = ref.msm(b, sc, grp)
                emit({"op": "cpu_msm_" + name, "log_n": lg, "threads": threads, "ms": t * 1e3, "cpu": cpu["model"],
                      "impl": "libff::multi_exp<BDLO12_signed, special> (oracle/_ref)"})
                del b
            for flags, name in ((0, "fft"), (3, "icoset_fft")):
                _, t = ref.ntt(sc, flags, gq)
                emit({"op": "cpu_ntt_" + name, "log_m": lg, "threads": threads, "ms": t * 1e3, "cpu": cpu["model"],
                      "impl": "libfqfft::basic_radix2_domain (oracle/_ref)"})
