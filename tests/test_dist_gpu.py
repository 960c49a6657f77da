"""The distributed 4-step NTT kernels and the sharded MSM building blocks on ONE GPU: the ranks of a
world of 1, 2, 4 or 8 are emulated one after the other on the same device (separate buffers; the
all-to-all is a set of tensor copies), which exercises exactly the index maps, twiddle offsets and
fused scalings the multi-GPU run uses. The NCCL exchange itself is covered by scripts/dist_check.py
(torchrun, --gpus N) and by the gloo tests."""
import numpy as np
import pytest
import torch

import bw6_constants as C
from zecale_b200 import dist as D

pytestmark = pytest.mark.gpu


def _emulated(gpu, x, log_m, flags, g, world):
    """x: global (m, 6) uint64. Returns the global result assembled from the ranks' output layout."""
    m = 1 << log_m
    ia = [D.layout_a_indices(log_m, world, r) for r in range(world)]
    ib = [D.layout_b_indices(log_m, world, r) for r in range(world)]
    src_idx, dst_idx = (ib, ia) if flags & 1 else (ia, ib)
    n_loc = m // world
    bufs = [torch.from_numpy(x[src_idx[r]].view(np.int64).copy()).cuda() for r in range(world)]
    send = [torch.empty_like(b) for b in bufs]
    for r in range(world):
        gpu.ntt_dist_phase(bufs[r].data_ptr(), send[r].data_ptr(), log_m, flags, g, world, r, 0)
    chunk = n_loc // world
    recv = [torch.empty_like(b) for b in bufs]
    for r in range(world):          # all-to-all: chunk s of recv[r] = chunk r of send[s]
        for s in range(world):
            recv[r][s * chunk:(s + 1) * chunk] = send[s][r * chunk:(r + 1) * chunk]
    torch.cuda.synchronize()
    out = np.zeros((m, 6), dtype=np.uint64)
    for r in range(world):
        gpu.ntt_dist_phase(recv[r].data_ptr(), bufs[r].data_ptr(), log_m, flags, g, world, r, 1)
        out[dst_idx[r]] = bufs[r].cpu().numpy().view(np.uint64)
    return out


@pytest.mark.parametrize("log_m,world", [(8, 1), (8, 2), (13, 4), (16, 8), (18, 2)])
def test_dist_ntt_matches_reference(gpu, ref, log_m, world):
    x = ref.fr_rng(17 * log_m + world, 1 << log_m)
    g = ref.fr_from_ints([C.FR_GENERATOR])[0]
    for flags in range(4):
        exp, _ = ref.ntt(x, flags, g)
        got = _emulated(gpu, x, log_m, flags, g if flags & 2 else None, world)
        assert np.array_equal(got, exp), (log_m, world, flags)


@pytest.mark.parametrize("log_m,world", [(22, 8), (24, 8)])
def test_dist_ntt_large_matches_libfqfft(gpu, ref, log_m, world):
    """BASELINE config 5 sizes: the 8-rank 4-step result equals libfqfft's basic_radix2_domain
    (basic_radix2_domain.tcc:55-88) bit for bit."""
    m = 1 << log_m
    x = ref.fr_rng(0xD157 + log_m, m)
    g = ref.fr_from_ints([C.FR_GENERATOR])[0]
    for flags in (0, 3):
        exp, _ = ref.ntt(x, flags, g)
        got = _emulated(gpu, x, log_m, flags, g if flags & 2 else None, world)
        assert np.array_equal(got, exp), (log_m, flags)
        del exp, got


def test_point_sum_and_sharded_msm(gpu, ref):
    n, world = 5000, 4
    bases = ref.bases_mul(ref.fr_rng(3, n), 1)
    scalars = ref.fr_rng(4, n)
    exp, _ = ref.msm(bases, scalars, 1)
    parts = []
    for r in range(world):
        lo, hi = D.msm_chunk_range(n, world, r)
        parts.append(gpu.msm_g1(bases[lo:hi], scalars[lo:hi]))
    assert np.array_equal(gpu.point_sum(np.stack(parts)), exp)
    # projective (non-special) inputs, infinity among them, and the empty sum
    dbl = ref.group_op(2, bases[:6], None, 1)
    inf = np.zeros(36, dtype=np.uint64)
    inf[12:24] = bases[0][24:36]
    pts = np.concatenate([dbl, inf[None, :]])
    acc = inf[None, :].copy()
    for i in range(pts.shape[0]):
        acc = ref.group_op(0, acc, pts[i:i + 1], 1)
    assert np.array_equal(gpu.point_sum(pts), ref.group_op(3, acc, None, 1)[0])
    assert np.array_equal(gpu.point_sum(pts[:0]), inf)


def test_dist_qap_h_emulated(gpu, ref):
    """The H polynomial through the distributed pipeline (world 2 emulated) equals libfqfft's."""
    log_m, world = 12, 2
    m = 1 << log_m
    g = ref.fr_from_ints([C.FR_GENERATOR])[0]
    vec = [ref.fr_rng(31 * k, m) for k in (1, 2, 3)]
    exp, _ = ref.qap_h(*vec)
    # iFFT (B->A), cosetFFT (A->B) per vector, pointwise per rank in layout B, icosetFFT (B->A)
    cur = []
    for v in vec:
        t = _emulated(gpu, v, log_m, 1, None, world)
        cur.append(_emulated(gpu, t, log_m, 2, g, world))
    d = [torch.from_numpy(c.view(np.int64).copy()).cuda() for c in cur]
    gpu.qap_pointwise_dev(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), log_m, g, m)
    h = _emulated(gpu, d[0].cpu().numpy().view(np.uint64), log_m, 3, g, world)
    assert np.array_equal(h, exp[:m])
