"""World-size-2 tests of the multi-GPU host layer (zecale_b200/dist.py) on CPU: gloo process
group, the C oracle as the arithmetic behind the same driver the GPU path uses. Checks the MSM
chunking + all-gather + ordered sum and the 4-step layouts + all-to-all against the golden
fixtures recorded from the unmodified reference."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests.golden import load as G


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, fn):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        globals()[fn](rank, world)
    finally:
        dist.destroy_process_group()


def _run(fn, world=2):
    mp.spawn(_worker, args=(world, _free_port(), fn), nprocs=world, join=True)


def _exchange(dst, src):
    dist.all_to_all_single(dst.view(-1), src.view(-1))


def _gather_layout(local, idx, m, world):
    """Assemble the global vector from every rank's (local rows, global indices)."""
    rows = [torch.empty_like(local) for _ in range(world)]
    ids = [torch.empty_like(idx) for _ in range(world)]
    dist.all_gather(rows, local)
    dist.all_gather(ids, idx)
    out = np.zeros((m, 6), dtype=np.uint64)
    for r, i in zip(rows, ids):
        out[i.numpy()] = r.numpy().view(np.uint64)
    return out


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64).copy())


def _dist_ntt(rank, world):
    from tests import dist_cpu_backend as B
    from zecale_b200 import dist as D
    g = G.load("ntt")
    log_m, m = 8, 256
    drv = D.DistNtt(B.phase_fn, _exchange, world, rank)
    ia, ib = D.layout_a_indices(log_m, world, rank), D.layout_b_indices(log_m, world, rank)
    # every index is owned exactly once in either layout
    own = _gather_layout(_t(np.zeros((ia.size, 6), dtype=np.uint64)), torch.from_numpy(ia), m, world)
    assert own.shape[0] == m and sorted(np.concatenate([D.layout_a_indices(log_m, world, r) for r in range(world)])) == list(range(m))
    assert sorted(np.concatenate([D.layout_b_indices(log_m, world, r) for r in range(world)])) == list(range(m))
    x = g["in_8"]
    for flags in range(4):
        src_idx, dst_idx = (ib, ia) if flags & 1 else (ia, ib)
        buf = _t(x[src_idx])
        scratch = torch.empty_like(buf)
        res, _ = drv.transform(buf, scratch, log_m, flags, g["coset_g"] if flags & 2 else None)
        got = _gather_layout(res, torch.from_numpy(dst_idx), m, world)
        assert np.array_equal(got, g["out_8_%d" % flags]), flags


def _dist_qap_h(rank, world):
    from tests import dist_cpu_backend as B
    from zecale_b200 import dist as D
    q = G.load("qap_h")
    log_m, m = 8, 256
    drv = D.DistNtt(B.phase_fn, _exchange, world, rank)
    ia, ib = D.layout_a_indices(log_m, world, rank), D.layout_b_indices(log_m, world, rank)
    aA, aB, aC = (_t(q[k][ib]) for k in ("aA", "aB", "aC"))
    scratch = torch.empty_like(aA)
    h = drv.qap_h(aA, aB, aC, scratch, log_m, q["coset_g"], B.pointwise_fn(log_m, q["coset_g"]))
    got = _gather_layout(h, torch.from_numpy(ia), m, world)
    assert np.array_equal(got, q["h"][:m])
    assert not q["h"][m].any()


def _dist_msm(rank, world):
    from oracle import port as P
    from zecale_b200 import dist as D
    port = P.Port()
    g = G.load("msm_g1_small")
    n = g["bases"].shape[0]
    lo, hi = D.msm_chunk_range(n, world, rank)
    assert (lo, hi) == ((0, n // 2) if rank == 0 else (n // 2, n))

    def local():
        return port.msm(g["bases"][lo:hi], g["scalars"][lo:hi])

    def psum(points):
        acc = points[0:1].copy()
        for i in range(1, points.shape[0]):
            acc = port.group_op(0, acc, points[i:i + 1])
        return acc[0]

    got = D.msm_sharded(local, psum, world)
    assert np.array_equal(got, g["result"])
    # sparse (kc) walk: shard the index list, scalars stay whole
    k = G.load("msm_kc_sparse")
    nnz = k["values"].shape[0]
    lo, hi = D.msm_chunk_range(nnz, world, rank)
    got = D.msm_sharded(lambda: port.msm(k["values"][lo:hi], k["scalars"], index=k["indices"][lo:hi]), psum, world)
    assert np.array_equal(got, k["result"][:36])
    # fewer points than ranks: rank 0 takes everything (multiexp.tcc:658-662)
    assert D.msm_chunk_range(1, world, 0) == (0, 1) and D.msm_chunk_range(1, world, 1) == (1, 1)


def test_msm_chunk_ranges_cover():
    from zecale_b200 import dist as D
    for n in (0, 1, 7, 8, 9, 1000, 237801):
        for world in (1, 2, 4, 8):
            r = [D.msm_chunk_range(n, world, k) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n or n < world
            assert sum(b - a for a, b in r) == n
            for k in range(world - 1):
                assert r[k][1] == r[k + 1][0] or n < world


def test_four_step_shape_errors():
    from zecale_b200 import dist as D
    with pytest.raises(ValueError):
        D.four_step_shape(8, 3)
    with pytest.raises(ValueError):
        D.four_step_shape(4, 4)
    assert D.four_step_shape(24, 8) == {"l1": 12, "l2": 12, "m1": 4096, "m2": 4096, "c1": 512, "c2": 512}


def test_dist_ntt_world2_gloo():
    _run("_dist_ntt")


def test_dist_qap_h_world2_gloo():
    _run("_dist_qap_h")


def test_dist_msm_world2_gloo():
    _run("_dist_msm")
