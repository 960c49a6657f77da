"""Multi-device contexts (zecale_gpu_init with n_devices > 1): pinned MSMs sharded by points over the
devices - libff::multi_exp's chunks (multiexp.tcc:663-685) - must give the reference's normalised point,
through the blocking entry point and through the prover-shaped asynchronous lanes (witness map on
devices[0], H coefficients sliced out to the other devices). With >= 2 GPUs (`gpurun --gpus 2`) the shards sit on
different devices and the partial sums / H slices travel by peer copies; on a one-GPU box the same device is listed
three times (ZECALE_GPU_TEST_REPEAT_DEVICES=1), which still runs shard_bounds -> per-shard Pippenger -> gather ->
sum -> normalise."""
import os

import numpy as np
import pytest

import bw6_constants as C

pytestmark = pytest.mark.gpu


def _device_count():
    import torch
    return torch.cuda.device_count()


@pytest.fixture(scope="module")
def multi():
    import zecale_b200 as z
    n = _device_count()
    if n < 2:
        os.environ["ZECALE_GPU_TEST_REPEAT_DEVICES"] = "1"
        try:
            ctx = z.Context([0, 0, 0])
        finally:
            del os.environ["ZECALE_GPU_TEST_REPEAT_DEVICES"]
        assert ctx.device_count() == 3
    else:
        ctx = z.Context(list(range(min(n, 8))))
        assert ctx.device_count() == min(n, 8)
    yield ctx
    ctx.close()


def test_repeated_devices_need_the_test_switch():
    import zecale_b200 as z
    assert "ZECALE_GPU_TEST_REPEAT_DEVICES" not in os.environ
    with pytest.raises(z.ZecaleGpuError):
        z.Context([0, 0])


@pytest.mark.parametrize("group", [1, 2])
def test_sharded_msm_vs_libff(multi, ref, group):
    n = 1 << 14
    bases = ref.bases_seq(n, group)
    scalars = ref.fr_rng(77 + group, n)
    scalars[5] = 0
    scalars[6] = ref.fr_from_ints([1])[0]
    exp, _ = ref.msm(bases, scalars, group)
    pin = multi.pin_bases(bases, group)
    try:
        assert np.array_equal(multi.msm_pinned(pin, scalars), exp)
        for k in (1, 3, 1000, n - 1):   # prefixes: shards beyond the prefix contribute nothing
            e2, _ = ref.msm(bases[:k], scalars[:k], group)
            assert np.array_equal(multi.msm_pinned(pin, scalars, n=k), e2), k
        idx = np.sort(np.random.default_rng(3).choice(n, 5000, replace=False)).astype(np.uint64)
        e3, _ = ref.msm(bases[:5000], np.ascontiguousarray(scalars[idx]), group)
        assert np.array_equal(multi.msm_pinned(pin, scalars, index=idx, n=5000), e3)
    finally:
        pin.release()


def test_sharded_prover_lanes(multi, ref):
    """The fused prover's calls on a multi-device context: identical to libff / libfqfft."""
    m, n = 1 << 12, 3000
    g = ref.fr_from_ints([C.FR_GENERATOR])[0]
    assert ref.session_synthetic(3000, 10, 400, 5) == 0
    ref.session_setup(4)
    zs = ref.session_export()
    nv, ni, mm = zs["num_variables"], zs["num_inputs"], zs["m"]
    w = zs["assignment"]
    r1cs = multi.r1cs_load(zs["num_constraints"], ni, nv, [zs["csr_a"], zs["csr_b"], zs["csr_c"]])
    pa, ph, pl = multi.pin_bases(zs["A_query"]), multi.pin_bases(zs["H_query"]), multi.pin_bases(zs["L_query"])
    pb2 = multi.pin_bases(zs["B_values"], 2)
    pb1 = multi.pin_bases(zs["B_values"], 1, base_offset_bytes=288)
    try:
        r, s = ref.fr_rng(1, 1)[0], ref.fr_rng(2, 1)[0]
        exp = ref.session_prove(r, s)
        for _ in range(2):
            out = [np.zeros(36, dtype=np.uint64) for _ in range(5)]
            h_out = np.zeros((mm + 1, 6), dtype=np.uint64)
            multi.qap_h_from_witness_async(0, r1cs, w, g, out=h_out)
            multi.msm_pinned_async(1, pa, w, out[0])
            multi.msm_pinned_async(2, pl, w[ni + 1:], out[4])
            multi.msm_pinned_async(3, pb2, w, out[1], index=zs["B_indices"])
            multi.msm_pinned_async(4, pb1, w, out[2], index=zs["B_indices"])
            multi.msm_pinned_async(0, ph, None, out[3], n=mm - 1, from_qap=True)
            multi.flush()
            multi.sync()
            assert np.array_equal(h_out, exp["h"])
            for k in range(5):
                assert np.array_equal(out[k], exp["msm5"][k]), k
    finally:
        for p in (pa, ph, pl, pb2, pb1):
            p.release()
        multi.r1cs_free(r1cs)


def test_two_witness_maps_in_one_epoch(multi, gpu, ref):
    """Two witness maps on two lanes between syncs, each followed by its own H MSM (ZECALE_GPU_SCALARS_FROM_QAP): every
    sharded H MSM must wait for the witness map of ITS lane (a per-lane event), not for whichever was recorded last.
    Expected values: coefficients_for_H of each assignment from the single-device context (itself compared with
    libsnark in test_prove_gpu.py) and libff::multi_exp over the H query with those coefficients."""
    g = ref.fr_from_ints([C.FR_GENERATOR])[0]
    assert ref.session_synthetic(3000, 10, 400, 5) == 0
    ref.session_setup(4)
    zs = ref.session_export()
    nv, ni, mm = zs["num_variables"], zs["num_inputs"], zs["m"]
    w0 = zs["assignment"]
    w1 = w0.copy()
    w1[ni + 1:] = ref.fr_rng(909, nv - ni)   # another (unsatisfying) assignment: the map is defined for any vector
    csr = [zs["csr_a"], zs["csr_b"], zs["csr_c"]]
    exp_h, exp_p = [], []
    r1 = gpu.r1cs_load(zs["num_constraints"], ni, nv, csr)
    try:
        for w in (w0, w1):
            h = np.zeros((mm + 1, 6), dtype=np.uint64)
            gpu.qap_h_from_witness_async(0, r1, w, g, out=h)
            gpu.sync()
            exp_h.append(h)
            exp_p.append(ref.msm(zs["H_query"][:mm - 1], np.ascontiguousarray(h[:mm - 1]), 1)[0])
    finally:
        gpu.r1cs_free(r1)
    assert not np.array_equal(exp_h[0], exp_h[1])
    r1cs = multi.r1cs_load(zs["num_constraints"], ni, nv, csr)
    ph = multi.pin_bases(zs["H_query"])
    try:
        for _ in range(3):
            out = [np.zeros(36, dtype=np.uint64) for _ in range(2)]
            h_out = [np.zeros((mm + 1, 6), dtype=np.uint64) for _ in range(2)]
            multi.qap_h_from_witness_async(0, r1cs, w0, g, out=h_out[0])
            multi.qap_h_from_witness_async(1, r1cs, w1, g, out=h_out[1])
            multi.msm_pinned_async(0, ph, None, out[0], n=mm - 1, from_qap=True)
            multi.msm_pinned_async(1, ph, None, out[1], n=mm - 1, from_qap=True)
            multi.flush()
            multi.sync()
            for k in range(2):
                assert np.array_equal(h_out[k], exp_h[k]), k
                assert np.array_equal(out[k], exp_p[k]), k
    finally:
        ph.release()
        multi.r1cs_free(r1cs)


def test_sharded_handle_is_refused_by_a_single_device_context(multi, gpu, ref):
    """A sharded image owns no device array of its own; a single-device context must refuse it (not fault)."""
    import zecale_b200 as z
    n = 256
    bases = ref.bases_seq(n, 1)
    scalars = ref.fr_rng(11, n)
    pin = multi.pin_bases(bases, 1)
    try:
        foreign = z.PinnedBases(gpu, pin.handle, n, 1)
        with pytest.raises(z.ZecaleGpuError) as e:
            gpu.msm_pinned(foreign, scalars)
        assert e.value.code == 1 and "sharded" in str(e.value)
        with pytest.raises(z.ZecaleGpuError):
            gpu.msm_pinned_async(0, foreign, scalars, np.zeros(36, dtype=np.uint64))
        assert np.array_equal(multi.msm_pinned(pin, scalars), ref.msm(bases, scalars, 1)[0])  # still usable by its maker
    finally:
        pin.release()


def test_kernel_ms_counts_the_workers(multi, ref):
    """zecale_gpu_kernel_ms on a multi-device context adds the shard MSMs the workers timed (class 0 = accumulate)."""
    n = 1 << 12
    bases = ref.bases_seq(n, 1)
    scalars = ref.fr_rng(12, n)
    pin = multi.pin_bases(bases, 1)
    try:
        multi.kernel_ms(0, reset=True)
        multi.msm_pinned(pin, scalars)
        ms, cnt = multi.kernel_ms(0, reset=True)
        assert cnt == multi.device_count() and ms > 0.0
        assert multi.kernel_ms(0)[1] == 0
    finally:
        pin.release()
