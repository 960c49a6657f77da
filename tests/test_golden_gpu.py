"""CUDA path (through the C ABI) against the committed golden fixtures — the outputs of the
unmodified reference recorded by tests/golden/make_golden.py. Needs no oracle library on the box."""
import numpy as np
import pytest

from tests.golden import load as G

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["fq", "fr"])
def test_field_ops(gpu, name):
    g = G.load("field_" + name)
    op = gpu.test_fq if name == "fq" else gpu.test_fr
    for k, nm in enumerate(["add", "sub", "mul", "sqr", "inv", "neg", "from_mont"]):
        assert np.array_equal(op(k, g["a"], g["b"]), g[nm]), nm


@pytest.mark.parametrize("grp", [1, 2])
def test_group_ops(gpu, grp):
    g = G.load("group_g%d" % grp)
    for k, nm in ((0, "add"), (1, "mixed_add"), (2, "dbl"), (3, "to_special")):
        assert np.array_equal(gpu.test_group(k, g["a"], g["b"]), g[nm]), nm
    assert np.array_equal(gpu.test_group(1, g["a_proj"], g["b"]), g["proj_mixed_add"])
    assert np.array_equal(gpu.test_group(0, g["a_proj"], g["b"]), g["proj_mixed_add"])


@pytest.mark.parametrize("grp", [1, 2])
def test_msm(gpu, grp):
    g = G.load("msm_g%d_small" % grp)
    f = gpu.msm_g1 if grp == 1 else gpu.msm_g2
    assert np.array_equal(f(g["bases"], g["scalars"]), g["result"])
    assert np.array_equal(f(g["bases"], g["scalars"], n=7), g["result_first7"])
    pin = gpu.pin_bases(g["bases"], grp)
    try:
        assert np.array_equal(gpu.msm_pinned(pin, g["scalars"]), g["result"])
        assert np.array_equal(gpu.msm_pinned(pin, g["scalars"], n=7), g["result_first7"])
    finally:
        pin.release()


def test_msm_kc_sparse(gpu):
    g = G.load("msm_kc_sparse")
    assert np.array_equal(gpu.msm_g2(g["values"], g["scalars"], index=g["indices"]), g["result"][:36])
    assert np.array_equal(gpu.msm_g1(g["values"], g["scalars"], index=g["indices"], base_offset_bytes=288), g["result"][36:])


def test_ntt(gpu):
    g = G.load("ntt")
    for lg in (1, 4, 8, 11):
        for fl in range(4):
            got = gpu.ntt(g["in_%d" % lg].copy(), fl, g["coset_g"] if fl & 2 else None)
            assert np.array_equal(got, g["out_%d_%d" % (lg, fl)]), (lg, fl)


def test_qap_h(gpu):
    g = G.load("qap_h")
    assert np.array_equal(gpu.qap_h(g["aA"], g["aB"], g["aC"], g["coset_g"]), g["h"])


def test_groth16_small(gpu):
    """Every device-side result of a complete small prove equals the recorded libsnark values."""
    g = G.load("groth16_small")
    nv, ni, m = g["sizes"].tolist()
    w = g["assignment"]
    h = gpu.qap_h(g["aA"], g["aB"], g["aC"], g["coset_g"])
    assert np.array_equal(h, g["h"])
    got = [gpu.msm_g1(g["A_query"][: nv + 1], w),
           gpu.msm_g2(g["B_values"], w, index=g["B_indices"]),
           gpu.msm_g1(g["B_values"], w, index=g["B_indices"], base_offset_bytes=288),
           gpu.msm_g1(g["H_query"][: m - 1], h[: m - 1]),
           gpu.msm_g1(g["L_query"], np.ascontiguousarray(w[ni + 1:]))]
    for k in range(5):
        assert np.array_equal(got[k], g["msm5"][k]), k


@pytest.mark.parametrize("grp", [1, 2])
def test_batch_exp(gpu, grp):
    """libff::batch_exp(_with_coeff) + batch_to_special, recorded by make_golden_setup.py."""
    g = G.load("setup_batch_exp")
    base, sc = g["base_g%d" % grp], g["scalars_g%d" % grp]
    assert np.array_equal(gpu.batch_exp(base, sc, grp), g["exp_g%d" % grp])
    assert np.array_equal(gpu.batch_exp(base, sc, grp, coeff=g["coeff_g%d" % grp]), g["exp_coeff_g%d" % grp])


def test_keyfile_bytes(gpu):
    """key-file bytes written by libzeth's codec -> device image; MSMs equal libff::multi_exp."""
    g = G.load("keyfile_bytes")
    for grp in (1, 2):
        raw, sc = np.ascontiguousarray(g["bytes_g%d" % grp]), g["scalars_g%d" % grp]
        pin = gpu.pin_bases_bytes(raw, 8, 192, sc.shape[0], grp)
        try:
            assert np.array_equal(gpu.msm_pinned(pin, sc), g["msm_g%d" % grp])
        finally:
            pin.release()
    raw = np.ascontiguousarray(g["kc_bytes"])
    sc = np.ascontiguousarray(g["kc_witness"][g["kc_indices"].astype(np.int64)])
    for grp, off in ((2, 24), (1, 216)):
        pin = gpu.pin_bases_bytes(raw, off, 392, sc.shape[0], grp)
        try:
            assert np.array_equal(gpu.msm_pinned(pin, sc), g["kc_msm_g%d" % grp])
        finally:
            pin.release()
