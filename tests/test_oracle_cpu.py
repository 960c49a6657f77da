"""The plain-C restatement (oracle/zecale_oracle.c) against the golden fixtures produced by the
unmodified reference (tests/golden/make_golden.py), and — where oracle/_ref is built — against
the reference itself on fresh random inputs. Pins the oracle (no GPU needed)."""
import numpy as np
import pytest

import bw6_constants as C
from oracle import port as P
from tests.golden import load as G


@pytest.fixture(scope="module")
def port():
    if not P.available():
        import subprocess, os
        subprocess.run(["make", "-C", os.path.dirname(P.PORT_SO), "port"], check=True, capture_output=True)
    return P.Port()


FIELD_OPS = ["add", "sub", "mul", "sqr", "inv", "neg", "from_mont"]


@pytest.mark.parametrize("name", ["fq", "fr"])
def test_field_ops_golden(port, name):
    g = G.load("field_" + name)
    op = port.fq_op if name == "fq" else port.fr_op
    for k, nm in enumerate(FIELD_OPS):
        assert np.array_equal(op(k, g["a"], g["b"]), g[nm]), nm


def test_signed_digits_golden(port):
    g = G.load("signed_digits")
    for i in range(g["plain"].shape[0]):
        for j, c in enumerate(g["cs"].tolist()):
            for k in range((379 + c - 1) // c):  # the rounds multi_exp_inner can ask for (multiexp.tcc:584-586)
                assert port.signed_digit(g["plain"][i], c, k) == g["digits"][i, j, k], (i, c, k)
    for n, c in g["optimal_c"].tolist():
        assert port.optimal_c(n) == c


@pytest.mark.parametrize("grp", [1, 2])
def test_group_ops_golden(port, grp):
    g = G.load("group_g%d" % grp)
    for k, nm in ((0, "add"), (1, "mixed_add"), (2, "dbl"), (3, "to_special"), (4, "neg")):
        assert np.array_equal(port.group_op(k, g["a"], g["b"]), g[nm]), nm
    assert np.array_equal(port.group_op(1, g["a_proj"], g["b"]), g["proj_mixed_add"])
    assert np.array_equal(port.group_op(3, g["a_proj"]), g["dbl"])


@pytest.mark.parametrize("grp", [1, 2])
@pytest.mark.parametrize("chunks", [1, 3, 16])
def test_msm_golden(port, grp, chunks):
    g = G.load("msm_g%d_small" % grp)
    assert np.array_equal(port.msm(g["bases"], g["scalars"], chunks=chunks), g["result"])
    assert np.array_equal(port.msm(g["bases"][:7], g["scalars"][:7], chunks=chunks), g["result_first7"])


def test_msm_empty(port):
    g = G.load("msm_g1_small")
    z = np.zeros(36, dtype=np.uint64)
    z[12:24] = g["bases"][0][24:36]
    assert np.array_equal(port.msm(g["bases"], g["scalars"], n=0), z)


def test_msm_kc_sparse_golden(port):
    g = G.load("msm_kc_sparse")
    for chunks in (1, 4):
        g2 = port.msm(g["values"], g["scalars"], index=g["indices"], chunks=chunks)
        g1 = port.msm(g["values"], g["scalars"], index=g["indices"], chunks=chunks, base_offset_bytes=288)
        assert np.array_equal(g2, g["result"][:36])
        assert np.array_equal(g1, g["result"][36:])


def test_ntt_golden(port):
    g = G.load("ntt")
    for lg in (1, 4, 8, 11):
        for fl in range(4):
            got = port.ntt(g["in_%d" % lg], fl, g["coset_g"] if fl & 2 else None)
            assert np.array_equal(got, g["out_%d_%d" % (lg, fl)]), (lg, fl)


def test_ntt_roundtrips(port):
    g = G.load("ntt")
    a = g["in_8"]
    assert np.array_equal(port.ntt(port.ntt(a, 0), 1), a)
    assert np.array_equal(port.ntt(port.ntt(a, 2, g["coset_g"]), 3, g["coset_g"]), a)


def test_qap_h_golden(port):
    g = G.load("qap_h")
    assert np.array_equal(port.qap_h(g["aA"], g["aB"], g["aC"], g["coset_g"]), g["h"])


def test_groth16_small_golden(port):
    """H polynomial and the five MSMs of a complete small Groth16 prove (fixed r, s)."""
    g = G.load("groth16_small")
    nv, ni, m = g["sizes"].tolist()
    w = g["assignment"]
    assert np.array_equal(port.qap_h(g["aA"], g["aB"], g["aC"], g["coset_g"]), g["h"])
    got = [port.msm(g["A_query"][: nv + 1], w, chunks=2),
           port.msm(g["B_values"], w, index=g["B_indices"]),
           port.msm(g["B_values"], w, index=g["B_indices"], base_offset_bytes=288),
           port.msm(g["H_query"][: m - 1], g["h"][: m - 1], chunks=4),
           port.msm(g["L_query"], np.ascontiguousarray(w[ni + 1:]))]
    for k in range(5):
        assert np.array_equal(got[k], g["msm5"][k]), k


# ---- against the reference itself, fresh inputs (only where oracle/_ref is built)
def test_field_vs_ref_random(port, ref):
    for name, p, seed in (("fq", C.Q, 5), ("fr", C.R, 6)):
        n64 = 12 if name == "fq" else 6
        rng = np.random.default_rng(seed)
        from oracle.ref import ints_to_limbs
        vals = [int.from_bytes(rng.bytes(8 * n64), "little") % p for _ in range(2000)]
        rop, pop = (ref.fq_op, port.fq_op) if name == "fq" else (ref.fr_op, port.fr_op)
        a = rop(7, ints_to_limbs(vals, n64))
        b = rop(7, ints_to_limbs(vals[1:] + vals[:1], n64))
        assert np.array_equal(pop(7, ints_to_limbs(vals, n64)), a)
        for k in (0, 1, 2, 3, 5, 6):
            assert np.array_equal(pop(k, a, b), rop(k, a, b)), (name, k)
        assert np.array_equal(pop(4, a[:50]), rop(4, a[:50]))


@pytest.mark.parametrize("grp", [1, 2])
def test_msm_vs_ref_random(port, ref, grp):
    n = 1500
    bases = ref.bases_mul(ref.fr_rng(99 + grp, n), grp)
    scalars = ref.fr_rng(7000 + grp, n)
    exp, _ = ref.msm(bases, scalars, grp)
    assert np.array_equal(port.msm(bases, scalars, chunks=ref.threads), exp)


def test_ntt_vs_ref_random(port, ref):
    g15 = ref.fr_from_ints([C.FR_GENERATOR])[0]
    a = ref.fr_rng(31337, 1 << 12)
    for fl in range(4):
        exp, _ = ref.ntt(a, fl, g15 if fl & 2 else None)
        assert np.array_equal(port.ntt(a, fl, g15 if fl & 2 else None), exp), fl


@pytest.mark.parametrize("grp", [1, 2])
def test_batch_exp_golden(port, grp):
    """SURVEY 8(f3) on the CPU: (coeff * s_i) * g restated as one-point sums of the C oracle, against
    libff::batch_exp(_with_coeff) + batch_to_special recorded by tests/golden/make_golden_setup.py."""
    g = G.load("setup_batch_exp")
    base, sc, coeff = g["base_g%d" % grp][None, :], g["scalars_g%d" % grp], g["coeff_g%d" % grp]
    scaled = port.fr_op(2, sc, np.repeat(coeff, sc.shape[0], axis=0))      # Montgomery product in Fr
    for i in range(sc.shape[0]):
        assert np.array_equal(port.msm(base, sc[i:i + 1]), g["exp_g%d" % grp][i]), i
        assert np.array_equal(port.msm(base, scaled[i:i + 1]), g["exp_coeff_g%d" % grp][i]), i


def _decode_points(port, raw, offset, stride, n, one):
    """Key-file points (X || Y, plain big-endian, libzeth group_element_write_bytes) -> libff special form:
    the CPU restatement of k_import_bytes (byte reversal, to-Montgomery, (0, 1) = zero)."""
    pts = np.zeros((n, 36), dtype=np.uint64)
    for i in range(n):
        rec = raw[offset + i * stride: offset + i * stride + 192]
        x = np.frombuffer(rec[:96][::-1].tobytes(), dtype=np.uint64)       # little-endian limbs of the plain integer
        y = np.frombuffer(rec[96:][::-1].tobytes(), dtype=np.uint64)
        if not x.any() and y[0] == 1 and not y[1:].any():
            pts[i, 12:24] = one
            continue
        pts[i, :12] = port.fq_op(7, x[None, :].copy())[0]
        pts[i, 12:24] = port.fq_op(7, y[None, :].copy())[0]
        pts[i, 24:36] = one
    return pts


def test_keyfile_bytes_golden(port):
    """SURVEY 8(f4) on the CPU: bytes written by libzeth's codec decode to the in-memory points, and the
    oracle's MSM over the decoded points equals libff::multi_exp over the originals."""
    g = G.load("keyfile_bytes")
    one = g["bases_g1"][0, 24:36]
    for grp in (1, 2):
        raw = g["bytes_g%d" % grp]
        n = int(raw[:8].view(np.uint64)[0])
        pts = _decode_points(port, raw, 8, 192, n, one)
        assert np.array_equal(pts, g["bases_g%d" % grp])
        assert np.array_equal(port.msm(pts, g["scalars_g%d" % grp]), g["msm_g%d" % grp])
    raw = g["kc_bytes"]
    nnz = int(raw[8:16].view(np.uint64)[0])
    idx = np.array([int(raw[16 + 392 * i: 24 + 392 * i].view(np.uint64)[0]) for i in range(nnz)], dtype=np.uint64)
    assert np.array_equal(idx, g["kc_indices"])
    sc = np.ascontiguousarray(g["kc_witness"][idx.astype(np.int64)])
    for grp, off in ((2, 24), (1, 216)):
        pts = _decode_points(port, raw, off, 392, nnz, one)
        assert np.array_equal(port.msm(pts, sc), g["kc_msm_g%d" % grp])
