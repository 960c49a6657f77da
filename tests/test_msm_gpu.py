"""zecale_gpu_msm_g1/g2 through the C ABI vs libff::multi_exp (BDLO12_signed, special bases).

Mirrors libff's test_multiexp.cpp (sizes 1..257, :204-283; edge case :337-383) for BW6-761,
which the reference itself never instantiates."""
import numpy as np
import pytest

import bw6_constants as C
from oracle.ref import ints_to_limbs

pytestmark = pytest.mark.gpu


@pytest.fixture(params=["xyzz", "tree"], autouse=True)
def acc_mode(request, monkeypatch):
    """Every MSM case runs under both bucket-accumulation kernels (XYZZ chains = default; batched affine
    pair tree = ZECALE_GPU_ACC=tree, read by the library at each call)."""
    monkeypatch.setenv("ZECALE_GPU_ACC", request.param)
    return request.param


def _inf(one_limbs):
    z = np.zeros(36, dtype=np.uint64)
    z[12:24] = one_limbs
    return z


@pytest.mark.parametrize("group", [1, 2])
@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 256, 257, 4096])
def test_msm_sizes(gpu, ref, group, n):
    bases = ref.bases_seq(n, group)
    scalars = ref.fr_rng(1000 * n, n)
    exp, _ = ref.msm(bases, scalars, group)
    f = gpu.msm_g1 if group == 1 else gpu.msm_g2
    assert np.array_equal(f(bases, scalars), exp)


def test_msm_empty(gpu, ref):
    bases = ref.bases_seq(1, 1)
    got = gpu.msm_g1(bases[:0], ref.fr_rng(0, 1)[:0], n=0)
    assert np.array_equal(got, _inf(bases[0][24:36]))


@pytest.mark.parametrize("group", [1, 2])
def test_msm_edge_cases(gpu, ref, group):
    n = 600
    bases = ref.bases_mul(ref.fr_rng(77, n), group)
    one = bases[0][24:36].copy()
    r = C.R
    vals = [0, 1, r - 1, r - 2, 2, (1 << 376), (1 << 377) - 1 - ((1 << 377) - 1 - (r - 1)), (r - 1) // 2, (r + 1) // 2]
    # top window digit negative + carry pattern (test_multiexp.cpp:337-383): all windows = 2^(c-1) for several c
    for c in (13, 14, 15, 16):
        v = sum((1 << (c - 1)) << (c * k) for k in range(377 // c + 1)) % r
        vals.append(v)
        vals.append(sum(((1 << c) - 1) << (c * k) for k in range(377 // c + 1)) % r)
    scal_int = [vals[i % len(vals)] if i < 200 else None for i in range(n)]
    rnd = ref.fr_rng(4242, n)
    scalars = rnd.copy()
    fixed = ref.fr_from_ints([v for v in scal_int if v is not None])
    scalars[: fixed.shape[0]] = fixed
    # zero points, duplicates, P / -P pairs with equal scalars (bucket P + (-P) and P + P)
    bases[5] = _inf(one)
    bases[6] = _inf(one)
    bases[300] = bases[301]
    scalars[300] = scalars[301]
    bases[302] = ref.group_op(4, bases[303:304], None, group)[0]
    scalars[302] = scalars[303]
    bases[310:330] = bases[310]
    scalars[310:330] = scalars[310]
    exp, _ = ref.msm(bases, scalars, group)
    f = gpu.msm_g1 if group == 1 else gpu.msm_g2
    assert np.array_equal(f(bases, scalars), exp)
    # filter_one_zero front-end of the prover gives the same element (multiexp.tcc:695-757)
    exp2, _ = ref.msm(bases, scalars, group, method=1)
    assert np.array_equal(exp2, exp)


@pytest.mark.parametrize("group", [1, 2])
@pytest.mark.parametrize("shape", ["PP", "P-P", "P-PQ", "PPP", "PQ-Q", "infPP"])
def test_msm_first_pair_special_cases(gpu, ref, group, shape):
    """The first two points of a bucket chain take the affine-accumulator addition (VM program madd_aff): equal points
    (-> doubling), opposite points (-> infinity, then the chain goes on), three equal points, a base at infinity in front.
    Equal scalars put the points into the same bucket of every window (cf. bw6_761_g1.cpp:160-358 special cases)."""
    pts = ref.bases_mul(ref.fr_rng(606, 2), group)
    P, Q = pts[0], pts[1]
    neg = lambda x: ref.group_op(4, x[None, :], None, group)[0]
    rows = {"PP": [P, P], "P-P": [P, neg(P)], "P-PQ": [P, neg(P), Q], "PPP": [P, P, P], "PQ-Q": [P, Q, neg(Q)],
            "infPP": [_inf(P[24:36]), P, P]}[shape]
    bases = np.stack(rows)
    s = ref.fr_rng(909, 1)
    scalars = np.repeat(s, len(rows), axis=0)
    exp, _ = ref.msm(bases, scalars, group)
    f = gpu.msm_g1 if group == 1 else gpu.msm_g2
    assert np.array_equal(f(bases, scalars), exp)
    pin = gpu.pin_bases(bases, group=group)
    try:
        assert np.array_equal(gpu.msm_pinned(pin, scalars), exp)
    finally:
        pin.release()


def test_msm_heavy_buckets(gpu, ref):
    """Aggregator-like scalars: many zeros / ones / small values -> very uneven buckets."""
    n = 5000
    bases = ref.bases_mul(ref.fr_rng(9, n), 1)
    scalars = ref.fr_rng(31337, n)
    small = ref.fr_from_ints([0, 1, 1, 1, 2, 3, 65535, 1, 0, 1])
    for i in range(0, 3000):
        scalars[i] = small[i % 10]
    exp, _ = ref.msm(bases, scalars, 1, method=1)
    assert np.array_equal(gpu.msm_g1(bases, scalars), exp)


def test_msm_kc_sparse(gpu, ref):
    """B-query shape: sparse (G2,G1) pairs, stride 576, scalars picked through indices
    (kc_multi_exp_with_mixed_addition, kc_multiexp.tcc:31-113)."""
    n_scalars, nnz = 3000, 1100
    rng = np.random.default_rng(3)
    idx = np.sort(rng.choice(n_scalars, nnz, replace=False)).astype(np.uint64)
    s = ref.fr_rng(555, nnz)
    vals = np.zeros((nnz, 72), dtype=np.uint64)
    vals[:, :36] = ref.bases_mul(s, 2)
    vals[:, 36:] = ref.bases_mul(ref.fr_rng(556, nnz), 1)
    scalars = ref.fr_rng(8080, n_scalars)
    scalars[idx[3]] = 0
    scalars[idx[4]] = ref.fr_from_ints([1])[0]
    exp, _ = ref.msm_kc(vals, idx, n_scalars, scalars)
    got_g2 = gpu.msm_g2(vals, scalars, index=idx)
    got_g1 = gpu.msm_g1(vals, scalars, index=idx, base_offset_bytes=288)
    assert np.array_equal(got_g2, exp[:36])
    assert np.array_equal(got_g1, exp[36:])


def test_msm_pinned_and_2_16(gpu, ref):
    n = 1 << 16
    bases = ref.bases_seq(n, 1)
    scalars = ref.fr_rng(1 << 32, n)
    exp, t = ref.msm(bases, scalars, 1)
    pinned = gpu.pin_bases(bases)
    try:
        assert np.array_equal(gpu.msm_pinned(pinned, scalars), exp)
        # prefix of the pinned set
        exp2, _ = ref.msm(bases[:1000], scalars[:1000], 1)
        assert np.array_equal(gpu.msm_pinned(pinned, scalars, n=1000), exp2)
    finally:
        pinned.release()


def test_msm_rejects_non_special(gpu, ref):
    import zecale_b200 as z
    bases = ref.bases_seq(8, 1)
    bad = ref.group_op(2, bases, None, 1)  # doubled: Z != one
    with pytest.raises(z.ZecaleGpuError) as e:
        gpu.msm_g1(bad, ref.fr_rng(1, 8))
    assert e.value.code == 3


def test_async_lanes_match_blocking_calls(gpu, ref):
    """The prover-shaped use of the lanes: H polynomial + H MSM on one lane, witness MSMs on others
    (shared witness uploaded once), results identical to the blocking calls and to libff."""
    import bw6_constants as C
    m, n = 1 << 12, 3000
    g = ref.fr_from_ints([C.FR_GENERATOR])[0]
    aA, aB, aC = (ref.fr_rng(31 * k, m) for k in (1, 2, 3))
    w = ref.fr_rng(99, n)
    bases = {k: ref.bases_mul(ref.fr_rng(1000 + k, n if k < 2 else m - 1), 1) for k in range(3)}
    pins = {k: gpu.pin_bases(v) for k, v in bases.items()}
    idx = np.sort(np.random.default_rng(1).choice(n, 700, replace=False)).astype(np.uint64)
    try:
        h_exp, _ = ref.qap_h(aA, aB, aC)
        exp = [ref.msm(bases[0], w, 1)[0], ref.msm(bases[1][10:], w[10:], 1)[0],
               ref.msm(bases[2], h_exp[: m - 1], 1)[0], ref.msm(bases[0][:700], np.ascontiguousarray(w[idx]), 1)[0]]
        out = [np.zeros(36, dtype=np.uint64) for _ in range(4)]
        h_out = np.zeros((m + 1, 6), dtype=np.uint64)
        for _ in range(2):  # second round reuses the staging buffers
            gpu.qap_h_async(0, aA, aB, aC, g, out=h_out)
            gpu.msm_pinned_async(1, pins[0], w, out[0])
            p1 = gpu.pin_bases(bases[1][10:])
            gpu.msm_pinned_async(2, p1, w[10:], out[1])
            gpu.msm_pinned_async(3, pins[0], w, out[3], n=700, index=idx)
            gpu.msm_pinned_async(0, pins[2], None, out[2], from_qap=True)
            gpu.sync()
            p1.release()
            assert np.array_equal(h_out, h_exp)
            for k in range(4):
                assert np.array_equal(out[k], exp[k]), k
    finally:
        for p in pins.values():
            p.release()


@pytest.mark.parametrize("group,log_n", [(1, 20), (2, 18)])
def test_msm_full_size_properties(gpu, ref, group, log_n):
    """BASELINE config 2 / the headline MSM size (2^20), where libff takes too long to be the checker:
    size-independent properties instead — linearity in the scalars, msm(a) + msm(b) == msm(a + b), and
    additivity over a split of the points (pinned prefix + un-pinned tail == whole), with the group
    additions done by the reference on the CPU."""
    n = 1 << log_n
    bases = gpu.synth_bases(n, group, seed=2024 + group)
    a, b = ref.fr_rng(5 << 32, n), ref.fr_rng(6 << 32, n)
    c = ref.fr_op(0, a, b)

    def add(p, q):
        s = ref.group_op(0, p[None, :].copy(), q[None, :].copy(), group)
        return ref.group_op(3, s, None, group)[0]

    pinned = gpu.pin_bases(bases, group)
    try:
        pa, pb, pc = (gpu.msm_pinned(pinned, s) for s in (a, b, c))
        assert np.array_equal(add(pa, pb), pc)
        half = n // 2 + 12345
        head = gpu.msm_pinned(pinned, a, n=half)
    finally:
        pinned.release()
    f = gpu.msm_g1 if group == 1 else gpu.msm_g2
    tail = f(np.ascontiguousarray(bases[half:]), np.ascontiguousarray(a[half:]))
    assert np.array_equal(add(head, tail), pa)


@pytest.mark.parametrize("group,log_n", [(1, 18), (2, 18), (1, 20), (2, 20)])
def test_msm_full_size_vs_libff(gpu, ref, group, log_n):
    """BASELINE config 2 at 2^18 and at the headline size 2^20, G1 and G2, bit-exact against
    libff::multi_exp<BDLO12_signed, special> (multiexp.tcc:648-688) on the reference's own bases
    P_i = (i+1) G and SHA512_rng scalars: the pinned (expanded image, one bucket set) and the un-pinned
    (one bucket set per window + doubling chain) entry points must both give libff's normalised point."""
    n = 1 << log_n
    bases = ref.bases_seq(n, group)
    scalars = ref.fr_rng((group << 40) + n, n)
    exp, _ = ref.msm(bases, scalars, group)
    pinned = gpu.pin_bases(bases, group)
    try:
        assert np.array_equal(gpu.msm_pinned(pinned, scalars), exp), "pinned MSM differs from libff"
        st = gpu.msm_last_stats()
        assert st["bucket_sets"] == 1 and st["base_copies"] > 1
    finally:
        pinned.release()
    f = gpu.msm_g1 if group == 1 else gpu.msm_g2
    assert np.array_equal(f(bases, scalars), exp), "un-pinned MSM differs from libff"
