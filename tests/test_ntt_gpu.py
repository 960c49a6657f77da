"""zecale_gpu_ntt through the C ABI vs libfqfft basic_radix2_domain<bw6_761_Fr>
(EvaluationDomainTest patterns, libfqfft/tests/evaluation_domain_test.cpp:40-154)."""
import numpy as np
import pytest

import bw6_constants as C
import zecale_b200 as z
from oracle.ref import limbs_to_ints, ints_to_limbs

pytestmark = pytest.mark.gpu


# 17 and 19 are the real batch-2/4 and batch-16 aggregator domains; odd sizes take the uneven R1 * R2 split
@pytest.mark.parametrize("log_m", [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19])
def test_ntt_all_modes(gpu, ref, log_m):
    m = 1 << log_m
    a = ref.fr_rng(7 * m, m)
    g = ref.fr_from_ints([C.FR_GENERATOR])[0]
    for flags in (0, 1, 2, 3):
        exp, _ = ref.ntt(a, flags, g)
        got = gpu.ntt(a.copy(), flags, g if flags & 2 else None)
        assert np.array_equal(got, exp), (log_m, flags)


def test_ntt_naive_small(gpu, ref):
    """FFT == Horner evaluation at the domain points (evaluation_domain_test.cpp:40-80)."""
    log_m, m = 4, 16
    a = ref.fr_rng(5, m)
    plain = limbs_to_ints(ref.fr_op(6, a))
    w = pow(C.FR_ROOT_OF_UNITY, 1 << (C.FR_S - log_m), C.R)
    exp = [sum(c * pow(w, i * j, C.R) for j, c in enumerate(plain)) % C.R for i in range(m)]
    got = limbs_to_ints(ref.fr_op(6, gpu.ntt(a.copy(), 0)))
    assert got == exp


@pytest.mark.parametrize("log_m", [20, 22, 23, 24])
def test_ntt_roundtrip_large(gpu, ref, log_m):
    """Size-independent properties at the BASELINE sweep sizes: iFFT(FFT(a)) = a,
    icosetFFT(cosetFFT(a)) = a, and linearity on a sampled prefix."""
    m = 1 << log_m
    base = ref.fr_rng(99, 1 << 16)
    a = np.tile(base, (m >> 16, 1))
    a[:, 0] ^= np.arange(m, dtype=np.uint64)  # make rows distinct, still < r (low limb only)
    g = ref.fr_from_ints([C.FR_GENERATOR])[0]
    x = a.copy()
    gpu.ntt(x, 0)
    assert not np.array_equal(x[:64], a[:64])
    gpu.ntt(x, z.NTT_INVERSE)
    assert np.array_equal(x, a)
    gpu.ntt(x, z.NTT_COSET, g)
    gpu.ntt(x, z.NTT_COSET | z.NTT_INVERSE, g)
    assert np.array_equal(x, a)


@pytest.mark.parametrize("log_m,flag_set", [(21, (0, 3)), (22, (0, 1, 2, 3)), (23, (0, 3)), (24, (0, 1, 2, 3))])
def test_ntt_large_vs_libfqfft(gpu, ref, log_m, flag_set):
    """BASELINE config 3 at its full sizes, bit for bit against libfqfft's basic_radix2_domain
    (basic_radix2_domain.tcc:55-88): 2^21 / 2^22 are the largest two-pass transforms, 2^23 / 2^24 run the
    three-pass factorisation (ntt.cuh, m = R1*R2*R3), which no smaller size reaches."""
    m = 1 << log_m
    a = ref.fr_rng(0xA5A5 + log_m, m)
    g = ref.fr_from_ints([C.FR_GENERATOR])[0]
    for flags in flag_set:
        exp, _ = ref.ntt(a, flags, g)
        got = gpu.ntt(a.copy(), flags, g if flags & 2 else None)
        assert np.array_equal(got, exp), (log_m, flags)
        del exp, got


def test_ntt_2_20_vs_libfqfft(gpu, ref):
    m = 1 << 20
    a = ref.fr_rng(123456, m)
    exp, _ = ref.ntt(a, 3, ref.fr_from_ints([C.FR_GENERATOR])[0])
    got = gpu.ntt(a.copy(), 3, ref.fr_from_ints([C.FR_GENERATOR])[0])
    assert np.array_equal(got, exp)


def test_ntt_errors(gpu):
    with pytest.raises(ValueError):
        gpu.ntt(np.zeros((12, 6), dtype=np.uint64))
    with pytest.raises(z.ZecaleGpuError):
        gpu.ntt(np.zeros((4, 6), dtype=np.uint64), z.NTT_COSET, None)


@pytest.mark.parametrize("log_m", [4, 12, 16])
def test_qap_h_vs_reference(gpu, ref, log_m):
    """zecale_gpu_qap_h == the transform sequence of r1cs_to_qap_witness_map (r1cs_to_qap.tcc:252-330)."""
    m = 1 << log_m
    aA, aB, aC = (ref.fr_rng(s * 1000003, m) for s in (1, 2, 3))
    g = ref.fr_from_ints([C.FR_GENERATOR])[0]
    exp, _ = ref.qap_h(aA, aB, aC)
    keep = (aA.copy(), aB.copy(), aC.copy())
    got = gpu.qap_h(aA, aB, aC, g)
    assert np.array_equal(got, exp)
    assert all(np.array_equal(x, y) for x, y in zip(keep, (aA, aB, aC)))  # inputs untouched
    assert not got[m].any()
