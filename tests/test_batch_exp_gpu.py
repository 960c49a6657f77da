"""zecale_gpu_batch_exp vs libff::batch_exp / batch_exp_with_coeff + batch_to_special — the fixed-base
exponentiations of r1cs_gg_ppzksnark_generator. Bit-exact special-form points."""
import numpy as np
import pytest

import bw6_constants as C

pytestmark = pytest.mark.gpu


def _needs(ref):
    if not hasattr(ref.L, "zref_batch_exp"):
        pytest.skip("oracle/_ref built before zref_batch_exp was added")


@pytest.mark.parametrize("group", [1, 2])
@pytest.mark.parametrize("n", [1, 7, 300, 5000])
def test_batch_exp_matches_libff(gpu, ref, group, n):
    _needs(ref)
    base = ref.bases_mul(ref.fr_rng(9, 1), group)[0]
    scalars = ref.fr_rng(31 * n + group, n)
    exp = ref.batch_exp(base, scalars, group)
    assert np.array_equal(gpu.batch_exp(base, scalars, group), exp)


@pytest.mark.parametrize("group", [1, 2])
def test_batch_exp_with_coeff_and_edge_scalars(gpu, ref, group):
    _needs(ref)
    base = ref.bases_seq(3, group)[2]
    r = C.R
    vals = [0, 1, 2, r - 1, r - 2, (r - 1) // 2, 1 << 376, (1 << 12) - 1, 1 << 12, (1 << 372) + 5, 0, 0]
    scalars = np.concatenate([ref.fr_from_ints(vals), ref.fr_rng(5, 200)])
    coeff = ref.fr_rng(77, 1)
    exp = ref.batch_exp(base, scalars, group, coeff=coeff)
    assert np.array_equal(gpu.batch_exp(base, scalars, group, coeff=coeff), exp)
    # coeff = 0: every output is the zero point (0, one, 0)
    zero = ref.fr_from_ints([0])
    exp0 = ref.batch_exp(base, scalars[:20], group, coeff=zero)
    assert np.array_equal(gpu.batch_exp(base, scalars[:20], group, coeff=zero), exp0)


def test_batch_exp_zero_base_and_generator(gpu, ref):
    _needs(ref)
    scalars = ref.fr_rng(123, 64)
    gen = ref.bases_seq(1, 1)[0]
    zero = gen.copy()
    zero[:12] = 0
    zero[24:36] = 0  # (0, one, 0); Y of a special-form point with x... keep Y = one
    zero[12:24] = gen[24:36]
    assert np.array_equal(gpu.batch_exp(zero, scalars, 1), ref.batch_exp(zero, scalars, 1))
    # s * G for the generator equals the harness's own scalar multiplications
    assert np.array_equal(gpu.batch_exp(gen, scalars, 1), ref.bases_mul(scalars, 1))


@pytest.mark.parametrize("log_m", [1, 5, 12, 16, 18])
def test_lagrange_vs_libfqfft(gpu, ref, log_m):
    """zecale_gpu_lagrange == basic_radix2_domain::evaluate_all_lagrange_polynomials
    (basic_radix2_domain_aux.tcc:186-239), the m inversions of the setup's QAP instance evaluation."""
    import zecale_b200 as z
    t = ref.fr_rng(0x7A6 + log_m, 1)[0]
    exp, _ = ref.lagrange(log_m, t)
    assert np.array_equal(gpu.lagrange(log_m, t), exp)
    # t on the domain: the unit-vector case stays with the reference's code
    e1 = np.zeros((1 << log_m, 6), dtype=np.uint64)
    e1[1] = ref.fr_from_ints([1])[0]
    w = ref.ntt(e1, 0)[0][1]          # FFT(e_1) = (1, w, w^2, ...): w is a point of the domain
    with pytest.raises(z.ZecaleGpuError) as e:
        gpu.lagrange(log_m, w)
    assert e.value.code == 4
