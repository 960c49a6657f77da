"""Generates the golden fixtures in this directory from oracle/_ref, i.e. from the UNMODIFIED
reference (libff / libfqfft / libsnark of /root/reference) compiled by oracle/Makefile.
Run in the build container (the only place /root/reference exists):

    make -C oracle ref && python tests/golden/make_golden.py

The fixtures are small on purpose (< 1 MB in total); they pin the C restatement
(oracle/zecale_oracle.c) and the CUDA path where oracle/_ref is not available.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bw6_constants as C  # noqa: E402
from oracle import ref as R  # noqa: E402
from oracle.ref import ints_to_limbs  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def save(name, **arrs):
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **arrs)
    print(name, {k: v.shape for k, v in arrs.items()})


def main():
    ref = R.Ref()
    q, r = C.Q, C.R
    # ---- fields: random pairs + edge values, all seven ops (test_fields.cpp:25-61 pattern)
    for name, p, n64, op, rng_seed in (("fq", q, 12, ref.fq_op, 1), ("fr", r, 6, ref.fr_op, 2)):
        rng = np.random.default_rng(rng_seed)
        edge = [0, 1, 2, p - 1, p - 2, (p - 1) // 2, (1 << (64 * n64)) % p, (1 << 64) - 1, 1 << 64]
        vals = edge + [int.from_bytes(rng.bytes(8 * n64), "little") % p for _ in range(55)]
        a = op(7, ints_to_limbs(vals, n64))                       # to Montgomery form
        b = op(7, ints_to_limbs(vals[::-1], n64))
        out = {"a": a, "b": b}
        for k, nm in enumerate(["add", "sub", "mul", "sqr", "inv", "neg", "from_mont"]):
            out[nm] = op(k, a, b)
        out["inv"][0] = 0                                          # inverse of zero is 0 by convention here
        save("field_" + name, **out)
    # ---- signed digits (test_fields.cpp:347-395 pattern): scalar x window size x index
    sc = [0, 1, r - 1, r - 2, (1 << 376) - 1, (r - 1) // 2] + \
         [sum((1 << (c - 1)) << (c * k) for k in range(377 // c + 1)) % r for c in (5, 13, 16)] + \
         [sum(((1 << c) - 1) << (c * k) for k in range(377 // c + 1)) % r for c in (5, 13, 16)]
    plain = ints_to_limbs(sc, 6)
    cs = [2, 5, 12, 13, 16, 17, 20]
    dig = np.zeros((len(sc), len(cs), 200), dtype=np.int64)
    for i in range(len(sc)):
        for j, c in enumerate(cs):
            for k in range((379 + c - 1) // c):  # the rounds multi_exp_inner can ask for (multiexp.tcc:584-586)
                dig[i, j, k] = ref.signed_digit(plain[i], c, k)
    save("signed_digits", plain=plain, cs=np.array(cs), digits=dig,
         optimal_c=np.array([[n, ref.optimal_c(n)] for n in (1, 2, 100, 4096, 8192, 1 << 16, 237801, 1 << 20)]))
    # ---- groups (test_groups.cpp:387-399 pattern): add / mixed_add / dbl / to_special / neg
    for g in (1, 2):
        n = 24
        a = ref.bases_mul(ref.fr_rng(10 + g, n), g)
        b = ref.bases_mul(ref.fr_rng(20 + g, n), g)
        zero = np.zeros(36, dtype=np.uint64)
        zero[12:24] = a[0][24:36]
        b[1] = a[1]                                   # P + P
        b[2] = ref.group_op(4, a[2:3], None, g)[0]    # P + (-P)
        a[3] = zero
        b[4] = zero
        a[5] = zero
        b[5] = zero
        a2 = ref.group_op(2, a, None, g)              # non-special left operands: 2P, normalised later by op 3
        out = {"a": a, "b": b}
        for k, nm in ((0, "add"), (1, "mixed_add"), (2, "dbl"), (3, "to_special"), (4, "neg")):
            out[nm] = ref.group_op(3, ref.group_op(k, a, b, g), None, g)
        out["a_proj"] = a2
        out["proj_mixed_add"] = ref.group_op(3, ref.group_op(1, a2, b, g), None, g)
        save("group_g%d" % g, **out)
    # ---- MSM: small dense cases with the edge scalars / points, both groups, and a sparse kc case
    for g in (1, 2):
        n = 300
        bases = ref.bases_mul(ref.fr_rng(70 + g, n), g)
        scalars = ref.fr_rng(4242 + g, n)
        fixed = ref.fr_from_ints([0, 1, r - 1, 2, (1 << 376), (r - 1) // 2] +
                                 [sum((1 << (c - 1)) << (c * k) for k in range(377 // c + 1)) % r for c in (8, 9, 13)])
        scalars[: fixed.shape[0]] = fixed
        zero = np.zeros(36, dtype=np.uint64)
        zero[12:24] = bases[0][24:36]
        bases[20] = zero
        bases[30] = bases[31]
        scalars[30] = scalars[31]
        bases[32] = ref.group_op(4, bases[33:34], None, g)[0]
        scalars[32] = scalars[33]
        bases[40:60] = bases[40]
        scalars[40:60] = scalars[40]
        res, _ = ref.msm(bases, scalars, g)
        res1, _ = ref.msm(bases, scalars, g, chunks=1)
        assert np.array_equal(res, res1)
        pre, _ = ref.msm(bases[:7], scalars[:7], g)
        save("msm_g%d_small" % g, bases=bases, scalars=scalars, result=res, result_first7=pre)
    n_sc, nnz = 500, 180
    rng = np.random.default_rng(3)
    idx = np.sort(rng.choice(n_sc, nnz, replace=False)).astype(np.uint64)
    vals = np.zeros((nnz, 72), dtype=np.uint64)
    vals[:, :36] = ref.bases_mul(ref.fr_rng(555, nnz), 2)
    vals[:, 36:] = ref.bases_mul(ref.fr_rng(556, nnz), 1)
    scalars = ref.fr_rng(8080, n_sc)
    scalars[idx[3]] = 0
    scalars[idx[4]] = ref.fr_from_ints([1])[0]
    res, _ = ref.msm_kc(vals, idx, n_sc, scalars)
    save("msm_kc_sparse", values=vals, indices=idx, scalars=scalars, result=res)
    # ---- NTT: the four transforms at m = 2, 2^4, 2^8, 2^11 (evaluation_domain_test.cpp:40-154 pattern), H polynomial
    g15 = ref.fr_from_ints([C.FR_GENERATOR])[0]
    out = {"coset_g": g15}
    for lg in (1, 4, 8, 11):
        a = ref.fr_rng(900 + lg, 1 << lg)
        out["in_%d" % lg] = a
        for fl in range(4):
            out["out_%d_%d" % (lg, fl)], _ = ref.ntt(a, fl, g15 if fl & 2 else None)
    save("ntt", **out)
    m = 1 << 8
    aA, aB, aC = (ref.fr_rng(31 * k, m) for k in (1, 2, 3))
    h, _ = ref.qap_h(aA, aB, aC)
    save("qap_h", aA=aA, aB=aB, aC=aC, coset_g=g15, h=h)
    # ---- a complete small Groth16 instance: pk queries, witness, evaluations, (r, s), the five MSM
    # results, coefficients_for_H and the 576 proof bytes (r1cs_gg_ppzksnark.tcc:464-603, groth16_snark.tcc:160-166)
    assert ref.session_synthetic(120, 4, 30, 7) == 0
    ref.session_setup(3)
    z = ref.session_export()
    rr, ss = ref.fr_rng(0xA11CE, 1)[0], ref.fr_rng(0xB0B, 1)[0]
    pr = ref.session_prove(rr, ss)
    assert pr["verified"]
    aA, aB, aC = ref.session_abc()
    save("groth16_small", A_query=z["A_query"], B_values=z["B_values"], B_indices=z["B_indices"], H_query=z["H_query"],
         L_query=z["L_query"], assignment=z["assignment"], aA=aA, aB=aB, aC=aC, coset_g=g15, r=rr, s=ss, h=pr["h"],
         msm5=pr["msm5"], proof=pr["proof"], sizes=np.array([z["num_variables"], z["num_inputs"], z["m"]]),
         csr_a_ptr=z["csr_a"][0], csr_a_col=z["csr_a"][1], csr_a_val=z["csr_a"][2],
         csr_b_ptr=z["csr_b"][0], csr_b_col=z["csr_b"][1], csr_b_val=z["csr_b"][2],
         csr_c_ptr=z["csr_c"][0], csr_c_col=z["csr_c"][1], csr_c_val=z["csr_c"][2])


if __name__ == "__main__":
    main()
