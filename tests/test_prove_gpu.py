"""End-to-end parity of the prove hot path: with the proving key, witness and the prover
randomness (r, s) fixed, the H polynomial, the five MSM results (normalised affine) and the
576 serialised proof bytes must equal libsnark's (r1cs_gg_ppzksnark.tcc:464-603 with r, s
injected; proof codec groth16_snark.tcc:160-166). The MSMs / NTTs run on the GPU through the
C ABI exactly as the hooked call sites issue them; assembly and serialisation stay on the
reference's CPU code, as in INTEGRATION.md."""
import numpy as np
import pytest

import bw6_constants as C

pytestmark = pytest.mark.gpu


def _gpu_prove(gpu, ref, r, s, pinned=False):
    z = ref.session_export()
    nv, ni, m = z["num_variables"], z["num_inputs"], z["m"]
    g = ref.fr_from_ints([C.FR_GENERATOR])[0]
    aA, aB, aC = ref.session_abc()
    h = gpu.qap_h(aA, aB, aC, g)
    w = z["assignment"]
    assert w.shape[0] == nv + 1
    if pinned:
        pa = gpu.pin_bases(z["A_query"][: nv + 1])
        ph = gpu.pin_bases(z["H_query"][: m - 1])
        try:
            At = gpu.msm_pinned(pa, w)
            Ht = gpu.msm_pinned(ph, h[: m - 1])
        finally:
            pa.release()
            ph.release()
    else:
        At = gpu.msm_g1(z["A_query"][: nv + 1], w)
        Ht = gpu.msm_g1(z["H_query"][: m - 1], h[: m - 1])
    Bg = gpu.msm_g2(z["B_values"], w, index=z["B_indices"])
    Bh = gpu.msm_g1(z["B_values"], w, index=z["B_indices"], base_offset_bytes=288)
    Lt = gpu.msm_g1(z["L_query"], np.ascontiguousarray(w[ni + 1:]))
    msm5 = np.stack([At, Bg, Bh, Ht, Lt])
    proof, ok = ref.session_assemble(r, s, msm5)
    return h, msm5, proof, ok


def _check(gpu, ref, pinned):
    r = ref.fr_rng(0xA11CE, 1)[0]
    s = ref.fr_rng(0xB0B, 1)[0]
    exp = ref.session_prove(r, s)
    assert exp["verified"]
    h, msm5, proof, ok = _gpu_prove(gpu, ref, r, s, pinned)
    assert np.array_equal(h, exp["h"]), "coefficients_for_H differ"
    for k, name in enumerate(["A", "B.g2", "B.g1", "H", "L"]):
        assert np.array_equal(msm5[k], exp["msm5"][k]), "MSM %s differs" % name
    assert ok, "assembled proof does not verify"
    assert proof.tobytes() == exp["proof"].tobytes(), "proof bytes differ"
    assert len(proof.tobytes()) == 576


@pytest.mark.parametrize("pinned", [False, True])
def test_synthetic_groth16_proof_bytes(gpu, ref, pinned):
    """1000-constraint synthetic R1CS with 100 boolean variables (cf. run_r1cs_gg_ppzksnark.tcc:66-127)."""
    assert ref.session_synthetic(1000, 10, 100, 7) == 0
    ref.session_setup(3)
    _check(gpu, ref, pinned)


def test_aggregator_batch1_proof_bytes(gpu):
    """BASELINE config 1: the real aggregator circuit wrapping one BLS12-377 dummy_app proof
    (aggregator_dummy_test.cpp:201-211), m = 2^16, proved with the GPU hot path."""
    from oracle import ref as R
    if not R.available(aggregator=True):
        pytest.skip("oracle/_ref/libzecale_ref_agg.so not built")
    agg = R.Ref(aggregator=True)
    assert agg.session_aggregator(1, 11) == 0
    agg.session_setup(5)
    sz = agg.session_sizes()
    assert sz["m"] == 1 << 16 and sz["num_inputs"] == 3
    _check(gpu, agg, pinned=True)


def test_h_polynomial_from_witness_on_device(gpu, ref):
    """the sparse mat-vecs <A_i,w>, <B_i,w>, <C_i,w> and the whole of
    r1cs_to_qap_witness_map on the device (zecale_gpu_r1cs_load + zecale_gpu_qap_h_from_witness_async),
    fused with the five MSMs of the prover on the asynchronous lanes; (r, s) fixed -> same proof bytes."""
    assert ref.session_synthetic(5000, 12, 700, 11) == 0
    ref.session_setup(4)
    z = ref.session_export()
    nv, ni, m = z["num_variables"], z["num_inputs"], z["m"]
    g = ref.fr_from_ints([C.FR_GENERATOR])[0]
    r, s = ref.fr_rng(1, 1)[0], ref.fr_rng(2, 1)[0]
    exp = ref.session_prove(r, s)
    w = z["assignment"]
    r1cs = gpu.r1cs_load(z["num_constraints"], ni, nv, [z["csr_a"], z["csr_b"], z["csr_c"]])
    assert (1 << gpu.r1cs_log_domain(r1cs)) == m
    pins = [gpu.pin_bases(z["A_query"][: nv + 1]), gpu.pin_bases(z["B_values"], 2), gpu.pin_bases(z["B_values"], 1, base_offset_bytes=288),
            gpu.pin_bases(z["H_query"][: m - 1]), gpu.pin_bases(z["L_query"])]
    try:
        for rnd in range(2):
            out = [np.zeros(36, dtype=np.uint64) for _ in range(5)]
            h = np.zeros((m + 1, 6), dtype=np.uint64)
            gpu.qap_h_from_witness_async(0, r1cs, w, g, out=h if rnd == 0 else None)
            gpu.msm_pinned_async(1, pins[0], w, out[0])
            gpu.msm_pinned_async(2, pins[1], w, out[1], index=z["B_indices"], n=z["nnz_B"])
            gpu.msm_pinned_async(3, pins[2], w, out[2], index=z["B_indices"], n=z["nnz_B"])
            gpu.msm_pinned_async(0, pins[3], None, out[3], from_qap=True)
            gpu.msm_pinned_async(4, pins[4], w[ni + 1:], out[4])
            gpu.sync()
            if rnd == 0:
                assert np.array_equal(h, exp["h"])
            assert np.array_equal(np.stack(out), exp["msm5"])
            proof, ok = ref.session_assemble(r, s, np.stack(out))
            assert ok and proof.tobytes() == exp["proof"].tobytes()
    finally:
        for p in pins:
            p.release()
        gpu.r1cs_free(r1cs)
