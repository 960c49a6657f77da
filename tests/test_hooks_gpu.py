"""Drop-in demonstration: the reference's own prover code, compiled with
`-include zecale_b200/hooks/zecale_gpu_hooks.hpp` and not edited anywhere (oracle/Makefile
ref-hooked), runs its five multi-exponentiations and seven transforms on the GPU and produces the
same proof bytes as the plain CPU build for the same key, witness and (r, s). Also runs the truly
unmodified libsnark::r1cs_gg_ppzksnark_prover (random r, s) on the hooked build and checks that its
proof verifies."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pair():
    from oracle import ref as R
    if not (R.available(aggregator=True) and R.available(hooked=True)):
        pytest.skip("oracle/_ref hooked / agg builds not present")
    cpu, hooked = R.Ref(aggregator=True), R.Ref(hooked=True)
    assert hooked.is_hooked() and not cpu.is_hooked()
    return cpu, hooked


SETUP_SECONDS = {}


def _same_session(cpu, hooked, make):
    import time
    for name, lib in (("cpu", cpu), ("hooked", hooked)):
        assert make(lib) == 0
        t0 = time.perf_counter()
        lib.session_setup(9)   # r1cs_gg_ppzksnark_generator_from_secrets; the hooked build runs its batch_exp calls on the GPU
        SETUP_SECONDS[name] = time.perf_counter() - t0
    # same statement in both builds (nested-proof generation is randomised inside the reference)
    assert hooked.session_set_assignment(cpu.session_assignment()) == 0
    r = cpu.fr_rng(0xC0FFEE, 1)[0]
    s = cpu.fr_rng(0xBEEF, 1)[0]
    exp = cpu.session_prove(r, s)
    assert exp["verified"]
    return r, s, exp


def _check_fused(hooked, r, s, exp):
    """The specialised libsnark::r1cs_gg_ppzksnark_prover body (fused device path when the key is resident)."""
    got = hooked.session_prove_fused(r, s)
    assert got["verified"], got["rc"]
    assert np.array_equal(got["msm5"], exp["msm5"]), "MSM results differ (fused prover)"
    assert got["proof"].tobytes() == exp["proof"].tobytes(), "proof bytes differ (fused prover)"
    return got["seconds"]


def _check(hooked, r, s, exp):
    got = hooked.session_prove(r, s)
    assert got["verified"]
    assert np.array_equal(got["h"], exp["h"]), "coefficients_for_H differ"
    assert np.array_equal(got["msm5"], exp["msm5"]), "MSM results differ"
    assert got["proof"].tobytes() == exp["proof"].tobytes(), "proof bytes differ"


def test_hooked_reference_prover_synthetic(pair):
    cpu, hooked = pair
    r, s, exp = _same_session(cpu, hooked, lambda lib: lib.session_synthetic(3000, 10, 400, 21))
    _check(hooked, r, s, exp)                       # un-pinned: bases uploaded per call
    _check_fused(hooked, r, s, exp)                 # un-pinned key: the sequential body
    assert hooked.hooks_pin_pk() == 0
    try:
        _check(hooked, r, s, exp)                   # pinned proving key
        _check_fused(hooked, r, s, exp)             # resident key: mat-vecs + transforms + MSMs fused on the device
        assert hooked.session_prove_stock() > 0     # unmodified prover, random r/s: proof verifies
    finally:
        hooked.hooks_unpin()


def test_hooked_reference_prover_aggregator_batch1(pair):
    """BASELINE config 1 through the hooks: aggregator circuit over one nested BLS12-377 proof."""
    cpu, hooked = pair
    r, s, exp = _same_session(cpu, hooked, lambda lib: lib.session_aggregator(1, 11))
    assert hooked.hooks_pin_pk() == 0
    try:
        _check(hooked, r, s, exp)
        _check_fused(hooked, r, s, exp)
        t_gpu = hooked.session_prove_stock()
        assert t_gpu > 0
        t_cpu = exp["timings"]["total"]
        print("aggregator batch-1 prove: reference CPU %.2f s, hooked (GPU) %.2f s" % (t_cpu, t_gpu))
    finally:
        hooked.hooks_unpin()


def test_hooked_reference_prover_aggregator_batch8(pair, capsys):
    """BASELINE config 4: batch-8 aggregation (m = 2^18) proved by the reference's code with the
    hooks, bit-identical to the CPU build; prints both prove times (same box, same process)."""
    cpu, hooked = pair
    r, s, exp = _same_session(cpu, hooked, lambda lib: lib.session_aggregator(8, 3))
    sz = cpu.session_sizes()
    assert sz["m"] == 1 << 18
    assert hooked.hooks_pin_pk() == 0
    try:
        _check(hooked, r, s, exp)
        got = hooked.session_prove(r, s)      # second run: tables and key resident
        _check_fused(hooked, r, s, exp)
        t_fused = min(_check_fused(hooked, r, s, exp) for _ in range(3))
        t_stock = min(hooked.session_prove_stock() for _ in range(3))
        assert t_stock > 0
        with capsys.disabled():
            print("\n[batch-8 aggregator prove, %d constraints] reference CPU (%d threads): %.2f s (witness map %.2f, MSM %.2f);"
                  " call-site hooks on the GPU: %.3f s (witness map %.3f, MSM %.3f); specialised prover (fused device path): %.3f s;"
                  " snark::generate_proof entry, hooked, random r/s: %.3f s"
                  % (sz["num_constraints"], cpu.threads, exp["timings"]["total"], exp["timings"]["witness_map"],
                     sum(exp["timings"][k] for k in ("msm_a", "msm_b", "msm_h", "msm_l")),
                     got["timings"]["total"], got["timings"]["witness_map"],
                     sum(got["timings"][k] for k in ("msm_a", "msm_b", "msm_h", "msm_l")), t_fused, t_stock))
            print("[batch-8 aggregator setup (generator_from_secrets)] reference CPU: %.2f s; hooked (A/B/H/L queries and gamma_ABC"
                  " as device batch exponentiations): %.2f s; proofs from both keys are byte-identical"
                  % (SETUP_SECONDS["cpu"], SETUP_SECONDS["hooked"]))
    finally:
        hooked.hooks_unpin()
