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
        # --- the drop-in at pipelined throughput: K proofs back to back through prove_async / collect,
        # alternating between two (r, s) pairs whose bytes the CPU build fixes
        r2, s2 = cpu.fr_rng(0xD00D, 1)[0], cpu.fr_rng(0xF00D, 1)[0]
        exp2 = cpu.session_prove(r2, s2)
        assert exp2["verified"]
        want = [exp["proof"].tobytes(), exp2["proof"].tobytes()]
        rs = np.stack([np.stack([r, s]), np.stack([r2, s2])])
        hooked.session_prove_pipeline(3, rs)          # every proof slot takes one proof first (tables, scratch)
        k = 8
        proofs, secs = hooked.session_prove_pipeline(k, rs)
        for i in range(k):
            assert proofs[i].tobytes() == want[i % 2], "pipelined proof %d differs" % i
        # --- two host threads proving at the same time (the aggregator server's gRPC workers)
        tproofs, tsecs = hooked.session_prove_threads(2, 3, r, s)
        for i in range(6):
            assert tproofs[i].tobytes() == want[0], "concurrent proof %d differs" % i
        with capsys.disabled():
            print("[batch-8 aggregator, hooked prover pipelined] %d proofs in %.3f s = %.1f proofs/s (prove_async/collect, real statement, "
                  "host assembly + serialisation included); 2 threads x 3 proofs concurrently: %.3f s = %.1f proofs/s; witness generation "
                  "(aggregator_circuit::prove's generate_r1cs_witness calls, reference CPU code, 1 thread): %.3f s"
                  % (k, secs, k / secs, tsecs, 6 / tsecs, cpu.session_witness_gen_seconds()))
            tpar, nthr = cpu.session_witness_gen_parallel()
            print("[batch-8 aggregator, SURVEY 8(f2)] witness generation, one nested proof per host thread (%d threads): %.3f s, same "
                  "assignment; so a whole aggregator_circuit::prove = witness generation %.3f s (serial) / %.3f s (parallel) + hooked prover %.3f s"
                  % (nthr, tpar, cpu.session_witness_gen_seconds(), tpar, t_fused))
        assert k / secs > 15.0
        assert cpu.session_witness_gen_parallel()[0] > 0, "parallel witness generation did not reproduce the serial assignment"
    finally:
        hooked.hooks_unpin()


def test_hooked_reference_prover_aggregator_batch32(pair, capsys, tmp_path):
    """BASELINE config 5's circuit for real: aggregator over 32 nested BLS12-377 proofs (920 718 constraints,
    m = 2^20; aggregator_circuit.tcc:119-170). The hooked build generates the key (device batch
    exponentiations) and writes it in libzeth's on-disk format; the plain CPU build reads that file and
    proves; the hooked prover must return the same 576 bytes."""
    import os
    import time
    cpu, hooked = pair
    assert hooked.session_aggregator(32, 5) == 0
    t0 = time.perf_counter()
    hooked.session_setup(9)
    t_setup = time.perf_counter() - t0
    path = os.path.join(str(tmp_path), "agg32.keypair")
    t_write = hooked.session_keypair_write(path)
    assert cpu.session_aggregator(32, 5) == 0
    t_read = cpu.session_keypair_read(path)
    os.remove(path)
    assert hooked.session_set_assignment(cpu.session_assignment()) == 0
    sz = cpu.session_sizes()
    assert sz["m"] == 1 << 20 and sz["num_constraints"] > 900000
    r, s = cpu.fr_rng(0x3232, 1)[0], cpu.fr_rng(0x2323, 1)[0]
    exp = cpu.session_prove(r, s)
    assert exp["verified"]
    assert hooked.hooks_pin_pk() == 0
    try:
        _check_fused(hooked, r, s, exp)
        t_fused = min(_check_fused(hooked, r, s, exp) for _ in range(2))
        with capsys.disabled():
            print("\n[batch-32 aggregator, %d constraints, m = 2^20, MSM sizes A %d / B %d / H %d / L %d] hooked setup %.1f s "
                  "(key file %.1f s to write, %.1f s to read); reference CPU prove (%d threads) %.2f s; specialised prover on the GPU %.3f s; "
                  "identical proof bytes" % (sz["num_constraints"], sz["n_A"], sz["nnz_B"], sz["n_H"], sz["n_L"], t_setup, t_write, t_read,
                                             cpu.threads, exp["timings"]["total"], t_fused))
    finally:
        hooked.hooks_unpin()


def test_default_build_keeps_libffs_window_table(pair):
    """The hooks WITHOUT -DZECALE_GPU_SLIM_WINDOW_TABLE (what `-include` alone gives a maintainer): libff's own
    get_window_table builds its full table on the CPU and the specialised batch_exp / kc_batch_exp readers take the base
    point table[0][1] from it. Setup from the same secrets and the proof must equal the CPU build's, as they do for the
    slim-table build above."""
    from oracle import ref as R
    if not R.available(hooked="fulltable"):
        pytest.skip("oracle/_ref/libzecale_ref_hooked_fulltable.so not present")
    cpu, _ = pair
    full = R.Ref(hooked="fulltable")
    assert full.is_hooked()
    r, s, exp = _same_session(cpu, full, lambda lib: lib.session_synthetic(3000, 10, 400, 21))
    _check(full, r, s, exp)
    assert full.hooks_pin_pk() == 0
    try:
        _check_fused(full, r, s, exp)
    finally:
        full.hooks_unpin()
