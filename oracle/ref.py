"""ctypes binding of oracle/_ref/libzecale_ref*.so — the UNMODIFIED reference
(libff / libfqfft / libsnark / libzeth / libzecale) compiled by oracle/Makefile.

TEST INFRASTRUCTURE ONLY. May be imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs — never by zecale_b200/.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(_HERE, "_ref", "libzecale_ref.so")
REF_AGG_SO = os.path.join(_HERE, "_ref", "libzecale_ref_agg.so")
# ref-agg compiled with -include zecale_gpu_hooks.hpp: the reference prover running on the GPU (drop-in demo)
REF_HOOKED_SO = os.path.join(_HERE, "_ref", "libzecale_ref_hooked.so")
REF_HOOKED_FULLTABLE_SO = os.path.join(_HERE, "_ref", "libzecale_ref_hooked_fulltable.so")  # hooked="fulltable"
# ref-agg with -DDEBUG (the reference's shipped default: is_satisfied + a QAP re-evaluation inside every prove)
REF_DEBUG_SO = os.path.join(_HERE, "_ref", "libzecale_ref_agg_debug.so")

_vp, _sz, _u64, _d, _i = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint64, ctypes.c_double, ctypes.c_int


def available(aggregator=False, hooked=False, debug=False):
    return os.path.exists(REF_DEBUG_SO if debug else REF_HOOKED_FULLTABLE_SO if hooked == "fulltable" else REF_HOOKED_SO if hooked
                          else REF_AGG_SO if aggregator else REF_SO)


def _p(a):
    return None if a is None else a.ctypes.data_as(_vp)


class Ref:
    """One loaded copy of the reference. Use aggregator=True for the zecale circuit."""

    def __init__(self, aggregator=False, hooked=False, debug=False):
        path = (REF_DEBUG_SO if debug else REF_HOOKED_FULLTABLE_SO if hooked == "fulltable" else REF_HOOKED_SO if hooked
                else REF_AGG_SO if aggregator else REF_SO)
        if not os.path.exists(path):
            raise FileNotFoundError(path + " (run `make -C oracle ref ref-agg` where /root/reference exists)")
        L = ctypes.CDLL(path)
        for f in ("zref_msm_g1", "zref_msm_g2", "zref_msm_g1_naive", "zref_msm_kc", "zref_ntt", "zref_qap_h",
                  "zref_session_setup", "zref_session_prove_stock", "zref_session_witness_map",
                  "zref_session_witness_gen_seconds", "zref_session_prove_pipeline", "zref_session_prove_threads",
                  "zref_session_keypair_write", "zref_session_keypair_read", "zref_lagrange",
                  "zref_session_witness_gen_parallel_seconds"):
            getattr(L, f).restype = _d
        L.zref_fr_signed_digit.restype = ctypes.c_long
        L.zref_session_abc.restype = ctypes.c_long
        L.zref_optimal_c.restype = _sz
        self.L = L
        self.threads = L.zref_init()

    def set_threads(self, n):
        """OpenMP threads of the reference from here on (overrides OMP_NUM_THREADS, which torchrun sets to 1)."""
        self.L.zref_set_threads(int(n))
        self.threads = int(self.L.zref_threads())
        return self.threads

    # ---- fields -------------------------------------------------------------------
    def fq_op(self, op, a, b=None):
        out = np.zeros_like(a)
        self.L.zref_fq_op(op, _p(a), _p(b if b is not None else a), _p(out), _sz(a.shape[0]))
        return out

    def fr_op(self, op, a, b=None):
        out = np.zeros_like(a)
        self.L.zref_fr_op(op, _p(a), _p(b if b is not None else a), _p(out), _sz(a.shape[0]))
        return out

    def fr_rng(self, seed, n):
        out = np.zeros((n, 6), dtype=np.uint64)
        self.L.zref_fr_rng(_u64(seed), _sz(n), _p(out))
        return out

    def fr_from_ints(self, vals):
        a = ints_to_limbs(vals, 6)
        return self.fr_op(7, a)

    def fq_from_ints(self, vals):
        a = ints_to_limbs(vals, 12)
        return self.fq_op(7, a)

    def signed_digit(self, plain_limbs, c, idx):
        return int(self.L.zref_fr_signed_digit(_p(np.ascontiguousarray(plain_limbs)), _sz(c), _sz(idx)))

    def optimal_c(self, n):
        return int(self.L.zref_optimal_c(_sz(n)))

    # ---- groups -------------------------------------------------------------------
    def bases_seq(self, n, group=1):
        out = np.zeros((n, 36), dtype=np.uint64)
        (self.L.zref_g1_bases_seq if group == 1 else self.L.zref_g2_bases_seq)(_sz(n), _p(out))
        return out

    def bases_mul(self, scalars, group=1):
        out = np.zeros((scalars.shape[0], 36), dtype=np.uint64)
        (self.L.zref_g1_bases_mul if group == 1 else self.L.zref_g2_bases_mul)(_p(scalars), _sz(scalars.shape[0]), _p(out))
        return out

    def batch_exp(self, base, scalars, group=1, coeff=None):
        """libff::batch_exp(_with_coeff) + batch_to_special: out[i] = (coeff * scalars[i]) * base."""
        out = np.zeros((scalars.shape[0], 36), dtype=np.uint64)
        self.L.zref_batch_exp(int(group), _p(base), _p(scalars), _sz(scalars.shape[0]),
                              _p(coeff) if coeff is not None else None, _p(out))
        return out

    def points_write_bytes(self, pts, group=1):
        """libzeth::group_elements_write_bytes of a vector of points (the on-disk form of a pk query)."""
        cap = 8 + 192 * pts.shape[0]
        out = np.zeros(cap, dtype=np.uint8)
        self.L.zref_points_write_bytes.restype = ctypes.c_size_t
        n = self.L.zref_points_write_bytes(int(group), _p(pts), _sz(pts.shape[0]), _p(out), _sz(cap))
        assert n == cap, n
        return out

    def kc_write_bytes(self, values, indices, domain):
        """libzeth::knowledge_commitment_vector_write_bytes of a sparse G2/G1 query (values: n x 72 u64)."""
        cap = 16 + 392 * values.shape[0]
        out = np.zeros(cap, dtype=np.uint8)
        self.L.zref_kc_write_bytes.restype = ctypes.c_size_t
        idx = np.ascontiguousarray(indices, dtype=np.uint64)
        n = self.L.zref_kc_write_bytes(_p(values), _p(idx), _sz(values.shape[0]), _sz(domain), _p(out), _sz(cap))
        assert n == cap, n
        return out

    def group_op(self, op, a, b=None, group=1):
        out = np.zeros_like(a)
        (self.L.zref_g1_op if group == 1 else self.L.zref_g2_op)(op, _p(a), _p(b if b is not None else a), _p(out), _sz(a.shape[0]))
        return out

    # ---- MSM ----------------------------------------------------------------------
    def msm(self, bases, scalars, group=1, chunks=0, method=0):
        out = np.zeros(36, dtype=np.uint64)
        b = np.ascontiguousarray(bases)
        s = np.ascontiguousarray(scalars)
        f = self.L.zref_msm_g1 if group == 1 else self.L.zref_msm_g2
        t = f(_p(b), _p(s), _sz(b.shape[0]), _sz(chunks), method, _p(out))
        return out, t

    def msm_naive_g1(self, bases, scalars):
        out = np.zeros(36, dtype=np.uint64)
        t = self.L.zref_msm_g1_naive(_p(bases), _p(scalars), _sz(bases.shape[0]), _p(out))
        return out, t

    def msm_kc(self, values, indices, domain_size, scalars, chunks=0):
        out = np.zeros(72, dtype=np.uint64)
        idx = np.ascontiguousarray(indices, dtype=np.uint64)
        t = self.L.zref_msm_kc(_p(values), _p(idx), _sz(values.shape[0]), _sz(domain_size), _p(scalars),
                               _sz(scalars.shape[0]), _sz(chunks), _p(out))
        return out, t

    # ---- NTT ----------------------------------------------------------------------
    def ntt(self, a, flags=0, coset_g=None):
        """Returns (result, seconds); input is not modified."""
        out = np.ascontiguousarray(a).copy()
        log_m = out.shape[0].bit_length() - 1
        t = self.L.zref_ntt(_p(out), ctypes.c_uint(log_m), ctypes.c_uint(flags), _p(coset_g))
        if t < 0:
            raise RuntimeError("reference NTT threw")
        return out, t

    def lagrange(self, log_m, t):
        """libfqfft evaluate_all_lagrange_polynomials(t) on the size-2^log_m domain; (u (m, 6), seconds)."""
        out = np.zeros((1 << log_m, 6), dtype=np.uint64)
        secs = self.L.zref_lagrange(ctypes.c_uint(log_m), _p(np.ascontiguousarray(t)), _p(out))
        return out, secs

    def qap_h(self, aA, aB, aC):
        """coefficients_for_H (m+1, 6) by the reference's own transform sequence; returns (h, seconds)."""
        m = aA.shape[0]
        h = np.zeros((m + 1, 6), dtype=np.uint64)
        t = self.L.zref_qap_h(_p(np.ascontiguousarray(aA)), _p(np.ascontiguousarray(aB)), _p(np.ascontiguousarray(aC)),
                              ctypes.c_uint(m.bit_length() - 1), _p(h))
        return h, t

    # ---- Groth16 sessions -----------------------------------------------------------
    def session_synthetic(self, num_constraints, num_inputs, n_bits, seed):
        return self.L.zref_session_synthetic(_sz(num_constraints), _sz(num_inputs), _sz(n_bits), _u64(seed))

    def session_aggregator(self, batch, seed=0):
        return self.L.zref_session_aggregator(_sz(batch), _u64(seed))

    def session_setup(self, seed):
        return self.L.zref_session_setup(_u64(seed))

    def session_keypair_write(self, path):
        t = self.L.zref_session_keypair_write(path.encode())
        if t < 0:
            raise RuntimeError("keypair_write failed: %r" % t)
        return t

    def session_keypair_read(self, path):
        t = self.L.zref_session_keypair_read(path.encode())
        if t < 0:
            raise RuntimeError("keypair_read failed: %r" % t)
        return t

    def session_from_csr(self, n_constraints, n_inputs, n_variables, csr, full):
        """Session over a constraint system given as three (row_ptr u64, col u32, val (nnz, 6) u64) CSR
        matrices and the assignment [1, primary, aux]."""
        keep = [(np.ascontiguousarray(rp, dtype=np.uint64), np.ascontiguousarray(c, dtype=np.uint32),
                 np.ascontiguousarray(v, dtype=np.uint64)) for rp, c, v in csr]
        arr = lambda k: (_vp * 3)(*[t[k].ctypes.data for t in keep])
        return self.L.zref_session_from_csr(_sz(n_constraints), _sz(n_inputs), _sz(n_variables), arr(0), arr(1), arr(2),
                                            _p(np.ascontiguousarray(full)))

    def session_witness_map(self, m=None):
        """libsnark::r1cs_to_qap_witness_map (mat-vecs + 7 transforms) on the session; returns
        (coefficients_for_H (m + 1, 6) or None when m is not given, seconds)."""
        h = np.zeros((m + 1, 6), dtype=np.uint64) if m else None
        t = self.L.zref_session_witness_map(_p(h))
        if t < 0:
            raise RuntimeError("no session")
        return h, t

    def session_witness_gen_seconds(self):
        return float(self.L.zref_session_witness_gen_seconds())

    def session_witness_gen_parallel(self):
        """(seconds, threads) of the per-nested-proof parallel witness generation prototype of the last
        aggregator session; seconds < 0 means it did not reproduce the serial assignment."""
        n = ctypes.c_int()
        t = self.L.zref_session_witness_gen_parallel_seconds(ctypes.byref(n))
        return float(t), int(n.value)

    def session_sizes(self):
        sz = np.zeros(16, dtype=np.uint64)
        rc = self.L.zref_session_sizes(_p(sz))
        assert rc == 0
        keys = ["num_constraints", "num_variables", "num_inputs", "m", "n_A", "nnz_B", "B_domain", "n_H", "n_L",
                "nnz_a", "nnz_b", "nnz_c"]
        return {k: int(sz[i]) for i, k in enumerate(keys)}

    def session_export(self):
        """All arrays of the current session (pk queries, CSR matrices, assignment)."""
        z = self.session_sizes()
        out = dict(z)

        def get(which, shape, dtype=np.uint64):
            a = np.zeros(shape, dtype=dtype)
            assert self.L.zref_session_get(which, _p(a)) == 0
            return a

        out["A_query"] = get(0, (z["n_A"], 36))
        out["B_values"] = get(1, (z["nnz_B"], 72))
        out["B_indices"] = get(2, (z["nnz_B"],))
        out["H_query"] = get(3, (z["n_H"], 36))
        out["L_query"] = get(4, (z["n_L"], 36))
        out["g1_consts"] = get(5, (3, 36))
        out["g2_consts"] = get(6, (2, 36))
        out["assignment"] = get(7, (z["num_variables"] + 1, 6))
        for k, name in enumerate("abc"):
            nnz = z["nnz_" + name]
            rp = np.zeros(z["num_constraints"] + 1, dtype=np.uint64)
            col = np.zeros(nnz, dtype=np.uint32)
            val = np.zeros((nnz, 6), dtype=np.uint64)
            assert self.L.zref_session_csr(k, _p(rp), _p(col), _p(val)) == 0
            out["csr_" + name] = (rp, col, val)
        return out

    def session_abc(self):
        """(aA, aB, aC): the evaluation vectors of r1cs_to_qap_witness_map before its first transform."""
        m = self.session_sizes()["m"]
        v = [np.zeros((m, 6), dtype=np.uint64) for _ in range(3)]
        got = self.L.zref_session_abc(_p(v[0]), _p(v[1]), _p(v[2]))
        assert got == m, (got, m)
        return v

    def session_prove(self, r, s, chunks=0):
        z = self.session_sizes()
        proof = np.zeros(576, dtype=np.uint8)
        msm5 = np.zeros((5, 36), dtype=np.uint64)
        h = np.zeros((z["m"] + 1, 6), dtype=np.uint64)
        tm = np.zeros(6, dtype=np.float64)
        rc = self.L.zref_session_prove(_p(r), _p(s), _sz(chunks), _p(proof), _p(msm5), _p(h), _p(tm))
        return {"verified": rc == 0, "proof": proof, "msm5": msm5, "h": h,
                "timings": dict(zip(["witness_map", "msm_a", "msm_b", "msm_h", "msm_l", "total"], tm.tolist()))}

    def session_assemble(self, r, s, msm5):
        proof = np.zeros(576, dtype=np.uint8)
        rc = self.L.zref_session_assemble(_p(r), _p(s), _p(np.ascontiguousarray(msm5)), _p(proof))
        return proof, rc == 0

    def session_assignment(self):
        n = self.session_sizes()["num_variables"] + 1
        a = np.zeros((n, 6), dtype=np.uint64)
        assert self.L.zref_session_get(7, _p(a)) == 0
        return a

    def session_set_assignment(self, full):
        full = np.ascontiguousarray(full)
        return self.L.zref_session_set_assignment(_p(full), _sz(full.shape[0]))

    def session_prove_stock(self):
        return self.L.zref_session_prove_stock()

    # ---- hooked build only ------------------------------------------------------------
    def is_hooked(self):
        return bool(self.L.zref_is_hooked())

    def hooks_pin_pk(self):
        return self.L.zref_hooks_pin_pk()

    def session_prove_fused(self, r, s):
        """Hooked build: the specialised prover body with (r, s) given (fused device path if pinned)."""
        proof = np.zeros(576, dtype=np.uint8)
        msm5 = np.zeros((5, 36), dtype=np.uint64)
        tm = np.zeros(1, dtype=np.float64)
        rc = self.L.zref_session_prove_fused(_p(r), _p(s), _p(proof), _p(msm5), _p(tm))
        return {"verified": rc == 0, "rc": rc, "proof": proof, "msm5": msm5, "seconds": float(tm[0])}

    def hooks_unpin(self):
        return self.L.zref_hooks_unpin()

    def session_prove_pipeline(self, count, rs_pairs):
        """Hooked build: `count` proofs back to back through prove_async / collect; rs_pairs: (k, 2, 6) u64.
        Returns (proofs (count, 576) u8, seconds)."""
        rs = np.ascontiguousarray(rs_pairs, dtype=np.uint64)
        proofs = np.zeros((count, 576), dtype=np.uint8)
        t = self.L.zref_session_prove_pipeline(_sz(count), _p(rs), _sz(rs.shape[0]), _p(proofs))
        if t < 0:
            raise RuntimeError("zref_session_prove_pipeline failed: %r" % t)
        return proofs, t

    def session_prove_threads(self, n_threads, per_thread, r, s):
        proofs = np.zeros((n_threads * per_thread, 576), dtype=np.uint8)
        t = self.L.zref_session_prove_threads(_sz(n_threads), _sz(per_thread), _p(r), _p(s), _p(proofs))
        if t < 0:
            raise RuntimeError("zref_session_prove_threads failed: %r" % t)
        return proofs, t


def ints_to_limbs(vals, n64):
    a = np.zeros((len(vals), n64), dtype=np.uint64)
    for i, v in enumerate(vals):
        for k in range(n64):
            a[i, k] = (v >> (64 * k)) & 0xFFFFFFFFFFFFFFFF
    return a


def limbs_to_ints(a):
    a = np.atleast_2d(a)
    return [sum(int(a[i, k]) << (64 * k) for k in range(a.shape[1])) for i in range(a.shape[0])]
