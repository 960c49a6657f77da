"""ctypes binding of oracle/libzecale_oracle.so — the plain-C restatement (zecale_oracle.c).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline leg; never from zecale_b200/. Built by `make -C oracle port` (gcc only)."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(_HERE, "libzecale_oracle.so")
_vp, _sz = ctypes.c_void_p, ctypes.c_size_t


def available():
    return os.path.exists(PORT_SO)


def _p(a):
    return None if a is None else a.ctypes.data_as(_vp)


class Port:
    def __init__(self):
        if not available():
            raise FileNotFoundError(PORT_SO + " (run `make -C oracle port`)")
        L = ctypes.CDLL(PORT_SO)
        L.zor_fr_signed_digit.restype = ctypes.c_long
        L.zor_optimal_c.restype = _sz
        self.L = L

    def fq_op(self, op, a, b=None):
        out = np.zeros_like(a)
        self.L.zor_fq_op(op, _p(a), _p(b), _p(out), _sz(a.shape[0]))
        return out

    def fr_op(self, op, a, b=None):
        out = np.zeros_like(a)
        self.L.zor_fr_op(op, _p(a), _p(b), _p(out), _sz(a.shape[0]))
        return out

    def group_op(self, op, a, b=None):
        out = np.zeros_like(a)
        self.L.zor_group_op(op, _p(a), _p(b), _p(out), _sz(a.shape[0]))
        return out

    def signed_digit(self, plain_limbs, c, idx):
        return int(self.L.zor_fr_signed_digit(_p(np.ascontiguousarray(plain_limbs)), _sz(c), _sz(idx)))

    def optimal_c(self, n):
        return int(self.L.zor_optimal_c(_sz(n)))

    def msm(self, bases, scalars, index=None, chunks=1, base_offset_bytes=0, n=None):
        out = np.zeros(36, dtype=np.uint64)
        n = bases.shape[0] if n is None else n
        idx = None if index is None else np.ascontiguousarray(index, dtype=np.uint64)
        self.L.zor_msm(_vp(bases.ctypes.data + base_offset_bytes), _sz(bases.strides[0]), _p(scalars), _sz(scalars.strides[0]),
                       _p(idx), _sz(n), _sz(chunks), _p(out))
        return out

    def ntt(self, a, flags=0, coset_g=None):
        out = np.ascontiguousarray(a).copy()
        rc = self.L.zor_ntt(_p(out), ctypes.c_uint(out.shape[0].bit_length() - 1), ctypes.c_uint(flags), _p(coset_g))
        if rc:
            raise ValueError("oracle NTT: bad domain size")
        return out

    def qap_h(self, aA, aB, aC, coset_g):
        m = aA.shape[0]
        h = np.zeros((m + 1, 6), dtype=np.uint64)
        rc = self.L.zor_qap_h(_p(np.ascontiguousarray(aA)), _p(np.ascontiguousarray(aB)), _p(np.ascontiguousarray(aC)),
                              ctypes.c_uint(m.bit_length() - 1), _p(coset_g), _p(h))
        assert rc == 0
        return h
