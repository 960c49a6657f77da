"""zecale_b200 — B200 (sm_100a) implementation of zecale's Groth16/BW6-761 prove hot path.

Only what the path needs lives here: csrc/ (CUDA kernels + the C ABI of
include/zecale_gpu.h), hooks/ (the C++ template specialisations that route the
reference's libff/libfqfft/libsnark call sites to the C ABI) and a thin ctypes
binding used by tests and bench.py.
"""
from .capi import (  # noqa: F401
    Context,
    PinnedBases,
    ZecaleGpuError,
    msm_window_bits,
    ntt_static,
    NTT_STATIC_MAX_LOG_M,
    NTT_INVERSE,
    NTT_COSET,
    G1,
    G2,
    LIB_PATH,
    EXPORTS,
)
