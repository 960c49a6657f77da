"""ctypes binding of the C ABI in include/zecale_gpu.h.

The product path is the CUDA library only: if libzecale_gpu.so is missing this
module raises at import, and every call raises ZecaleGpuError on a non-zero
status. There is no CPU fallback anywhere in this package.

Array conventions (numpy, little-endian host = libff's in-memory layout):
  Fr vector   : uint64 array of shape (n, 6)     Montgomery form
  points      : uint64 array of shape (n, 36)    {X, Y, Z}, 12 limbs each, special form
  kc pairs    : uint64 array of shape (n, 72)    {G2 point, G1 point}
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# ZECALE_GPU_LIB selects another build of the SAME library (the range-checked one, libzecale_gpu_checked.so)
LIB_PATH = os.environ.get("ZECALE_GPU_LIB") or os.path.join(_HERE, "libzecale_gpu.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        "zecale_b200: %s not found — build it with `python -m zecale_b200.build` "
        "(nvcc, sm_100a). There is no CPU fallback." % LIB_PATH)

_lib = ctypes.CDLL(LIB_PATH)

_vp = ctypes.c_void_p
_sz = ctypes.c_size_t
_u = ctypes.c_uint
_i = ctypes.c_int

NTT_INVERSE = 1
NTT_COSET = 2
G1 = 1
G2 = 2

EXPORTS = {
    "zecale_gpu_init": (_i, [ctypes.POINTER(_i), _i, ctypes.POINTER(_vp)]),
    "zecale_gpu_device_count": (_i, [_vp]),
    "zecale_gpu_shutdown": (None, [_vp]),
    "zecale_gpu_last_error": (ctypes.c_char_p, [_vp]),
    "zecale_gpu_msm_g1": (_i, [_vp, _vp, _sz, _vp, _sz, _vp, _sz, _vp]),
    "zecale_gpu_msm_g2": (_i, [_vp, _vp, _sz, _vp, _sz, _vp, _sz, _vp]),
    "zecale_gpu_ntt": (_i, [_vp, _vp, _u, _u, _vp]),
    "zecale_gpu_pin_bases": (_i, [_vp, _vp, _sz, _sz, _i, ctypes.POINTER(_vp)]),
    "zecale_gpu_pin_bases_ex": (_i, [_vp, _vp, _sz, _sz, _i, _i, ctypes.POINTER(_vp)]),
    "zecale_gpu_msm_last_stats": (_i, [_vp, ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64),
                                  ctypes.POINTER(_u), ctypes.POINTER(_u), ctypes.POINTER(_u)]),
    "zecale_gpu_unpin_bases": (_i, [_vp, _vp]),
    "zecale_gpu_msm_pinned": (_i, [_vp, _vp, _vp, _sz, _vp, _sz, _vp]),
    "zecale_gpu_msm_pinned_dev": (_i, [_vp, _vp, _vp, _sz, _vp, _sz, _vp, ctypes.POINTER(ctypes.c_float)]),
    "zecale_gpu_ntt_dev": (_i, [_vp, _vp, _u, _u, _vp, ctypes.POINTER(ctypes.c_float)]),
    "zecale_gpu_ntt_static": (_i, [_vp, _vp, _u, _u, _vp, ctypes.c_char_p, _sz]),
    "zecale_gpu_qap_h": (_i, [_vp, _vp, _vp, _vp, _u, _vp, _vp]),
    "zecale_gpu_qap_h_dev": (_i, [_vp, _vp, _vp, _vp, _u, _vp, ctypes.POINTER(ctypes.c_float)]),
    "zecale_gpu_point_sum": (_i, [_vp, _vp, _sz, _vp]),
    "zecale_gpu_ntt_dist_phase": (_i, [_vp, _vp, _vp, _u, _u, _vp, _i, _i, _i, _i]),
    "zecale_gpu_qap_pointwise_dev": (_i, [_vp, _vp, _vp, _vp, _u, _vp, _sz]),
    "zecale_gpu_ntt_dist_phase_on": (_i, [_vp, _vp, _vp, _vp, _u, _u, _vp, _i, _i, _i]),
    "zecale_gpu_qap_pointwise_dev_on": (_i, [_vp, _vp, _vp, _vp, _vp, _u, _vp, _sz]),
    "zecale_gpu_r1cs_load": (_i, [_vp, _sz, _sz, _sz, ctypes.POINTER(_vp), ctypes.POINTER(_vp), ctypes.POINTER(_vp),
                                  ctypes.POINTER(_vp)]),
    "zecale_gpu_r1cs_free": (_i, [_vp, _vp]),
    "zecale_gpu_r1cs_log_domain": (_u, [_vp]),
    "zecale_gpu_qap_h_from_witness_async": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _u]),
    "zecale_gpu_lane_count": (_i, []),
    "zecale_gpu_msm_pinned_async": (_i, [_vp, _i, _vp, _vp, _sz, _vp, _sz, _vp, _u]),
    "zecale_gpu_qap_h_async": (_i, [_vp, _i, _vp, _vp, _vp, _u, _vp, _vp]),
    "zecale_gpu_sync": (_i, [_vp]),
    "zecale_gpu_flush": (_i, [_vp]),
    "zecale_gpu_host_alloc": (_vp, [_sz]),
    "zecale_gpu_host_free": (None, [_vp]),
    "zecale_gpu_synth_bases": (_i, [_vp, _i, ctypes.c_uint64, _sz, _vp]),
    "zecale_gpu_pin_bases_bytes": (_i, [_vp, _vp, _sz, _sz, _i, _i, ctypes.POINTER(_vp)]),
    "zecale_gpu_batch_exp": (_i, [_vp, _i, _vp, _vp, _sz, _sz, _vp, _vp]),
    "zecale_gpu_lagrange": (_i, [_vp, _u, _vp, _vp]),
    "zecale_gpu_kernel_ms": (_i, [_vp, _i, _i, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_uint64)]),
    "zecale_gpu_launch_count": (ctypes.c_uint64, [_vp]),
    "zecale_gpu_msm_window_bits": (_u, [_sz]),
    "zecale_gpu_msm_window_bits_pinned": (_u, [_sz]),
    "zecale_gpu_test_fq": (_i, [_vp, _i, _vp, _vp, _vp, _sz]),
    "zecale_gpu_test_fr": (_i, [_vp, _i, _vp, _vp, _vp, _sz]),
    "zecale_gpu_test_group": (_i, [_vp, _i, _vp, _vp, _vp, _sz]),
    "zecale_gpu_bench_fq_mul": (_i, [_vp, _u, _u, _u, ctypes.POINTER(ctypes.c_float)]),
    "zecale_gpu_bench_imad": (_i, [_vp, _u, _u, _u, ctypes.POINTER(ctypes.c_float)]),
}

for _name, (_res, _args) in EXPORTS.items():
    _f = getattr(_lib, _name)  # AttributeError here == header/library mismatch
    _f.restype = _res
    _f.argtypes = _args


class ZecaleGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("zecale_gpu error %d: %s" % (code, msg))
        self.code = code


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, int):
        return ctypes.c_void_p(a)
    assert a.flags["C_CONTIGUOUS"], "array must be C-contiguous"
    return a.ctypes.data_as(ctypes.c_void_p)


NTT_STATIC_MAX_LOG_M = 16


def ntt_static(d_ptr, log_m, flags=0, coset_g=None, stream=0):
    """zecale_gpu_ntt_static: the transform of Context.ntt_dev for callers that may not allocate device memory
    (no context, working memory inside the library image, log_m <= NTT_STATIC_MAX_LOG_M). d_ptr: device address of
    2^log_m Fr, transformed in place on `stream` (a cudaStream_t value, 0 = default stream); not synchronised."""
    err = ctypes.create_string_buffer(512)
    rc = _lib.zecale_gpu_ntt_static(ctypes.c_void_p(stream), ctypes.c_void_p(d_ptr), log_m, flags, _ptr(coset_g), err, len(err))
    if rc != 0:
        raise ZecaleGpuError(rc, err.value.decode())


class PinnedBases:
    def __init__(self, ctx, handle, n, group):
        self.ctx, self.handle, self.n, self.group = ctx, handle, n, group

    def release(self):
        if self.handle is not None:
            _lib.zecale_gpu_unpin_bases(self.ctx._h, self.handle)
            self.handle = None


class Context:
    """One zecale_gpu_ctx (one device, one stream)."""

    def __init__(self, device=None):
        """device: None (current), an int, or a list of ints (multi-device context: pinned MSMs are sharded)."""
        h = ctypes.c_void_p()
        if device is None:
            rc = _lib.zecale_gpu_init(None, 0, ctypes.byref(h))
        else:
            devs = list(device) if isinstance(device, (list, tuple)) else [device]
            arr = (ctypes.c_int * len(devs))(*devs)
            rc = _lib.zecale_gpu_init(arr, len(devs), ctypes.byref(h))
        if rc != 0:
            raise ZecaleGpuError(rc, "zecale_gpu_init failed (no usable sm_100 device?)")
        self._h = h

    def device_count(self):
        return int(_lib.zecale_gpu_device_count(self._h))

    def close(self):
        if self._h is not None:
            _lib.zecale_gpu_shutdown(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc != 0:
            raise ZecaleGpuError(rc, _lib.zecale_gpu_last_error(self._h).decode())

    # -- MSM ------------------------------------------------------------------------
    def _msm(self, fn, bases, scalars, index, n, base_offset_bytes):
        out = np.zeros(36, dtype=np.uint64)
        n = int(bases.shape[0] if n is None else n)
        bstride = bases.strides[0] if bases.ndim == 2 else 288
        sstride = scalars.strides[0] if scalars.ndim == 2 else 48
        bptr = bases.ctypes.data + base_offset_bytes if n else 0
        idx = None
        if index is not None:
            idx = np.ascontiguousarray(index, dtype=np.uint64)
        self._ck(fn(self._h, ctypes.c_void_p(bptr), bstride, _ptr(scalars) if n else None, sstride, _ptr(idx), n, _ptr(out)))
        return out

    def msm_g1(self, bases, scalars, index=None, n=None, base_offset_bytes=0):
        return self._msm(_lib.zecale_gpu_msm_g1, bases, scalars, index, n, base_offset_bytes)

    def msm_g2(self, bases, scalars, index=None, n=None, base_offset_bytes=0):
        return self._msm(_lib.zecale_gpu_msm_g2, bases, scalars, index, n, base_offset_bytes)

    def pin_bases(self, bases, group=G1, base_offset_bytes=0, expand=True):
        h = ctypes.c_void_p()
        n = int(bases.shape[0])
        self._ck(_lib.zecale_gpu_pin_bases_ex(self._h, ctypes.c_void_p(bases.ctypes.data + base_offset_bytes),
                                              bases.strides[0], n, group, 1 if expand else 0, ctypes.byref(h)))
        return PinnedBases(self, h, n, group)

    def msm_last_stats(self):
        e, t = ctypes.c_uint64(), ctypes.c_uint64()
        c, v, g = _u(), _u(), _u()
        self._ck(_lib.zecale_gpu_msm_last_stats(self._h, ctypes.byref(e), ctypes.byref(t), ctypes.byref(c),
                                                ctypes.byref(v), ctypes.byref(g)))
        return {"entries": int(e.value), "tasks": int(t.value), "c": int(c.value), "bucket_sets": int(v.value),
                "base_copies": int(g.value)}

    def msm_pinned(self, pinned, scalars, index=None, n=None):
        out = np.zeros(36, dtype=np.uint64)
        n = pinned.n if n is None else int(n)
        idx = None if index is None else np.ascontiguousarray(index, dtype=np.uint64)
        self._ck(_lib.zecale_gpu_msm_pinned(self._h, pinned.handle, _ptr(scalars), scalars.strides[0], _ptr(idx), n, _ptr(out)))
        return out

    def msm_pinned_dev(self, pinned, d_scalars_ptr, n, scalar_stride=48, d_index_ptr=None):
        """Scalars already on the device (raw pointer). Returns (point, kernel ms)."""
        out = np.zeros(36, dtype=np.uint64)
        ms = ctypes.c_float()
        self._ck(_lib.zecale_gpu_msm_pinned_dev(self._h, pinned.handle, ctypes.c_void_p(d_scalars_ptr), scalar_stride,
                                                ctypes.c_void_p(d_index_ptr) if d_index_ptr else None, int(n), _ptr(out),
                                                ctypes.byref(ms)))
        return out, ms.value

    # -- NTT ------------------------------------------------------------------------
    def ntt(self, a, flags=0, coset_g=None):
        """In place on a host (m, 6) uint64 array; m must be a power of two."""
        m = a.shape[0]
        log_m = m.bit_length() - 1
        if m == 0 or (1 << log_m) != m:
            # libfqfft: basic_radix2_domain throws DomainSizeException (basic_radix2_domain.tcc:42-57)
            raise ValueError("NTT size must be a power of two")
        self._ck(_lib.zecale_gpu_ntt(self._h, _ptr(a), log_m, flags, _ptr(coset_g)))
        return a

    def ntt_dev(self, d_ptr, log_m, flags=0, coset_g=None):
        ms = ctypes.c_float()
        self._ck(_lib.zecale_gpu_ntt_dev(self._h, ctypes.c_void_p(d_ptr), log_m, flags, _ptr(coset_g), ctypes.byref(ms)))
        return ms.value

    def qap_h(self, aA, aB, aC, coset_g):
        """coefficients_for_H (m+1, 6) from the three evaluation vectors (host arrays, not modified)."""
        m = aA.shape[0]
        log_m = m.bit_length() - 1
        if m == 0 or (1 << log_m) != m or aB.shape[0] != m or aC.shape[0] != m:
            raise ValueError("qap_h: vectors must have the same power-of-two length")
        out = np.zeros((m + 1, 6), dtype=np.uint64)
        self._ck(_lib.zecale_gpu_qap_h(self._h, _ptr(aA), _ptr(aB), _ptr(aC), log_m, _ptr(coset_g), _ptr(out)))
        return out

    def qap_h_dev(self, d_aA, d_aB, d_aC, log_m, coset_g):
        ms = ctypes.c_float()
        self._ck(_lib.zecale_gpu_qap_h_dev(self._h, ctypes.c_void_p(d_aA), ctypes.c_void_p(d_aB), ctypes.c_void_p(d_aC),
                                           log_m, _ptr(coset_g), ctypes.byref(ms)))
        return ms.value

    # -- R1CS image / H polynomial from the witness ---------------------------------------
    def r1cs_load(self, n_constraints, n_inputs, n_variables, csr):
        """csr: three (row_ptr uint64, col uint32, val (nnz, 6) uint64) tuples for A, B, C."""
        keep = [(np.ascontiguousarray(rp, dtype=np.uint64), np.ascontiguousarray(c, dtype=np.uint32),
                 np.ascontiguousarray(v, dtype=np.uint64)) for rp, c, v in csr]
        arr = lambda k: (_vp * 3)(*[t[k].ctypes.data for t in keep])
        h = ctypes.c_void_p()
        self._ck(_lib.zecale_gpu_r1cs_load(self._h, n_constraints, n_inputs, n_variables, arr(0), arr(1), arr(2), ctypes.byref(h)))
        return h

    def r1cs_free(self, handle):
        _lib.zecale_gpu_r1cs_free(self._h, handle)

    def r1cs_log_domain(self, handle):
        return int(_lib.zecale_gpu_r1cs_log_domain(handle))

    def qap_h_from_witness_async(self, lane, r1cs, assignment, coset_g, out=None, device=False):
        a = ctypes.c_void_p(assignment) if device else _ptr(assignment)
        self._ck(_lib.zecale_gpu_qap_h_from_witness_async(self._h, lane, r1cs, a, _ptr(coset_g), _ptr(out), 1 if device else 0))

    # -- multi-GPU building blocks ---------------------------------------------------------
    def point_sum(self, points):
        """Sum of (n, 36) projective points, special form out."""
        out = np.zeros(36, dtype=np.uint64)
        pts = np.ascontiguousarray(points)
        self._ck(_lib.zecale_gpu_point_sum(self._h, _ptr(pts) if pts.shape[0] else None, pts.shape[0], _ptr(out)))
        return out

    def ntt_dist_phase(self, d_in, d_out, log_m, flags, coset_g, world, rank, phase, sync=True):
        self._ck(_lib.zecale_gpu_ntt_dist_phase(self._h, ctypes.c_void_p(d_in), ctypes.c_void_p(d_out), log_m, flags,
                                                _ptr(coset_g), world, rank, phase, 1 if sync else 0))

    def ntt_dist_phase_on(self, stream, d_in, d_out, log_m, flags, coset_g, world, rank, phase):
        """The phase on the caller's CUDA stream (raw handle), not synchronised."""
        self._ck(_lib.zecale_gpu_ntt_dist_phase_on(self._h, ctypes.c_void_p(stream), ctypes.c_void_p(d_in), ctypes.c_void_p(d_out),
                                                   log_m, flags, _ptr(coset_g), world, rank, phase))

    def qap_pointwise_dev_on(self, stream, d_aA, d_aB, d_aC, log_m, coset_g, count):
        self._ck(_lib.zecale_gpu_qap_pointwise_dev_on(self._h, ctypes.c_void_p(stream), ctypes.c_void_p(d_aA), ctypes.c_void_p(d_aB),
                                                      ctypes.c_void_p(d_aC), log_m, _ptr(coset_g), count))

    def qap_pointwise_dev(self, d_aA, d_aB, d_aC, log_m, coset_g, count):
        self._ck(_lib.zecale_gpu_qap_pointwise_dev(self._h, ctypes.c_void_p(d_aA), ctypes.c_void_p(d_aB), ctypes.c_void_p(d_aC),
                                                   log_m, _ptr(coset_g), count))

    # -- asynchronous lanes -----------------------------------------------------------
    def msm_pinned_async(self, lane, pinned, scalars, out, n=None, index=None, device=False, from_qap=False):
        """Enqueue an MSM on `lane`; `out` (uint64[36]) is filled by sync(). scalars/index: numpy arrays
        (kept alive by the caller until sync) or raw device pointers when device=True."""
        n = pinned.n if n is None else int(n)
        flags = (1 if device else 0) | (2 if from_qap else 0)
        if from_qap:
            sp, ip, stride = None, None, 48
        elif device:
            sp, ip, stride = ctypes.c_void_p(scalars), (ctypes.c_void_p(index) if index else None), 48
        else:
            sp, ip, stride = _ptr(scalars), _ptr(index), scalars.strides[0]
        self._ck(_lib.zecale_gpu_msm_pinned_async(self._h, lane, pinned.handle, sp, stride, ip, n, _ptr(out), flags))

    def qap_h_async(self, lane, aA, aB, aC, coset_g, out=None, log_m=None):
        """aA/aB/aC: numpy arrays, or raw pointers (host or device) with log_m given."""
        if log_m is None:
            log_m = aA.shape[0].bit_length() - 1
        self._ck(_lib.zecale_gpu_qap_h_async(self._h, lane, _ptr(aA), _ptr(aB), _ptr(aC), log_m, _ptr(coset_g), _ptr(out)))

    def sync(self):
        self._ck(_lib.zecale_gpu_sync(self._h))

    def flush(self):
        """Enqueue the deferred MSM tails of the async calls so far, without waiting."""
        self._ck(_lib.zecale_gpu_flush(self._h))

    def synth_bases(self, n, group=G1, seed=0):
        out = np.zeros((n, 36), dtype=np.uint64)
        self._ck(_lib.zecale_gpu_synth_bases(self._h, group, seed, n, _ptr(out)))
        return out

    def pin_bases_bytes(self, buf, offset, stride, n, group=G1, expand=True):
        """Pin n points from key-file bytes (numpy uint8 buffer): X || Y big-endian plain, `stride` apart."""
        h = ctypes.c_void_p()
        self._ck(_lib.zecale_gpu_pin_bases_bytes(self._h, ctypes.c_void_p(buf.ctypes.data + offset), stride, n, group,
                                                 1 if expand else 0, ctypes.byref(h)))
        return PinnedBases(self, h, n, group)

    def batch_exp(self, base, scalars, group=G1, coeff=None):
        """out[i] = (coeff * scalars[i]) * base (libff::batch_exp + batch_to_special)."""
        scalars = np.ascontiguousarray(scalars)
        n = scalars.shape[0]
        out = np.zeros((n, 36), dtype=np.uint64)
        base = np.ascontiguousarray(base)
        cf = np.ascontiguousarray(coeff) if coeff is not None else None
        self._ck(_lib.zecale_gpu_batch_exp(self._h, group, _ptr(base), _ptr(scalars) if n else None, 48, n,
                                           _ptr(cf) if cf is not None else None, _ptr(out)))
        return out

    def lagrange(self, log_m, t):
        """All Lagrange basis polynomials of the size-2^log_m domain evaluated at t: (m, 6) uint64."""
        out = np.zeros((1 << log_m, 6), dtype=np.uint64)
        self._ck(_lib.zecale_gpu_lagrange(self._h, log_m, _ptr(np.ascontiguousarray(t)), _ptr(out)))
        return out

    def kernel_ms(self, kernel_class, reset=True):
        ms = ctypes.c_double()
        cnt = ctypes.c_uint64()
        self._ck(_lib.zecale_gpu_kernel_ms(self._h, kernel_class, 1 if reset else 0, ctypes.byref(ms), ctypes.byref(cnt)))
        return ms.value, int(cnt.value)

    # -- misc -----------------------------------------------------------------------
    def launch_count(self):
        return int(_lib.zecale_gpu_launch_count(self._h))

    def test_fq(self, op, a, b=None):
        out = np.zeros_like(a)
        self._ck(_lib.zecale_gpu_test_fq(self._h, op, _ptr(a), _ptr(b), _ptr(out), a.shape[0]))
        return out

    def test_fr(self, op, a, b=None):
        out = np.zeros_like(a)
        self._ck(_lib.zecale_gpu_test_fr(self._h, op, _ptr(a), _ptr(b), _ptr(out), a.shape[0]))
        return out

    def test_group(self, op, a, b=None):
        out = np.zeros_like(a)
        self._ck(_lib.zecale_gpu_test_group(self._h, op, _ptr(a), _ptr(b), _ptr(out), a.shape[0]))
        return out

    def bench_fq_mul(self, blocks, threads_per_block, iters):
        ms = ctypes.c_float()
        self._ck(_lib.zecale_gpu_bench_fq_mul(self._h, blocks, threads_per_block, iters, ctypes.byref(ms)))
        return ms.value

    def bench_imad(self, blocks, threads_per_block, iters):
        """ms for blocks * threads_per_block threads x 8 * iters dependent IMAD.WIDE.U32 (8 independent chains per thread)."""
        ms = ctypes.c_float()
        self._ck(_lib.zecale_gpu_bench_imad(self._h, blocks, threads_per_block, iters, ctypes.byref(ms)))
        return ms.value


def msm_window_bits(n, pinned=False):
    return int((_lib.zecale_gpu_msm_window_bits_pinned if pinned else _lib.zecale_gpu_msm_window_bits)(n))
