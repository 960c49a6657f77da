"""fq_inv.cuh (the device's batched-step binary-GCD inverse) compiled for the HOST with g++ and checked
against Python's pow(y, -1, q): the same text runs inside the CUDA library (VOP_INVP of the Fq VM).
Reference behaviour replaced: Fp_model::invert's mpn_gcdext (libff/algebra/fields/fp.tcc:680-727)."""
import ctypes
import os
import random
import subprocess

import numpy as np
import pytest

import bw6_constants as C

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = r'''
#include "fq_inv.cuh"
extern "C" void inv_host(const uint32_t *y, const uint32_t *P, uint32_t pinv32, uint32_t *r, int n)
{
    for (int j = 0; j < n; j++) {
        uint32_t yy[24], rr[24];
        for (int i = 0; i < 24; i++) yy[i] = y[24 * j + i];
        zecale::fq_inv_bingcd(rr, yy, P, pinv32);
        for (int i = 0; i < 24; i++) r[24 * j + i] = rr[i];
    }
}
'''


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    d = tmp_path_factory.mktemp("inv")
    src = d / "inv_host.cpp"
    src.write_text(SHIM)
    so = d / "libinv_host.so"
    subprocess.run(["/usr/bin/g++", "-O2", "-shared", "-fPIC", "-I", os.path.join(ROOT, "zecale_b200", "csrc"),
                    "-o", str(so), str(src)], check=True)
    return ctypes.CDLL(str(so))


def _limbs(x):
    return [(x >> (32 * i)) & 0xFFFFFFFF for i in range(24)]


def _val(a):
    return sum(int(a[i]) << (32 * i) for i in range(24))


def test_bingcd_inverse_matches_pow(lib):
    q = C.Q
    rnd = random.Random(7)
    ys = [0, 1, 2, 3, q - 1, q - 2, (q + 1) // 2, 1 << 760, (1 << 760) - 1, 1 << 384, 0xFFFFFFFF, 1 << 30, (1 << 30) + 1,
          (1 << 64) - 1, 1 << 64, (1 << 96) - 1]
    ys += [rnd.randrange(1, q) for _ in range(1500)]
    ys += [rnd.randrange(1, 1 << rnd.randrange(1, 761)) for _ in range(1000)]   # every bit length
    y = np.array([_limbs(v) for v in ys], dtype=np.uint32)
    p = np.array(_limbs(q), dtype=np.uint32)
    r = np.zeros_like(y)
    vp = ctypes.c_void_p
    lib.inv_host(y.ctypes.data_as(vp), p.ctypes.data_as(vp), ctypes.c_uint32(C.FQ["inv32"]), r.ctypes.data_as(vp), len(ys))
    for i, v in enumerate(ys):
        assert _val(r[i]) == (pow(v, -1, q) if v else 0), hex(v)
