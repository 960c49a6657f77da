"""The C-ABI library loads and exports every symbol include/zecale_gpu.h declares, the ctypes
binding covers exactly that set, and without a GPU every entry point fails loudly (no CPU
fallback). No compute calls are made here."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "zecale_gpu.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(zecale_gpu_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_north_star_entry_points():
    names = declared_functions()
    for must in ("zecale_gpu_msm_g1", "zecale_gpu_msm_g2", "zecale_gpu_ntt", "zecale_gpu_init", "zecale_gpu_shutdown",
                 "zecale_gpu_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    import zecale_b200 as z
    lib = ctypes.CDLL(z.LIB_PATH)
    for name in declared_functions():
        assert hasattr(lib, name), "libzecale_gpu.so does not export " + name


def test_binding_covers_the_header():
    import zecale_b200 as z
    assert sorted(z.EXPORTS) == declared_functions()


def test_header_is_plain_c():
    """The boundary is a C ABI: the header must compile as C (no C++ / torch types)."""
    import subprocess
    r = subprocess.run(["gcc", "-std=c99", "-fsyntax-only", "-Wall", "-Werror", "-x", "c", HEADER], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_no_cpu_fallback_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import zecale_b200 as z
    with pytest.raises(z.ZecaleGpuError) as e:
        z.Context()
    assert e.value.code == 2  # ZECALE_GPU_E_CUDA


def test_product_does_not_import_the_oracle():
    """zecale_b200/ must never reach into oracle/ (the oracle is test infrastructure)."""
    pkg = os.path.join(ROOT, "zecale_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle" not in txt.replace("# oracle-free", ""), os.path.join(dp, f)
