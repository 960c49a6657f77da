"""Build libzecale_gpu.so in-tree with nvcc for sm_100a (no JIT cache, no torch extension)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "zecale_gpu.cu")
LIB = os.path.join(HERE, "libzecale_gpu.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-diag-suppress", "1160",
    "-shared", "-Xcompiler", "-fPIC",
]


def sources():
    d = os.path.join(HERE, "csrc")
    inc = os.path.join(HERE, "..", "include", "zecale_gpu.h")
    return [os.path.join(d, f) for f in sorted(os.listdir(d)) if f.endswith((".cu", ".cuh"))] + [inc]


def is_stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(s) > t for s in sources())


LIB_CHECKED = os.path.join(HERE, "libzecale_gpu_checked.so")


def build(force=False, verbose=False, checked=False):
    """Compile the CUDA library if it is missing or older than its sources. checked=True builds
    libzecale_gpu_checked.so with -DZECALE_GPU_CHECKED (device-side range checks on every index, fqvm.cuh ZG_CHECK);
    load it with ZECALE_GPU_LIB=<path> (zecale_b200/capi.py)."""
    if checked:
        cmd = [NVCC] + FLAGS + ["-DZECALE_GPU_CHECKED", "-o", LIB_CHECKED, SRC]
        env = dict(os.environ)
        env.pop("CXX", None)
        env.pop("CC", None)
        r = subprocess.run(cmd, capture_output=True, text=True, env=env)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
        return LIB_CHECKED
    if not force and not is_stale():
        return LIB
    cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB, SRC]
    env = dict(os.environ)
    env.pop("CXX", None)
    env.pop("CC", None)
    r = subprocess.run(cmd, capture_output=True, text=True, env=env)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, checked="--checked" in sys.argv))
