"""zecale_gpu_ntt_static: the NTT for host programs that are not allowed to allocate device memory.

Who needs it: a prover living inside a service whose own pool grabbed all of the GPU's memory at start (a library's
cudaMalloc then fails), or a request path that bans cudaMalloc / cudaFree because they stall the whole device. Such
a caller cannot build a zecale_gpu_ctx either, so the entry point takes no context and keeps its twiddle table,
scale table and second vector in a fixed array inside the library image (sizes up to 2^16 only).

What is checked, through the C ABI: the result equals, bit for bit (field arithmetic is exact, so the tolerance of
every NTT test here is zero), what the existing context entry point zecale_gpu_ntt gives on the same input, in all
four modes; on two calls in a row, because every call shares the one array and the second call finds the first
one's tables in it; when a large transform is followed by a small one of another mode; that a size above the limit
is refused with a message; and that a call leaves the device's free memory where it was.

This file is named to be collected last: `pytest -x` stops at the first failure, and a problem in this extra must
not hide the parity tests. It needs nothing the other tests leave behind and passes alone in a fresh process.
"""
import numpy as np
import pytest

import bw6_constants as C
import zecale_b200 as z

pytestmark = pytest.mark.gpu


def fr_vector(seed, m):
    """m pseudo-random Fr in libff's in-memory form (6 x u64, Montgomery); top limb kept below the modulus' top limb."""
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 1 << 63, size=(m, 6), dtype=np.uint64)
    a[:, 5] %= np.uint64((C.R >> 320) - 1)
    return a


def generator_montgomery():
    x = (C.FR_GENERATOR << 384) % C.R
    return np.array([(x >> (64 * i)) & ((1 << 64) - 1) for i in range(6)], dtype=np.uint64)


def run_static(a, flags, g):
    """Upload a, transform it in place with the allocation-free entry point on torch's current stream, download."""
    import torch
    d = torch.from_numpy(a.view(np.int64).copy()).cuda()
    log_m = a.shape[0].bit_length() - 1
    z.ntt_static(d.data_ptr(), log_m, flags, g if flags & 2 else None, stream=torch.cuda.current_stream().cuda_stream)
    return d.cpu().numpy().view(np.uint64)  # the copy back is ordered after the transform on the same stream


@pytest.mark.parametrize("log_m", [1, 3, 8, 11, 12, 13, 16])
def test_static_ntt_equals_context_ntt_twice_in_a_row(gpu, log_m):
    m = 1 << log_m
    g = generator_montgomery()
    for flags in (0, 1, 2, 3):
        first, second = fr_vector(1000 * log_m + flags, m), fr_vector(7777 * log_m + flags, m)
        exp_first = gpu.ntt(first.copy(), flags, g if flags & 2 else None)
        exp_second = gpu.ntt(second.copy(), flags, g if flags & 2 else None)
        assert np.array_equal(run_static(first, flags, g), exp_first), (log_m, flags, "first call")
        assert np.array_equal(run_static(second, flags, g), exp_second), (log_m, flags, "second call")


def test_static_ntt_does_not_depend_on_what_the_last_call_left():
    """A 2^16 icosetFFT fills the whole array; the 2^5 FFT, 2^12 iFFT and 2^16 cosetFFT that follow must not see it.
    Expected values come from the context path, created here so that the test also runs on its own."""
    g = generator_montgomery()
    ctx = z.Context()
    try:
        for log_m, flags in ((16, 3), (5, 0), (12, 1), (16, 2), (16, 2)):
            a = fr_vector(31 * log_m + flags, 1 << log_m)
            exp = ctx.ntt(a.copy(), flags, g if flags & 2 else None)
            assert np.array_equal(run_static(a, flags, g), exp), (log_m, flags)
    finally:
        ctx.close()


def test_static_ntt_refuses_sizes_beyond_its_array():
    import torch
    d = torch.zeros(8, dtype=torch.int64, device="cuda")
    with pytest.raises(z.ZecaleGpuError) as e:
        z.ntt_static(d.data_ptr(), z.NTT_STATIC_MAX_LOG_M + 1, 0)
    assert e.value.code == 4  # ZECALE_GPU_E_DOMAIN
    assert "2^16" in str(e.value)
    with pytest.raises(z.ZecaleGpuError) as e:
        z.ntt_static(d.data_ptr(), 4, z.NTT_COSET, None)  # coset mode without the generator
    assert e.value.code == 1  # ZECALE_GPU_E_ARG
    with pytest.raises(z.ZecaleGpuError):
        z.ntt_static(0, 4, 0)


def test_static_ntt_takes_no_device_memory():
    import torch
    a = fr_vector(5, 1 << 14)
    g = generator_montgomery()
    d = torch.from_numpy(a.view(np.int64).copy()).cuda()
    z.ntt_static(d.data_ptr(), 14, 3, g)  # first call on this device: set-up (address look-up, kernel loading)
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    z.ntt_static(d.data_ptr(), 14, 2, g)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before
    assert np.array_equal(d.cpu().numpy().view(np.uint64), a)  # icosetFFT then cosetFFT: back to the input
