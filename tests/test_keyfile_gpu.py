"""the proving key's on-disk bytes -> device image (zecale_gpu_pin_bases_bytes).
The bytes are written by the reference's own codec (libzeth group_elements_write_bytes /
knowledge_commitment_vector_write_bytes, groth16_snark.tcc:93-110) and the MSMs over the image made
from them must equal libff::multi_exp over the in-memory points, bit for bit."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _needs(ref):
    if not hasattr(ref.L, "zref_points_write_bytes"):
        pytest.skip("oracle/_ref built before the key-file codec exports were added")


def _inf_like(p):
    z = np.zeros(36, dtype=np.uint64)
    z[12:24] = p[24:36]
    return z


@pytest.mark.parametrize("group", [1, 2])
def test_query_bytes_to_device_image(gpu, ref, group):
    _needs(ref)
    n = 3000
    bases = ref.bases_mul(ref.fr_rng(55 + group, n), group)
    bases[7] = _inf_like(bases[0])          # zero points are written as (0, 1)
    bases[2999] = _inf_like(bases[0])
    scalars = ref.fr_rng(999, n)
    exp, _ = ref.msm(bases, scalars, group)
    raw = ref.points_write_bytes(bases, group)
    assert int(raw[:8].view(np.uint64)[0]) == n
    for expand in (True, False):
        pinned = gpu.pin_bases_bytes(raw, 8, 192, n, group, expand=expand)
        try:
            assert np.array_equal(gpu.msm_pinned(pinned, scalars), exp)
        finally:
            pinned.release()
    # the staging copy must not depend on the buffer's alignment in the file
    shifted = np.zeros(raw.size + 3, dtype=np.uint8)
    shifted[3:] = raw
    pinned = gpu.pin_bases_bytes(shifted, 11, 192, n, group)
    try:
        assert np.array_equal(gpu.msm_pinned(pinned, scalars), exp)
    finally:
        pinned.release()


def test_sparse_b_query_bytes(gpu, ref):
    """B query: entries of index | g (G2) | h (G1); both halves pinned from the same byte range with stride 392."""
    _needs(ref)
    nnz, domain = 1500, 4000
    g2 = ref.bases_mul(ref.fr_rng(71, nnz), 2)
    g1 = ref.bases_mul(ref.fr_rng(72, nnz), 1)
    values = np.ascontiguousarray(np.concatenate([g2, g1], axis=1))      # knowledge_commitment<G2, G1>: 576 B
    idx = np.sort(np.random.default_rng(3).choice(domain, nnz, replace=False)).astype(np.uint64)
    raw = ref.kc_write_bytes(values, idx, domain)
    hdr = raw[:16].view(np.uint64)
    assert int(hdr[0]) == domain and int(hdr[1]) == nnz
    # indices as the file holds them
    ent = raw[16:].reshape(nnz, 392)
    file_idx = np.ascontiguousarray(ent[:, :8]).view(np.uint64).reshape(nnz)
    assert np.array_equal(file_idx, idx)
    w = ref.fr_rng(4321, domain)
    sc = np.ascontiguousarray(w[idx.astype(np.int64)])
    for group, off, pts in ((2, 24, g2), (1, 216, g1)):
        exp, _ = ref.msm(pts, sc, group)
        pinned = gpu.pin_bases_bytes(raw, off, 392, nnz, group)
        try:
            assert np.array_equal(gpu.msm_pinned(pinned, sc), exp)
        finally:
            pinned.release()
