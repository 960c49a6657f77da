"""Golden fixtures for the SURVEY 8(f3)/(f4) rows, recorded from oracle/_ref (the UNMODIFIED reference):
libff::batch_exp / batch_exp_with_coeff + batch_to_special, and the proving-key byte codec of libzeth
(group_elements_write_bytes, knowledge_commitment_vector_write_bytes) with libff::multi_exp results over
the same points. Run in the build container:

    make -C oracle ref && python tests/golden/make_golden_setup.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bw6_constants as C  # noqa: E402
from oracle import ref as R  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    ref = R.Ref()
    r = C.R
    out = {}
    edge = [0, 1, 2, r - 1, r - 2, (r - 1) // 2, 1 << 376, (1 << 12) - 1, 1 << 12, (1 << 372) + 5]
    for g in (1, 2):
        base = ref.bases_mul(ref.fr_rng(900 + g, 1), g)[0]
        scalars = np.concatenate([ref.fr_from_ints(edge), ref.fr_rng(40 + g, 38)])
        coeff = ref.fr_rng(77 + g, 1)
        out["base_g%d" % g] = base
        out["scalars_g%d" % g] = scalars
        out["coeff_g%d" % g] = coeff
        out["exp_g%d" % g] = ref.batch_exp(base, scalars, g)
        out["exp_coeff_g%d" % g] = ref.batch_exp(base, scalars, g, coeff=coeff)
    np.savez_compressed(os.path.join(HERE, "setup_batch_exp.npz"), **out)
    print("setup_batch_exp", {k: v.shape for k, v in out.items()})

    out = {}
    n = 40
    for g in (1, 2):
        bases = ref.bases_mul(ref.fr_rng(300 + g, n), g)
        bases[5, :12] = 0                      # a zero point: (0, one, 0)
        bases[5, 12:24] = bases[0, 24:36]
        bases[5, 24:36] = 0
        scalars = ref.fr_rng(310 + g, n)
        out["bases_g%d" % g] = bases
        out["scalars_g%d" % g] = scalars
        out["bytes_g%d" % g] = ref.points_write_bytes(bases, g)
        out["msm_g%d" % g] = ref.msm(bases, scalars, g)[0]
    nnz, domain = 24, 64
    g2 = ref.bases_mul(ref.fr_rng(320, nnz), 2)
    g1 = ref.bases_mul(ref.fr_rng(321, nnz), 1)
    values = np.ascontiguousarray(np.concatenate([g2, g1], axis=1))
    idx = np.sort(np.random.default_rng(9).choice(domain, nnz, replace=False)).astype(np.uint64)
    w = ref.fr_rng(322, domain)
    sc = np.ascontiguousarray(w[idx.astype(np.int64)])
    out.update(kc_bytes=ref.kc_write_bytes(values, idx, domain), kc_indices=idx, kc_witness=w,
               kc_msm_g2=ref.msm(g2, sc, 2)[0], kc_msm_g1=ref.msm(g1, sc, 1)[0])
    np.savez_compressed(os.path.join(HERE, "keyfile_bytes.npz"), **out)
    print("keyfile_bytes", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
