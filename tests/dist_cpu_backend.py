"""CPU stand-in for zecale_gpu_ntt_dist_phase, used only by the gloo tests: an independent
restatement of the two local phases of the distributed 4-step transform (layouts as documented in
include/zecale_gpu.h) on top of the C oracle's small NTTs. TEST INFRASTRUCTURE."""
import numpy as np

from oracle import port as P
from zecale_b200 import dist as D

_port = None


def port():
    global _port
    if _port is None:
        _port = P.Port()
    return _port


def powers_of_omega(log_m):
    """w_m^k for k < m: the forward transform of the unit vector e_1."""
    m = 1 << log_m
    e1 = np.zeros((m, 6), dtype=np.uint64)
    e1[1] = port().fr_op(7, np.array([[1, 0, 0, 0, 0, 0]], dtype=np.uint64))[0]
    return port().ntt(e1, 0)


def coset_powers(g, m, inverse):
    """g^i (or g^-i) for i < m, Montgomery."""
    pt = port()
    base = g.reshape(1, 6)
    if inverse:
        base = pt.fr_op(4, base)
    out = np.zeros((m, 6), dtype=np.uint64)
    out[0] = pt.fr_op(7, np.array([[1, 0, 0, 0, 0, 0]], dtype=np.uint64))[0]
    for i in range(1, m):
        out[i] = pt.fr_op(2, out[i - 1:i], base)[0]
    return out


def phase_fn(src, dst, log_m, flags, coset_g, world, rank, ph):
    pt = port()
    s = D.four_step_shape(log_m, world)
    m, m1, m2, c1, c2 = 1 << log_m, s["m1"], s["m2"], s["c1"], s["c2"]
    inverse, coset = flags & 1, (flags >> 1) & 1
    a = src.numpy().view(np.uint64).reshape(-1, 6)
    w = powers_of_omega(log_m)
    out = np.zeros_like(a)
    mul = lambda x, y: pt.fr_op(2, np.ascontiguousarray(x), np.ascontiguousarray(y))
    if not inverse and ph == 0:
        x = a.reshape(m1, c2, 6).copy()
        if coset:
            gp = coset_powers(coset_g, m, False)
            for j2l in range(c2):
                j = np.arange(m1) * m2 + rank * c2 + j2l
                x[:, j2l] = mul(x[:, j2l], gp[j])
        y = np.zeros_like(x)
        for j2l in range(c2):
            col = pt.ntt(np.ascontiguousarray(x[:, j2l]), 0)            # over j1 -> k1
            j2 = rank * c2 + j2l
            y[:, j2l] = mul(col, w[(j2 * np.arange(m1)) % m])
        out[:] = y.reshape(-1, 6)                                      # [k1][j2l]: chunk d = rows of rank d
    elif not inverse:
        x = a.reshape(world, c1, c2, 6)                                # [sender][k1l][j2l]
        z = np.zeros((c1, m2, 6), dtype=np.uint64)
        for k1l in range(c1):
            row = np.concatenate([x[g, k1l] for g in range(world)])    # j2 = g*c2 + j2l
            z[k1l] = pt.ntt(np.ascontiguousarray(row), 0)              # over j2 -> k2
        out[:] = z.reshape(-1, 6)                                      # layout B
    elif ph == 0:
        x = a.reshape(c1, m2, 6)                                       # layout B [k1l][k2]
        t = np.zeros((world, c1, c2, 6), dtype=np.uint64)
        for k1l in range(c1):
            k1 = rank * c1 + k1l
            row = pt.ntt(np.ascontiguousarray(x[k1l]), 1)              # inverse over k2 -> j2 (includes 1/m2)
            row = mul(row, w[(-(k1 * np.arange(m2))) % m])
            t[:, k1l] = row.reshape(world, c2, 6)
        out[:] = t.reshape(-1, 6)                                      # chunk d = columns of rank d
    else:
        x = a.reshape(world, c1, c2, 6)                                # [sender h][k1l][j2l]
        y = np.zeros((m1, c2, 6), dtype=np.uint64)
        gp = coset_powers(coset_g, m, True) if coset else None
        for j2l in range(c2):
            col = np.concatenate([x[h, :, j2l] for h in range(world)])  # k1 = h*c1 + k1l
            col = pt.ntt(np.ascontiguousarray(col), 1)                 # inverse over k1 -> j1 (includes 1/m1)
            if coset:
                j = np.arange(m1) * m2 + rank * c2 + j2l
                col = mul(col, gp[j])
            y[:, j2l] = col
        out[:] = y.reshape(-1, 6)                                      # layout A
    dst.numpy().view(np.uint64).reshape(-1, 6)[:] = out


def pointwise_fn(log_m, coset_g):
    """(aA*aB - aC) / (g^m - 1) on locally held elements."""
    pt = port()
    x = coset_g.reshape(1, 6).copy()
    for _ in range(log_m):
        x = pt.fr_op(3, x)
    one = pt.fr_op(7, np.array([[1, 0, 0, 0, 0, 0]], dtype=np.uint64))
    zinv = pt.fr_op(4, pt.fr_op(1, x, one))

    def f(aA, aB, aC):
        a, b, c = (t.numpy().view(np.uint64).reshape(-1, 6) for t in (aA, aB, aC))
        r = pt.fr_op(1, pt.fr_op(2, a, b), c)
        a[:] = pt.fr_op(2, r, np.repeat(zinv, r.shape[0], axis=0))
    return f
