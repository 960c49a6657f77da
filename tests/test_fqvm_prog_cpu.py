"""The Fq-VM micro-programs (tools/gen_fqvm_prog.py -> csrc/fqvm_prog.cuh) interpreted on Python integers and compared with
the affine group law of y^2 = x^3 + b over Fq: mixed addition in its three variants (madd, madd_acc, madd_aff), general
addition, doubling, normalisation, with the special cases libff's bw6_761_G1/G2::mixed_add / add / dbl handle
(bw6_761_g1.cpp:111-358: an operand at infinity, P == Q -> doubling, P == -Q -> infinity). The interpreter decodes the SAME
code words the device executes (op | d << 8 | imm16 << 16); field values are plain residues, which is what the Montgomery
form is up to the fixed factor R."""
import random

import pytest

import bw6_constants as C
import gen_fqvm_prog as G

Q = C.Q
CODE, LABELS = G.assemble(G.PROGRAM)
OPS = {v: k for k, v in G.OPS.items()}
S = G.SLOTS


def run(pc, slots, max_steps=200000):
    """Runs until a HOOK (returns its id) or HALT. `slots`: dict slot -> int."""
    ret = cnt = 0
    for _ in range(max_steps):
        u = CODE[pc]
        pc += 1
        op, d, imm = OPS[u & 0xFF], (u >> 8) & 0xFF, u >> 16
        a, b = imm & 0xFF, imm >> 8
        g = lambda s: slots.get(s, 0)
        if op == "MUL":
            slots[d] = g(a) * g(b) % Q
        elif op == "ADD":
            slots[d] = (g(a) + g(b)) % Q
        elif op == "SUB":
            slots[d] = (g(a) - g(b)) % Q
        elif op == "MULW":
            slots["wide"] = g(a) * g(b)          # unreduced, kept apart (the W0/W1 slot pair)
        elif op == "MULFIN":
            slots[d] = (g(a) * g(b) - slots.pop("wide")) % Q
        elif op == "MOV":
            slots[d] = g(imm)
        elif op == "ONE":
            slots[d] = 1
        elif op == "ZERO":
            slots[d] = 0
        elif op == "JZ":
            if g(d) == 0:
                pc = imm
        elif op == "JMP":
            pc = imm
        elif op == "CALL":
            ret, pc = pc, imm
        elif op == "RET":
            pc = ret
        elif op == "INVP":
            slots[d] = pow(g(imm), -1, Q) if g(imm) else 0   # the device returns a^-1 R^-1; LDR3 + MUL restore the form
        elif op == "LDR3":
            slots[d] = 1                                     # ... so R^3 is 1 on plain residues
        elif op == "SETCNT":
            cnt = imm
        elif op == "DECJNZ":
            if cnt:
                cnt -= 1
                pc = imm
        elif op == "JBIT":
            if ((Q - 2) >> cnt) & 1:
                pc = imm
        elif op == "HOOK":
            return d
        else:
            return -1
    raise AssertionError("program did not finish")


# ---- reference: affine group law, None = infinity
def aff_add(P, Qp):
    if P is None:
        return Qp
    if Qp is None:
        return P
    (x1, y1), (x2, y2) = P, Qp
    if x1 == x2:
        if (y1 + y2) % Q == 0:
            return None
        lam = 3 * x1 * x1 * pow(2 * y1, -1, Q) % Q
    else:
        lam = (y2 - y1) * pow(x2 - x1, -1, Q) % Q
    x3 = (lam * lam - x1 - x2) % Q
    return x3, (lam * (x1 - x3) - y1) % Q


def aff_mul(k, P):
    R = None
    while k:
        if k & 1:
            R = aff_add(R, P)
        P = aff_add(P, P)
        k >>= 1
    return R


def to_xyzz(P, rnd):
    """A random XYZZ representative of an affine point (infinity: ZZ = ZZZ = 0, Y = 1 as vs_acc_inf leaves it)."""
    if P is None:
        return 0, 1, 0, 0
    z = rnd.randrange(1, Q)
    return P[0] * z * z % Q, P[1] * z * z * z % Q, z * z % Q, z * z * z % Q


def from_xyzz(x, y, zz, zzz):
    if zz == 0:
        return None
    return x * pow(zz, -1, Q) % Q, y * pow(zzz, -1, Q) % Q


def acc_of(slots, names=("X", "Y", "ZZ", "ZZZ")):
    return from_xyzz(*(slots.get(S[n], 0) for n in names))


def set_acc(slots, P, rnd, names=("X", "Y", "ZZ", "ZZZ")):
    for n, v in zip(names, to_xyzz(P, rnd)):
        slots[S[n]] = v


@pytest.fixture(scope="module", params=[C.G1_GEN, C.G2_GEN], ids=["G1", "G2"])
def points(request):
    g = request.param
    rnd = random.Random(11)
    return [aff_mul(rnd.randrange(1, 1 << 64), g) for _ in range(4)]


@pytest.mark.parametrize("prog", ["madd", "madd_acc"])
def test_mixed_addition_programs(points, prog):
    rnd = random.Random(5)
    P, Qp = points[0], points[1]
    neg = lambda p: (p[0], (-p[1]) % Q)
    for acc, pt in [(P, Qp), (None, Qp), (P, P), (P, neg(P)), (aff_add(P, Qp), Qp), (aff_add(P, P), P)]:
        slots = {S["QX"]: pt[0], S["QY"]: pt[1]}
        set_acc(slots, acc, rnd)
        assert run(LABELS[prog], slots) == 0
        assert acc_of(slots) == aff_add(acc, pt), (prog, acc is None, acc == pt)


def test_affine_accumulator_program(points):
    """madd_aff: the accumulator is (X, Y) with ZZ = ZZZ = 1, as madd_accinf leaves it after the first point of a chain."""
    P, Qp = points[0], points[1]
    neg = lambda p: (p[0], (-p[1]) % Q)
    for pt in (Qp, P, neg(P)):
        slots = {S["QX"]: P[0], S["QY"]: P[1], S["X"]: 0, S["Y"]: 1, S["ZZ"]: 0, S["ZZZ"]: 0}
        assert run(LABELS["madd"], slots) == 0                  # first point: madd_accinf
        assert (slots[S["ZZ"]], slots[S["ZZZ"]]) == (1, 1)
        slots[S["QX"]], slots[S["QY"]] = pt
        assert run(LABELS["madd_aff"], slots) == 0
        assert acc_of(slots) == aff_add(P, pt)


def test_a_chain_like_k_accumulate(points):
    """first point -> madd (accinf), second -> madd_aff, the rest -> madd_acc: the sequence k_accumulate's hook issues."""
    rnd = random.Random(9)
    chain = [points[i % 4] for i in range(3)] + [points[0], (points[1][0], (-points[1][1]) % Q)] + [aff_mul(rnd.randrange(2, 99), points[2])
                                                                                                       for _ in range(4)]
    slots = {S["X"]: 0, S["Y"]: 1, S["ZZ"]: 0, S["ZZZ"]: 0}
    exp = None
    for i, pt in enumerate(chain):
        slots[S["QX"]], slots[S["QY"]] = pt
        assert run(LABELS["madd" if i == 0 else "madd_aff" if i == 1 else "madd_acc"], slots) == 0
        exp = aff_add(exp, pt)
        assert acc_of(slots) == exp


@pytest.mark.parametrize("prog,a,b", [("add", ("X", "Y", "ZZ", "ZZZ"), ("QX", "QY", "QZZ", "QZZZ")),
                                      ("add_r", ("RX", "RY", "RZZ", "RZZZ"), ("X", "Y", "ZZ", "ZZZ"))])
def test_general_addition_programs(points, prog, a, b):
    rnd = random.Random(6)
    P, Qp = points[2], points[3]
    neg = lambda p: (p[0], (-p[1]) % Q)
    for x, y in [(P, Qp), (None, Qp), (P, None), (None, None), (P, P), (P, neg(P))]:
        slots = {}
        set_acc(slots, x, rnd, a)
        set_acc(slots, y, rnd, b)
        assert run(LABELS[prog], slots) == 0
        assert acc_of(slots, a) == aff_add(x, y)


def test_doubling_and_normalisation(points):
    rnd = random.Random(7)
    P = points[1]
    for prog, names in (("dbl", ("X", "Y", "ZZ", "ZZZ")), ("dbl_r", ("RX", "RY", "RZZ", "RZZZ"))):
        for x in (P, None):
            slots = {}
            set_acc(slots, x, rnd, names)
            assert run(LABELS[prog], slots) == 0
            assert acc_of(slots, names) == aff_add(x, x)
    slots = {S["QX"]: P[0], S["QY"]: P[1]}
    assert run(LABELS["mdbl"], slots) == 0
    assert acc_of(slots) == aff_add(P, P)
    slots = {}
    set_acc(slots, P, rnd)
    assert run(LABELS["toaff"], slots) == 1
    assert (slots[S["QX"]], slots[S["QY"]]) == P
    slots = {}
    set_acc(slots, None, rnd)
    assert run(LABELS["toaff"], slots) == 2
