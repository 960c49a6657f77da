#!/usr/bin/env python3
"""Assemble the Fq-VM micro-programs (zecale_b200/csrc/fqvm_prog.cuh).

The EC layer of the CUDA library is written for a tiny register machine whose
registers ("slots") are 761-bit Fq elements held in per-thread local memory and
whose instructions are whole field operations (see fqvm.cuh for why). This script
holds the point-arithmetic formulas as assembly text and resolves labels.

Formulas (EFD, short Weierstrass a = 0, XYZZ coordinates x = X/ZZ, y = Y/ZZZ):
  madd-2008-s, mdbl-2008-s-1, add-2008-s, dbl-2008-s-1.
They replace libff's homogeneous-projective bw6_761_G1/G2::mixed_add / operator+ / dbl
(bw6_761_g1.cpp:111-358, bw6_761_g2.cpp:116-364) including their special cases
(either operand zero, P == Q -> doubling, P == -Q -> zero); the coordinate system is
free because parity is checked on the normalised affine point, which is unique.
"""
import os
import re

SLOTS = {"X": 0, "Y": 1, "ZZ": 2, "ZZZ": 3, "QX": 4, "QY": 5, "QZZ": 6, "QZZZ": 7,
         "T0": 8, "T1": 9, "T2": 10, "T3": 11, "T4": 12, "RX": 13, "RY": 14, "RZZ": 15, "RZZZ": 16,
         "K0": 17, "K1": 18, "W0": 19, "W1": 20}
NSLOT = 21

# MULW c, e ; MULFIN d, a, b: d = a*b - c*e with ONE Montgomery reduction. MULW leaves the unreduced 1536-bit product
# c*e in the wide slot pair (W0, W1); MULFIN subtracts it from a*b before reducing (fq_mulw_regs, fq_wsubk_regs,
# fq_redc_regs in field_asm.cuh).
OPS = {"MUL": 0, "ADD": 1, "SUB": 2, "MULW": 3, "MULFIN": 4, "MOV": 5, "ONE": 6, "ZERO": 7, "JZ": 8, "JMP": 9, "JBIT": 10,
       "DECJNZ": 11, "SETCNT": 12, "CALL": 13, "RET": 14, "HOOK": 15, "HALT": 16, "INVP": 17, "LDR3": 18}

PROGRAM = r"""
; ---------------------------------------------------------------- acc += (QX, QY) affine, not infinity
madd:
    JZ   ZZ, madd_accinf
    MUL  T0, QX, ZZ         ; U2
    MUL  T1, QY, ZZZ        ; S2
    SUB  T0, T0, X          ; P = U2 - X1
    SUB  T1, T1, Y          ; R = S2 - Y1
    JZ   T0, madd_samex
    MUL  T2, T0, T0         ; PP
    MUL  T0, T0, T2         ; PPP
    MUL  T3, X, T2          ; Q = X1*PP
    MUL  ZZ, ZZ, T2         ; ZZ3
    MUL  ZZZ, ZZZ, T0       ; ZZZ3
    MUL  T2, T1, T1         ; R^2
    SUB  T2, T2, T0
    SUB  T2, T2, T3
    SUB  X, T2, T3          ; X3 = R^2 - PPP - 2Q
    SUB  T3, T3, X
    MULW Y, T0
    MULFIN Y, T1, T3        ; Y3 = R(Q - X3) - Y1*PPP
    HOOK 0
madd_accinf:
    MOV  X, QX
    MOV  Y, QY
    ONE  ZZ
    ONE  ZZZ
    HOOK 0
; ---------------------------------------------------------------- k_accumulate's variants of madd: the point slots double as
; temporaries (QX <- U2, P, PPP; QY <- S2, R), so a chain touches 8 slots + the wide pair instead of 10 + 2 and the per-thread
; state that has to stay in L2 shrinks by a fifth. The point is gone afterwards: P == Q doubles the accumulator (dbl), which
; is the same group element as doubling the point.
madd_acc:
    JZ   ZZ, madd_accinf
    MUL  QX, QX, ZZ         ; U2
    MUL  QY, QY, ZZZ        ; S2
    SUB  QX, QX, X          ; P = U2 - X1
    SUB  QY, QY, Y          ; R = S2 - Y1
    JZ   QX, madd_acc_samex
    MUL  T2, QX, QX         ; PP
    MUL  QX, QX, T2         ; PPP
    MUL  T3, X, T2          ; Q = X1*PP
    MUL  ZZ, ZZ, T2         ; ZZ3
    MUL  ZZZ, ZZZ, QX       ; ZZZ3
    MUL  T2, QY, QY         ; R^2
    SUB  T2, T2, QX
    SUB  T2, T2, T3
    SUB  X, T2, T3          ; X3 = R^2 - PPP - 2Q
    SUB  T3, T3, X
    MULW Y, QX
    MULFIN Y, QY, T3        ; Y3 = R(Q - X3) - Y1*PPP
    HOOK 0
; the same with ZZ = ZZZ = 1 (what madd_accinf leaves: the first point of a chain): U2 = QX, S2 = QY, ZZ3 = PP, ZZZ3 = PPP -
; 6 products and 5 reductions instead of 10 and 9
madd_aff:
    SUB  QX, QX, X          ; P
    SUB  QY, QY, Y          ; R
    JZ   QX, madd_acc_samex
    MUL  ZZ, QX, QX         ; PP = ZZ3
    MUL  ZZZ, QX, ZZ        ; PPP = ZZZ3
    MUL  T3, X, ZZ          ; Q = X1*PP
    MUL  T2, QY, QY         ; R^2
    SUB  T2, T2, ZZZ
    SUB  T2, T2, T3
    SUB  X, T2, T3          ; X3 = R^2 - PPP - 2Q
    SUB  T3, T3, X
    MULW Y, ZZZ
    MULFIN Y, QY, T3        ; Y3 = R(Q - X3) - Y1*PPP
    HOOK 0
madd_acc_samex:
    JZ   QY, dbl
    JMP  set_inf
madd_samex:
    JZ   T1, mdbl
    JMP  set_inf
mdbl:                       ; acc = 2*(QX, QY)
    ADD  T0, QY, QY         ; U = 2 Y1
    MUL  T1, T0, T0         ; V = U^2
    MUL  T2, T0, T1         ; W = U V
    MUL  T3, QX, T1         ; S = X1 V
    MUL  T4, QX, QX
    ADD  T0, T4, T4
    ADD  T0, T0, T4         ; M = 3 X1^2
    MUL  T4, T0, T0
    SUB  T4, T4, T3
    SUB  X, T4, T3          ; X3 = M^2 - 2S
    SUB  T4, T3, X
    MULW T2, QY
    MULFIN Y, T0, T4        ; Y3 = M(S - X3) - W Y1
    MOV  ZZ, T1
    MOV  ZZZ, T2
    HOOK 0
; ---------------------------------------------------------------- acc -> affine (QX, QY); HOOK 1, or HOOK 2 if infinity
toaff:
    JZ   ZZ, toaff_inf
    MOV  T1, ZZZ
    CALL inv                ; T0 = 1/ZZZ
    MUL  QY, Y, T0          ; y = Y/ZZZ
    MUL  T2, T0, ZZ         ; ZZ/ZZZ = 1/Z
    MUL  T2, T2, T2         ; 1/ZZ
    MUL  QX, X, T2          ; x = X/ZZ
    HOOK 1
toaff_inf:
    HOOK 2
; T0 = 1/T1 (0 -> 0). INVP is the binary extended Euclid on the plain integer aR, giving
; a^-1 R^-1; one Montgomery product with R^3 restores a^-1 R — the same two steps as libff's
; Fp_model::invert (mpn_gcdext then mul_reduce(Rcubed), fp.tcc:680-727).
inv:
    INVP T0, T1
    LDR3 K0
    MUL  T0, T0, K0
    RET
; T0 = T1^(q-2): Fermat ladder, kept as an independent check of INVP (tests only)
inv_fermat:
    MOV  T0, T1
    SETCNT 759              ; q-2 has 761 bits; the top one is the MOV
inv_loop:
    MUL  T0, T0, T0
    JBIT inv_mul
inv_next:
    DECJNZ inv_loop
    HOOK 0
inv_mul:
    MUL  T0, T0, T1
    JMP  inv_next
; ---------------------------------------------------------------- Q = libff projective (QX, QY, QZZ = Z) -> XYZZ in Q
fromproj:
    JZ   QZZ, fromproj_inf
    MUL  T0, QZZ, QZZ       ; Z^2
    MUL  QX, QX, QZZ        ; X Z
    MUL  QY, QY, T0         ; Y Z^2
    MUL  QZZZ, T0, QZZ      ; Z^3
    MOV  QZZ, T0
    HOOK 0
fromproj_inf:
    ZERO QX
    ONE  QY
    ZERO QZZ
    ZERO QZZZ
    HOOK 0
; ---------------------------------------------------------------- acc = Q
movqa:
    MOV  X, QX
    MOV  Y, QY
    MOV  ZZ, QZZ
    MOV  ZZZ, QZZZ
    HOOK 0
; ---------------------------------------------------------------- batched inversion helpers (precompute normalisation)
; K1 = running product (a slot no point program touches); step: K1 *= ZZZ
prodstep:
    MUL  K1, K1, ZZZ
    HOOK 0
; T0 = 1/K1
invprod:
    MOV  T1, K1
    CALL inv
    HOOK 0
; given T0 = 1/(prefix * ZZZ), T3 = prefix: T4 = T0*T3 = 1/ZZZ ; T0 = T0*ZZZ = 1/prefix ; then affine into QX,QY
unstep:
    MUL  T4, T0, T3         ; 1/ZZZ
    MUL  T0, T0, ZZZ        ; inverse of the shorter prefix
    MUL  QY, Y, T4
    MUL  T4, T4, ZZ         ; 1/Z
    MUL  T4, T4, T4         ; 1/ZZ
    MUL  QX, X, T4
    HOOK 0
; ---------------------------------------------------------------- batched affine pair additions (k_tree_round)
; A thread adds K independent pairs P_j + Q_j in affine coordinates with ONE inversion (Montgomery's trick
; over the denominators d_j = qx_j - px_j). P in (X, Y), Q in (QX, QY), relative slot @0 = product of the
; denominators before pair j, K1 = running product, T4 = running inverse.
; forward step:  d = qx - px ; @0 = K1 ; K1 *= d          HOOK 0 = pair joins the inversion
;   same x: equal points -> tangent (d = 2 py, HOOK 2 marks it); opposite points / 2-torsion -> HOOK 1 (sum = infinity)
tfwd:
    SUB  T0, QX, X
    JZ   T0, tfwd_samex
tfwd_go:
    MOV  @0, K1
    MUL  K1, K1, T0
    HOOK 0
tfwd_samex:
    HOOK 3                  ; the forward pass only loads x: fetch py, qy now
    SUB  T1, QY, Y
    JZ   T1, tfwd_dbl
    HOOK 1
tfwd_dbl:
    ADD  T0, Y, Y
    JZ   T0, tfwd_inf
    HOOK 2
tfwd_inf:
    HOOK 1
; T4 = 1 / K1
tinv:
    MOV  T1, K1
    CALL inv
    MOV  T4, T0
    HOOK 0
; backward step (pairs in reverse order): 1/d = T4 * @0 ; T4 *= d ; lambda = (qy - py)/d ;
;   x3 = lambda^2 - px - qx -> T2 ; y3 = lambda (px - x3) - py -> T3
tbwd:
    SUB  T0, QX, X
    MUL  T1, T4, @0
    MUL  T4, T4, T0
    SUB  T2, QY, Y
tbwd_fin:
    MUL  T1, T1, T2
    MUL  T2, T1, T1
    SUB  T2, T2, X
    SUB  T2, T2, QX
    SUB  T3, X, T2
    MUL  T3, T1, T3
    SUB  T3, T3, Y
    HOOK 0
tbwd_dbl:                   ; Q == P: lambda = 3 px^2 / (2 py)
    ADD  T0, Y, Y
    MUL  T1, T4, @0
    MUL  T4, T4, T0
    MUL  T2, X, X
    ADD  T3, T2, T2
    ADD  T2, T3, T2
    JMP  tbwd_fin
"""

ADD_TEMPLATE = r"""
; ---------------------------------------------------------------- {AX}.. += {BX}..   (general XYZZ addition)
add{S}:
    JZ   {BZZ}, add_done{S}
    JZ   {AZZ}, add_accinf{S}
    MUL  T0, {AX}, {BZZ}      ; U1
    MUL  T1, {BX}, {AZZ}      ; U2
    MUL  T2, {AY}, {BZZZ}     ; S1
    MUL  T3, {BY}, {AZZZ}     ; S2
    SUB  T1, T1, T0         ; P
    SUB  T3, T3, T2         ; R
    JZ   T1, add_samex{S}
    MUL  T4, T1, T1         ; PP
    MUL  T1, T1, T4         ; PPP
    MUL  T0, T0, T4         ; Q = U1 PP
    MUL  {AZZ}, {AZZ}, {BZZ}
    MUL  {AZZ}, {AZZ}, T4       ; ZZ3
    MUL  {AZZZ}, {AZZZ}, {BZZZ}
    MUL  {AZZZ}, {AZZZ}, T1     ; ZZZ3
    MUL  T4, T3, T3         ; R^2
    SUB  T4, T4, T1
    SUB  T4, T4, T0
    SUB  {AX}, T4, T0         ; X3
    SUB  T0, T0, {AX}
    MULW T2, T1
    MULFIN {AY}, T3, T0       ; Y3 = R(Q - X3) - S1 PPP
add_done{S}:
    HOOK 0
add_accinf{S}:
    MOV  {AX}, {BX}
    MOV  {AY}, {BY}
    MOV  {AZZ}, {BZZ}
    MOV  {AZZZ}, {BZZZ}
    HOOK 0
add_samex{S}:
    JZ   T3, dbl{S}
set_inf{S}:
    ZERO {AX}
    ONE  {AY}
    ZERO {AZZ}
    ZERO {AZZZ}
    HOOK 0
; ---------------------------------------------------------------- {AX}.. = 2 {AX}..
dbl{S}:
    JZ   {AZZ}, dbl_done{S}
    ADD  T0, {AY}, {AY}         ; U
    MUL  T1, T0, T0         ; V
    MUL  T2, T0, T1         ; W
    MUL  T3, {AX}, T1         ; S
    MUL  T4, {AX}, {AX}
    ADD  T0, T4, T4
    ADD  T0, T0, T4         ; M
    MUL  T4, T0, T0
    SUB  T4, T4, T3
    SUB  {AX}, T4, T3         ; X3
    SUB  T4, T3, {AX}
    MULW T2, {AY}
    MULFIN {AY}, T0, T4       ; Y3 = M(S - X3) - W Y1
    MUL  {AZZ}, T1, {AZZ}
    MUL  {AZZZ}, T2, {AZZZ}
dbl_done{S}:
    HOOK 0
"""

ACC = dict(AX="X", AY="Y", AZZ="ZZ", AZZZ="ZZZ")
PROGRAM += ADD_TEMPLATE.format(S="", BX="QX", BY="QY", BZZ="QZZ", BZZZ="QZZZ", **ACC)
# second accumulator R += acc (running-sum bucket reduction)
PROGRAM += ADD_TEMPLATE.format(S="_r", AX="RX", AY="RY", AZZ="RZZ", AZZZ="RZZZ", BX="X", BY="Y", BZZ="ZZ", BZZZ="ZZZ")

ENTRY_POINTS = ["tfwd", "tfwd_go", "tinv", "tbwd", "tbwd_dbl", "madd", "madd_acc", "madd_aff", "mdbl", "add", "dbl", "toaff", "fromproj", "movqa", "set_inf", "add_r", "dbl_r", "prodstep", "invprod", "unstep", "inv_fermat"]


def slot(x):
    """Absolute slot name, or @k = relative slot k (bit 7 set; see fqvm_run)."""
    if x.startswith("@"):
        return 0x80 | int(x[1:])
    return SLOTS[x]


def assemble(text):
    lines = []
    labels = {}
    addr = 0
    for raw in text.splitlines():
        ln = raw.split(";")[0].strip()
        if not ln:
            continue
        m = re.match(r"^([A-Za-z_][A-Za-z_0-9]*):$", ln)
        if m:
            labels[m.group(1)] = addr
            continue
        lines.append(ln)
        addr += 1
    code = []
    for ln in lines:
        parts = ln.replace(",", " ").split()
        op = parts[0]
        args = parts[1:]
        o = OPS[op]
        d = a = b = 0
        if op == "MULW":
            a, b = (slot(x) for x in args)
            imm = a | (b << 8)
        elif op in ("MUL", "ADD", "SUB", "MULFIN"):
            d, a, b = (slot(x) for x in args)
            imm = a | (b << 8)
        elif op in ("MOV", "INVP"):
            d, a = slot(args[0]), slot(args[1])
            imm = a
        elif op in ("ONE", "ZERO", "LDR3"):
            d = slot(args[0])
            imm = 0
        elif op == "JZ":
            d = slot(args[0])
            imm = labels[args[1]]
        elif op in ("JMP", "JBIT", "DECJNZ", "CALL"):
            imm = labels[args[0]]
        elif op == "SETCNT":
            imm = int(args[0])
        elif op == "HOOK":
            d = int(args[0])
            imm = 0
        elif op in ("RET", "HALT"):
            imm = 0
        else:
            raise ValueError(ln)
        assert 0 <= imm < 65536 and 0 <= d < 256
        code.append(o | (d << 8) | (imm << 16))
    return code, labels


def main():
    code, labels = assemble(PROGRAM)
    o = "// GENERATED by tools/gen_fqvm_prog.py — do not edit by hand.\n"
    o += "// Micro-programs of the Fq VM (fqvm.cuh): word = op | d << 8 | imm16 << 16.\n"
    o += "#pragma once\n#include <cstdint>\nnamespace zecale {\n"
    o += "enum : uint32_t {\n"
    for k, v in OPS.items():
        o += "    VOP_%s = %d,\n" % (k, v)
    o += "};\nenum : uint32_t {\n"
    for k, v in SLOTS.items():
        o += "    VS_%s = %d,\n" % (k, v)
    o += "    VS_COUNT = %d\n};\nenum : uint32_t {\n" % NSLOT
    for k in ENTRY_POINTS:
        o += "    VPC_%s = %d,\n" % (k.upper(), labels[k])
    o += "};\n"
    o += "static constexpr int FQVM_CODE_LEN = %d;\n" % len(code)
    o += "__device__ __constant__ uint32_t FQVM_CODE[%d] = {\n" % len(code)
    for i in range(0, len(code), 8):
        o += "    " + ", ".join("0x%08xu" % w for w in code[i:i + 8]) + ",\n"
    o += "};\n} // namespace zecale\n"
    dst = os.path.join(os.path.dirname(__file__), "..", "zecale_b200", "csrc", "fqvm_prog.cuh")
    with open(dst, "w") as f:
        f.write(o)
    print("wrote", os.path.abspath(dst), len(code), "uops")


if __name__ == "__main__":
    main()
