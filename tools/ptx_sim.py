"""Tiny interpreter for the PTX subset emitted by gen_field_asm.py (CPU-side check of the
carry-chain code without a GPU). Test infrastructure only."""
M32 = 0xFFFFFFFF


def run(lines, regs):
    cc = 0
    pred = {}

    def val(tok):
        tok = tok.strip()
        if tok.startswith("0x"):
            return int(tok, 16)
        if tok.isdigit():
            return int(tok)
        return regs[tok]

    for ln in lines:
        ln = ln.strip().rstrip(";")
        op, rest = ln.split(" ", 1)
        args = [a.strip() for a in rest.split(",")]
        p = op.split(".")
        base = p[0]
        sets_cc = "cc" in p
        if base in ("mul",):
            a, b = val(args[1]), val(args[2])
            pr = a * b
            regs[args[0]] = (pr & M32) if "lo" in p else (pr >> 32)
        elif base in ("mad", "madc"):
            a, b, c = val(args[1]), val(args[2]), val(args[3])
            pr = a * b
            part = (pr & M32) if "lo" in p else (pr >> 32)
            s = part + c + (cc if base == "madc" else 0)
            regs[args[0]] = s & M32
            if sets_cc:
                cc = s >> 32
        elif base in ("add", "addc"):
            s = val(args[1]) + val(args[2]) + (cc if base == "addc" else 0)
            regs[args[0]] = s & M32
            if sets_cc:
                cc = s >> 32
        elif base in ("sub", "subc"):
            s = val(args[1]) - val(args[2]) - (cc if base == "subc" else 0)
            regs[args[0]] = s & M32
            if sets_cc:
                cc = 1 if s < 0 else 0
        elif base == "shf":   # shf.l.wrap.b32 d, lo, hi, c
            lo, hi, c = val(args[1]), val(args[2]), val(args[3]) & 31
            regs[args[0]] = (((hi << 32) | lo) << c >> 32) & M32
        elif base == "shr":
            regs[args[0]] = val(args[1]) >> (val(args[2]) & 31)
        elif base == "shl":
            regs[args[0]] = (val(args[1]) << (val(args[2]) & 31)) & M32
        elif base == "and":
            regs[args[0]] = val(args[1]) & val(args[2])
        elif base == "xor":
            regs[args[0]] = val(args[1]) ^ val(args[2])
        elif base == "lop3":  # lop3.b32 d, a, b, c, immLut
            a, b, c, lut = val(args[1]), val(args[2]), val(args[3]), val(args[4])
            r = 0
            for bit in range(8):
                if (lut >> bit) & 1:
                    ta = a if bit & 4 else ~a
                    tb = b if bit & 2 else ~b
                    tc = c if bit & 1 else ~c
                    r |= ta & tb & tc
            regs[args[0]] = r & M32
        elif base == "mov":
            regs[args[0]] = val(args[1])
        elif base == "not":
            regs[args[0]] = ~val(args[1]) & M32
        elif base == "setp":
            pred[args[0]] = val(args[1]) != val(args[2])
        elif base == "selp":
            regs[args[0]] = val(args[1]) if pred[args[3]] else val(args[2])
        else:
            raise ValueError(ln)
    return regs
