#!/usr/bin/env python3
"""Per-kernel SASS evidence of the shipped library: instruction mnemonic counts (cuobjdump -sass) and
registers / stack (cuobjdump -res-usage) -> a markdown table under profiles/.

  python tools/sass_summary.py zecale_b200/libzecale_gpu.so profiles/r02_sass_summary.md
"""
import collections
import re
import subprocess
import sys

CUOBJDUMP = "/usr/local/cuda/bin/cuobjdump"
COLS = ["IMAD.WIDE", "IMAD", "IADD3", "LDG", "STG", "LDL", "STL", "LDS", "STS", "LDC", "SHFL", "BAR", "ATOM", "RED"]


def demangle(name):
    m = re.match(r"_ZN6zecale(\d+)", name)
    if m:
        k = int(m.group(1))
        start = m.end()
        base = name[start:start + k]
        t = re.search(r"ILi(\d+)E", name[start + k:])
        return base + ("<%s>" % t.group(1) if t else "")
    m = re.match(r"_ZN\d+_GLOBAL__N__[0-9a-f_]+zecale_gpu_cu_[0-9a-f]+_\d+(\d\d)(k_[a-z_0-9]+)", name)
    return name


def main():
    lib, dst = sys.argv[1], sys.argv[2]
    res = subprocess.run([CUOBJDUMP, "-res-usage", lib], capture_output=True, text=True).stdout
    usage = {}
    cur = None
    for ln in res.splitlines():
        m = re.search(r"Function (\S+):", ln)
        if m:
            cur = m.group(1)
            continue
        m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", ln)
        if m and cur:
            usage[cur] = tuple(int(x) for x in m.groups())
            cur = None
    sass = subprocess.run([CUOBJDUMP, "-sass", lib], capture_output=True, text=True).stdout
    counts = collections.OrderedDict()
    cur = None
    for ln in sass.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if not m:
            continue
        op = m.group(1)
        counts[cur]["total"] += 1
        if op.startswith("IMAD.WIDE"):
            counts[cur]["IMAD.WIDE"] += 1
        elif op.startswith("IMAD"):
            counts[cur]["IMAD"] += 1
        else:
            base = op.split(".")[0]
            for c in COLS:
                if base == c or (c in ("ATOM", "RED") and base.startswith(c)):
                    counts[cur][c] += 1
        if op.startswith(("UTMA", "UTCMMA", "TCGEN", "HMMA", "IMMA", "UTCHMMA")):
            counts[cur]["tensor/TMA"] += 1
    rows = []
    for fn, c in counts.items():
        short = fn
        m = re.search(r"(k_[a-z_0-9]+)", fn)
        if m:
            short = m.group(1)
            t = re.search(r"ILi(\d+)E", fn)
            if t:
                short += "<%s>" % t.group(1)
        u = usage.get(fn, (0, 0, 0, 0))
        rows.append((short, u, c))
    rows.sort(key=lambda r: -r[2]["total"])
    with open(dst, "w") as f:
        f.write("# SASS summary of zecale_b200/libzecale_gpu.so (sm_100a), per kernel\n\n")
        f.write("Source: `cuobjdump -sass` / `cuobjdump -res-usage` on the shipped library; counts are static instructions.\n")
        f.write("One Fq Montgomery product = 432 IMAD.WIDE (Karatsuba product, fq_mulw_regs) + 576 (reduction, fq_redc_regs: 552 IMAD.WIDE + 24 "
                "IMAD.HI for the first column) + 24 IMAD for m = t0*inv; one Fr body = 288 (+12). The IMAD column counts every "
                "other IMAD.* form (IMAD, IMAD.HI, IMAD.X, IMAD.MOV, IMAD.IADD): all of them occupy the multiply pipe.\n")
        f.write("No tensor-core or TMA instruction appears in any kernel (the path is carry-chain integer arithmetic).\n\n")
        f.write("| kernel | regs | stack B | total | " + " | ".join(COLS) + " | tensor/TMA |\n")
        f.write("|---|---|---|---|" + "---|" * (len(COLS) + 1) + "\n")
        for short, u, c in rows:
            f.write("| %s | %d | %d | %d | %s | %d |\n" % (short, u[0], u[1], c["total"], " | ".join(str(c[k]) for k in COLS), c["tensor/TMA"]))
    print("wrote", dst, len(rows), "kernels")


if __name__ == "__main__":
    main()
