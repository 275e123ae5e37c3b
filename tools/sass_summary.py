#!/usr/bin/env python
"""Per-kernel SASS evidence of the shipped libdipole.so (run in the build container; writes profiles/sass_summary.txt):
whole-function counts of the instructions that show what the kernel is built from -- packed FP32 (FFMA2 / FMUL2 /
FADD2), scalar FP32 (FFMA / FMUL / FADD), MUFU, uniform-datapath constant loads (LDCU) vs vector constant loads (LDC),
shared-memory loads, TMA bulk copies (UBLKCP), cluster barriers / DSMEM stores, REDUX -- plus the opcode histogram of the
innermost MUFU-bearing loop of the all-pairs kernels."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "dipole-normal-prop_b200", "libdipole.so")
OUT = os.path.join(ROOT, "profiles", "sass_summary.txt")
WATCH = ["FFMA2", "FMUL2", "FADD2", "FFMA", "FMUL", "FADD", "MUFU.RSQ", "MUFU.RCP", "MUFU.SQRT", "FMNMX", "LDCU", "LDC", "LDS",
         "LDG", "UBLKCP", "SYNCS", "UCGABAR", "BAR", "REDUX", "ATOMG", "ST", "DFMA", "DADD", "DMUL"]


def demangle(names):
    try:
        out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
        out = [o.replace("(anonymous namespace)::", "").replace("void ", "") for o in out]
        return [re.sub(r"\((?!anonymous).*", "", o) for o in out]
    except OSError:
        return names


def main():
    sass = subprocess.check_output(["cuobjdump", "-sass", LIB], text=True)
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
    funcs = []
    for fn in re.split(r"\n\s+Function : ", sass)[1:]:
        name = fn.split("\n")[0].strip()
        ins = [(int(m.group(1), 16), re.sub(r"^@!?U?P\d\s+", "", m.group(2).strip()))
               for m in re.finditer(r"/\*([0-9a-f]{4,5})\*/\s+(.*?);", fn)]
        ops = collections.Counter()
        for _, t in ins:
            op = t.split()[0]
            for w in WATCH:
                if op == w or op.startswith(w + "."):
                    ops[w] += 1
                    break
        loops = []
        for a, t in ins:
            m = re.search(r"BRA\S*\s+(?:\S+,\s+)?0x([0-9a-f]+)", t)
            if m and int(m.group(1), 16) < a:
                loops.append((int(m.group(1), 16), a))
        inner = [l for l in loops if any("MUFU" in t for a, t in ins if l[0] <= a <= l[1])]
        hist = None
        if inner:
            lo, hi = min(inner, key=lambda l: l[1] - l[0])
            hist = collections.Counter(t.split()[0] for a, t in ins if lo <= a <= hi)
        funcs.append((name, len(ins), ops, hist))
    names = demangle([f[0] for f in funcs])
    with open(OUT, "w") as f:
        f.write(f"# tools/sass_summary.py: cuobjdump -sass dipole-normal-prop_b200/libdipole.so ({', '.join(arch)} cubins only)\n")
        f.write("# whole-function instruction counts; then, for kernels with a MUFU-bearing loop, that loop's histogram\n")
        f.write("kernel,instructions," + ",".join(WATCH) + "\n")
        for (name, n, ops, hist), pretty in zip(funcs, names):
            f.write(f"\"{pretty}\",{n}," + ",".join(str(ops.get(w, 0)) for w in WATCH) + "\n")
        f.write("\n# ---- innermost MUFU-bearing loops\n")
        for (name, n, ops, hist), pretty in zip(funcs, names):
            if hist and ("pairs" in pretty or "lattice" in pretty or "edge_" in pretty or "loop" in pretty):
                f.write(f"{pretty}: {sum(hist.values())} instructions: " +
                        ", ".join(f"{k} {v}" for k, v in hist.most_common()) + "\n")
        tot = collections.Counter()
        for _, _, ops, _ in funcs:
            tot.update(ops)
        f.write("\n# ---- library totals: " + ", ".join(f"{w} {tot.get(w, 0)}" for w in WATCH) + "\n")
        f.write("# no UTCxMMA / UTMALDG / HMMA: none of these kernels is a dense contraction (the weight 1/(|R|^3+eps) is per pair)\n")
    print("wrote", OUT, len(funcs), "kernels")


if __name__ == "__main__":
    main()
