"""Per-kernel SASS instruction-class counts of the shipped library (cuobjdump -sass / -res-usage over csrc/_obj/*.o):
packed fp32 (FFMA2 / FMUL2 / FADD2), TMA bulk copies (UBLKCP), mbarrier ops (SYNCS), MATCH, SHFL, MUFU, multimem (in-switch
reduce), local-memory traffic (LDL / STL = spills), registers, shared memory.  usage: python tools/sass_summary.py > profiles/rN_sass_summary.md"""
import glob
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "gaussiansplatting_b200", "csrc", "_obj")
CLASSES = [("FFMA2", r"\bFFMA2\b"), ("FMUL2", r"\bFMUL2\b"), ("FADD2", r"\bFADD2\b"), ("FFMA", r"\bFFMA\b"), ("MUFU", r"\bMUFU"),
           ("SHFL", r"\bSHFL"), ("MATCH", r"\bMATCH"), ("VOTE", r"\bVOTE"), ("UBLKCP", r"\bUBLKCP"), ("SYNCS", r"\bSYNCS"),
           ("ATOM/RED", r"\b(ATOM|ATOMG|RED|ATOMS)\b"), ("MULTIMEM", r"MULTIMEM|\.MMLD|\.MMST|LDGMC|STGMC|REDG?MC"), ("LDL", r"\bLDL\b"), ("STL", r"\bSTL\b")]


def demangle(names):
    try:
        out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
        return dict(zip(names, out))
    except OSError:
        return {n: n for n in names}


def short(name):
    name = re.sub(r"\(anonymous namespace\)::", "", name)
    name = re.sub(r"\(.*", "", name)
    return re.sub(r"^void ", "", name)


def main():
    rows = []
    for obj in sorted(glob.glob(os.path.join(OBJ, "*.o"))):
        sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
        res = subprocess.run(["cuobjdump", "-res-usage", obj], capture_output=True, text=True).stdout
        usage = {}
        for m in re.finditer(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", res):
            usage[m.group(1)] = (int(m.group(2)), int(m.group(3)), int(m.group(4)), int(m.group(5)))
        parts = re.split(r"\n\s*Function : ", sass)[1:]
        for p in parts:
            fname = p.split("\n", 1)[0].strip()
            body = [l for l in p.split("\n") if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l)]
            total = len(body)
            text = "\n".join(body)
            counts = {k: len(re.findall(rx, text)) for k, rx in CLASSES}
            rows.append((os.path.basename(obj)[:-2], fname, total, counts, usage.get(fname)))
    names = demangle([r[1] for r in rows])
    print("# SASS summary of libgs_b200.so (sm_100a) — instruction-class counts per kernel\n")
    print("Static counts from `cuobjdump -sass` of `csrc/_obj/*.o` (what the shared library links); registers / stack / static shared memory from "
          "`cuobjdump -res-usage`. LDL / STL are local-memory loads / stores: a non-zero count with STACK 0 would be a spill.\n")
    print("| file | kernel | SASS instr | regs | stack B | smem B | " + " | ".join(k for k, _ in CLASSES) + " |")
    print("|---|---|---|---|---|---|" + "---|" * len(CLASSES))
    for f, fn, total, c, u in rows:
        reg, stack, shared, _ = u if u else ("?", "?", "?", "?")
        print(f"| {f} | `{short(names.get(fn, fn))}` | {total} | {reg} | {stack} | {shared} | " + " | ".join(str(c[k]) for k, _ in CLASSES) + " |")
    tot = {k: sum(r[3][k] for r in rows) for k, _ in CLASSES}
    print(f"\nTotals: " + ", ".join(f"{k} {v}" for k, v in tot.items()) + f"; kernels {len(rows)}; kernels with local-memory traffic: "
          + (", ".join(sorted({short(names.get(r[1], r[1])) for r in rows if r[3]['LDL'] or r[3]['STL']})) or "none"))


if __name__ == "__main__":
    main()
