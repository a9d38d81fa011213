"""Opcode histogram per kernel of libb200_w8a8.so (cuobjdump -sass), written to profiles/sass_summary.txt.

    python tools/sass_summary.py [out_path]

Lists, per kernel: instruction count, registers are in the -Xptxas -v build log (not here), and the counts of the
Blackwell-specific opcodes that prove the design (UTCIMMA / UTCQMMA = tcgen05.mma kind::i8 / kind::f8f6f4,
UTMALDG / UTMASTG = TMA loads / stores, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, FFMA2 = packed fp32 FMA,
SYNCS = mbarrier ops, MULTIMEM / RED for the NVSwitch reduction).  Runs on the CPU-only container.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "sglang-diffusion_b200", "libb200_w8a8.so")
KEY = ("UTCIMMA", "UTCQMMA", "UTCHMMA", "UTMALDG", "UTMASTG", "UTMAPF", "LDTM", "STTM", "UTCBAR", "UTCATOM", "FFMA2", "FMUL2",
       "SYNCS", "MULTIMEM", "ACQBULK", "UCGABAR", "CCTL", "LDGSTS", "LDG", "STG", "LDS", "STS", "F2FP", "F2I", "I2F",
       "SHFL", "MUFU", "FFMA", "FMNMX", "BAR", "RED", "ATOM", "ST.E", "LD.E")


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "sass_summary.txt")
    sass = subprocess.run(["cuobjdump", "-sass", LIB], check=True, capture_output=True, text=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*(?:\.[A-Za-z0-9_]+)*)", line)
        if m and cur:
            kernels[cur][m.group(1)] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    with open(out_path, "w") as f:
        f.write(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)} -- opcode histogram per kernel ({len(kernels)} kernels)\n")
        f.write("# key opcodes: base mnemonic -> count (all modifiers folded); 'full' lists the tcgen05/TMA forms verbatim\n\n")
        for (name, ops), dm in zip(kernels.items(), demangle):
            total = sum(ops.values())
            base = collections.Counter()
            for op, n in ops.items():
                base[op.split(".")[0]] += n
            f.write(f"{dm}\n  instructions: {total}\n")
            keys = [(k, base[k]) for k in KEY if base.get(k)]
            f.write("  key: " + ", ".join(f"{k}={n}" for k, n in keys) + "\n")
            full = sorted((op, n) for op, n in ops.items()
                          if op.split(".")[0] in ("UTCIMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "LDTM", "UTCBAR", "MULTIMEM", "FFMA2", "UTCATOM"))
            if full:
                f.write("  full: " + ", ".join(f"{op}={n}" for op, n in full) + "\n")
            top = ", ".join(f"{op}={n}" for op, n in base.most_common(12))
            f.write(f"  top: {top}\n\n")
    print(out_path, len(kernels), "kernels")


if __name__ == "__main__":
    main()
