#!/usr/bin/env python
"""Opcode counts per kernel from `cuobjdump -sass` of the built library (no GPU needed).

    python profiles/sass_summary.py > profiles/sass_summary.md
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "mm_masking_b200", "libmm_masking_b200.so")
WATCH = ["UBLKCP", "SYNCS", "FFMA2", "FMNMX3", "CREDUX", "UCGABAR_ARV", "ACQBULK", "PREEXIT", "BAR", "SHFL", "VOTE", "LDG", "LDS",
         "STG", "STS", "LDL", "STL", "ATOMG", "ATOMS", "REDG", "DFMA", "DADD", "DMUL", "MUFU", "UTCHMMA", "LDTM", "HMMA"]


def main():
    sys.path.insert(0, ROOT)
    import bench
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", out)))
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and cur:
            kernels[cur][m.group(1)] += 1
    demangle = subprocess.run(["c++filt"] + list(kernels), capture_output=True, text=True).stdout.splitlines()
    total = collections.Counter()
    print("# SASS opcode counts per kernel (`cuobjdump -sass mm_masking_b200/libmm_masking_b200.so`)\n")
    print(f"Architectures in the library: {', '.join(arch)}.  CUDA sources sha256[:16] = `{bench.kernel_source_sha()}`.")
    print("Static instruction counts (not executed counts).  No tensor-core opcodes (`UTC*MMA`, `LDTM`, `HMMA`) are expected: the "
          "hot path is exact nearest-neighbour search and fp64 reductions on the CUDA cores (DESIGN.md section 3).\n")
    names = [n for n in WATCH]
    print("| kernel | instr | " + " | ".join(names) + " |")
    print("|---|---|" + "---|" * len(names))
    for (mangled, c), name in zip(kernels.items(), demangle):
        short = re.sub(r"\(.*", "", name).replace("void ", "").replace("mmb::", "")
        total.update(c)
        if sum(c.values()) < 40:
            continue
        print(f"| `{short}` | {sum(c.values())} | " + " | ".join(str(c.get(n, 0)) for n in names) + " |")
    print("\n| whole library | " + str(sum(total.values())) + " | " + " | ".join(str(total.get(n, 0)) for n in names) + " |")


if __name__ == "__main__":
    main()
