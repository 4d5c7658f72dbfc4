#!/usr/bin/env python
"""Per-kernel evidence from the SHIPPED library, no GPU needed: registers / static shared memory (cuobjdump -res-usage)
and counts of the SASS mnemonics that show how a kernel moves data and reduces (UBLKCP = cp.async.bulk / TMA bulk copy,
SYNCS = mbarrier ops, REDG = global reductions incl. vector F32x4, SHFL, MUFU.EX2, LDGMC = multimem.ld_reduce, BAR).
    python scripts/sass_summary.py > profiles/r02_sass_summary.md"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "splat-slam_b200", "diff_gaussian_rasterization", "libsplat_raster.so")
MNEMONICS = ("UBLKCP", "SYNCS", "REDG", "REDG.E.ADD.F32x4", "ATOMG", "ATOMS", "SHFL", "VOTE", "MUFU.EX2", "MUFU.RCP",
             "BAR.SYNC", "LDGMC", "STG.E.128", "LDS.128")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return [re.sub(r"\(.*", "", n).replace("splat::", "") for n in out]


def main():
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    usage = {}
    arch = set(re.findall(r"arch = (sm_\w+)", res))
    for m in re.finditer(r"Function (\S+):\n\s+REG:(\d+) STACK:(\d+) SHARED:(\d+)", res):
        usage[m.group(1)] = (int(m.group(2)), int(m.group(3)), int(m.group(4)))
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts, cur = {}, None
    for line in sass.splitlines():
        f = re.match(r"\s+Function : (\S+)", line)
        if f:
            cur = f.group(1)
            counts[cur] = {k: 0 for k in MNEMONICS}
            counts[cur]["instructions"] = 0
            continue
        if cur and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
            counts[cur]["instructions"] += 1
            for k in MNEMONICS:
                if re.search(r"\b" + re.escape(k) + r"\b", line):
                    counts[cur][k] += 1
    names = sorted(counts)
    pretty = dict(zip(names, demangle(names)))
    print("# SASS / resource summary of the shipped libsplat_raster.so (`python scripts/sass_summary.py`)\n")
    print(f"cubin architectures in the library: {', '.join(sorted(arch)) or 'unknown'} (no PTX fallback is relied on)\n")
    cols = ["instructions"] + list(MNEMONICS)
    print("| kernel | regs | static smem (B) | " + " | ".join(cols) + " |")
    print("|---|---|---|" + "---|" * len(cols))
    for n in sorted(names, key=lambda x: pretty[x]):
        r = usage.get(n, (0, 0, 0))
        print(f"| `{pretty[n]}` | {r[0]} | {r[2]} | " + " | ".join(str(counts[n][c]) for c in cols) + " |")
    print("\nDynamic shared memory is not in the static column: render_backward_kernel<8> launches with 35 KB, <16> with 53 KB; "
          "preprocess_backward_kernel stages SH gradient rows in up to 64 KB when M > 1.")


if __name__ == "__main__":
    sys.exit(main())
