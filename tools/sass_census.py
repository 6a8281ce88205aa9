#!/usr/bin/env python
"""Count the sm_100a SASS mnemonics that matter per kernel of libcppre.so.

CPU-only (cuobjdump). Writes one line per kernel: registers are not in SASS, so this
lists instruction counts only: TMA loads (UTMALDG), mbarrier traffic (SYNCS), packed
fp32x2 math (FFMA2 / FMUL2 / FADD2), scalar FFMA, warp reductions (REDUX), named
barriers (BAR), shared / global loads and stores, local-memory spills (LDL / STL).

    python tools/sass_census.py > profiles/r01_sass_census.txt
"""
import collections
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "..", "cp_pre_b200", "libcppre.so")
COLS = ["UTMALDG", "SYNCS", "FFMA2", "FMUL2", "FADD2", "FFMA", "REDUX", "BAR",
        "LDS", "STS", "LDG", "STG", "LDL", "STL"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout
    return out.splitlines()


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts = collections.OrderedDict()
    cur = None
    op = re.compile(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)")
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = counts.setdefault(m.group(1), collections.Counter())
            continue
        if cur is None:
            continue
        m = op.match(line)
        if m:
            cur[m.group(1).split(".")[0]] += 1
            cur["_total"] += 1
    names = demangle(list(counts))
    print("# cuobjdump -sass cp_pre_b200/libcppre.so, instruction counts per kernel")
    print("# " + " ".join(f"{c:>7}" for c in ["total"] + COLS) + "  kernel")
    for (mangled, c), name in zip(counts.items(), names):
        short = re.sub(r"\s+", " ", name)
        if len(short) > 150:
            short = short[:147] + "..."
        print("  " + " ".join(f"{c[k]:>7}" for k in ["_total"] + COLS) + "  " + short)


if __name__ == "__main__":
    sys.exit(main())
