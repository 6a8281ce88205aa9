"""`cuobjdump -res-usage` of every star / t-innermost kernel in libcppre.so (CPU only): registers and the
per-thread STACK (local-memory spill frame).  Kernel names: star_kernel<KIND, BATCH_ROWS, RX, MASKS, MODE, OPER>
(KIND: 3 ns_momentum, 4 mhd_continuity, 5 mhd_momentum, 6 mhd_energy, 7 mhd_induction, 8 mhd_gauss; MODE 0 STORE /
1 SCORE / 2 COVER; OPER -1 = run-time operand set, 0 none, 2 weight only), tin_kernel<KIND, RX, MASKS, MODE>.
MASKS 101382 = the MHD operator set the reference's constructors build (D_t, D_x, D_y == D_t), 253856774 = the NS
momentum set; 2097151 / 268435455 = every arm of every slot (operator sets no reference script builds).

    python tools/res_usage.py > profiles/r02_res_usage.txt
"""
import os
import re
import subprocess

LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "cp_pre_b200", "libcppre.so")
SHIPPED = {"101382ul", "253856774ul"}


def main():
    txt = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
    rows, seen = [], set()
    for m in re.finditer(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", txt):
        if "star_kernel" not in m.group(1) and "tin_kernel" not in m.group(1):
            continue
        tu = re.search(r"_(stencil_t[a-z]+_[a-z_0-9]+?)_cu_", m.group(1))
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        k = re.search(r"(star_kernel|tin_kernel)<([^>]*)>", name)
        args = [a.strip() for a in k.group(2).split(",")]
        mask = args[3] if k.group(1) == "star_kernel" else args[2]
        row = (tu.group(1) if tu else "?", f"{k.group(1)}<{k.group(2)}>", mask, int(m.group(2)), int(m.group(3)))
        if row[:2] not in seen:
            seen.add(row[:2])
            rows.append(row)
    hot = ("stencil_tma_store_plain", "stencil_tma_score_plain", "stencil_tma_score_w", "stencil_tma_cover_w",
           "stencil_tma_store_dyn", "stencil_tin_store", "stencil_tin_cover")
    print(f"{len(rows)} kernels; {sum(1 for r in rows if r[4] == 0)} with STACK 0")
    print("\n== translation units the benchmark configurations launch (fp32; plain / weight-only operand sets, STORE with a "
          "reference), shipped operator sets ==")
    for r in rows:
        if r[0] in hot and r[2] in SHIPPED:
            print(f"{r[0]:28s} {r[1]:60s} REG {r[3]:3d} STACK {r[4]}")
    print("\n== every shipped-operator kernel with a spill frame (run-time operand sets, bf16 staging, decode fold) ==")
    for r in rows:
        if r[4] > 0 and r[2] in SHIPPED:
            print(f"{r[0]:28s} {r[1]:60s} REG {r[3]:3d} STACK {r[4]}")
    print("\n== all-arms fallbacks (operator sets no reference script builds) with a spill frame: "
          f"{sum(1 for r in rows if r[4] > 0 and r[2] not in SHIPPED)} kernels, STACK up to "
          f"{max([r[4] for r in rows if r[2] not in SHIPPED] + [0])} bytes ==")


if __name__ == "__main__":
    main()
