"""ncu raw CSV (`ncu -i X.ncu-rep --page raw --csv`) -> profiles/k1_traffic_<workload>[_<equation>].json, the file
bench.py's `roofline.traffic` is read from.  The record carries the fingerprint of the K1 sources it was
captured from (bench.kernel_source_sha): bench.py reports the traffic only while the kernel is unchanged.

    python tools/k1_traffic.py gpurun_out/r2g_prof_raw.csv ns "" 256
    python tools/k1_traffic.py gpurun_out/r2g_prof_mhd_raw.csv mhd momentum 64
"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def to_bytes(v, unit):
    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]
    return float(v.replace(",", "")) * mult


def main(path, workload, equation=None, samples=None):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    w = bench.WORKLOADS[workload]
    nf = bench.MHD_EQUATIONS[equation] if workload == "mhd" else w["nf"]
    rec = cover = None
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        if "star_kernel" not in name:
            continue
        rd = to_bytes(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
        wr = to_bytes(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
        if wr < 0.05 * rd:          # a COVER / SCORE launch: nothing written
            if cover is None:
                cover = {"kernel": name[:160], "dram_bytes_read": rd, "dram_bytes_write": wr}
            continue
        if rec is not None:
            continue
        grid = int(r[col["launch__grid_size"]].replace(",", ""))
        ms = float(r[col["gpu__time_duration.sum"]].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(
            units[col["gpu__time_duration.sum"]], 1e-6)
        rec = {"kernel": name[:160], "workload": workload, "grid": grid, "dram_bytes_read": rd, "dram_bytes_write": wr,
               "ncu_duration_ms": ms}
    assert rec is not None, "no STORE launch of star_kernel in the capture"
    # algorithmic bytes of the captured launch: (4F + 4) B per point; points from the written bytes (4 B per point)
    import numpy as np
    cells = int(np.prod(w["grid"]))
    # samples per launch: given, or whole samples from the written bytes (4 B per point; ncu sees a little less than
    # the algorithmic write volume because the last lines are still in L2 when the kernel ends)
    samples = int(samples) if samples else round(rec["dram_bytes_write"] / 4 / cells)
    points = samples * cells
    rec["algorithmic_bytes"] = (4 * nf + 4) * points
    rec["points_per_launch"] = points
    rec["samples_per_launch"] = samples
    rec["traffic_over_algorithmic"] = (rec["dram_bytes_read"] + rec["dram_bytes_write"]) / rec["algorithmic_bytes"]
    if cover is not None:
        # the COVER launch of the same step (validation half, as many samples as the STORE launch in bench.py's split):
        # reads the fields + the per-cell band tile, writes per-sample counters only
        cover["algorithmic_bytes"] = 4 * nf * points
        cover["traffic_over_algorithmic"] = (cover["dram_bytes_read"] + cover["dram_bytes_write"]) / cover["algorithmic_bytes"]
        rec["cover_launch"] = cover
    rec["kernel_source_sha"] = bench.kernel_source_sha()
    rec["source"] = f"ncu --set full --clock-control none, {os.path.basename(path)} (first STORE launch of star_kernel)"
    name = f"k1_traffic_{workload}" + (f"_{equation}" if workload == "mhd" else "") + ".json"
    json.dump(rec, open(os.path.join(ROOT, "profiles", name), "w"), indent=1)
    print(json.dumps(rec))


if __name__ == "__main__":
    a = sys.argv[1:]
    main(a[0], a[1], (a[2] or None) if len(a) > 2 else None, a[3] if len(a) > 3 else None)
