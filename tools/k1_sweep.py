"""Times the K1 residual kernels (CUDA events, median of reps) over programs / epilogues and the
CPPRE_TMA_RX / CPPRE_TMA_STAGES tuning knobs.  Run on the GPU box:  python tools/k1_sweep.py
Each configuration runs in a fresh subprocess because the knobs are read once per process."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def child(which):
    import numpy as np
    import torch
    import cp_pre_b200
    from cp_pre_b200 import _lib as L, programs
    from oracle import synth
    cp_pre_b200.install()
    import ConvOps_1d as C1
    import Utils.ConvOps_2d as C2
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(0)
    D = dict(t=C2.ConvOperator("t", 1), x=C2.ConvOperator("x", 1), y=C2.ConvOperator("y", 1), l=C2.ConvOperator(("x", "y"), 2))
    c = synth.NS
    if which == "ns":
        shape, nf = (128, 3, 50, 256, 256), 3
        prog = programs.ns_momentum(D["t"], D["x"], D["y"], D["l"], c["dx"], c["dy"], c["dt"], c["nu"])
    elif which == "wave":
        shape, nf = (256, 80, 128, 128), 1
        W = C2.ConvOperator()
        W.kernel = C2.ConvOperator("t", 2).kernel - 0.25 * C2.ConvOperator(("x", "y"), 2).kernel
        prog = programs.linear(W)
    elif which == "burgers":
        shape, nf = (4096, 100, 1024), 1
        b = synth.BURGERS
        prog = programs.burgers(C1.ConvOperator("t", 1), C1.ConvOperator("x", 1), C1.ConvOperator("x", 2), b["dx"], b["dt"], b["nu"])
    elif which.startswith("mhd"):
        shape, nf = (48, 6, 100, 256, 256), {"mhd_momentum": 6, "mhd_energy": 6, "mhd_induction": 4, "mhd_continuity": 3, "mhd_gauss": 2}[which]
        prog = programs.mhd(which[4:], D["t"], D["x"], D["y"])
    v = torch.rand(shape, device=dev, generator=g) + 0.5
    pts = v.numel() // (shape[1] if len(shape) == 5 else 1)
    cells = shape[-3:] if len(shape) >= 4 else shape[-2:]
    w = torch.rand(cells, device=dev, generator=g) + 0.5

    def timeit(fn, reps=7):
        for _ in range(2):
            fn()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return float(np.median(ts))

    out = {"which": which, "rx": os.environ.get("CPPRE_TMA_RX", "auto"), "stages": os.environ.get("CPPRE_TMA_STAGES", "auto")}
    for name, fn, bpp in (("store", lambda: prog.residual(v, boundary=True), 4 * nf + 4),
                          ("score", lambda: prog.joint_scores(v, w), 4 * nf)):
        ms = timeit(fn)
        out[name] = {"ms": round(ms, 3), "path": L.last_path(), "gpts": round(pts / ms / 1e6, 1), "GBs": round(bpp * pts / ms / 1e6, 0)}
    prog.path = L.PATH_GENERIC
    if os.environ.get("SWEEP_GENERIC"):
        ms = timeit(lambda: prog.residual(v, boundary=True), 3)
        out["generic_store"] = {"ms": round(ms, 3), "gpts": round(pts / ms / 1e6, 1), "GBs": round((4 * nf + 4) * pts / ms / 1e6, 0)}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--child":
        child(sys.argv[2])
        sys.exit(0)
    progs = sys.argv[1].split(",") if len(sys.argv) > 1 else ["ns", "wave", "burgers", "mhd_momentum"]
    for which in progs:
        for rx in ("1", "2"):
            for st in ("3", "4", "6"):
                env = dict(os.environ, CPPRE_TMA_RX=rx, CPPRE_TMA_STAGES=st)
                if rx == "1" and st == "4":
                    env["SWEEP_GENERIC"] = "1"
                subprocess.run([sys.executable, __file__, "--child", which], env=env)
