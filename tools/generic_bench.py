"""Throughput of the requests the TMA star kernels do NOT take (missing 7): Taylor-4 / Taylor-6 operators
(k = 5 / 7), ragged rows (N0 % 4 != 0, e.g. config 1's Nx = 197), the JOREK coefficient-field programs, the valid
convolution of ConvOps_Spatial -- all served by K1g (stencil_generic.cu) -- with the star kernel on the nearest
eligible shape beside each, so the gap is a number.  One JSON line.  Run on the GPU box: python tools/generic_bench.py"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cp_pre_b200  # noqa: E402
from cp_pre_b200 import _lib as L, programs  # noqa: E402
from oracle import synth  # noqa: E402


def timed(fn, reps=7):
    fn(); fn()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def main():
    cp_pre_b200.install()
    import ConvOps_1d as C1
    import Utils.ConvOps_2d as C2
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(0)
    out = {}

    def record(name, prog, v, nf, note=""):
        pts = v.numel() // (nf if v.dim() == 5 else 1)
        ms = timed(lambda: prog.residual(v, boundary=True))
        out[name] = {"path": L.last_path(), "shape": list(v.shape), "ms": round(ms, 3), "gpoints_s": round(pts / ms / 1e6, 1),
                     "GBs": round((4 * nf + 4) * pts / ms / 1e6, 0), "note": note}

    u = torch.randn((64, 50, 256, 256), device=dev, generator=g)
    for taylor, k in ((2, 3), (4, 5), (6, 7)):
        lap = C2.ConvOperator(("x", "y"), 2, taylor_order=taylor)
        record(f"laplacian_taylor{taylor}_k{k}", programs.linear(lap), u, 1, f"{k}x{k}x{k} kernel, plus-shaped taps")
    W = C2.ConvOperator()
    W.kernel = C2.ConvOperator("t", 2).kernel - 0.25 * C2.ConvOperator(("x", "y"), 2).kernel
    record("wave_k3_aligned", programs.linear(W), u, 1, "the star kernel's case")
    record("wave_k3_ragged_N0_254", programs.linear(W), u[..., :254].contiguous(), 1, "N0 % 4 != 0 -> flat-line star kernel (generic kernel before it existed)")
    c = synth.NS
    D = [C2.ConvOperator("t", 1), C2.ConvOperator("x", 1), C2.ConvOperator("y", 1), C2.ConvOperator(("x", "y"), 2)]
    ns = programs.ns_momentum(*D, c["dx"], c["dy"], c["dt"], c["nu"])
    v3 = torch.randn((32, 3, 50, 256, 256), device=dev, generator=g)
    record("ns_momentum_aligned", ns, v3, 3, "the star kernel's case")
    record("ns_momentum_ragged_N0_254", ns, v3[..., :254].contiguous(), 3, "N0 % 4 != 0 -> flat-line star kernel (generic kernel before it existed)")
    ns.path = L.PATH_GENERIC
    record("ns_momentum_aligned_forced_generic", ns, v3, 3, "same request, generic kernel forced")
    # config 1's shape: [1000, 10, 197] advection (1-D + time, ragged)
    a = synth.ADVECTION
    A = C1.ConvOperator()
    A.kernel = C1.ConvOperator("t", 1).kernel + (a["v"] * 2 * (a["t_end"] / a["nt"]) / ((a["x_max"] - a["x_min"]) / a["nx"])) * C1.ConvOperator("x", 1).kernel
    record("advection_config1_1000x10x197", programs.linear(A), torch.randn((1000, 10, 197), device=dev, generator=g), 1,
           "7.9 MB: launch-latency bound")
    # JOREK reduced-MHD programs (R rides along as a stride-0 field)
    gam = torch.tensor(5 / 3, dtype=torch.float32)
    DJ = [C2.ConvOperator(domain="t", order=1, scale=1), C2.ConvOperator(domain="x", order=1, scale=1),
          C2.ConvOperator(domain="y", order=1, scale=1), C2.ConvOperator(domain="x", order=2, scale=gam),
          C2.ConvOperator(domain="y", order=2, scale=gam)]
    R = torch.linspace(1.0, 2.0, 256)
    vj = torch.rand((32, 3, 50, 256, 256), device=dev, generator=g) + 0.5
    for eq, coef in (("continuity", 3.4), ("temperature", 2.25e-7)):
        record(f"jorek_{eq}", programs.jorek(eq, *DJ, R, coef, gam), vj, 3, "coefficient-field program")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
