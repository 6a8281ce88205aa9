"""What a reference script gains UNCHANGED from the drop-in.

The reference's residual functions are Python expressions over ~10 ConvOperator calls plus ~30 torch
elementwise ops on whatever tensors the script holds -- CPU tensors by default.  This
times `residual_momentum` of Joint/NS_Residuals_CP.py:231-240, written exactly as in the script, on a
[BS,3,50,256,256] batch in four settings:

  cpu_reference   the script's own arithmetic on CPU tensors: F.conv3d per operator (oracle/torch_port.py)
  dropin_cpu      the script unchanged, `from Utils.ConvOps_2d import ConvOperator` resolved to this repo:
                  every D_k(u) stages its CPU input to the GPU, runs K1, brings the result back; the
                  elementwise algebra stays on the host like in the script
  dropin_cuda     the script with `vars = vars.cuda()` in front (one line changed): K1 per operator, the
                  elementwise algebra as torch CUDA ops
  fused           programs.ns_momentum(D_t, D_x, D_y, D_xx_yy, ...)(vars): the whole expression in one K1 launch

    python tools/dropin_script_bench.py [BS]     -> one JSON line
"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cp_pre_b200  # noqa: E402
from cp_pre_b200 import programs  # noqa: E402
from oracle import synth, torch_port as TP  # noqa: E402

cp_pre_b200.install()
from Utils.ConvOps_2d import ConvOperator  # noqa: E402

c = synth.NS
dx, dy, dt, nu = c["dx"], c["dy"], c["dt"], c["nu"]
D_t = ConvOperator(domain='t', order=1)
D_x = ConvOperator(domain='x', order=1)
D_y = ConvOperator(domain='y', order=1)
D_xx_yy = ConvOperator(domain=('x', 'y'), order=2)


def residual_momentum(vars, boundary=False):          # verbatim shape of Joint/NS_Residuals_CP.py:231-240
    u, v, p = vars[:, 0], vars[:, 1], vars[:, 2]
    res_x = D_t(u)*dx*dy + u*D_x(u)*dt*dy + v*D_y(u)*dt*dx - nu*D_xx_yy(u)*dt + D_x(p)*dt*dy
    res_y = D_t(v)*dx*dy + u*D_x(v)*dt*dx + v*D_y(v)*dt*dy - nu*D_xx_yy(v)*dt + D_y(p)*dt*dx
    if boundary:
        return res_x + res_y
    return res_x[..., 1:-1, 1:-1, 1:-1] + res_y[..., 1:-1, 1:-1, 1:-1]


def timed(fn, reps, sync):
    fn()
    if sync:
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    if sync:
        torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps, out


def main():
    bs = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    torch.set_num_threads(os.cpu_count())
    v = torch.from_numpy(synth.fields(3, 3, (bs, 50, 256, 256)))
    pts = bs * 50 * 256 * 256
    t_ref, r_ref = timed(lambda: TP.ns_momentum(v, dx, dy, dt, nu), 1, False)
    t_cpu, r_cpu = timed(lambda: residual_momentum(v), 2, True)
    vc = v.cuda()
    t_cuda, r_cuda = timed(lambda: residual_momentum(vc), 5, True)
    prog = programs.ns_momentum(D_t, D_x, D_y, D_xx_yy, dx, dy, dt, nu)
    t_fused, r_fused = timed(lambda: prog.residual(vc), 10, True)
    scale = float(r_ref.abs().max())
    err = lambda r: float((r.cpu() - r_ref).abs().max()) / scale
    print(json.dumps({
        "what": "Joint/NS_Residuals_CP.py residual_momentum on [%d,3,50,256,256]" % bs, "points": pts, "host_threads": os.cpu_count(),
        "gpoints_s": {"cpu_reference": pts / t_ref / 1e9, "dropin_cpu_tensors": pts / t_cpu / 1e9,
                      "dropin_cuda_tensors": pts / t_cuda / 1e9, "fused_program": pts / t_fused / 1e9},
        "seconds": {"cpu_reference": t_ref, "dropin_cpu_tensors": t_cpu, "dropin_cuda_tensors": t_cuda, "fused_program": t_fused},
        "max_rel_err_vs_cpu_reference": {"dropin_cpu_tensors": err(r_cpu), "dropin_cuda_tensors": err(r_cuda), "fused_program": err(r_fused)},
    }))


if __name__ == "__main__":
    main()
