"""K1 on the t-innermost layout FNO surrogates emit (pred.permute(0,1,4,2,3) of a [BS,F,Nx,Ny,Nt] tensor,
SURVEY fact 6): processed in memory order, no .contiguous() copy.  Prints Gpoints/s per layout."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cp_pre_b200  # noqa: E402
from cp_pre_b200 import _lib, programs  # noqa: E402
from oracle import synth  # noqa: E402


def timed(fn, reps=5):
    fn(); fn()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    for i in range(reps):
        e[i].record(); fn()
    e[reps].record(); torch.cuda.synchronize()
    return min(e[i].elapsed_time(e[i + 1]) for i in range(reps))


def main():
    cp_pre_b200.install()
    import Utils.ConvOps_2d as C2
    c = synth.NS
    D = [C2.ConvOperator("t", 1), C2.ConvOperator("x", 1), C2.ConvOperator("y", 1), C2.ConvOperator(("x", "y"), 2)]
    prog = programs.ns_momentum(*D, c["dx"], c["dy"], c["dt"], c["nu"])
    out = {}
    for nt in (50, 48):
        bs = 64
        raw = torch.randn((bs, 3, 256, 256, nt), device="cuda")           # what the surrogate emits: t innermost
        view = raw.permute(0, 1, 4, 2, 3)                                  # [BS,F,Nt,Nx,Ny] view the scripts pass on
        contig = view.contiguous()
        pts = bs * nt * 256 * 256
        ms_v = timed(lambda: prog.residual(view, boundary=True)); path_v = _lib.last_path()
        ms_c = timed(lambda: prog.residual(contig, boundary=True)); path_c = _lib.last_path()
        ms_copy = timed(lambda: view.contiguous())
        same = bool(torch.equal(prog.residual(view, boundary=True), prog.residual(contig, boundary=True)))
        out[f"Nt={nt}"] = {"view_path": path_v, "view_gpoints_s": round(pts / ms_v / 1e6, 1), "contiguous_path": path_c,
                           "contiguous_gpoints_s": round(pts / ms_c / 1e6, 1),
                           "copy_then_contiguous_gpoints_s": round(pts / (ms_c + ms_copy) / 1e6, 1), "bit_equal": same}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
