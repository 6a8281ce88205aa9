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
    from cp_pre_b200 import ops
    from Neural_PDE.UQ.inductive_cp import calibrate_rank
    for nt in (50, 48):
        bs = 128
        raw = torch.randn((bs, 3, 256, 256, nt), device="cuda")           # what the surrogate emits: t innermost
        view = raw.permute(0, 1, 4, 2, 3)                                  # [BS,F,Nt,Nx,Ny] view the scripts pass on
        contig = view.contiguous()
        pts = bs * nt * 256 * 256
        ms_v = timed(lambda: prog.residual(view, boundary=True)); path_v = _lib.last_path()
        ms_c = timed(lambda: prog.residual(contig, boundary=True)); path_c = _lib.last_path()
        ms_copy = timed(lambda: view.contiguous())
        same = bool(torch.equal(prog.residual(view, boundary=True), prog.residual(contig, boundary=True)))
        rec = {"view_path": path_v, "view_gpoints_s": round(pts / ms_v / 1e6, 1), "contiguous_path": path_c,
               "contiguous_gpoints_s": round(pts / ms_c / 1e6, 1),
               "copy_then_contiguous_gpoints_s": round(pts / (ms_c + ms_copy) / 1e6, 1), "bit_equal": same}
        # the whole marginal-CP step (STORE |R| -> per-cell quantile -> COVER) with every tensor t-innermost
        n_cal = bs // 2
        k = calibrate_rank(n_cal, n_cal, 0.1)
        sc_t = torch.empty((n_cal, 256, 256, nt), device="cuda")
        sc_c = torch.empty((n_cal, nt, 256, 256), device="cuda")

        def step(v, sc, tin):
            prog.abs_scores(v[:n_cal], boundary=True, out=sc.permute(0, 3, 1, 2) if tin else sc)
            q = ops.select_ranks(sc, [k])[0]
            return q, prog.coverage_counts(v[n_cal:], 1.0, half=q.permute(2, 0, 1) if tin else q)

        ms_st = timed(lambda: step(view, sc_t, True)); path_st = _lib.last_path()
        ms_sc = timed(lambda: step(contig, sc_c, False))
        qt, ct = step(view, sc_t, True)
        qc, cc = step(contig, sc_c, False)
        rec["cp_step"] = {"view_path": path_st, "view_ms": round(ms_st, 3), "contiguous_ms": round(ms_sc, 3),
                          "view_gpoints_s": round(pts / ms_st / 1e6, 1), "contiguous_gpoints_s": round(pts / ms_sc / 1e6, 1),
                          "qhat_bit_equal": bool(torch.equal(qt.permute(2, 0, 1), qc)),
                          "coverage_counts_equal": bool(torch.equal(ct, cc))}
        out[f"Nt={nt}"] = rec
        del raw, view, contig, sc_t, sc_c
    # config 5's kernel on the same layout (MHD momentum, 6 fields, 128x128x20 grid)
    import bench
    mp = bench.make_program("mhd", "momentum")
    raw = (torch.rand((512, 6, 128, 128, 20), device="cuda") + 0.5)
    view, pts = raw.permute(0, 1, 4, 2, 3), 512 * 20 * 128 * 128
    contig = view.contiguous()
    ms_v = timed(lambda: mp.residual(view, boundary=True)); path_v = _lib.last_path()
    ms_c = timed(lambda: mp.residual(contig, boundary=True))
    out["mhd_momentum"] = {"view_path": path_v, "view_gpoints_s": round(pts / ms_v / 1e6, 1),
                           "contiguous_gpoints_s": round(pts / ms_c / 1e6, 1),
                           "bit_equal": bool(torch.equal(mp.residual(view, boundary=True), mp.residual(contig, boundary=True)))}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
