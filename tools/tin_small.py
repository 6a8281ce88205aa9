"""Small t-innermost cases for a sanitizer / bounds-check pass: every mode, ragged tiles, both Nt%4 classes."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from cp_pre_b200 import _lib  # noqa: E402


def tin(t):
    nd = t.dim()
    return t.permute(list(range(nd - 3)) + [nd - 2, nd - 1, nd - 3]).contiguous().permute(
        list(range(nd - 3)) + [nd - 1, nd - 3, nd - 2])


def main():
    for wl, eq, nf in (("ns", None, 3), ("mhd", "momentum", 6), ("wave", None, 1)):
        prog = bench.make_program(wl, eq) if eq else bench.make_program(wl)
        for (bs, nt, nx, ny) in ((2, 50, 5, 44), (1, 20, 4, 130), (2, 6, 3, 4), (1, 258, 3, 10)):
            shape = (bs, nf, nt, nx, ny) if nf > 1 else (bs, nt, nx, ny)
            v = torch.rand(shape, device="cuda") + 0.5
            vt = tin(v)
            a = prog.residual(v, boundary=True)
            b = prog.residual(vt, boundary=True)
            assert _lib.last_path() == "tin" and torch.equal(a, b), (wl, shape)
            w = torch.rand(a.shape[1:], device="cuda") + 0.1
            assert torch.equal(prog.joint_scores(v, inv_modulation=w), prog.joint_scores(vt, inv_modulation=w))
            assert torch.equal(prog.coverage_counts(v, 0.5, half=w, centre=a * 0.9),
                               prog.coverage_counts(vt, 0.5, half=w, centre=b * 0.9))
            assert _lib.last_path() == "tin"
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
