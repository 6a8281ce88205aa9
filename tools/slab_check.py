"""Nx-slab partition on real GPUs (torchrun, NCCL point-to-point): every rank owns an x-slab of the same
seeded field, exchanges one ghost row per side and evaluates the NS momentum residual on its padded slab;
the result must equal its slice of the undivided residual bit for bit."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cp_pre_b200  # noqa: E402
from cp_pre_b200 import dist as D, programs  # noqa: E402
from oracle import synth  # noqa: E402


def main():
    rank, world, dev = D.init_from_env()
    cp_pre_b200.install()
    import Utils.ConvOps_2d as C2
    c = synth.NS
    ops = [C2.ConvOperator("t", 1), C2.ConvOperator("x", 1), C2.ConvOperator("y", 1), C2.ConvOperator(("x", "y"), 2)]
    prog = programs.ns_momentum(*ops, c["dx"], c["dy"], c["dt"], c["nu"])
    g = torch.Generator(device=dev).manual_seed(11)                       # same field on every rank
    v = torch.randn((8, 3, 20, 256, 256), device=dev, generator=g)
    want = prog.residual(v, boundary=True)
    lo, hi = D.shard_range(256, rank, world)
    padded = torch.zeros((8, 3, 20, hi - lo + 2, 256), device=dev)
    padded[..., 1:-1, :] = v[..., lo:hi, :]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    D.slab_residual(prog, padded)                                          # warm-up (communicator set-up)
    padded[..., 0, :] = 0
    padded[..., -1, :] = 0
    e0.record()
    mine = D.slab_residual(prog, padded)
    e1.record()
    torch.cuda.synchronize()
    ok = torch.tensor([int(torch.equal(mine, want[..., lo:hi, :]))], device=dev)
    torch.distributed.all_reduce(ok, op=torch.distributed.ReduceOp.MIN)
    if rank == 0:
        print(json.dumps({"world": world, "slab_equals_undivided": bool(ok.item()), "ms": round(e0.elapsed_time(e1), 3)}))
    torch.distributed.destroy_process_group()
    assert bool(ok.item())


if __name__ == "__main__":
    main()
