"""Per-phase CUDA-event timing of the multi-GPU marginal quantile (run under torchrun on the GPU box):
python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/dist_profile.py"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cp_pre_b200 import dist as D, ops  # noqa: E402


class Timed:
    def __init__(self):
        self.t = {}

    def wrap(self, name, fn):
        def inner(*a, **k):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn(*a, **k)
            e1.record()
            self.t.setdefault(name, []).append((e0, e1))
            return out
        return inner

    def summary(self, skip):
        torch.cuda.synchronize()
        return {k: round(sum(a.elapsed_time(b) for a, b in v[skip:]) / max(len(v) - skip, 1), 3) for k, v in self.t.items()}


def main():
    rank, world, dev = D.init_from_env()
    n, cells = 256, 50 * 256 * 256
    g = torch.Generator(device=dev).manual_seed(rank)
    scores = torch.randn((n, cells), device=dev, generator=g).abs_()
    T = Timed()
    kern = {k: T.wrap(k, f) for k, f in D._cuda_kernels().items()}
    for name in ("all_reduce", "all_to_all_single", "all_gather"):
        setattr(dist, name, T.wrap(name, getattr(dist, name)))
    k = int(0.9 * n * world)
    reps, skip = 6, 2
    tot = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); dist.barrier()
        e0.record()
        D.marginal_quantiles(scores, [k], kernels=kern)
        e1.record(); torch.cuda.synchronize()
        tot.append(e0.elapsed_time(e1))
    out = T.summary(skip)
    # calls per quantile: all_reduce x3, all_gather x2(+2 on overflow)
    out = {k: v for k, v in out.items()}
    out["calls_per_quantile"] = {k: len(v) // reps for k, v in T.t.items()}
    out["total_ms"] = round(sum(tot[skip:]) / (reps - skip), 3)
    out["world"] = world
    out["stats"] = dict(D.LAST_STATS)
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
