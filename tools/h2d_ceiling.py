"""Host -> device copy ceiling of the box (run under torchrun with N ranks, one per GPU):
every rank copies a pinned 2 GiB host buffer to its GPU with plain cudaMemcpyAsync, first alone (the other
ranks idle), then all ranks together.  Prints per-rank and aggregate GB/s: the denominator against which the
e2e line of bench.py (which is H2D-bound) has to be read.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/h2d_ceiling.py
"""
import json
import os
import subprocess
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cp_pre_b200 import dist as D  # noqa: E402


def rate(host, dev_buf, reps=4):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        dev_buf.copy_(host, non_blocking=True)
    b.record()
    torch.cuda.synchronize()
    return reps * host.numel() * 4 / (a.elapsed_time(b) * 1e-3) / 1e9


def main():
    rank, world, dev = D.init_from_env()
    host = torch.empty(1 << 29, dtype=torch.float32, pin_memory=True)      # 2 GiB
    host.fill_(1.0)
    buf = torch.empty_like(host, device=dev)
    rate(host, buf, 1)
    alone = torch.zeros(world, dtype=torch.float64, device=dev)
    for r in range(world):                       # one rank at a time
        if world > 1:
            dist.barrier()
        if r == rank:
            alone[r] = rate(host, buf)
    together = torch.zeros(world, dtype=torch.float64, device=dev)
    if world > 1:
        dist.barrier()
    together[rank] = rate(host, buf, 8)
    if world > 1:
        dist.all_reduce(alone)
        dist.all_reduce(together)
    if rank == 0:
        topo = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout
        print(json.dumps({"world": world, "h2d_gbs_alone": [round(float(v), 1) for v in alone],
                          "h2d_gbs_together": [round(float(v), 1) for v in together],
                          "aggregate_together_gbs": round(float(together.sum()), 1),
                          "host_cpus": os.cpu_count(), "topo": topo.splitlines()[:12]}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
