"""K3x kernels at full size with the ranks emulated on ONE GPU (plain device buffers stand in for the
peers' symmetric memory): CUDA-event time per kernel.  python tools/k3x_profile.py [world] [slots]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cp_pre_b200 import dist as D, ops  # noqa: E402


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    e[0].record()
    for i in range(reps):
        fn()
        e[i + 1].record()
    torch.cuda.synchronize()
    return round(min(e[i].elapsed_time(e[i + 1]) for i in range(reps)), 4)


def main():
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    slots = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    n, cells = 256, 50 * 256 * 256
    g = torch.Generator(device="cuda").manual_seed(0)
    s = torch.randn((n, cells), device="cuda", generator=g).abs_()     # every emulated rank scans this shard
    N = world * n
    kt = N - int(0.9 * N)
    J = -(-kt // world)
    slab = -(-cells // world)
    cells_pad = (slab * world + 3) // 4 * 4
    keys = torch.zeros((world, cells_pad), dtype=torch.int32, device="cuda")
    recv = [torch.zeros((world, 1 + slots, slab), device="cuda") for _ in range(world)]
    qbuf = [torch.zeros((cells_pad,), device="cuda") for _ in range(world)]
    ov = torch.zeros(1, dtype=torch.int32, device="cuda")
    ovc = torch.full((8192,), -1, dtype=torch.int32, device="cuda")
    out = {"world": world, "slots": slots}
    a = ops.select_ranks(s, [n - J])[0]
    out["select"] = timed(lambda: ops.select_ranks(s, [n - J]))
    # the other ranks' statistics: this shard's, jittered by a few ulps so the bracket has a width
    for r in range(world):
        ar = a if r == 0 else a * (1.0 + 0.02 * torch.randn_like(a))
        tmp = torch.zeros((world, cells_pad), dtype=torch.int32, device="cuda")
        ops.peer_bcast_keys(ar, cells_pad, [tmp.data_ptr()], 0)
        keys[r] = tmp[0]
    out["bcast_keys"] = timed(lambda: ops.peer_bcast_keys(a, cells_pad, [keys.data_ptr()] * world, 0))
    ops.peer_bcast_keys(a, cells_pad, [keys.data_ptr()], 0)     # row 0 back to this shard's own statistic
    ptrs = [t.data_ptr() for t in recv]
    out["window_x"] = timed(lambda: ops.window_exchange(s, keys, slots, slab, ptrs, 0, ov, ovc))
    for me in range(1, world):
        ops.window_exchange(s, keys, slots, slab, ptrs, me, ov, ovc)
    hdr = recv[0][:, 0].view(torch.int32)
    out["mean_inside"] = float(((hdr >> 24) & 0x7f).float().mean())
    out["owner"] = timed(lambda: ops.owner_select(recv[0], slab, kt, D.owner_network(world, slots), 0,
                                                  [qbuf[0].data_ptr()], ov, ovc))
    out["peer_copy"] = timed(lambda: ops.peer_copy(qbuf[0][:slab], 0, [t.data_ptr() for t in qbuf[1:]]))
    out["overflow"] = int(ov.item())
    print(json.dumps(out))


if __name__ == "__main__":
    main()
