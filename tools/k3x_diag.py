"""Where do the overflowing cells of the fused multi-GPU quantile come from?  World emulated on one GPU with
independent shards (|N(0,1)| scores).  python tools/k3x_diag.py [world] [slots] [network]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cp_pre_b200 import dist as D, ops  # noqa: E402


def main():
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    slots = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    n, cells = 256, 50 * 256 * 256
    network = int(sys.argv[3]) if len(sys.argv) > 3 else D.owner_network(world, slots)
    N = world * n
    kt = N - int(0.9 * N)
    J = -(-kt // world)
    slab = -(-cells // world)
    cells_pad = (slab * world + 3) // 4 * 4
    keys = torch.zeros((world, cells_pad), dtype=torch.int32, device="cuda")
    recv = [torch.zeros((world, 1 + slots, slab), device="cuda") for _ in range(world)]
    qbuf = [torch.zeros((cells_pad,), device="cuda") for _ in range(world)]
    ov = torch.zeros(1, dtype=torch.int32, device="cuda")
    ovc = torch.full((65536,), -1, dtype=torch.int32, device="cuda")
    shards = []
    for r in range(world):
        g = torch.Generator(device="cuda").manual_seed(r)
        shards.append(torch.randn((n, cells), device="cuda", generator=g).abs_())
        ops.peer_bcast_keys(ops.select_ranks(shards[r], [n - J])[0], cells_pad, [keys.data_ptr()] * world, r)
    for r in range(world):
        ops.window_exchange(shards[r], keys, slots, slab, [t.data_ptr() for t in recv], r, ov, ovc)
    out = {"world": world, "slots": slots, "network": network, "overflow_after_window": int(ov.item())}
    hdr = torch.stack([t[:, 0].view(torch.int32) for t in recv])          # [owner, rank, slab]
    inside = ((hdr >> 24) & 0x7f).float()
    tot = inside.sum(1)
    out["inside_per_rank_mean"] = float(inside.mean())
    out["inside_per_rank_max"] = float(inside.max())
    out["total_mean"] = float(tot.mean())
    out["total_q999"] = float(torch.quantile(tot.flatten()[:4_000_000], 0.999))
    out["total_max"] = float(tot.max())
    out["cells_total_gt_network"] = int((tot > network).sum())
    for me in range(world):
        ops.owner_select(recv[me], slab, kt, network, me * slab, [qbuf[me].data_ptr()], ov, ovc)
    out["overflow_after_owner"] = int(ov.item())
    print(json.dumps(out))


if __name__ == "__main__":
    main()
