"""Times the conformal kernels (K2 moments, K3 select, K5 coverage) with CUDA events on
score tensors of the benchmark shapes.  Run on the GPU box:  python tools/k3_sweep.py"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cp_pre_b200 import ops  # noqa: E402


def timeit(fn, reps=7):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def main():
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(0)
    cells = 50 * 256 * 256
    ns = [int(v) for v in sys.argv[1].split(',')] if len(sys.argv) > 1 else (32, 64, 128, 256, 512, 1024, 2048)
    for n in ns:
        c = cells if n <= 512 else cells // 4
        s = torch.rand((n, c), device=dev, generator=g)
        for nr, ranks in ((1, [int(0.9 * n)]), (10, [int(q * (n - 1)) for q in np.linspace(0.05, 0.95, 10)])):
            ms = timeit(lambda: ops.select_ranks(s, ranks))
            print(json.dumps({"op": "select_ranks", "n": n, "cells": c, "ranks": nr, "ms": round(ms, 3),
                              "Gscores_s": round(n * c / ms / 1e6, 1), "GBs": round(4 * n * c / ms / 1e6, 0)}), flush=True)
        if n == 256 and len(sys.argv) <= 1:
            q = ops.select_ranks(s, [230])[0]
            ms = timeit(lambda: ops.inside_count_band(s, None, q, 1.0))
            print(json.dumps({"op": "inside_count_band", "n": n, "cells": c, "ms": round(ms, 3),
                              "GBs": round(4 * n * c / ms / 1e6, 0)}), flush=True)
            ms = timeit(lambda: ops.cell_moments(s))
            print(json.dumps({"op": "cell_moments", "n": n, "cells": c, "ms": round(ms, 3),
                              "GBs": round(4 * n * c / ms / 1e6, 0)}), flush=True)
            w = torch.rand(c, device=dev, generator=g) + 0.5
            ms = timeit(lambda: ops.joint_score(s, None, w))
            print(json.dumps({"op": "joint_score", "n": n, "cells": c, "ms": round(ms, 3),
                              "GBs": round(4 * n * c / ms / 1e6, 0)}), flush=True)
        del s


if __name__ == "__main__":
    main()
