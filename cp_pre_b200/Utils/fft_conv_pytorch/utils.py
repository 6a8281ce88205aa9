"""Timing / comparison helpers of the vendored fft-conv package
[ref: Utils/fft_conv_pytorch/utils.py] (unused by the hot path; kept for API parity)."""
from timeit import Timer
from typing import Callable, NamedTuple

import numpy as np
import torch
from torch import Tensor


class Benchmark(NamedTuple):
    mean: float
    std: float

    def __repr__(self):
        return f"BenchmarkResult(mean: {self.mean:.3e}, std: {self.std:.3e})"

    def __str__(self):
        return f"({self.mean:.3e} ± {self.std:.3e}) s"


def benchmark(fn: Callable, *args, num_iterations: int = 10, **kwargs) -> Benchmark:
    """Wall-clock mean/std of ``fn`` (first run discarded; CUDA work is synchronised)."""
    def run():
        fn(*args, **kwargs)
        if torch.cuda.is_available():
            torch.cuda.synchronize()
    times = Timer(run).repeat(number=1, repeat=num_iterations + 1)[1:]
    return Benchmark(float(np.mean(times)), float(np.std(times)))


def _assert_almost_equal(x: Tensor, y: Tensor) -> bool:
    err = torch.abs(x - y)
    assert err.mean().item() < 5e-5
    assert err.max().item() < 1e-4
    return True


def _gcd(x: int, y: int) -> int:
    while y:
        x, y = y, x % y
    return x
