"""Batch-sharded multi-GPU calibration (one process per GPU, ``torch.distributed``; NCCL over
NVLink 5 / NVSwitch on the B200 box, gloo on CPU for the host-logic tests).

PRE evaluation, scoring and coverage act within one sample, so the batch axis shards with no
halo and no data-path collective.  Only calibration couples samples:

  Joint     per-sample scores are rank-local; the global alpha-quantile is found either by
            all-gathering the <= BS floats (``joint_quantile_gather``) or -- the form that scales
            to any BS -- by four 8-bit radix passes whose 256-bin histograms are all-reduced
            (``joint_quantile_hist``, kernel ``cppre_radix_hist``).  The per-cell modulation (std
            over the GLOBAL batch) is merged Chan-style from per-rank (count, mean, M2) with two
            all-reduces (``global_cell_std``).
  Marginal  per-cell quantiles over a batch split across ranks: scores are transposed
            batch-shard -> cell-shard with one all-to-all, each rank selects on its cell slab,
            the slabs of qhat are all-gathered (``marginal_quantiles``).

The kernels are injected (``hist_fn`` / ``select_fn`` / ``moments_fn``) so the orchestration can
be tested on CPU with the oracle's NumPy implementations; the defaults are the CUDA kernels and
there is no CPU fallback in the product path.
"""
from __future__ import annotations

import os
from typing import Callable, Optional, Sequence

import torch
import torch.distributed as dist


def init_from_env(backend: Optional[str] = None):
    """Reads RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* (torchrun) and initialises the process
    group.  Returns (rank, world, device)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    use_cuda = torch.cuda.is_available()
    device = torch.device("cuda", local) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(device)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        kw = {"device_id": device} if use_cuda else {}
        dist.init_process_group(backend or ("nccl" if use_cuda else "gloo"), rank=rank, world_size=world, **kw)
    return rank, world, device


def shard_range(total: int, rank: int, world: int):
    """Contiguous slab [lo, hi) of ``total`` units for ``rank`` (first ``total % world`` ranks
    get one extra)."""
    base, extra = divmod(total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def _world(group=None) -> int:
    return dist.get_world_size(group) if dist.is_initialized() else 1


# ---------------------------------------------------------------------------------------
# Joint: modulation over the global batch
# ---------------------------------------------------------------------------------------
def merge_moments(mean: torch.Tensor, m2: torch.Tensor, count: float, group=None):
    """Chan et al. pairwise merge of per-rank (count, mean, M2) -> global (mean, M2, count) with
    two all-reduces of [cells] fp64:  mean = sum(n_r mean_r)/N;  M2 = sum(M2_r + n_r (mean_r - mean)^2)."""
    if _world(group) == 1:
        return mean, m2, count
    n = torch.tensor([float(count)], dtype=torch.float64, device=mean.device)
    s = mean * float(count)
    dist.all_reduce(n, group=group)
    dist.all_reduce(s, group=group)
    total = float(n.item())
    gmean = s / total
    g2 = m2 + float(count) * (mean - gmean) ** 2
    dist.all_reduce(g2, group=group)
    return gmean, g2, total


def global_cell_std(a: torch.Tensor, b: Optional[torch.Tensor] = None, eps: float = 0.0, group=None,
                    moments_fn: Optional[Callable] = None, std_fn: Optional[Callable] = None):
    """Population std over the GLOBAL sample axis of a-b, per cell; (std, 1/std) shaped like a[0]."""
    if moments_fn is None or std_fn is None:
        from . import ops
        moments_fn, std_fn = ops.cell_moments, ops.cell_std_from_m2
    mean, m2 = moments_fn(a, b)
    _, g2, total = merge_moments(mean, m2, a.shape[0], group)
    sd, inv = std_fn(g2, total, eps, True)
    return sd.reshape(a.shape[1:]), inv.reshape(a.shape[1:])


# ---------------------------------------------------------------------------------------
# Joint: global quantile of per-sample scores
# ---------------------------------------------------------------------------------------
def joint_quantile_gather(scores: torch.Tensor, ranks: Sequence[int], group=None,
                          select_fn: Optional[Callable] = None) -> torch.Tensor:
    """All-gather the rank-local scores (equal counts per rank) and select on every rank."""
    if select_fn is None:
        from . import ops
        select_fn = ops.select_ranks
    w = _world(group)
    if w > 1:
        parts = [torch.empty_like(scores) for _ in range(w)]
        dist.all_gather(parts, scores.contiguous(), group=group)
        scores = torch.cat(parts)
    return select_fn(scores, ranks)


def _key_to_float(key: int) -> float:
    import struct
    bits = (key & 0x7FFFFFFF) if key & 0x80000000 else (~key & 0xFFFFFFFF)
    return struct.unpack("<f", struct.pack("<I", bits))[0]


def joint_quantile_hist(scores: torch.Tensor, rank_k: int, group=None,
                        hist_fn: Optional[Callable] = None) -> float:
    """k-th smallest (0-based, over ALL ranks' scores) by radix select on order-preserving
    uint32 keys: 4 passes x (local 256-bin histogram kernel -> all-reduce -> pick digit).
    Exchanges 4 KiB per rank in total, whatever the batch size.  NaN anywhere -> NaN (np.quantile)."""
    if hist_fn is None:
        from . import ops
        hist_fn = ops.radix_hist
    nan = torch.isnan(scores).any().to(torch.int32).reshape(1)
    if _world(group) > 1:
        dist.all_reduce(nan, op=dist.ReduceOp.MAX, group=group)
    prefix = mask = 0
    k = int(rank_k)
    for shift in (24, 16, 8, 0):
        hist = torch.zeros(256, dtype=torch.int32, device=scores.device)
        hist_fn(scores, prefix, mask, shift, hist)
        if _world(group) > 1:
            dist.all_reduce(hist, group=group)
        c = torch.cumsum(hist.to(torch.int64), 0).cpu()
        if k >= int(c[-1]):
            raise ValueError(f"rank {rank_k} out of range for {int(c[-1])} scores")
        d = int(torch.searchsorted(c, torch.tensor(k), right=True))
        k -= int(c[d - 1]) if d else 0
        prefix |= d << shift
        mask |= 255 << shift
    return float("nan") if int(nan.item()) else _key_to_float(prefix)


# ---------------------------------------------------------------------------------------
# Marginal: per-cell quantiles over a batch split across ranks
# ---------------------------------------------------------------------------------------
def marginal_quantiles(scores: torch.Tensor, ranks: Sequence[int], group=None,
                       select_fn: Optional[Callable] = None) -> torch.Tensor:
    """scores: rank-local [n_local, cells...] (same n_local on every rank).  Returns
    [len(ranks), cells...] quantiles over the global n_local*world samples, on every rank.
    Cells are padded to a multiple of the world size for the exchange."""
    if select_fn is None:
        from . import ops
        select_fn = ops.select_ranks
    w = _world(group)
    if w == 1:
        return select_fn(scores, ranks)
    n = scores.shape[0]
    cell_shape = tuple(scores.shape[1:])
    flat = scores.reshape(n, -1)
    cells = flat.shape[1]
    slab = (cells + w - 1) // w
    if slab * w != cells:
        flat = torch.nn.functional.pad(flat, (0, slab * w - cells))
    # batch-shard -> cell-shard: chunk r of the send buffer is this rank's samples of slab r
    send = flat.reshape(n, w, slab).permute(1, 0, 2).contiguous()
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=group)          # recv[r] = rank r's samples of MY slab
    q_slab = select_fn(recv.reshape(w * n, slab), ranks).contiguous()    # [R, slab]
    parts = [torch.empty_like(q_slab) for _ in range(w)]
    dist.all_gather(parts, q_slab, group=group)
    q = torch.cat(parts, dim=1)[:, :cells]
    return q.reshape((len(ranks),) + cell_shape)


def allreduce_counts(*values: int, device=None, group=None):
    """Sums a few integer counters (coverage numerators / denominators) over ranks."""
    t = torch.tensor([int(v) for v in values], dtype=torch.int64, device=device)
    if _world(group) > 1:
        dist.all_reduce(t, group=group)
    return [int(v) for v in t.tolist()]
