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
  Marginal  per-cell quantiles over a batch split across ranks (``marginal_quantiles``): every
            rank selects one local order statistic per cell; their all-reduced min / max bracket
            the global quantile; one pass over the local scores collects the few samples inside
            the bracket (K3w); only those candidates cross NVLink (all-to-all, batch-shard ->
            cell-shard) and the owner of a cell slab picks the answer.  Exact; a cell whose
            bracket holds more candidates than the window falls back to transposing the whole
            score tensor.

  Nx-slab   a single field too large for one GPU is split along x instead: one ghost row per
            side travels (``exchange_x_halo``), everything else -- PRE, per-cell calibration --
            is rank-local (``slab_residual``).

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
_SIDE_STREAMS = {}


def _rank_counts(n_local: int, group) -> list:
    """Every rank's local sample count, as one all-reduce that EVERY rank issues on EVERY call.

    There is deliberately no cache here: whether a rank may skip the collective cannot be decided from
    its own ``n_local`` (the global total is not a function of one rank's count), so a rank-local cache
    lets ranks disagree on hit vs miss -- one raises from a stale entry while its peer waits alone in
    the all-reduce.  Under NCCL the collective and the host read run on a side stream that has nothing
    else queued, so they do not wait for (or stall) the K1 work already enqueued on the caller's stream."""
    w, me = _world(group), dist.get_rank(group)
    mine = [0] * w
    mine[me] = int(n_local)
    if dist.get_backend(group) != "nccl":
        t = torch.tensor(mine, dtype=torch.int64)
        dist.all_reduce(t, group=group)
        return t.tolist()
    dev = torch.device("cuda", torch.cuda.current_device())
    side = _SIDE_STREAMS.get(dev.index)
    if side is None:
        side = _SIDE_STREAMS[dev.index] = torch.cuda.Stream(dev)
    with torch.cuda.stream(side):
        t = torch.tensor(mine, dtype=torch.int64, device=dev)
        dist.all_reduce(t, group=group)
        return t.tolist()                       # synchronises the side stream only


def global_count(n_local: int, group=None, require_equal: bool = False) -> int:
    """Sum of the ranks' local sample counts (one small all-reduce per call, see ``_rank_counts``).
    ``require_equal`` raises -- on every rank alike -- when the shards are uneven (``shard_range``
    produces such shards whenever BS % world != 0; the windowed marginal quantile and the equal-count
    gathers assume equal shards and would otherwise return a wrong quantile or hang in NCCL)."""
    if _world(group) == 1:
        return int(n_local)
    counts = _rank_counts(n_local, group)
    n_min, n_max = min(counts), max(counts)
    if require_equal and n_max != n_min:
        raise ValueError(f"uneven batch shards (between {n_min} and {n_max} samples per rank): this exchange needs the "
                         "same number of samples on every rank -- pad or trim the batch to a multiple of the world size")
    return int(sum(counts))


def merge_moments(mean: torch.Tensor, m2: torch.Tensor, count: float, group=None):
    """Chan et al. pairwise merge of per-rank (count, mean, M2) -> global (mean, M2, count) with
    two all-reduces of [cells] fp64:  mean = sum(n_r mean_r)/N;  M2 = sum(M2_r + n_r (mean_r - mean)^2).
    The global count comes from ``global_count`` before the first of them, so no host round trip
    sits between the two all-reduces."""
    if _world(group) == 1:
        return mean, m2, count
    total = float(global_count(int(count), group))
    s = mean * float(count)
    dist.all_reduce(s, group=group)
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
        global_count(scores.shape[0], group, require_equal=True)
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


def joint_quantile_device(scores: torch.Tensor, rank_k: int, group=None, select_fn: Optional[Callable] = None) -> torch.Tensor:
    """``joint_quantile_hist`` with the radix state on the device (kernels K3d): four passes of
    (local 257-counter histogram -> all-reduce -> pick digit on the device) enqueued back to back
    on the current stream.  Returns a one-element float32 device tensor that the coverage launch
    takes as ``band_q`` -- no host round trip anywhere between calibration and coverage.
    ``select_fn(scores, k, allreduce)`` replaces the CUDA kernels in the CPU tests."""
    if select_fn is None:
        from . import ops
        select_fn = ops.radix_select
    n_total = global_count(scores.numel(), group)
    if not 0 <= int(rank_k) < n_total:
        raise ValueError(f"rank {rank_k} out of range for {n_total} scores")
    ar = (lambda h: dist.all_reduce(h, group=group)) if _world(group) > 1 else None
    return select_fn(scores, int(rank_k), ar)


# ---------------------------------------------------------------------------------------
# Marginal: per-cell quantiles over a batch split across ranks
# ---------------------------------------------------------------------------------------
def _cuda_kernels():
    from . import ops
    return {"select": ops.select_ranks, "keys": ops.ordered_keys, "window": ops.window_compact,
            "top": ops.select_from_top, "peer_exchange": True, "bcast_keys": ops.peer_bcast_keys,
            "window_x": ops.window_exchange, "owner": ops.owner_select, "peer_copy": ops.peer_copy}


_SYMM = {}
PENDING = []      # overflow counters of deferred fused quantiles (device tensors)
LAST_STATS = {}   # diagnostics of the last fused quantile (overflowing cells, network size)
_TUNE = {}     # self-tuning of the owner's network size, keyed by problem shape


def _symm_buffer(tag: str, shape, dtype, device, group):
    """A buffer of CUDA symmetric memory, allocated once per (group, shape) and mapped into every
    process of the group: (local tensor, every rank's base pointer as seen from this process)."""
    import torch.distributed._symmetric_memory as symm
    name = (group or dist.group.WORLD).group_name
    key = (tag, name, tuple(shape), dtype, device.index)
    if key not in _SYMM:
        t = symm.empty(tuple(shape), dtype=dtype, device=device)
        hdl = symm.rendezvous(t, name)
        _SYMM[key] = (t, hdl, [int(p) for p in hdl.buffer_ptrs])
    t, _, ptrs = _SYMM[key]
    return t, ptrs


def _peer_recv_buffer(world: int, slots: int, slab: int, device, group):
    """Every rank's candidate receive buffer [world, slots, slab] (K3w stores into the peers'
    buffers directly over NVLink, no all-to-all call)."""
    return _symm_buffer("recv", (world, slots, slab), torch.float32, device, group)


def _keys_by_allreduce(world: int) -> bool:
    """Bracket exchange: NCCL MAX all-reduce (True) or peer-store broadcast of the keys (False).
    CPPRE_KEYS_ALLREDUCE=0/1 overrides; default: all-reduce from 4 ranks up, where the broadcast's
    world x 13 MB per rank starts to dominate."""
    env = os.environ.get("CPPRE_KEYS_ALLREDUCE")
    return (env != "0") if env is not None else world >= 4


def _barrier(flag: torch.Tensor, group):
    """Orders the peer stores of the kernel just enqueued before the peers' next kernel: a
    one-word all-reduce on the same stream (nothing else travels)."""
    dist.all_reduce(flag, group=group)


def pending_overflow() -> int:
    """Settles every deferred fused quantile since the last call: one host sync, returns how many
    cells could not be answered from their candidate windows (0 = every deferred result stands)."""
    total = 0
    while PENDING:
        total += int(PENDING.pop().sum().item())
    return total


def fused_slab(cells: int, world: int) -> int:
    """Cells owned by one rank in the fused exchange: ceil(cells / world) rounded up to whole 32-cell
    tiles (a warp of the sender packs the candidate lists of exactly one owner tile)."""
    return (-(-cells // world) + 31) // 32 * 32


def _fused_window_quantile(scores: torch.Tensor, rank_k: int, group, kern, slots: int, defer: bool = False):
    """K3x: `_window_quantile` with every exchange inside the kernels (include/cppre.h).  Per
    quantile: K3t local order statistic -> keys stored into every peer -> barrier -> window scan,
    headers + candidates stored into the owners -> barrier -> owner select, qhat stored into
    every peer -> all-gather of the overflow counters (the last barrier).  Same bracket argument
    and the same exact redo of overflowing cells as `_window_quantile`."""
    w = _world(group)
    me = dist.get_rank(group)
    n = scores.shape[0]
    global_count(n, group, require_equal=True)
    flat = scores.reshape(n, -1)
    cells = flat.shape[1]
    dev = scores.device
    kt = w * n - int(rank_k)
    J = -(-kt // w)
    slab = fused_slab(cells, w)
    cells_pad = (slab * w + 3) // 4 * 4
    # every rank's receive buffer: per sender and 32-cell tile one block [header row | dense candidate rows]
    recv, recv_ptrs = _symm_buffer("recvx", (w, slab // 32, 1 + slots, 32), torch.float32, dev, group)
    qbuf, q_ptrs = _symm_buffer("q", (cells_pad,), torch.float32, dev, group)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    over = torch.zeros(1, dtype=torch.int32, device=dev)
    over_cells = torch.full((OVERFLOW_CAP,), -1, dtype=torch.int32, device=dev)

    a = kern["select"](flat, [n - J])[0]
    if _keys_by_allreduce(w):
        # bracket by one MAX all-reduce of [~key, key] (2 x 13 MB through the switch, reduced in it
        # where NVLS is up) instead of world x 13 MB of peer stores + a barrier
        ak = kern["keys"](a).reshape(-1)
        both = torch.empty((2, cells_pad), dtype=torch.int32, device=dev)
        both[0, :cells] = ~ak
        both[1, :cells] = ak
        dist.all_reduce(both, op=dist.ReduceOp.MAX, group=group)
        kern["window_x"](flat, both, slots, slab, recv_ptrs, me, over, over_cells, 1)
    else:
        keys, key_ptrs = _symm_buffer("keys", (w, cells_pad), torch.int32, dev, group)
        kern["bcast_keys"](a, cells_pad, key_ptrs, me)
        _barrier(flag, group)
        kern["window_x"](flat, keys, slots, slab, recv_ptrs, me, over, over_cells)
    _barrier(flag, group)
    live = max(0, min(slab, cells - me * slab))
    tune = ("net", w, n, cells, slots)
    network = min(1024, owner_network(w, slots) << _TUNE.get(tune, 0))
    kern["owner"](recv, live, kt, network, me * slab, [q_ptrs[me]], over, over_cells)     # my slab, written locally
    if live > 0:                                                                          # ... then pushed to the peers
        kern["peer_copy"](qbuf[me * slab:me * slab + live], me * slab, [p for r, p in enumerate(q_ptrs) if r != me])
    counts = [torch.empty_like(over) for _ in range(w)]
    dist.all_gather(counts, over, group=group)                # also the barrier that orders the q stores
    q = qbuf[:cells].clone()
    if defer:
        # no host round trip here: the caller consumes q right away and settles the overflow
        # counters of all deferred calls later with `pending_overflow()`; a non-zero total means
        # those results must be recomputed with defer=False (which redoes the cells exactly)
        PENDING.append(torch.stack(counts))
        return q
    counts = torch.stack(counts).cpu()
    LAST_STATS.update(overflow_cells=int(counts.sum()), network=network, slots=slots)
    if int(counts.max()) > OVERFLOW_CAP // 2 and network < 1024:
        # close to the point where the whole tensor would be transposed instead: sort more keys per
        # cell from now on (a few thousand redone cells are cheaper than a network twice the size);
        # the counts are the same on every rank, so every rank widens alike
        _TUNE[tune] = _TUNE.get(tune, 0) + 1
    if int(counts.sum()) == 0:
        return q
    if int(counts.max()) > OVERFLOW_CAP:
        return None
    return _redo_cells(flat, q, over_cells, rank_k, group, kern["select"])


def _redo_cells(flat, q, over_cells, rank_k, group, select_fn):
    """Cells on some rank's overflow list, redone exactly from their gathered full columns."""
    w = _world(group)
    cells = flat.shape[1]
    lists = [torch.empty_like(over_cells) for _ in range(w)]
    dist.all_gather(lists, over_cells, group=group)
    ids = torch.unique(torch.cat(lists))
    ids = ids[(ids >= 0) & (ids < cells)].to(torch.int64)
    cols = flat[:, ids].contiguous()
    allcols = [torch.empty_like(cols) for _ in range(w)]
    dist.all_gather(allcols, cols, group=group)
    q[ids] = select_fn(torch.cat(allcols), [int(rank_k)])[0].to(q.dtype)
    return q


def _exchange_quantiles(scores: torch.Tensor, ranks: Sequence[int], group, select_fn) -> torch.Tensor:
    """Exact slow path: transpose the WHOLE score tensor batch-shard -> cell-shard with one
    all-to-all, select on [world*n, slab], all-gather the slabs of qhat."""
    w = _world(group)
    n = scores.shape[0]
    global_count(n, group, require_equal=True)
    flat = scores.reshape(n, -1)
    cells = flat.shape[1]
    slab = (cells + w - 1) // w
    if slab * w != cells:
        flat = torch.nn.functional.pad(flat, (0, slab * w - cells))
    # chunk r of the send buffer is this rank's samples of slab r
    send = flat.reshape(n, w, slab).permute(1, 0, 2).contiguous()
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=group)          # recv[r] = rank r's samples of MY slab
    q_slab = select_fn(recv.reshape(w * n, slab), ranks).contiguous()    # [R, slab]
    parts = [torch.empty_like(q_slab) for _ in range(w)]
    dist.all_gather(parts, q_slab, group=group)
    return torch.cat(parts, dim=1)[:, :cells]


def _window_quantile(scores: torch.Tensor, rank_k: int, group, kern, slots: int):
    """One global order statistic per cell, exchanging `slots` candidates per cell and rank
    instead of the n local samples (include/cppre.h, K3w).  Returns q [cells], or None when more
    cells overflowed the candidate window than the redo list holds (the caller then transposes
    the whole tensor).  The few cells whose bracket holds more than `slots` samples on some rank
    are redone exactly from their full columns (all-gathered).

    With kt = N - rank_k (the kt-th largest of the N = world*n global samples) and
    J = ceil(kt / world): every rank's J-th largest a_r satisfies min_r a_r <= g <= max_r a_r
    for the global answer g (>= : each rank holds J samples >= its a_r, so world*J >= kt samples
    are >= the minimum; <= : a sample above max_r a_r is among the J-1 largest of its rank, so
    fewer than kt samples exceed the maximum).  g is then the (kt - #{x > hi})-th largest of
    the samples inside [lo, hi], and only those travel."""
    w = _world(group)
    me = dist.get_rank(group)
    n = scores.shape[0]
    global_count(n, group, require_equal=True)
    flat = scores.reshape(n, -1)
    cells = flat.shape[1]
    kt = w * n - int(rank_k)
    J = -(-kt // w)
    a = kern["keys"](kern["select"](flat, [n - J])[0])        # int32 ordered keys, [cells]
    both = torch.stack([~a, a])                              # one MAX all-reduce: min(a) = ~max(~a)
    dist.all_reduce(both, op=dist.ReduceOp.MAX, group=group)
    lo, hi = ~both[0], both[1]
    slab = (cells + w - 1) // w
    dev = scores.device
    win_over = torch.zeros(1, dtype=torch.int32, device=dev)       # cells whose bracket exceeds `slots`
    over_cells = torch.full((OVERFLOW_CAP,), -1, dtype=torch.int32, device=dev)
    if kern.get("peer_exchange") and scores.is_cuda and os.environ.get("CPPRE_PEER_EXCHANGE", "1") != "0":
        # fused: K3w stores block r straight into rank r's receive buffer over NVLink; the
        # all-reduce of the counts below is also the barrier that orders those stores
        recv, ptrs = _peer_recv_buffer(w, slots, slab, dev, group)
        _, cnt = kern["window"](flat, lo, hi, slots, slab, win_over, over_cells, ptrs, me)
        dist.all_reduce(cnt, group=group)
    else:
        win, cnt = kern["window"](flat, lo, hi, slots, slab, win_over, over_cells)    # [w, slots, slab], [w*slab]
        dist.all_reduce(cnt, group=group)
        recv = torch.empty_like(win)
        dist.all_to_all_single(recv, win, group=group)       # recv[r] = rank r's candidates of MY slab
    q_slab = kern["top"](recv.reshape(w * slots, slab), cnt[me * slab:(me + 1) * slab], kt, win_over, over_cells,
                         me * slab, owner_network(w, slots))
    parts = [torch.empty_like(q_slab) for _ in range(w)]
    dist.all_gather(parts, q_slab.contiguous(), group=group)
    q = torch.cat(parts)[:cells]
    # overflow bookkeeping, identical on every rank: cells listed by K3w and by the owner's select
    counts = [torch.empty_like(win_over) for _ in range(w)]
    dist.all_gather(counts, win_over, group=group)
    counts = torch.stack(counts).cpu()
    if int(counts.sum()) == 0:
        return q
    if int(counts.max()) > OVERFLOW_CAP:
        return None                                          # too many: the caller transposes everything
    return _redo_cells(flat, q, over_cells, rank_k, group, kern["select"])


OVERFLOW_CAP = 8192


def owner_network(world: int, slots: int) -> int:
    """Keys the owner of a cell slab sorts per cell: a power of two above the number of real
    candidates the union of `world` brackets holds.  Measured on |N(0,1)| scores, alpha = 0.1,
    n = 256 per rank, 3.3 M cells (tools/k3x_diag.py): world 2 -> 12 per cell on average, 62 at
    most; world 8 -> 110 on average, 224 at the 99.9th percentile, 323 at most.  The fused owner
    kernel sorts N, N/2 or N/4 keys depending on the fullest cell of each 32-cell tile, so the
    capacity is sized for the maximum (64 per rank) and costs only where it is needed; a cell
    above it is redone exactly.  Never more than the candidate rows."""
    need = max(32, 64 * world)
    net = 32
    while net < need and net < world * slots:
        net *= 2
    return min(net, 1024)


def marginal_quantiles(scores: torch.Tensor, ranks: Sequence[int], group=None,
                       select_fn: Optional[Callable] = None, kernels: Optional[dict] = None,
                       slots: int = 60, mode: str = "auto", defer: bool = False) -> torch.Tensor:
    """scores: rank-local [n_local, cells...] (same n_local on every rank).  Returns
    [len(ranks), cells...] order statistics over the global n_local*world samples, on every rank.

    mode "auto": candidate-window exchange per rank (`slots` floats per cell and rank over
    NVLink instead of n_local); cells that overflow the window are redone exactly from their
    full columns, and past OVERFLOW_CAP such cells the whole tensor is transposed instead
    (decided collectively, so every process takes the same path); "window" raises instead of
    transposing; "exchange" always transposes the whole tensor.

    defer=True (fused CUDA path only) skips the host round trip that checks the overflow
    counters, so the next kernel can be enqueued while this one runs; the caller MUST then call
    `pending_overflow()` before trusting the results and recompute with defer=False when it
    returns non-zero."""
    w = _world(group)
    if kernels is None and select_fn is None:
        kernels = _cuda_kernels()
    if select_fn is None:
        select_fn = kernels["select"]
    if w == 1:
        return select_fn(scores, ranks)
    cell_shape = tuple(scores.shape[1:])
    if mode == "exchange" or kernels is None:
        q = _exchange_quantiles(scores, ranks, group, select_fn)
        return q.reshape((len(ranks),) + cell_shape)
    out = []
    for k in ranks:
        fused = ("owner" in kernels and scores.is_cuda and os.environ.get("CPPRE_FUSED_EXCHANGE", "1") != "0")
        if fused:
            q = _fused_window_quantile(scores, k, group, kernels, slots, defer)
        else:
            q = _window_quantile(scores, k, group, kernels, slots)
        if q is None:
            if mode == "window":
                raise RuntimeError(f"candidate window overflowed in more than {OVERFLOW_CAP} cells (slots={slots})")
            q = _exchange_quantiles(scores, [k], group, select_fn)[0]
        out.append(q)
    return torch.stack(out).reshape((len(ranks),) + cell_shape)


# ---------------------------------------------------------------------------------------
# Nx-slab partition: one field too large for one GPU
# ---------------------------------------------------------------------------------------
def exchange_x_halo(padded: torch.Tensor, axis: int = -2, group=None) -> torch.Tensor:
    """``padded`` is this rank's slab of a field split along ``axis`` (the x axis of
    [..., Nt, Nx, Ny] data), allocated WITH one ghost row on each side: [..., nx_loc + 2, Ny].
    Fills the ghost rows in place from the neighbouring ranks (rank-1 owns the rows below,
    rank+1 the rows above); the two global edges keep whatever they hold -- zeros reproduce the
    reference's zero "same" padding (F.conv3d padding=k//2, Utils/ConvOps_2d.py:135-150).
    One send + one receive per neighbour (NCCL point-to-point over NVLink, gloo on CPU); the
    star stencils of the hot path reach one row, so one ghost row is all that ever travels."""
    w = _world(group)
    if w == 1:
        return padded
    me = dist.get_rank(group)
    n = padded.shape[axis]
    lo_ghost, lo_edge = padded.select(axis, 0), padded.select(axis, 1)
    hi_edge, hi_ghost = padded.select(axis, n - 2), padded.select(axis, n - 1)
    ops_, bufs = [], []
    to_global = (lambda r: dist.get_global_rank(group, r)) if group is not None else (lambda r: r)
    if me > 0:
        send, recv = lo_edge.contiguous(), torch.empty_like(lo_ghost, memory_format=torch.contiguous_format)
        ops_ += [dist.P2POp(dist.isend, send, to_global(me - 1), group), dist.P2POp(dist.irecv, recv, to_global(me - 1), group)]
        bufs.append((lo_ghost, recv))
    if me < w - 1:
        send, recv = hi_edge.contiguous(), torch.empty_like(hi_ghost, memory_format=torch.contiguous_format)
        ops_ += [dist.P2POp(dist.isend, send, to_global(me + 1), group), dist.P2POp(dist.irecv, recv, to_global(me + 1), group)]
        bufs.append((hi_ghost, recv))
    for req in dist.batch_isend_irecv(ops_):
        req.wait()
    for ghost, recv in bufs:
        ghost.copy_(recv)
    return padded


def slab_residual(program, padded_vars: torch.Tensor, axis: int = -2, group=None, residual_fn: Optional[Callable] = None):
    """PRE of an x-slab: exchange the ghost rows, evaluate the residual program on the padded
    slab (one K1 launch, boundary rows included) and drop the two ghost rows of the result.
    Every kept cell sees exactly the neighbourhood it has in the undivided field, so the
    concatenation of the ranks' results equals the single-GPU residual bit for bit.  Per-cell
    (marginal) calibration of a slab needs no exchange at all: a rank owns all samples of its
    cells.  ``residual_fn(padded_vars)`` replaces the CUDA program in the CPU tests."""
    exchange_x_halo(padded_vars, axis, group)
    full = residual_fn(padded_vars) if residual_fn is not None else program.residual(padded_vars, boundary=True)
    n = full.shape[axis]
    return full.narrow(axis if axis >= 0 else full.dim() + axis, 1, n - 2)


def allreduce_counts(*values: int, device=None, group=None):
    """Sums a few integer counters (coverage numerators / denominators) over ranks."""
    t = torch.tensor([int(v) for v in values], dtype=torch.int64, device=device)
    if _world(group) > 1:
        dist.all_reduce(t, group=group)
    return [int(v) for v in t.tolist()]
