"""world_size-2 gloo tests (CPU) of the multi-GPU orchestration in cp_pre_b200/dist.py.  The CUDA
kernels are replaced by the oracle's NumPy implementations through the injection points, so only
the host logic (sharding, Chan merge, radix passes + all-reduce, all-to-all transposition) runs."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import cpu_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _np_hist(x, prefix, mask, shift, hist):
    keys = O.float_order_key(x.numpy().reshape(-1))
    sel = (keys & np.uint32(mask)) == np.uint32(prefix)
    h = np.bincount((keys[sel] >> np.uint32(shift)) & np.uint32(255), minlength=256)
    hist += torch.from_numpy(h.astype(np.int32))
    return hist


def _np_select(scores, ranks):
    s = scores.numpy()
    return torch.from_numpy(np.stack([O.select_rank(s, r) for r in ranks]))


def _np_keys(x):
    return torch.from_numpy(O.ordered_keys(x.numpy()))


def _np_window(scores, lo, hi, slots, slab, overflow, overflow_cells):
    win, cnt, ov = O.window_compact(scores.numpy(), lo.numpy(), hi.numpy(), slots, slab)
    for c in ov:
        if int(overflow) < overflow_cells.numel():
            overflow_cells[int(overflow)] = c
        overflow += 1
    return torch.from_numpy(win), torch.from_numpy(cnt)


def _np_top(cand, cnt, kt, overflow, overflow_cells, cell_offset, network):
    out, ov = O.select_from_top(cand.numpy(), cnt.numpy(), kt, network)
    for c in ov:
        if int(overflow) < overflow_cells.numel():
            overflow_cells[int(overflow)] = cell_offset + c
        overflow += 1
    return torch.from_numpy(out)


def _np_radix_select(scores, k, allreduce):
    """NumPy stand-in for ops.radix_select: the same four passes over (prefix, mask, k) state, the
    257-counter histogram summed over ranks by the injected all-reduce."""
    x = scores.numpy().reshape(-1)
    keys = O.float_order_key(x)
    prefix = mask = 0
    flags = 0
    for p in range(4):
        shift = 24 - 8 * p
        sel = (keys & np.uint32(mask)) == np.uint32(prefix)
        h = np.bincount((keys[sel] >> np.uint32(shift)) & np.uint32(255), minlength=256)
        hist = torch.from_numpy(np.concatenate([h, [np.isnan(x).sum() if p == 0 else 0]]).astype(np.int32))
        if allreduce is not None:
            allreduce(hist)
        c = np.cumsum(hist.numpy()[:256].astype(np.int64))
        if p == 0 and int(hist[256]):
            flags |= 1
        if k >= c[-1]:
            flags |= 2
            break
        d = int(np.searchsorted(c, k, side="right"))
        k -= int(c[d - 1]) if d else 0
        prefix |= d << shift
        mask |= 255 << shift
    from cp_pre_b200.dist import _key_to_float
    val = np.nan if flags else _key_to_float(prefix)
    return torch.tensor([val], dtype=torch.float32)


NP_KERNELS = {"select": _np_select, "keys": _np_keys, "window": _np_window, "top": _np_top}


def _np_moments(a, b):
    x = a.numpy().astype(np.float64) - (0 if b is None else b.numpy().astype(np.float64))
    x = x.reshape(x.shape[0], -1)
    return torch.from_numpy(x.mean(0)), torch.from_numpy(((x - x.mean(0)) ** 2).sum(0))


def _np_std(m2, count, eps, want_inv):
    sd = (torch.sqrt(m2 / count) + eps).float()
    return (sd, 1.0 / sd) if want_inv else sd


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from cp_pre_b200 import dist as D
    r, w, dev = D.init_from_env("gloo")
    assert (r, w, dev.type) == (rank, world, "cpu")
    rng = np.random.default_rng(123)                       # same stream on both ranks
    full = (rng.standard_normal((40, 6, 7)) * 2 + 1).astype(np.float32)
    lo, hi = D.shard_range(40, rank, world)
    mine = torch.from_numpy(full[lo:hi])
    res = {}
    # global modulation = std over the union of both shards
    sd, inv = D.global_cell_std(mine, None, 1e-6, moments_fn=_np_moments, std_fn=_np_std)
    res["std_err"] = float(np.max(np.abs(sd.numpy() - (np.std(full, axis=0) + 1e-6))))
    # joint quantile, both exchange forms
    scores = np.abs(full).reshape(40, -1).max(axis=1)
    my_scores = torch.from_numpy(scores[lo:hi])
    want = np.sort(scores)
    for k in (0, 13, 39):
        assert D.joint_quantile_hist(my_scores, k, hist_fn=_np_hist) == float(want[k])
        got = D.joint_quantile_gather(my_scores, [k], select_fn=_np_select)
        assert float(got[0]) == float(want[k])
    nan_scores = my_scores.clone()
    if rank == 1:
        nan_scores[0] = float("nan")
    assert np.isnan(D.joint_quantile_hist(nan_scores, 5, hist_fn=_np_hist))
    with pytest.raises(ValueError):
        D.joint_quantile_hist(my_scores, 40, hist_fn=_np_hist)
    # device-state form (K3d orchestration): same answers, result stays a one-element tensor
    for k in (0, 13, 39):
        got = D.joint_quantile_device(my_scores, k, select_fn=_np_radix_select)
        assert got.shape == (1,) and float(got[0]) == float(want[k])
    assert np.isnan(float(D.joint_quantile_device(nan_scores, 5, select_fn=_np_radix_select)[0]))
    with pytest.raises(ValueError):
        D.joint_quantile_device(my_scores, 40, select_fn=_np_radix_select)
    # uneven shards: the equal-count exchanges refuse instead of hanging / answering wrongly
    assert D.global_count(lo + 1 if world > 1 else 1) == sum(D.shard_range(40, r, world)[0] + 1 for r in range(world))
    # two calibration sets of different size in one process: shards (3, 2, ..) then (3, 3, ..) then back.  Rank 0
    # sees n_local = 3 every time; every rank must still take part in every count exchange and agree on the answer
    if world > 1:
        first = 3 if rank == 0 else 2
        assert D.global_count(first) == 3 + 2 * (world - 1)
        with pytest.raises(ValueError, match="uneven batch shards"):
            D.global_count(first, require_equal=True)
        assert D.global_count(3, require_equal=True) == 3 * world
        assert D.global_count(first) == 3 + 2 * (world - 1)
    uneven = torch.from_numpy(np.abs(full[: 3 + rank]))
    with pytest.raises(ValueError, match="uneven batch shards"):
        D.marginal_quantiles(uneven, [1], select_fn=_np_select)
    with pytest.raises(ValueError, match="uneven batch shards"):
        D.joint_quantile_gather(torch.from_numpy(scores[: 3 + rank]), [1], select_fn=_np_select)
    # marginal per-cell quantiles through the all-to-all (42 cells: exercises the padding to a
    # multiple of the world size when world == 4; divisible when world == 2)
    q = D.marginal_quantiles(torch.from_numpy(np.abs(full[lo:hi])), [5, 20, 36], select_fn=_np_select)
    ref = np.sort(np.abs(full), axis=0)[[5, 20, 36]]
    res["marginal_equal"] = bool(np.array_equal(q.numpy(), ref))
    odd = np.abs(full[:, :, :5])                           # 30 cells -> not divisible by 4 either way
    q = D.marginal_quantiles(torch.from_numpy(odd[lo:hi]).contiguous(), [7], select_fn=_np_select)
    res["marginal_odd_equal"] = bool(np.array_equal(q.numpy(), np.sort(odd, axis=0)[[7]]))
    # candidate-window exchange (K3w orchestration): every rank of interest, ties, +-inf, a NaN
    # column, cells not divisible by the world size; then a window too small for the bracket
    big = (rng.standard_normal((40 * world, 9, 5)) * 3).astype(np.float32)
    big[3, 0, 0] = np.inf
    big[5, 0, 1] = -np.inf
    big[:, 0, 2] = 1.25
    big[7, 0, 3] = np.nan
    n_loc = 40
    mine_big = torch.from_numpy(big[rank * n_loc:(rank + 1) * n_loc])
    ref = np.sort(big, axis=0)
    ref[:, np.isnan(ref).any(axis=0)] = np.nan
    N = 40 * world
    ks = [N - 1, N - 5, N - 17, int(0.9 * N), N // 2, 3, 0]
    q = D.marginal_quantiles(mine_big, ks, kernels=NP_KERNELS, mode="window", slots=40)
    res["window_equal"] = bool(np.array_equal(q.numpy(), ref[ks], equal_nan=True))
    # 6 slots: most mid-range brackets overflow and are redone from their full columns
    q = D.marginal_quantiles(mine_big, ks, kernels=NP_KERNELS, mode="window", slots=6)
    res["auto_equal"] = bool(np.array_equal(q.numpy(), ref[ks], equal_nan=True))
    cap, D.OVERFLOW_CAP = D.OVERFLOW_CAP, 3             # a redo list that is too short
    try:
        D.marginal_quantiles(mine_big, [N // 2], kernels=NP_KERNELS, mode="window", slots=1)
        res["overflow_raises"] = False
    except RuntimeError:
        res["overflow_raises"] = True
    q = D.marginal_quantiles(mine_big, [N // 2], kernels=NP_KERNELS, mode="auto", slots=1)
    res["fallback_equal"] = bool(np.array_equal(q.numpy(), ref[[N // 2]], equal_nan=True))
    D.OVERFLOW_CAP = cap
    # Nx-slab partition of one field: ghost rows exchanged, the oracle's wave residual
    # on the padded slab, ghost rows dropped -> the concatenation equals the residual of the whole field
    field = rng.standard_normal((3, 6, 8 * world + 3, 12)).astype(np.float32)       # x extent not divisible
    xlo, xhi = D.shard_range(field.shape[2], rank, world)
    padded = torch.zeros((3, 6, xhi - xlo + 2, 12))
    padded[:, :, 1:-1] = torch.from_numpy(field[:, :, xlo:xhi])
    kw = O.build_kernel("2d", "t", 2) - np.float32(0.25) * O.build_kernel("2d", ("x", "y"), 2)   # wave operator
    wave = lambda t: torch.from_numpy(O.correlate(t.numpy(), kw))
    mine_res = D.slab_residual(None, padded, residual_fn=wave)
    res["slab_equal"] = bool(np.array_equal(mine_res.numpy(), O.correlate(field, kw)[:, :, xlo:xhi]))
    res["counts"] = D.allreduce_counts(rank + 1, 10)
    out[rank] = res
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_calibration_orchestration(world):
    port = _free_port()
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
        assert len(out) == world
        for rank in range(world):
            r = out[rank]
            assert r["std_err"] < 1e-6
            assert r["marginal_equal"] and r["marginal_odd_equal"]
            assert r["window_equal"] and r["auto_equal"] and r["overflow_raises"] and r["fallback_equal"]
            assert r["slab_equal"]
            assert r["counts"] == [world * (world + 1) // 2, 10 * world]


def test_shard_range_partitions_exactly():
    from cp_pre_b200.dist import shard_range
    for total in (0, 1, 7, 512, 2048):
        for world in (1, 2, 3, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
