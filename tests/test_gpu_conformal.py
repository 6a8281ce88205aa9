"""GPU parity of the conformal half: K2 cell statistics, K3 order statistics, K5 coverage.
Bit-exact where the contract says so (quantile values given identical scores, coverage counts,
accept masks); fp32 tolerance for std / joint scores."""
import numpy as np
import pytest
import torch

import cp_pre_b200
from conftest import assert_close_rel, load_golden
from cp_pre_b200 import dist as D
from cp_pre_b200 import ops, programs
from oracle import cpu_oracle as O
from oracle import synth

cp_pre_b200.install()
pytestmark = pytest.mark.gpu

from Neural_PDE.UQ.inductive_cp import (calibrate, calibrate_many, emp_cov, emp_cov_joint,  # noqa: E402
                                        filter_sims_joint, filter_sims_within_bounds, modulation_func,
                                        ncf_metric_joint)

ALPHAS = np.arange(0.05, 0.95 + 0.1, 0.1)


def cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_helpers_against_reference_inline_twins_golden():
    g = load_golden("cp")
    cal, val = g["cal"], g["val"]
    mod = modulation_func(cal, np.zeros(cal.shape))
    assert isinstance(mod, np.ndarray) and mod.shape == cal.shape[1:]
    assert_close_rel(mod, g["modulation"], 1e-6, "modulation")
    assert_close_rel(modulation_func(np.zeros(cal.shape, np.float32) + 0, cal), g["modulation"], 1e-6)  # symmetric
    sc = ncf_metric_joint(cal, np.zeros(cal.shape), g["modulation"])
    assert_close_rel(sc, g["joint_scores"], 1e-6, "joint scores")
    for q in (1.5, 2.5, 3.5):
        sets = [-q * g["modulation"], q * g["modulation"]]
        np.testing.assert_array_equal(filter_sims_joint(sets, val), g[f"joint_mask|{q}"])
        assert emp_cov_joint(sets, val) == float(g[f"joint_cov|{q}"])
    for within in (False, True):
        np.testing.assert_array_equal(
            filter_sims_within_bounds(-g["modulation"], g["modulation"], val, 0.5, within),
            g[f"filter_bounds|{int(within)}"])


@pytest.mark.parametrize("n,cells", [(20, (7,)), (100, (5, 33)), (256, (3, 10, 20)), (500, (40,)), (1000, (9,)),
                                     (1, (4,)), (1500, (70,)), (2, (1,)), (31, (65,)), (32, (32,)), (33, (31,)),
                                     (64, (3, 11)), (65, (100,)), (128, (97,)), (129, (33,)), (257, (34,)),
                                     (512, (48,)), (513, (17,)), (1024, (33,)), (1025, (19,)), (2048, (21,)),
                                     (2049, (5,)), (3000, (3,))])
def test_calibrate_is_bit_exact_per_cell(n, cells):
    rng = np.random.default_rng(n)
    s = np.abs(rng.standard_normal((n,) + cells)).astype(np.float32)
    s[rng.integers(0, n, 5), ..., 0] = 0.25            # ties
    for alpha in ALPHAS:
        try:
            want = O.calibrate(s, n, alpha)
        except ValueError:
            with pytest.raises(ValueError, match="Quantiles must be in the range"):
                calibrate(s, n, alpha)
            continue
        got = calibrate(s, n, alpha)
        assert got.dtype == np.float32 and got.shape == want.shape
        np.testing.assert_array_equal(got, want, err_msg=f"n={n} alpha={alpha}")
    ok = [a for a in ALPHAS if 0 <= O.calibrate_quantile_level(n, a) <= 1]
    if ok:
        many = calibrate_many(cuda(s), n, ok)
        for i, a in enumerate(ok):
            np.testing.assert_array_equal(many[i].cpu().numpy(), O.calibrate(s, n, a))


@pytest.mark.parametrize("n", [7, 100, 256, 300, 2048])
def test_select_ranks_every_rank_signed_values_nan_and_strided_rows(n):
    """Every order statistic at once (one sort serves all ranks), signed values, +-inf, NaN
    columns, and a row stride larger than the cell count (a column slice of a wider tensor)."""
    rng = np.random.default_rng(100 + n)
    cells = 77
    wide = (rng.standard_normal((n, cells + 13)) * 5).astype(np.float32)
    wide[rng.integers(0, n, 3), 5] = np.inf
    wide[rng.integers(0, n, 3), 6] = -np.inf
    wide[rng.integers(0, n), 9] = np.nan
    wide[:, 11] = 1.5                                   # all ties
    wide[:, 12] = -0.0
    t = cuda(wide)[:, 3:3 + cells]
    ref = np.sort(wide[:, 3:3 + cells], axis=0)
    ranks = list(range(n)) if n <= 300 else [0, 1, 31, 32, 33, 1023, 1024, 1025, n - 2, n - 1]
    for r0 in range(0, len(ranks), 40):                 # > kMaxRanks per call exercises the chunking
        rr = ranks[r0:r0 + 40]
        got = ops.select_ranks(t, rr).cpu().numpy()
        want = ref[rr]
        want[:, np.isnan(ref).any(axis=0)] = np.nan
        np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("n", [65, 100, 256, 300, 1000, 3000])
def test_select_ranks_near_one_end_streaming_topk_windows(n):
    """Ranks within 8 / 16 / 32 / 64 of either end take the streaming top-K kernel (K3t): every window
    size, both ends, n not a multiple of the window, +-inf, ties, NaNs of either sign, strided rows."""
    rng = np.random.default_rng(7 * n)
    cells = 150
    wide = (rng.standard_normal((n, cells + 9)) * 3).astype(np.float32)
    wide[rng.integers(0, n, 2), 4] = np.inf
    wide[rng.integers(0, n, 2), 5] = -np.inf
    wide[rng.integers(0, n), 6] = np.nan
    wide[rng.integers(0, n), 7] = np.float32(np.nan) * np.float32(-1)          # sign bit set
    wide.view(np.uint32)[rng.integers(0, n), 8] = 0xffc00001                      # negative quiet NaN payload
    wide[:, 10] = 2.5
    wide[: n // 2, 11] = -1.0
    t = cuda(wide)[:, 2:2 + cells]
    ref = np.sort(wide[:, 2:2 + cells], axis=0)
    nan_cols = np.isnan(ref).any(axis=0)
    assert nan_cols.sum() == 3
    for w in (1, 8, 9, 16, 17, 32, 33, 52, 64):
        if w > n:
            continue
        for rr in (list(range(n - w, n)), list(range(w)), [n - w], [w - 1], [n - 1, n - w]):
            got = ops.select_ranks(t, rr).cpu().numpy()
            want = ref[rr]
            want[:, nan_cols] = np.nan
            np.testing.assert_array_equal(got, want, err_msg=f"n={n} window={w} ranks={rr[:3]}..")


def test_calibrate_scalar_joint_form_negative_values_and_nan():
    rng = np.random.default_rng(0)
    for n in (19, 512, 4096, 5000):
        s = rng.standard_normal(n).astype(np.float32) * 3         # negatives exercise the key order
        for alpha in (0.05, 0.5, 0.9):
            got = calibrate(s, n, alpha)
            assert np.ndim(got) == 0 and got == O.calibrate(s, n, alpha)
        t = calibrate(cuda(s), n, 0.1)
        assert t.is_cuda and t.dim() == 0 and float(t) == float(O.calibrate(s, n, 0.1))
    s = np.abs(rng.standard_normal((64, 6))).astype(np.float32)
    s[10, 2] = np.nan
    s[11, 4] = np.inf
    got, want = calibrate(s, 64, 0.1), O.calibrate(s, 64, 0.1)
    assert np.isnan(got[2]) and np.isnan(want[2])
    np.testing.assert_array_equal(np.delete(got, 2), np.delete(want, 2))
    # n argument different from the number of rows (scripts pass len(other_array))
    assert calibrate(s[:, 0], 80, 0.3) == O.calibrate(s[:, 0], 80, 0.3)


def test_modulation_and_joint_scores_against_oracle():
    rng = np.random.default_rng(4)
    a = (0.1 * rng.standard_normal((300, 12, 34))).astype(np.float32) + np.float32(5.0)   # mean >> std
    b = (0.1 * rng.standard_normal((300, 12, 34))).astype(np.float32)
    mod = modulation_func(a, b)
    assert_close_rel(mod, O.modulation_func(a, b), 2e-6, "std with large mean")
    sc = ncf_metric_joint(a, b, mod)
    assert_close_rel(sc, O.ncf_metric_joint(a, b, mod), 1e-6, "joint score")
    ta, tb = cuda(a), cuda(b)
    m2 = modulation_func(ta, tb)
    assert m2.is_cuda and torch.equal(m2.cpu(), torch.from_numpy(mod))
    # strided (interior-crop) views are accepted
    v = ta[:, 1:-1, 1:-1]
    assert_close_rel(modulation_func(v, tb[:, 1:-1, 1:-1]).cpu().numpy(), O.modulation_func(a, b)[1:-1, 1:-1], 2e-6)


def test_coverage_counts_are_exact():
    rng = np.random.default_rng(5)
    y = rng.standard_normal((200, 9, 31)).astype(np.float32)
    pred = (y + 0.3 * rng.standard_normal(y.shape)).astype(np.float32)
    q_cell = np.abs(rng.standard_normal(y.shape[1:])).astype(np.float32)
    for sets in ([pred - q_cell, pred + q_cell], [-q_cell, q_cell], [np.float32(-0.7), np.float32(0.7)],
                 [pred - np.float32(0.4), pred + np.float32(0.4)]):
        assert emp_cov(sets, y) == O.emp_cov(sets, y)
        assert emp_cov_joint(sets, y) == O.emp_cov_joint(sets, y)
        np.testing.assert_array_equal(filter_sims_joint(sets, y), O.filter_sims_joint(sets, y))
    big = [-10 * np.ones_like(q_cell), 10 * np.ones_like(q_cell)]
    assert emp_cov_joint(big, y) == 1.0 and emp_cov(big, y) == 1.0
    # values exactly on the bounds are inside (>=, <=)
    assert emp_cov([y, y], y) == 1.0
    # fused band form: [c - q*h, c + q*h] never materialised, same fp32 roundings as NumPy
    h = np.abs(rng.standard_normal(y.shape[1:])).astype(np.float32)
    q = np.float32(1.37)
    cnt = ops.inside_count_band(cuda(y), cuda(pred), cuda(h), float(q)).cpu().numpy()
    want = ((y >= pred - q * h) & (y <= pred + q * h)).sum(axis=(1, 2))
    np.testing.assert_array_equal(cnt, want)
    cnt = ops.inside_count_band(cuda(y), None, cuda(h), float(q)).cpu().numpy()
    np.testing.assert_array_equal(cnt, ((y >= -q * h) & (y <= q * h)).sum(axis=(1, 2)))
    for thr in (0.3, 0.8):
        for within in (False, True):
            np.testing.assert_array_equal(filter_sims_within_bounds(-h, h, y, thr, within),
                                          O.filter_sims_within_bounds(-h, h, y, thr, within))


def test_radix_histogram_passes_reproduce_the_joint_quantile():
    """The multi-GPU Joint path: 4 x 8-bit histogram passes (all-reduced between passes when the
    scores are sharded) find the same order statistic as a sort."""
    rng = np.random.default_rng(6)
    x = (rng.standard_normal(10000) * 4).astype(np.float32)
    keys = O.float_order_key(x)
    xc = cuda(x)
    for k in (0, 17, 5000, 9999):
        prefix = mask = 0
        kk = k
        for shift in (24, 16, 8, 0):
            hist = torch.zeros(256, dtype=torch.int32, device="cuda")
            ops.radix_hist(xc, prefix, mask, shift, hist)
            h = hist.cpu().numpy().astype(np.int64)
            sel = (keys & np.uint32(mask)) == np.uint32(prefix)
            np.testing.assert_array_equal(h, np.bincount((keys[sel] >> np.uint32(shift)) & np.uint32(255), minlength=256))
            c = np.cumsum(h)
            d = int(np.searchsorted(c, kk, side="right"))
            kk -= int(c[d - 1]) if d else 0
            prefix |= d << shift
            mask |= 255 << shift
        assert prefix == int(np.sort(keys)[k])


def test_joint_and_marginal_pipelines_end_to_end():
    """residual -> modulation -> scores -> qhat -> coverage, reference formulas on the oracle side
    (Joint/NS_Residuals_CP.py:289-337, Marginal/NS_Residuals_CP.py:289-337)."""
    import Utils.ConvOps_2d as C2
    n = synth.NS
    ko = O.ops_2d()
    D = dict(D_t=C2.ConvOperator("t", 1), D_x=C2.ConvOperator("x", 1), D_y=C2.ConvOperator("y", 1),
             D_xx_yy=C2.ConvOperator(("x", "y"), 2))
    prog = programs.ns_momentum(D["D_t"], D["D_x"], D["D_y"], D["D_xx_yy"], n["dx"], n["dy"], n["dt"], n["nu"])
    shape = (48, 8, 16, 128)
    pred = synth.fields(41, 3, shape)
    out = (pred + 0.02 * synth.normal_fields(42, 3, shape)).astype(np.float32)
    R = lambda v: O.residual_ns_momentum(v[:, 0], v[:, 1], v[:, 2], n["dx"], n["dy"], n["dt"], n["nu"], ko)
    inner = (slice(None), slice(1, -1), slice(1, -1), slice(1, -1))
    r_pred, r_out = R(pred)[inner], R(out)[inner]
    mod_o = O.modulation_func(r_out, r_pred)
    sc_o = O.ncf_metric_joint(r_out, r_pred, mod_o)
    mod, inv, sc = programs.joint_calibration(prog, cuda(pred), cuda(out))
    assert_close_rel(mod[1:-1, 1:-1, 1:-1].cpu().numpy(), mod_o, 2e-5, "pipeline modulation")
    assert_close_rel(sc.cpu().numpy(), sc_o, 5e-5, "pipeline joint scores")
    # identical score arrays -> identical qhat (bit exact), then coverage with the fused band kernel
    for alpha in (0.1, 0.5):
        qhat = calibrate(sc_o, len(sc_o), alpha)
        assert qhat == O.calibrate(sc_o, len(sc_o), alpha)
        sets = [r_pred - qhat * mod_o, r_pred + qhat * mod_o]
        assert emp_cov_joint(sets, r_out) == O.emp_cov_joint(sets, r_out)
    # marginal, data-free form: |res(pred)| per-cell quantiles for the whole alpha grid in one pass
    scores, qh = programs.marginal_calibration(prog, cuda(pred), None, alphas=(0.1, 0.3, 0.5))
    s_o = np.abs(r_pred)
    assert_close_rel(scores.cpu().numpy(), s_o, 1e-5, "marginal scores")
    mine = scores.cpu().numpy()
    for i, a in enumerate((0.1, 0.3, 0.5)):
        np.testing.assert_array_equal(qh[i].cpu().numpy(), O.calibrate(mine, len(mine), a))


def test_window_kernels_match_the_oracle():
    """K3w pieces against oracle/cpu_oracle.py: ordered keys, window compaction (counts, NaN flag,
    slot contents in row order, overflow count, blocked layout with a ragged last slab) and the
    per-cell select-from-top (NaN flags, out-of-window ranks)."""
    rng = np.random.default_rng(5)
    x = (rng.standard_normal(1000) * 4).astype(np.float32)
    x[:6] = [np.inf, -np.inf, np.nan, 0.0, -0.0, -np.float32(np.nan)]
    np.testing.assert_array_equal(ops.ordered_keys(cuda(x)).cpu().numpy(), O.ordered_keys(x))
    fin = x[np.isfinite(x)]
    assert (np.diff(fin[np.argsort(O.ordered_keys(fin), kind="stable")]) >= 0).all()   # key order = float order

    n, cells, slots, slab = 77, 301, 12, 128
    s = (rng.standard_normal((n, cells)) * 2).astype(np.float32)
    s[:, 1] = 0.5
    srt = np.sort(s, axis=0)
    lo = O.ordered_keys(srt[n - 14])
    hi = O.ordered_keys(srt[n - 6])
    lo[2], hi[2] = O.ordered_keys(srt[5, 2:3])[0], O.ordered_keys(srt[40, 2:3])[0]      # 36 inside: overflow
    hi[4] = O.ordered_keys(np.array([np.nan], dtype=np.float32))[0]                       # a rank saw a NaN
    want_win, want_cnt, want_ov = O.window_compact(s, lo, hi, slots, slab)
    ov = torch.zeros(1, dtype=torch.int32, device="cuda")
    ov_cells = torch.full((8,), -1, dtype=torch.int32, device="cuda")
    win, cnt = ops.window_compact(cuda(s), cuda(lo), cuda(hi), slots, slab, ov, ov_cells)
    np.testing.assert_array_equal(cnt.cpu().numpy(), want_cnt)
    np.testing.assert_array_equal(win.cpu().numpy(), want_win)
    assert int(ov.item()) == len(want_ov) == 2            # cell 2 and the all-ties cell 1
    assert sorted(ov_cells.cpu().numpy()[:2].tolist()) == want_ov == [1, 2]
    assert (cnt.cpu().numpy()[cells:] == 1 << 24).all()   # padding tail of the ragged last slab

    rows, c2 = 96, 210
    cand = (rng.standard_normal((rows, c2)) * 3).astype(np.float32)
    cand[rng.integers(0, rows, 40), rng.integers(0, c2, 40)] = -np.inf
    packed = rng.integers(0, 20, c2).astype(np.int32)
    packed[7] |= 1 << 24
    packed[9] = 0                                         # want = kt: fine
    packed[11] = 25                                       # want < 1: overflow
    kt = 21
    want_out, want_ov = O.select_from_top(cand, packed, kt)
    ov.zero_()
    ov_cells.fill_(-1)
    got = ops.select_from_top(cuda(cand), cuda(packed), kt, ov, ov_cells, 1000).cpu().numpy()
    np.testing.assert_array_equal(got, want_out)
    assert int(ov.item()) == len(want_ov) >= 1
    assert sorted(ov_cells.cpu().numpy()[:len(want_ov)].tolist()) == [1000 + c for c in want_ov]
    # packing form: blocks that are mostly -inf padding, a network smaller than the row count,
    # a column with more real candidates than the network, a rank that lands in the padding
    for rows, network in ((80, 32), (160, 64), (320, 128), (320, 256), (40, 32)):
        cand = np.full((rows, c2), -np.inf, dtype=np.float32)
        for c in range(c2):
            m = int(rng.integers(0, min(network, rows) + 1))
            cand[rng.choice(rows, m, replace=False), c] = (rng.standard_normal(m) * 2).astype(np.float32)
        cand[:, 3] = np.arange(rows, dtype=np.float32)                        # rows real candidates
        packed = rng.integers(0, 10, c2).astype(np.int32)
        packed[5] |= 1 << 24
        kt = 14
        want_out, want_ov = O.select_from_top(cand, packed, kt, network)
        ov.zero_()
        got = ops.select_from_top(cuda(cand), cuda(packed), kt, ov, ov_cells, 0, network).cpu().numpy()
        np.testing.assert_array_equal(got, want_out, err_msg=f"rows={rows} network={network}")
        assert int(ov.item()) == len(want_ov) == (1 if rows > network else 0)


@pytest.mark.parametrize("world,n_loc,alpha", [(2, 256, 0.1), (8, 64, 0.1), (4, 100, 0.05), (3, 50, 0.5)])
def test_window_protocol_emulated_ranks_equals_global_quantile(world, n_loc, alpha):
    """The multi-GPU marginal quantile with its CUDA kernels, the ranks emulated one after the
    other on this GPU (collectives done by hand): bit-equal to np.quantile over the union."""
    rng = np.random.default_rng(world * 1000 + n_loc)
    cells = 1000
    full = np.abs(rng.standard_normal((world * n_loc, cells))).astype(np.float32)
    full[rng.integers(0, world * n_loc, 4), 3] = 0.75
    N = world * n_loc
    k = O.calibrate_rank(N, N, alpha)
    kt, slots = N - k, 32
    J = -(-kt // world)
    shards = [cuda(full[r * n_loc:(r + 1) * n_loc]) for r in range(world)]
    keys = torch.stack([ops.ordered_keys(ops.select_ranks(s, [n_loc - J])[0]) for s in shards])
    lo, hi = keys.min(0).values, keys.max(0).values
    slab = -(-cells // world)
    ov = torch.zeros(1, dtype=torch.int32, device="cuda")
    cells_ov = torch.full((64,), -1, dtype=torch.int32, device="cuda")
    wins, cnts = zip(*[ops.window_compact(s, lo, hi, slots, slab, ov, cells_ov) for s in shards])
    cnt = torch.stack(cnts).sum(0).to(torch.int32)
    q = []
    for me in range(world):                                # all-to-all: block `me` of every rank
        cand = torch.cat([w[me] for w in wins])            # [world*slots, slab]
        q.append(ops.select_from_top(cand, cnt[me * slab:(me + 1) * slab], kt, ov, cells_ov, me * slab,
                                     D.owner_network(world, slots)))
    got = torch.cat(q)[:cells].cpu().numpy()
    if int(ov.item()) == 0:
        np.testing.assert_array_equal(got, O.calibrate(full, N, alpha))
    else:                                                  # a bracket wider than 32 slots is reported, never silent
        assert alpha == 0.5


@pytest.mark.parametrize("world,n_loc,alpha,cells", [(2, 256, 0.1, 1000), (8, 64, 0.1, 1003), (4, 100, 0.05, 517),
                                                     (3, 50, 0.5, 1000), (8, 256, 0.1, 4096)])
@pytest.mark.parametrize("key_mode", [0, 1])
def test_fused_exchange_protocol_emulated_ranks_equals_global_quantile(world, n_loc, alpha, cells, key_mode):
    """K3x (keys broadcast -> window scan with in-band headers -> owner select with the qhat
    broadcast), the ranks emulated one after the other on this GPU: plain device buffers stand in
    for the peers' symmetric memory.  Bit-equal to np.quantile over the union; NaN cells are NaN;
    cells that cannot be answered from the window are listed, never silently wrong."""
    rng = np.random.default_rng(world * 1000 + n_loc)
    full = np.abs(rng.standard_normal((world * n_loc, cells))).astype(np.float32)
    full[rng.integers(0, world * n_loc, 4), 3] = 0.75          # ties
    full[rng.integers(0, world * n_loc), 11] = np.nan          # one NaN -> that cell is NaN
    full[:, 17] = 1.25                                         # a constant column: every sample sits in the bracket
    N = world * n_loc
    k = O.calibrate_rank(N, N, alpha)
    kt, slots = N - k, 32
    J = -(-kt // world)
    slab = D.fused_slab(cells, world)
    cells_pad = (slab * world + 3) // 4 * 4
    shards = [cuda(full[r * n_loc:(r + 1) * n_loc]) for r in range(world)]
    keys = [torch.zeros((world, cells_pad), dtype=torch.int32, device="cuda") for _ in range(world)]
    recv = [torch.full((world, slab // 32, 1 + slots, 32), float("nan"), device="cuda") for _ in range(world)]   # stale garbage
    qbuf = [torch.full((cells_pad,), -7.0, device="cuda") for _ in range(world)]
    ov = torch.zeros(1, dtype=torch.int32, device="cuda")
    cells_ov = torch.full((256,), -1, dtype=torch.int32, device="cuda")
    stats = [ops.select_ranks(s, [n_loc - J])[0] for s in shards]
    if key_mode == 0:                                          # keys stored into every peer
        for me in range(world):
            ops.peer_bcast_keys(stats[me], cells_pad, [t.data_ptr() for t in keys], me)
    else:                                                      # [~min, max] of signed keys, as a MAX all-reduce leaves it
        ak = torch.stack([ops.ordered_keys(a) for a in stats])
        both = torch.zeros((2, cells_pad), dtype=torch.int32, device="cuda")
        both[0, :cells], both[1, :cells] = (~ak).max(0).values, ak.max(0).values
        keys = [both] * world
    for me, s in enumerate(shards):
        ops.window_exchange(s, keys[me], slots, slab, [t.data_ptr() for t in recv], me, ov, cells_ov, key_mode)
    for me in range(world):                                    # owner writes its slab locally, then pushes it
        live = max(0, min(slab, cells - me * slab))
        ops.owner_select(recv[me], live, kt, D.owner_network(world, slots), me * slab, [qbuf[me].data_ptr()],
                         ov, cells_ov)
        ops.peer_copy(qbuf[me][me * slab:me * slab + live], me * slab,
                      [t.data_ptr() for r, t in enumerate(qbuf) if r != me])
    want = O.calibrate(full, N, alpha)
    listed = set(int(c) for c in cells_ov.cpu().numpy() if c >= 0)
    assert int(ov.item()) <= 256
    assert 17 in listed or n_loc <= slots                      # the constant column overflows the window
    for r in range(world):                                     # every rank received the same, complete qhat
        got = qbuf[r][:cells].cpu().numpy()
        keep = np.array([c not in listed for c in range(cells)])
        np.testing.assert_array_equal(got[keep], want[keep])
        assert np.isnan(got[11]) and np.isnan(want[11])
    if alpha != 0.5:                                           # 32 slots: the odd wide bracket at world = 8 is listed
        assert len(listed - {17}) <= cells // 50


# ----------------------------------------------------------------------------------------------
# round 2: device-resident Joint flow (K3d radix select, fused modulation, COVER with a device q)
# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [1, 19, 256, 1000, 4096, 50000])
def test_device_radix_select_equals_sorted_order_statistic(n):
    """K3d: four radix passes with the state on the device return sorted(x)[k] bit for bit, for
    negative values, ties, +-inf, -0.0; NaN anywhere -> NaN; k out of range -> NaN."""
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(n) * 5).astype(np.float32)
    if n >= 19:
        x[:6] = [np.inf, -np.inf, 0.0, -0.0, 1.5, 1.5]
    want = np.sort(x)
    xc = cuda(x)
    for k in sorted({0, n // 3, n // 2, int(0.9 * (n - 1)), n - 1}):
        got = ops.radix_select(xc, k)                                     # one launch (single-CTA form)
        assert got.shape == (1,) and got.is_cuda
        assert float(got) == float(want[k])         # == : +0.0 and -0.0 tie in np.sort's order too
        multi = ops.radix_select(xc, k, allreduce=lambda h: None)         # four passes, state on the device
        assert torch.equal(multi, got) or (float(multi) == 0.0 and float(got) == 0.0)
    assert np.isnan(float(ops.radix_select(xc, n)))                       # rank out of range
    if n >= 19:
        y = x.copy()
        y[7] = np.nan
        assert np.isnan(float(ops.radix_select(cuda(y), n // 2)))
        y[7] = -np.float32(np.nan)                                        # NaN with the sign bit set
        assert np.isnan(float(ops.radix_select(cuda(y), n // 2)))


@pytest.mark.parametrize("world", [2, 8])
def test_device_radix_select_over_emulated_ranks(world):
    """The multi-GPU Joint quantile with the CUDA kernels: every rank's shard adds its 257 counters to
    the same histogram (what the NCCL all-reduce does between passes); the pick runs on the summed
    counters.  Equal to the order statistic of the concatenated scores; NaN on one rank -> NaN."""
    rng = np.random.default_rng(world)
    shards = [np.abs(rng.standard_normal(128 + 0 * r) * (1 + r)).astype(np.float32) for r in range(world)]
    allx = np.sort(np.concatenate(shards))
    cs = [cuda(s) for s in shards]
    for k in (0, len(allx) // 2, int(np.ceil((len(allx) + 1) * 0.9)) - 1, len(allx) - 1):
        got = ops.radix_select(cs, k)
        assert float(got) == float(allx[k])
    shards[world - 1][5] = np.nan
    assert np.isnan(float(ops.radix_select([cuda(s) for s in shards], 3)))


def test_fused_modulation_equals_two_kernel_path_and_chan_merge_of_emulated_ranks():
    """cppre_cell_modulation (one launch) == cell_moments + cell_std_from_m2 bit for bit, and the Chan
    merge of per-rank fp64 moments (the multi-GPU modulation, dist.merge_moments) reproduces the std
    over the whole batch: emulated ranks = batch shards on one GPU, CUDA kernels throughout."""
    rng = np.random.default_rng(11)
    a = (0.3 * rng.standard_normal((96, 7, 40))).astype(np.float32) + np.float32(2.0)
    ta = cuda(a)
    sd, inv = ops.cell_modulation(ta, None, 1e-6)
    sd2, inv2 = ops.cell_std(ta, None, 1e-6, want_inv=True)
    assert torch.equal(sd, sd2) and torch.equal(inv, inv2)
    assert_close_rel(sd.cpu().numpy(), O.modulation_func(a, np.zeros_like(a)) + np.float32(1e-6), 2e-6, "fused modulation")
    for world in (2, 3, 8):
        cuts = np.linspace(0, 96, world + 1).astype(int)
        parts = [ops.cell_moments(ta[lo:hi]) for lo, hi in zip(cuts[:-1], cuts[1:])]
        ns = [float(hi - lo) for lo, hi in zip(cuts[:-1], cuts[1:])]
        total = sum(ns)
        gmean = sum(m * n for (m, _), n in zip(parts, ns)) / total
        g2 = sum(m2 + n * (m - gmean) ** 2 for (m, m2), n in zip(parts, ns))
        merged = ops.cell_std_from_m2(g2, total, 1e-6).reshape(sd.shape)
        assert float(((merged - sd).abs() / sd).max()) <= 2e-7, world


def test_coverage_with_a_device_resident_q_equals_the_host_scalar_form():
    """COVER reads q from a device scalar (cppre_epilogue_t::d_band_q): same counts as q by value,
    on both kernel paths; and the whole Joint flow stays on the device."""
    import Utils.ConvOps_2d as C2
    from cp_pre_b200 import _lib as L
    n = synth.NS
    prog = programs.ns_momentum(C2.ConvOperator("t", 1), C2.ConvOperator("x", 1), C2.ConvOperator("y", 1),
                                C2.ConvOperator(("x", "y"), 2), n["dx"], n["dy"], n["dt"], n["nu"])
    v = cuda(synth.fields(3, 3, (40, 6, 16, 128)))
    mod, inv, scores = programs.joint_calibration(prog, v[:20])
    qd = programs.joint_quantile(scores, 20, 0.1)
    assert qd.is_cuda and qd.shape == (1,)
    assert float(qd) == float(O.calibrate(scores.cpu().numpy(), 20, 0.1))
    for path in (L.PATH_TMA, L.PATH_GENERIC):
        prog.path = path
        a = prog.coverage_counts(v[20:], qd, mod)
        b = prog.coverage_counts(v[20:], float(qd), mod)
        assert torch.equal(a, b) and int(a.sum()) > 0
    prog.path = L.PATH_AUTO
    with pytest.raises(ValueError):
        prog.coverage_counts(v[20:], torch.ones(2, device="cuda"), mod)


@pytest.mark.parametrize("equation", ["continuity", "momentum", "energy", "induction", "gauss"])
def test_mhd_joint_cp_flow_against_the_oracle(equation):
    """BASELINE config 5's flow (data-free Joint CP on an MHD residual, Joint/MHD_Residuals_CP.py:225-278,
    405-420) on a small slab against the oracle's formulas: residual 1e-5, modulation 2e-5, scores 5e-5;
    given the oracle's score array qhat is bit-equal, and the fused coverage counts (band qhat*modulation
    read from the device) equal NumPy's on the stored residual."""
    import Utils.ConvOps_2d as C2
    c = synth.MHD
    prog = programs.mhd(equation, C2.ConvOperator("t", 1), C2.ConvOperator("x", 1), C2.ConvOperator("y", 1), c["gamma"])
    shape = (28, 8, 12, 128)
    v = synth.fields(77, 6, shape, positive=(0, 3))
    n_cal = 16
    res_o = O.residual_mhd(equation, *[v[:, j] for j in range(6)], gamma=c["gamma"], ops=O.ops_2d())
    inner = (slice(None), slice(1, -1), slice(1, -1), slice(1, -1))
    cal_o, val_o = res_o[:n_cal][inner], res_o[n_cal:][inner]
    mod_o = O.modulation_func(cal_o, np.zeros_like(cal_o)) + np.float32(1e-6)
    sc_o = O.ncf_metric_joint(cal_o, np.zeros_like(cal_o), mod_o)
    vc = cuda(v)
    mod, inv, sc = programs.joint_calibration(prog, vc[:n_cal], eps=1e-6)
    from cp_pre_b200 import _lib as L
    assert L.last_path() != "", equation
    assert_close_rel(mod[1:-1, 1:-1, 1:-1].cpu().numpy(), mod_o, 2e-5, equation + " modulation")
    assert_close_rel(sc.cpu().numpy(), sc_o, 5e-5, equation + " joint scores")
    # identical scores -> identical qhat, bit for bit, through the device-side radix select
    qd = programs.joint_quantile(cuda(sc_o), n_cal, 0.1)
    assert float(qd) == float(O.calibrate(sc_o, n_cal, 0.1))
    # coverage: fused COVER with the library's own residual of the validation samples as ground truth
    mine = prog.residual(vc[n_cal:], boundary=True).cpu().numpy()
    q, m = np.float32(float(qd)), mod.cpu().numpy()
    want = ((mine >= -(q * m)) & (mine <= q * m))[inner].sum(axis=(1, 2, 3))
    got = prog.coverage_counts(vc[n_cal:], qd, mod).cpu().numpy()
    np.testing.assert_array_equal(got, want)
    cells = prog.interior_cells(vc.shape)
    assert O.emp_cov_joint([-(q * m)[inner[1:]], (q * m)[inner[1:]]], mine[inner]) == float((got == cells).mean())


@pytest.mark.parametrize("n,alpha", [(4096, 0.1), (3000, 0.1), (2500, 0.05), (8192, 0.02), (2048, 0.1), (1536, 0.1), (1100, 0.08)])
def test_segmented_select_for_long_columns_equals_numpy(n, alpha, monkeypatch):
    """n > 1024 (where the register sort is slow, and beyond 2048 where only the radix select is left): the top-end order statistic through the candidate-window kernels
    with row segments in the role of ranks (ops._segmented_select) -- bit-equal to np.quantile(method='higher')
    and to the radix kernel it replaces, incl. NaN cells, ties, a constant column and +-inf."""
    rng = np.random.default_rng(n)
    cells = 1500
    s = np.abs(rng.standard_normal((n, cells))).astype(np.float32)
    s[rng.integers(0, n, 50), 3] = 0.75                  # ties
    s[:, 4] = 1.25                                       # constant column
    s[7, 5] = np.nan                                     # NaN cell -> NaN
    s[rng.integers(0, n, 5), 6] = np.inf
    s[rng.integers(0, n, 5), 8] = -np.inf
    k = O.calibrate_rank(n, n, alpha)
    want = O.calibrate(s, n, alpha)
    t = cuda(s)
    assert ops._segmented_select(t, k) is not None       # the request fits the segmented path
    got = ops.select_ranks(t, [k])[0].cpu().numpy()
    np.testing.assert_array_equal(got, want)
    monkeypatch.setenv("CPPRE_SEGMENTED_SELECT", "0")
    np.testing.assert_array_equal(ops.select_ranks(t, [k])[0].cpu().numpy(), want)
    monkeypatch.delenv("CPPRE_SEGMENTED_SELECT")
    # a mid-column rank does not fit (the bracket would hold hundreds of samples): radix kernel, same answer
    assert ops._segmented_select(t, n // 2) is None
    np.testing.assert_array_equal(ops.select_ranks(t, [n // 2])[0].cpu().numpy(), np.sort(s, axis=0)[n // 2] if not np.isnan(s).any() else
                                  np.where(np.isnan(s).any(axis=0), np.nan, np.sort(s, axis=0)[n // 2]))


def test_calibrate_long_one_dimensional_scores_and_narrow_arrays():
    """The Joint form on long score vectors (configs[1]: 2 048 calibration scores; here 4 096) and per-cell arrays
    with only a few cells: too narrow for the segmented path, served by the library's own dispatch; bit-equal to
    NumPy either way."""
    rng = np.random.default_rng(12)
    for n, cells in ((4096, 1), (2048, 1), (3000, 7), (1500, 31), (1500, 33)):
        s = np.abs(rng.standard_normal((n, cells))).astype(np.float32)
        arr = s[:, 0] if cells == 1 else s
        for alpha in (0.1, 0.5):
            got = calibrate(arr, n, alpha)
            np.testing.assert_array_equal(np.asarray(got), O.calibrate(arr, n, alpha), err_msg=f"{n} {cells} {alpha}")
