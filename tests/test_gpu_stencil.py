"""GPU parity of K1 (the fused residual/stencil kernels) against the CPU oracle and the golden
vectors produced by the unmodified reference.  Every call goes through the C ABI
(cp_pre_b200.ops -> ctypes -> libcppre.so).  Tolerance:
max|new-ref| <= 1e-5 * max|ref| for residual fields; exact for nothing here (fp32 sums)."""
import numpy as np
import pytest
import torch

import cp_pre_b200
from conftest import assert_close_rel, load_golden
from cp_pre_b200 import _lib as L
from cp_pre_b200 import ops, programs
from oracle import cpu_oracle as O
from oracle import synth

cp_pre_b200.install()
pytestmark = pytest.mark.gpu

TOL = 1e-5


def cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _mods():
    import ConvOps_0d
    import ConvOps_1d
    import ConvOps_Spatial
    import Utils.ConvOps_2d as ConvOps_2d
    return {"2d": ConvOps_2d, "1d": ConvOps_1d, "0d": ConvOps_0d, "sp": ConvOps_Spatial}


@pytest.mark.parametrize("prefix,names", [
    ("2d", ["Dt", "Dx", "Dy", "Dtt", "Dxx", "I", "L3", "L5", "L7", "wave", "dense", "aniso"]),
    ("1d", ["Dt", "Dx", "Dxx", "I", "L3", "L5", "adv"]),
    ("0d", ["0_2", "1_2", "1_4", "2_2", "2_4", "2_6"]),
    ("sp", ["Dx", "L3", "L5"]),
])
def test_operator_call_matches_reference_golden(prefix, names):
    g = load_golden("conv")
    m = _mods()[prefix]
    u = g[f"{prefix}|u"]
    for nm in names:
        D = m.ConvOperator(1) if prefix == "0d" else m.ConvOperator("x", 1, device="cpu")
        D.kernel = torch.from_numpy(g[f"{prefix}|{nm}|k"])
        want = g[f"{prefix}|{nm}|y"]
        # host tensors in, host tensor out (the reference's default workflow) ...
        got = D(torch.from_numpy(u))
        assert got.device.type == "cpu" and got.dtype == torch.float32
        assert_close_rel(got.numpy(), want, TOL, f"{prefix}|{nm} (host)")
        # ... and device tensors in, device tensor out
        got = D(cuda(u))
        assert got.is_cuda
        assert_close_rel(got.cpu().numpy(), want, TOL, f"{prefix}|{nm} (device)")
        # convolution(field, kernel) overwrites .kernel permanently, like the reference
        E = m.ConvOperator(1) if prefix == "0d" else m.ConvOperator("x", 1, device="cpu")
        E.convolution(cuda(u), D.kernel)
        assert E.kernel is D.kernel


def test_permuted_view_is_processed_in_place_and_matches_reference():
    g = load_golden("conv")
    m = _mods()["2d"]
    W = m.ConvOperator()
    W.kernel = torch.from_numpy(g["2d|wave|k"])
    up = cuda(g["2d|perm|u"]).permute(0, 3, 1, 2)          # [BS,Nt,Nx,Ny] view, t innermost
    got = W(up)
    assert got.shape == up.shape
    assert_close_rel(got.cpu().numpy(), g["2d|perm|y"], TOL, "perm")


def _programs_2d():
    m = _mods()["2d"]
    C = m.ConvOperator
    return dict(D_t=C("t", 1), D_x=C("x", 1), D_y=C("y", 1), D_xx_yy=C(("x", "y"), 2))


def test_residual_programs_match_reference_golden():
    g = load_golden("residuals")
    m1 = _mods()["1d"]
    c = synth.BURGERS
    pb = programs.burgers(m1.ConvOperator("t", 1), m1.ConvOperator("x", 1), m1.ConvOperator("x", 2),
                          torch.tensor(c["dx"]), torch.tensor(c["dt"]), torch.tensor(c["nu"]))
    assert_close_rel(pb.residual(cuda(g["burgers|u"]), boundary=True).cpu().numpy(), g["burgers|res"], TOL, "burgers")
    o = _programs_2d()
    n = synth.NS
    v = cuda(g["ns|vars"])
    pc = programs.ns_continuity(o["D_x"], o["D_y"], np.asarray(n["dx"]), np.asarray(n["dy"]))
    assert_close_rel(pc.residual(v, True).cpu().numpy(), g["ns|cont"], TOL, "ns cont")
    pm = programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], n["dx"], n["dy"], n["dt"], n["nu"])
    assert_close_rel(pm.residual(v, True).cpu().numpy(), g["ns|mom"], TOL, "ns mom")
    # interior crop is a view of the full field, like the reference's [...,1:-1,1:-1,1:-1]
    assert_close_rel(pm.residual(v).cpu().numpy(), g["ns|mom"][..., 1:-1, 1:-1, 1:-1], TOL)
    w = cuda(g["mhd|vars"])
    for eq in ("continuity", "momentum", "energy", "induction", "gauss"):
        p = programs.mhd(eq, o["D_t"], o["D_x"], o["D_y"])
        assert_close_rel(p.residual(w, True).cpu().numpy(), g[f"mhd|{eq}"], TOL, eq)
    # class forms (Other_UQ/Evaluation/PRE_estimations.py)
    from Other_UQ.Evaluation.PRE_estimations import PRE_MHD, PRE_NS, PRE_Wave
    pw = PRE_Wave(dt=0.5, dx=1.0, c=1.0)
    np.testing.assert_array_equal(pw.D.kernel.numpy(), g["wave|k"])
    assert_close_rel(pw.residual(cuda(g["wave|u"]), boundary=True).cpu().numpy(), g["wave|res_full"], TOL)
    assert_close_rel(pw.residual(torch.from_numpy(g["wave|u"])).numpy(), g["wave|res"], TOL)
    assert_close_rel(PRE_NS(n["dt"], n["dx"], n["dy"]).residual(v, boundary=True).cpu().numpy(), g["ns|mom_class"], TOL)
    assert_close_rel(PRE_MHD(0.01, 1 / 128, 1 / 128).residual(w, boundary=True).cpu().numpy(),
                     g["mhd|energy_class"], TOL)


SHAPES_3D = [(3, 7, 19, 132), (2, 5, 16, 256), (1, 3, 9, 4), (2, 4, 40, 128)]


@pytest.mark.parametrize("shape", SHAPES_3D)
@pytest.mark.parametrize("path", [L.PATH_GENERIC, L.PATH_TMA])
def test_all_programs_against_oracle_on_both_kernel_paths(shape, path):
    """Seeded smooth + N(0,1) fields; the TMA star kernel and the generic kernel must both match
    the oracle, including tiles that straddle the column (128) and row (8/16) tile edges."""
    o = _programs_2d()
    ko = O.ops_2d()
    n = synth.NS
    for gen in ("smooth", "normal"):
        v = (synth.fields(5, 6, shape, positive=(0, 3)) if gen == "smooth" else
             np.abs(synth.normal_fields(6, 6, shape)) + np.float32(0.5))
        vc = cuda(v)
        f = [v[:, j] for j in range(6)]
        progs = {
            "wave": (programs.linear(_wave_op()), lambda: O.correlate(f[0], _wave_op().kernel.numpy()), vc[:, 0]),
            "ns_cont": (programs.ns_continuity(o["D_x"], o["D_y"], n["dx"], n["dy"]),
                        lambda: O.residual_ns_continuity(f[0], f[1], n["dx"], n["dy"], ko), vc),
            "ns_mom": (programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], n["dx"], n["dy"], n["dt"], n["nu"]),
                       lambda: O.residual_ns_momentum(f[0], f[1], f[2], n["dx"], n["dy"], n["dt"], n["nu"], ko), vc),
        }
        for eq in ("continuity", "momentum", "energy", "induction", "gauss"):
            progs["mhd_" + eq] = (programs.mhd(eq, o["D_t"], o["D_x"], o["D_y"]),
                                  (lambda eq=eq: O.residual_mhd(eq, *f, ops=ko)), vc)
        for name, (prog, oracle, inp) in progs.items():
            prog.path = path
            got = prog.residual(inp, boundary=True)
            assert L.last_path() == ("tma" if path == L.PATH_TMA else "generic"), name
            assert_close_rel(got.cpu().numpy(), oracle(), TOL, f"{name} {gen} {shape} path={path}")


def _wave_op():
    m = _mods()["2d"]
    W = m.ConvOperator()
    W.kernel = m.ConvOperator("t", 2).kernel - 0.25 * m.ConvOperator(("x", "y"), 2).kernel
    return W


@pytest.mark.parametrize("shape", [(5, 9, 260), (17, 6, 128), (3, 4, 8)])
@pytest.mark.parametrize("path", [L.PATH_GENERIC, L.PATH_TMA])
def test_burgers_and_advection_1d(shape, path):
    m1 = _mods()["1d"]
    c = synth.BURGERS
    u = synth.fields(9, 1, shape)[:, 0]
    pb = programs.burgers(m1.ConvOperator("t", 1), m1.ConvOperator("x", 1), m1.ConvOperator("x", 2),
                          c["dx"], c["dt"], c["nu"])
    pb.path = path
    assert_close_rel(pb.residual(cuda(u), True).cpu().numpy(), O.residual_burgers(u, c["dx"], c["dt"], c["nu"]),
                     TOL, f"burgers {shape}")
    A = m1.ConvOperator()
    A.kernel = m1.ConvOperator("t", 1).kernel + 0.5 * m1.ConvOperator("x", 1).kernel
    pa = programs.linear(A)
    pa.path = path
    assert_close_rel(pa.residual(cuda(u), True).cpu().numpy(), O.correlate(u, A.kernel.numpy()), TOL)
    assert L.last_path() == ("tma" if path == L.PATH_TMA else "generic")


def test_ragged_and_empty_inputs_take_the_generic_kernel():
    m = _mods()
    W = _wave_op()
    rng = np.random.default_rng(3)
    for shape, path in [((2, 5, 7, 197), "flat"), ((1, 1, 1, 1), "generic"), ((3, 2, 1, 5), "flat"), ((2, 80, 3, 2), "flat")]:
        u = rng.standard_normal(shape, dtype=np.float32)
        got = W(cuda(u))
        assert L.last_path() == path        # ragged rows: the flat-line star kernel; degenerate rows: the generic one
        assert_close_rel(got.cpu().numpy(), O.correlate(u, W.kernel.numpy()), TOL, str(shape))
    W5 = m["2d"].ConvOperator(("x", "y"), 2, taylor_order=4)            # 5x5x5 taps: generic kernel on any shape
    u = rng.standard_normal((2, 5, 7, 197), dtype=np.float32)
    got = W5(cuda(u))
    assert L.last_path() == "generic"
    assert_close_rel(got.cpu().numpy(), O.correlate(u, W5.kernel.numpy()), TOL)
    e = W(torch.zeros(0, 4, 4, 4, device="cuda"))
    assert e.shape == (0, 4, 4, 4)
    A = m["1d"].ConvOperator()                              # config 1: Nx = 197 (not a multiple of 4)
    A.kernel = m["1d"].ConvOperator("t", 1).kernel + 0.5 * m["1d"].ConvOperator("x", 1).kernel
    u = rng.standard_normal((100, 10, 197), dtype=np.float32)
    assert_close_rel(A(torch.from_numpy(u)).numpy(), O.correlate(u, A.kernel.numpy()), TOL)
    with pytest.raises(NotImplementedError):                # forcing TMA on an unqualified request fails loudly
        ops.run_program(L.PROG_LINEAR, [cuda(u).unsqueeze(2)], [A.kernel.numpy().reshape(3, 1, 3)], [(1, 0, 1)],
                        path=L.PATH_TMA)
    with pytest.raises(NotImplementedError):                # kernels beyond 7 taps per axis
        ops.run_program(L.PROG_LINEAR, [cuda(u).unsqueeze(2)], [np.ones((9, 1, 3), np.float32)], [(4, 0, 1)])


@pytest.mark.parametrize("path", [L.PATH_GENERIC, L.PATH_TMA])
@pytest.mark.parametrize("shape", [(6, 9, 20, 132), (4, 5, 16, 256)])
def test_fused_joint_score_epilogue(path, shape):
    """scores[b] = max over interior cells of |R(pred) - ref| * weight -- one pass, nothing
    but [BS] floats written -- against the oracle's residual + np.max."""
    o, ko, n = _programs_2d(), O.ops_2d(), synth.NS
    v = synth.fields(21, 3, shape)
    ref = 0.3 * O.residual_ns_momentum(*(synth.fields(22, 3, shape)[:, j] for j in range(3)),
                                       n["dx"], n["dy"], n["dt"], n["nu"], ko)
    rng = np.random.default_rng(1)
    w = rng.uniform(0.5, 2.0, shape[1:]).astype(np.float32)
    res = O.residual_ns_momentum(v[:, 0], v[:, 1], v[:, 2], n["dx"], n["dy"], n["dt"], n["nu"], ko)
    prog = programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], n["dx"], n["dy"], n["dt"], n["nu"])
    prog.path = path
    inner = (slice(None), slice(1, -1), slice(1, -1), slice(1, -1))
    for use_ref in (False, True):
        for use_w in (False, True):
            d = np.abs(res - ref) if use_ref else np.abs(res)
            if use_w:
                d = d * w
            got = prog.joint_scores(cuda(v), cuda(w) if use_w else None, cuda(ref) if use_ref else None)
            assert got.shape == (shape[0],)
            assert_close_rel(got.cpu().numpy(), d[inner].max(axis=(1, 2, 3)), 2e-5, f"max ref={use_ref} w={use_w}")
            got = prog.joint_scores(cuda(v), cuda(w) if use_w else None, cuda(ref) if use_ref else None,
                                    interior=False)
            assert_close_rel(got.cpu().numpy(), d.max(axis=(1, 2, 3)), 2e-5, "max, shell included")
            got = prog.joint_scores(cuda(v), cuda(w) if use_w else None, cuda(ref) if use_ref else None,
                                    reduce="sum")
            assert_close_rel(got.cpu().numpy(), d[inner].astype(np.float64).sum(axis=(1, 2, 3)), 1e-4, "sum")
    # marginal score tensor |R - ref|
    got = prog.abs_scores(cuda(v), ref=cuda(ref))
    assert_close_rel(got.cpu().numpy(), np.abs(res - ref)[inner], TOL, "abs_scores")
    # NaN anywhere in a sample's interior propagates to that sample's score only (np.max semantics)
    vn = v.copy()
    vn[1, 0, 3, 5, 7] = np.nan
    s = prog.joint_scores(cuda(vn)).cpu().numpy()
    assert np.isnan(s[1]) and not np.isnan(np.delete(s, 1)).any()


@pytest.mark.parametrize("path", [L.PATH_GENERIC, L.PATH_TMA])
def test_fused_coverage_epilogue_counts_are_exact(path):
    """COVER epilogue: per-sample count of cells with c - q*h <= R(vars) <= c + q*h, the sets and
    the residual never written.  Given the kernel's own residual (STORE epilogue, same arithmetic)
    the counts equal NumPy's evaluation of the reference expressions bit for bit
    [ref: Joint/NS_Residuals_CP.py:312,336; Marginal/NS_Residuals_CP.py:311-312]."""
    o, n = _programs_2d(), synth.NS
    shape = (10, 7, 24, 132)
    v = synth.fields(51, 3, shape)
    prog = programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], n["dx"], n["dy"], n["dt"], n["nu"])
    prog.path = path
    res = prog.residual(cuda(v), boundary=True).cpu().numpy()
    assert_close_rel(res, O.residual_ns_momentum(v[:, 0], v[:, 1], v[:, 2], n["dx"], n["dy"], n["dt"], n["nu"],
                                                 O.ops_2d()), TOL, "residual")
    rng = np.random.default_rng(3)
    inner = (slice(None), slice(1, -1), slice(1, -1), slice(1, -1))
    scale = np.float32(np.abs(res[inner]).std())
    half = (scale * rng.uniform(0.2, 2.0, shape[1:])).astype(np.float32)
    centre = (res + scale * rng.standard_normal(res.shape)).astype(np.float32)
    q = np.float32(0.83)
    for use_c in (False, True):
        for use_h in (False, True):
            c = centre if use_c else np.zeros_like(res)
            hh = (q * half).astype(np.float32) if use_h else np.full(half.shape, q, np.float32)
            lo, hi = (c - hh).astype(np.float32), (c + hh).astype(np.float32)
            ins = (res >= lo) & (res <= hi)
            got = prog.coverage_counts(cuda(v), float(q), cuda(half) if use_h else None, cuda(centre) if use_c else None)
            assert got.dtype == torch.int32 and got.shape == (shape[0],)
            np.testing.assert_array_equal(got.cpu().numpy(), ins[inner].sum(axis=(1, 2, 3)), f"c={use_c} h={use_h}")
            got = prog.coverage_counts(cuda(v), float(q), cuda(half) if use_h else None, cuda(centre) if use_c else None,
                                       interior=False)
            np.testing.assert_array_equal(got.cpu().numpy(), ins.sum(axis=(1, 2, 3)))
    assert prog.interior_cells(v.shape) == 5 * 22 * 130
    # marginal data-free form: emp_cov([-qhat, qhat], res) with a per-cell qhat, q = 1
    qhat = half
    got = prog.coverage_counts(cuda(v), 1.0, cuda(qhat)).cpu().numpy()
    np.testing.assert_array_equal(got, ((res >= -qhat) & (res <= qhat))[inner].sum(axis=(1, 2, 3)))
    # 1-D + time (batch-rows tiles)
    m1 = _mods()["1d"]
    cb = synth.BURGERS
    u = synth.fields(52, 1, (37, 12, 260))[:, 0]
    pb = programs.burgers(m1.ConvOperator("t", 1), m1.ConvOperator("x", 1), m1.ConvOperator("x", 2),
                          cb["dx"], cb["dt"], cb["nu"])
    pb.path = path
    rb = pb.residual(cuda(u), boundary=True).cpu().numpy()
    hb = (np.abs(rb).mean() * rng.uniform(0.5, 2.0, rb.shape[1:])).astype(np.float32)
    got = pb.coverage_counts(cuda(u), 1.0, cuda(hb)).cpu().numpy()
    np.testing.assert_array_equal(got, ((rb >= -hb) & (rb <= hb))[:, 1:-1, 1:-1].sum(axis=(1, 2)))


def test_fused_score_1d_batch_rows_mode():
    m1 = _mods()["1d"]
    c = synth.BURGERS
    shape = (37, 12, 260)
    u = synth.fields(31, 1, shape)[:, 0]
    res = O.residual_burgers(u, c["dx"], c["dt"], c["nu"])
    w = np.random.default_rng(2).uniform(0.5, 2.0, shape[1:]).astype(np.float32)
    for path in (L.PATH_GENERIC, L.PATH_TMA):
        pb = programs.burgers(m1.ConvOperator("t", 1), m1.ConvOperator("x", 1), m1.ConvOperator("x", 2),
                              c["dx"], c["dt"], c["nu"])
        pb.path = path
        got = pb.joint_scores(cuda(u), cuda(w))
        assert_close_rel(got.cpu().numpy(), (np.abs(res) * w)[:, 1:-1, 1:-1].max(axis=(1, 2)), 2e-5, f"path {path}")


def test_linearity_and_null_space_at_full_config_size():
    """Size-independent properties at BASELINE config 3 (wave, [1024,80,128,128]): the composite
    kernel annihilates constants away from the zero-padded shell, the operator is linear, and
    the TMA and generic kernels agree on a slab."""
    W = _wave_op()
    bs = 1024
    g = torch.Generator(device="cuda").manual_seed(0)
    u = torch.randn(bs, 80, 128, 128, device="cuda", generator=g)
    r = W(u)
    assert L.last_path() == "tma" and r.shape == u.shape
    c = W(torch.full((2, 80, 128, 128), 3.25, device="cuda"))
    assert float(c[:, 1:-1, 1:-1, 1:-1].abs().max()) < 1e-5
    v = torch.randn(8, 80, 128, 128, device="cuda", generator=g)
    lhs = W(u[:8] + 2.0 * v)
    rhs = r[:8] + 2.0 * W(v)
    assert float((lhs - rhs).abs().max()) <= 1e-5 * float(rhs.abs().max())
    pg = programs.linear(W)
    pg.path = L.PATH_GENERIC
    rg = pg.residual(u[:16], boundary=True)
    assert float((rg - r[:16]).abs().max()) <= 1e-6 * float(rg.abs().max())
    # fused joint score == max of the stored residual, bit for bit (same arithmetic, no weight)
    s = programs.linear(W).joint_scores(u[:64])
    assert torch.equal(s, r[:64, 1:-1, 1:-1, 1:-1].abs().amax(dim=(1, 2, 3)))


def test_autograd_through_operator_matches_torch_conv3d():
    """PRE as a training loss (Physics_Informed/Wave_FNO_PI.py): d(loss)/d(field) through D.
    Floating-point kernel -> checked against a plain PyTorch fp32 reference of the same op."""
    import torch.nn.functional as F
    W = _wave_op()
    u = torch.randn(2, 6, 12, 16, device="cuda", requires_grad=True)
    loss = (W(u) ** 2).sum()
    loss.backward()
    u2 = u.detach().clone().requires_grad_(True)
    k = W.kernel.cuda()
    ref = (F.conv3d(u2.unsqueeze(1), k[None, None], padding=1).squeeze(1) ** 2).sum()
    ref.backward()
    assert_close_rel(u.grad.cpu().numpy(), u2.grad.cpu().numpy(), 1e-5, "grad")


def test_burgers_joint_cp_at_full_config_size():
    """BASELINE config 2 at its full size ([4096,100,1024], 1-D Burgers + Joint CP) through
    size-independent properties: the fused SCORE and COVER epilogues reproduce, bit for bit, what
    the stored residual gives (same arithmetic, nothing but [BS] numbers written); the joint
    quantile is the order statistic np.quantile(method='higher') picks; and a slab of the residual
    agrees with the CPU oracle [ref: Joint/Burgers_Residuals_CP.py:182-187,212-260]."""
    from Neural_PDE.UQ.inductive_cp import calibrate
    m1, c = _mods()["1d"], synth.BURGERS
    pb = programs.burgers(m1.ConvOperator("t", 1), m1.ConvOperator("x", 1), m1.ConvOperator("x", 2),
                          c["dx"], c["dt"], c["nu"])
    bs, nt, nx = 4096, 100, 1024
    g = torch.Generator(device="cuda").manual_seed(2)
    x = torch.linspace(0, 1, nx, device="cuda")
    u = torch.sin(2 * np.pi * (x[None, None, :] + torch.rand(bs, 1, 1, device="cuda", generator=g))) \
        * torch.exp(-torch.linspace(0, 1, nt, device="cuda"))[None, :, None] \
        + 0.01 * torch.randn(bs, nt, nx, device="cuda", generator=g)
    res = pb.residual(u, boundary=True)
    assert L.last_path() == "tma" and res.shape == u.shape
    want = O.residual_burgers(u[:4].cpu().numpy(), c["dx"], c["dt"], c["nu"])
    assert_close_rel(res[:4, 1:-1, 1:-1].cpu().numpy(), want[:, 1:-1, 1:-1], TOL, "oracle slab")
    n_cal = bs // 2
    inner = res[:, 1:-1, 1:-1]
    mod = inner[:n_cal].std(dim=0, unbiased=False)                     # modulation over the calibration half
    inv = torch.zeros(nt, nx, device="cuda")
    inv[1:-1, 1:-1] = 1.0 / mod
    scores = pb.joint_scores(u, inv)                                   # one pass, [BS] floats written
    assert torch.equal(scores, (inner.abs() * inv[1:-1, 1:-1]).amax(dim=(1, 2)))
    qhat = calibrate(scores[:n_cal].cpu().numpy(), n_cal, 0.1)
    assert qhat == np.sort(scores[:n_cal].cpu().numpy())[O.calibrate_rank(n_cal, n_cal, 0.1)]
    half = torch.zeros(nt, nx, device="cuda")
    half[1:-1, 1:-1] = mod
    inside = pb.coverage_counts(u[n_cal:], float(qhat), half)          # -qhat*mod <= R <= qhat*mod, sets never built
    band = half[1:-1, 1:-1] * float(qhat)
    assert torch.equal(inside.to(torch.int64), (inner[n_cal:].abs() <= band).sum(dim=(1, 2)))
    joint_cov = float((inside == pb.interior_cells(u.shape)).float().mean())
    assert 0.85 <= joint_cov <= 0.95, joint_cov                        # exchangeable halves: ~1 - alpha


def test_mhd_programs_on_a_config5_per_gpu_slab():
    """BASELINE config 5 (2-D MHD, Nt=100, 256x256, 6 fields; 256 samples per GPU at 8 GPUs) on a
    32-sample slab of the per-GPU shape: every MHD program runs on the TMA kernel, its fused joint
    score equals the max of its stored residual bit for bit, and one sample agrees with the oracle."""
    o, c = _programs_2d(), synth.MHD
    bs, nt, n = 32, 100, 256
    g = torch.Generator(device="cuda").manual_seed(5)
    w = 0.1 * torch.randn(bs, 6, nt, n, n, device="cuda", generator=g)
    w[:, 0] += 1.0                                                      # rho, a divisor: strictly positive
    w[:, 3] += 1.0
    f1 = [w[:1, j].cpu().numpy() for j in range(6)]
    for eq in ("continuity", "momentum", "energy", "induction", "gauss"):
        p = programs.mhd(eq, o["D_t"], o["D_x"], o["D_y"], c["gamma"])
        r = p.residual(w, boundary=True)
        assert L.last_path() == "tma", eq
        s = p.joint_scores(w)
        assert torch.equal(s, r[:, 1:-1, 1:-1, 1:-1].abs().amax(dim=(1, 2, 3))), eq
        want = O.residual_mhd(eq, *f1, gamma=c["gamma"], ops=O.ops_2d())
        assert_close_rel(r[:1].cpu().numpy(), want, TOL, eq)


def test_x_slab_partition_equals_the_undivided_field():
    """Nx-slab partition (one field too large for one GPU), two slabs emulated on
    this GPU: ghost rows filled by hand the way dist.exchange_x_halo fills them, the NS momentum
    program on each padded slab, ghost rows dropped -> bit-identical to the undivided launch."""
    o, n = _programs_2d(), synth.NS
    prog = programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], n["dx"], n["dy"], n["dt"], n["nu"])
    v = cuda(synth.fields(61, 3, (4, 9, 40, 128)))                     # [BS, F, Nt, Nx, Ny]
    want = prog.residual(v, boundary=True)
    parts = []
    for lo, hi in ((0, 24), (24, 40)):
        padded = torch.zeros((4, 3, 9, hi - lo + 2, 128), device="cuda")
        padded[..., 1:-1, :] = v[..., lo:hi, :]
        if lo > 0:
            padded[..., 0, :] = v[..., lo - 1, :]
        if hi < 40:
            padded[..., -1, :] = v[..., hi, :]
        parts.append(prog.residual(padded, boundary=True)[..., 1:-1, :])
    assert torch.equal(torch.cat(parts, dim=-2), want)


def test_bf16_input_mode_equals_fp32_kernel_on_the_widened_input():
    """bf16-input mode (CPPRE_FLAG_IN_BF16; north_star: 'bf16-input mode stated separately'): the
    fields are staged as bf16 tiles and widened in the kernel, all arithmetic stays fp32.  Stated
    parity: (1) bit-identical to the fp32 kernel run on the widened input, for every epilogue;
    (2) within 1e-5 of the fp32 oracle evaluated on the bf16-rounded input."""
    o, n, c = _programs_2d(), synth.NS, synth.MHD
    shape = (6, 9, 20, 136)                                            # two column tiles, the second 8 wide
    rng = np.random.default_rng(4)
    w = rng.uniform(0.5, 2.0, shape[1:]).astype(np.float32)
    cases = {
        "ns_mom": (programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], n["dx"], n["dy"], n["dt"], n["nu"]),
                   cuda(synth.fields(71, 3, shape)),
                   lambda f: O.residual_ns_momentum(f[:, 0], f[:, 1], f[:, 2], n["dx"], n["dy"], n["dt"], n["nu"], O.ops_2d())),
        "wave": (programs.linear(_wave_op()), cuda(synth.fields(72, 1, shape))[:, 0],
                 lambda f: O.correlate(f, _wave_op().kernel.numpy())),
        "mhd_mom": (programs.mhd("momentum", o["D_t"], o["D_x"], o["D_y"], c["gamma"]),
                    cuda(synth.fields(73, 6, shape, positive=(0, 3))),
                    lambda f: O.residual_mhd("momentum", *(f[:, j] for j in range(6)), gamma=c["gamma"], ops=O.ops_2d())),
    }
    for name, (prog, v, oracle) in cases.items():
        vb = v.bfloat16()
        wide = vb.float()
        r = prog.residual(vb, boundary=True)
        assert L.last_path() == "tma" and r.dtype == torch.float32, name
        assert torch.equal(r, prog.residual(wide, boundary=True)), name
        assert_close_rel(r.cpu().numpy(), oracle(wide.cpu().numpy()), TOL, name + " vs oracle on rounded input")
        assert torch.equal(prog.joint_scores(vb, cuda(w)), prog.joint_scores(wide, cuda(w))), name
        assert torch.equal(prog.coverage_counts(vb, 0.7, cuda(w)), prog.coverage_counts(wide, 0.7, cuda(w))), name
        ref = 0.5 * r
        assert torch.equal(prog.abs_scores(vb, ref=ref, boundary=True), prog.abs_scores(wide, ref=ref, boundary=True)), name
    # 1-D + time (batch-rows tiles)
    m1, cb = _mods()["1d"], synth.BURGERS
    pb = programs.burgers(m1.ConvOperator("t", 1), m1.ConvOperator("x", 1), m1.ConvOperator("x", 2), cb["dx"], cb["dt"], cb["nu"])
    ub = cuda(synth.fields(74, 1, (37, 12, 264))[:, 0]).bfloat16()
    assert torch.equal(pb.residual(ub, boundary=True), pb.residual(ub.float(), boundary=True))
    assert L.last_path() == "tma"
    # rows that are not 16-byte multiples in bf16 cannot be staged by TMA: refused, never silently widened
    bad = cuda(synth.fields(75, 1, (2, 5, 8, 132)))[:, 0].bfloat16()
    with pytest.raises(NotImplementedError):
        programs.linear(_wave_op()).residual(bad, boundary=True)


def test_jorek_coefficient_field_programs_match_reference_golden():
    """the JOREK reduced-MHD residuals (R-dependent coefficients, second-order D_RR / D_ZZ
    scaled by the re-bound gamma) against tests/golden/jorek.npz, which the unmodified reference's
    operators produced; the fused score / coverage epilogues work on them like on every program."""
    g = load_golden("jorek")
    C = _mods()["2d"].ConvOperator
    gam = torch.tensor(5 / 3, dtype=torch.float32)
    D = [C(domain="t", order=1, scale=1), C(domain="x", order=1, scale=1), C(domain="y", order=1, scale=1),
         C(domain="x", order=2, scale=gam), C(domain="y", order=2, scale=gam)]
    np.testing.assert_array_equal(D[3].kernel.cpu().numpy(), g["k|D_RR"])
    v = cuda(g["vars"])
    for eq, coef in (("continuity", torch.tensor(3.4)), ("temperature", torch.tensor(2.25 * 1e-7))):
        p = programs.jorek(eq, *D, torch.from_numpy(g["R"]), coef, gam)
        r = p.residual(v, boundary=True)
        assert L.last_path() == "flat"                                 # R(x, y) staged as a stationary plane
        p.path = L.PATH_GENERIC
        assert_close_rel(p.residual(v, boundary=True).cpu().numpy(), g[eq], TOL, eq + " generic")
        assert L.last_path() == "generic"
        p.path = L.PATH_AUTO
        assert_close_rel(r.cpu().numpy(), g[eq], TOL, eq)
        assert_close_rel(p.residual(v).cpu().numpy(), g[eq][:, 1:-1, 1:-1, 1:-1], TOL, eq + " interior")
        s = p.joint_scores(v)
        assert torch.equal(s, r[:, 1:-1, 1:-1, 1:-1].abs().amax(dim=(1, 2, 3))), eq
    with pytest.raises(RuntimeError):                                  # R must match the last axis, like torch broadcasting
        programs.jorek("continuity", *D, torch.ones(5), 3.4).residual(v)


def test_fno_layout_views_are_transposed_by_the_library_and_match_the_contiguous_run():
    """The t-innermost tensors FNO surrogates emit, passed on as pred.permute(0,1,4,2,3):
    cppre_transpose_inner brings each field to [Nt][Nx*Ny] order (bit-equal to torch's copy, ragged tile
    edges included) and the residual equals the run on a contiguous copy bit for bit."""
    o, n = _programs_2d(), synth.NS
    prog = programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], n["dx"], n["dy"], n["dt"], n["nu"])
    g = torch.Generator(device="cuda").manual_seed(9)
    for nx, ny, nt in ((12, 20, 50), (16, 128, 7), (40, 132, 33)):
        raw = torch.randn((3, 3, nx, ny, nt), device="cuda", generator=g)      # [BS,F,Nx,Ny,Nt]
        view = raw.permute(0, 1, 4, 2, 3)                                      # [BS,F,Nt,Nx,Ny]
        for j in range(3):
            assert torch.equal(ops._to_logical_order(view[:, j], [1, 2, 0]), view[:, j].contiguous())
        want = prog.residual(view.contiguous(), boundary=True)
        got = prog.residual(view, boundary=True)
        assert got.shape == want.shape and torch.equal(got, want), (nx, ny, nt)


def _tin(t):
    """Same logical values, stored t-innermost: what pred.permute(0,1,4,2,3) of an FNO output looks like."""
    nd = t.dim()
    fwd = list(range(nd - 3)) + [nd - 2, nd - 1, nd - 3]
    inv = list(range(nd - 3)) + [nd - 1, nd - 3, nd - 2]
    return t.permute(fwd).contiguous().permute(inv)


def test_t_innermost_kernel_takes_fno_views_in_place_and_is_bit_equal_to_the_contiguous_run():
    """the t-innermost star kernel reads [BS,F,Nx,Ny,Nt] memory through its
    [BS,F,Nt,Nx,Ny] view with no transpose pass; STORE / SCORE / COVER results equal the contiguous
    run BIT FOR BIT (same arm accumulation order), for every shipped program family, ragged tiles,
    short and long t lines, and sample strides that are not the dense ones (field slices of a 5-D tensor)."""
    import bench
    o, n = _programs_2d(), synth.NS
    progs = {"ns": (programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], n["dx"], n["dy"], n["dt"],
                                         n["nu"]), 3),
             "cont": (programs.ns_continuity(o["D_x"], o["D_y"], n["dx"], n["dy"]), 3),
             "wave": (bench.make_program("wave"), 1)}
    for eq in ("continuity", "momentum", "energy", "induction", "gauss"):
        progs["mhd_" + eq] = (bench.make_program("mhd", eq), 6)
    g = torch.Generator(device="cuda").manual_seed(21)
    shapes = ((3, 50, 12, 20), (2, 4, 7, 9), (2, 100, 33, 17), (1, 80, 5, 130), (2, 20, 64, 64), (1, 512, 6, 10),
              (2, 6, 3, 4), (1, 258, 9, 10))
    for name, (prog, nf) in progs.items():
        for (bs, nt, nx, ny) in shapes:
            if name.startswith("mhd") and nt > 100:
                continue
            shape = (bs, nf, nt, nx, ny) if nf > 1 else (bs, nt, nx, ny)
            v = torch.randn(shape, device="cuda", generator=g)
            if name.startswith("mhd"):
                v = v.abs() + 0.5                                        # density / pressure away from 0
            vt = _tin(v)
            assert not vt.is_contiguous()
            want = prog.residual(v, boundary=True)
            got = prog.residual(vt, boundary=True)
            assert L.last_path() == "tin", (name, shape)
            assert torch.equal(got, want), (name, shape)
            assert got.stride(1) == 1                                   # result stays t-innermost
            ref = torch.randn(want.shape, device="cuda", generator=g)
            w = torch.rand(want.shape[1:], device="cuda", generator=g) + 0.1
            for interior in (True, False):
                for red in ("max", "sum"):
                    a = prog.joint_scores(v, inv_modulation=w, ref=ref, interior=interior, reduce=red)
                    b = prog.joint_scores(vt, inv_modulation=w, ref=_tin(ref), interior=interior, reduce=red)
                    assert L.last_path() == "tin"
                    if red == "max":
                        assert torch.equal(a, b), (name, shape, interior)
                    else:                                               # sums: different cell order
                        assert torch.allclose(a, b, rtol=2e-5, atol=1e-6), (name, shape, interior)
                a = prog.coverage_counts(v, 0.7, half=w, centre=ref, interior=interior)
                b = prog.coverage_counts(vt, 0.7, half=_tin(w), centre=ref, interior=interior)   # ref copied once
                assert L.last_path() == "tin"
                assert torch.equal(a, b), (name, shape, interior)
            sc = torch.empty_like(want)
            a = prog.abs_scores(v, ref=ref, boundary=True)
            outp = _tin(sc)
            b = prog.abs_scores(vt, ref=_tin(ref), boundary=True, out=outp)
            assert L.last_path() == "tin" and b.data_ptr() == outp.data_ptr() and torch.equal(a, b)


def test_t_innermost_kernel_odd_lines_unaligned_views_and_what_it_declines():
    """FNO views the plain t-innermost units cannot stage -- odd Nt (the y-neighbour is no longer an 8-byte shared
    load), a base that is not 16-byte aligned -- run on the flexible units (stencil_tinx_*: 4-byte cp.async staging,
    scalar epilogue) and equal the contiguous run bit for bit, every epilogue; Taylor-4 taps and bf16 input still fall
    back to the transpose + contiguous kernels and match; PATH_TIN on a contiguous tensor raises."""
    o, n = _programs_2d(), synth.NS
    prog = programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], n["dx"], n["dy"], n["dt"], n["nu"])
    g = torch.Generator(device="cuda").manual_seed(5)
    for shape, misalign in (((2, 3, 33, 12, 20), False), ((2, 3, 25, 9, 7), False), ((2, 3, 50, 12, 20), True),
                            ((1, 3, 101, 6, 10), True)):
        v = torch.randn(shape, device="cuda", generator=g)
        bs, nf, nt, nx, ny = shape
        buf = torch.empty(v.numel() + 1, device="cuda")
        raw = buf[1 if misalign else 0:][: v.numel()].view(bs, nf, nx, ny, nt)      # physical [BS,F,Nx,Ny,Nt]
        raw.copy_(v.permute(0, 1, 3, 4, 2))
        vt = raw.permute(0, 1, 4, 2, 3)
        want = prog.residual(v, boundary=True)
        got = prog.residual(vt, boundary=True)
        assert L.last_path() == "tin", (shape, misalign)
        assert torch.equal(got, want), (shape, misalign)
        ref = torch.randn(want.shape, device="cuda", generator=g)
        w = torch.rand(want.shape[1:], device="cuda", generator=g) + 0.1
        for interior in (True, False):
            a = prog.joint_scores(v, inv_modulation=w, ref=ref, interior=interior)
            b = prog.joint_scores(vt, inv_modulation=w, ref=_tin(ref), interior=interior)
            assert L.last_path() == "tin" and torch.equal(a, b), (shape, interior)
            a = prog.coverage_counts(v, 0.7, half=w, centre=ref, interior=interior)
            b = prog.coverage_counts(vt, 0.7, half=w, centre=ref, interior=interior)
            assert L.last_path() == "tin" and torch.equal(a, b), (shape, interior)
    prog.path = L.PATH_TIN
    with pytest.raises(NotImplementedError):
        prog.residual(torch.randn((2, 3, 16, 12, 20), device="cuda"), boundary=True)     # contiguous input is not t-innermost
    prog.path = L.PATH_AUTO
    lap4 = programs.linear(_mods()["2d"].ConvOperator(("x", "y"), 2, taylor_order=4))     # 5x5x5 taps
    u = torch.randn((2, 16, 12, 20), device="cuda", generator=g)
    assert torch.equal(lap4.residual(_tin(u), boundary=True), lap4.residual(u, boundary=True))
    assert L.last_path() != "tin"
    vb = torch.randn((2, 3, 16, 12, 128), device="cuda").to(torch.bfloat16)
    assert torch.equal(prog.residual(_tin(vb), boundary=True), prog.residual(vb, boundary=True))
    assert L.last_path() != "tin"


def test_ns_marginal_cp_at_full_config_size():
    """The HEADLINE configuration at its full size (BASELINE config 4: NS momentum PRE + marginal per-cell
    CP, [512,3,50,256,256], Marginal/NS_Residuals_CP.py:327-337), the exact pipeline bench.py times:
    K1 STORE on the TMA path -> K3 per-cell quantile -> K1 COVER.  A 2-sample slab of the residual against
    the oracle at 1e-5; qhat of 4096 random cells bit-equal to the oracle's calibrate() on the downloaded
    score columns; COVER counts of 4 validation samples bit-equal to NumPy on their stored residual."""
    import bench
    from Neural_PDE.UQ.inductive_cp import calibrate_rank
    bs, nt, n = 512, 50, 256
    n_cal = bs // 2
    c = synth.NS
    prog = bench.make_program("ns")
    v = bench.synth_fields_gpu(bs, 3, (nt, n, n), torch.device("cuda"), seed=1234)
    scores = torch.empty((n_cal, nt, n, n), device="cuda")
    prog.abs_scores(v[:n_cal], boundary=True, out=scores)
    assert L.last_path() == "tma"
    f = v[:2].cpu().numpy()
    want = O.residual_ns_momentum(f[:, 0], f[:, 1], f[:, 2], c["dx"], c["dy"], c["dt"], c["nu"], O.ops_2d())
    assert_close_rel(scores[:2].cpu().numpy(), np.abs(want), TOL, "|R| slab vs oracle")
    k = calibrate_rank(n_cal, n_cal, 0.1)
    qhat = ops.select_ranks(scores, [k])[0]
    ids = torch.randint(0, nt * n * n, (4096,), generator=torch.Generator().manual_seed(3)).cuda()
    cols = scores.reshape(n_cal, -1)[:, ids].cpu().numpy()
    np.testing.assert_array_equal(qhat.reshape(-1)[ids].cpu().numpy(), O.calibrate(cols, n_cal, 0.1))
    inside = prog.coverage_counts(v[n_cal:], 1.0, half=qhat)
    assert L.last_path() == "tma" and inside.shape == (bs - n_cal,)
    r4 = prog.residual(v[n_cal:n_cal + 4], boundary=True).cpu().numpy()
    q = qhat.cpu().numpy()
    want_cnt = ((r4 >= -q) & (r4 <= q))[:, 1:-1, 1:-1, 1:-1].sum(axis=(1, 2, 3))
    np.testing.assert_array_equal(inside[:4].cpu().numpy(), want_cnt)
    cov = float(inside.to(torch.int64).sum()) / ((bs - n_cal) * prog.interior_cells(v.shape))
    assert 0.88 <= cov <= 0.92, cov                                    # exchangeable halves: ~1 - alpha
    del v, scores
    torch.cuda.empty_cache()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_tma_path_on_two_devices_in_one_process():
    """Function attributes (the dynamic shared memory opt-in of the TMA ring, ~88 KB for NS momentum) are
    per device: the same process must be able to run the TMA kernel on cuda:0 and then on cuda:1."""
    prog = __import__("bench").make_program("ns")
    outs = []
    for d in (0, 1, 0):
        v = cuda(synth.fields(9, 3, (4, 6, 16, 128))).to(f"cuda:{d}")
        outs.append(prog.residual(v, boundary=True).cpu())
        assert L.last_path() == "tma"
        s = ops.select_ranks(outs[-1].abs().to(f"cuda:{d}"), [2])      # K3 opt-in path on the same device
        assert s.device.index == d
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])


def test_autograd_through_nonlinear_programs_kernel_tensor_and_spectral_paths():
    """SURVEY 8 f-2 / ADVICE r1: a call whose input, kernel or weight requires grad must either produce
    the gradient or raise -- never detach silently.  Nonlinear residual programs (product rule), d/d(kernel)
    (D.kernel.requires_grad = True, Physics_Informed/Wave_FNO_PI.py:206), the spectral path and the
    _FFTConv layers all backpropagate; checked against torch.autograd through plain PyTorch fp32
    expressions of the same ops (oracle/torch_port.py)."""
    import torch.nn.functional as F
    from oracle import torch_port as TP
    from fft_conv_pytorch.fft_conv import FFTConv2d, fft_conv
    o, c = _programs_2d(), synth.NS
    # NS momentum as a PRE loss: d(loss)/d(vars)
    prog = programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], c["dx"], c["dy"], c["dt"], c["nu"])
    v = cuda(synth.fields(5, 3, (2, 6, 12, 16))).requires_grad_(True)
    r = prog.residual(v)
    assert r.requires_grad
    (r ** 2).sum().backward()
    v2 = v.detach().cpu().clone().requires_grad_(True)
    (TP.ns_momentum(v2, c["dx"], c["dy"], c["dt"], c["nu"]) ** 2).sum().backward()
    assert_close_rel(v.grad.cpu().numpy(), v2.grad.numpy(), 2e-5, "d loss / d vars (NS momentum)")
    # the fused forward of the same call (no grad) agrees with the autograd forward
    assert_close_rel(prog.residual(v.detach()).cpu().numpy(), r.detach().cpu().numpy(), TOL, "fused vs autograd forward")
    # MHD momentum (a division by rho inside the product rule) through the fused score epilogue's twin
    m = synth.MHD
    pm = programs.mhd("momentum", o["D_t"], o["D_x"], o["D_y"], m["gamma"])
    w = cuda(synth.fields(6, 6, (2, 5, 10, 16), positive=(0, 3))).requires_grad_(True)
    pm.joint_scores(w).sum().backward()
    w2 = w.detach().cpu().clone().requires_grad_(True)
    TP.mhd_momentum(w2).abs().flatten(1).amax(dim=1).sum().backward()
    assert_close_rel(w.grad.cpu().numpy(), w2.grad.numpy(), 5e-5, "d score / d vars (MHD momentum)")
    # d/d(kernel): a trainable operator
    W = _wave_op()
    W.kernel = W.kernel.clone().cuda().requires_grad_(True)
    u = torch.randn(3, 6, 12, 16, device="cuda")
    (W(u) ** 2).sum().backward()
    k2 = W.kernel.detach().clone().requires_grad_(True)
    (F.conv3d(u.unsqueeze(1), k2[None, None], padding=1).squeeze(1) ** 2).sum().backward()
    assert_close_rel(W.kernel.grad.cpu().numpy(), k2.grad.cpu().numpy(), 1e-4, "d loss / d kernel")   # fp32 sums of 3456 products
    # spectral path and the FFTConv layer: gradients w.r.t. signal, weight and bias
    sig = torch.randn(2, 1, 16, 18, device="cuda", requires_grad=True)
    ker = torch.randn(1, 1, 3, 3, device="cuda", requires_grad=True)
    fft_conv(sig, ker, padding=1).sum().backward()
    s2, k2 = sig.detach().clone().requires_grad_(True), ker.detach().clone().requires_grad_(True)
    F.conv2d(s2, k2, padding=1).sum().backward()
    assert_close_rel(sig.grad.cpu().numpy(), s2.grad.cpu().numpy(), 1e-4, "fft_conv d/d signal")
    assert_close_rel(ker.grad.cpu().numpy(), k2.grad.cpu().numpy(), 1e-4, "fft_conv d/d kernel")
    layer = FFTConv2d(1, 1, 3, padding=1).cuda()
    layer(torch.randn(2, 1, 8, 8, device="cuda")).sum().backward()
    assert layer.weight.grad is not None and layer.bias.grad is not None and float(layer.weight.grad.abs().sum()) > 0
    # differentiate(): the manual spectral form keeps its graph too
    x = torch.randn(2, 6, 8, 10, device="cuda", requires_grad=True)
    o["D_x"].differentiate(x, correlation=True).sum().backward()
    assert x.grad is not None and float(x.grad.abs().sum()) > 0


def _every_program():
    """One instance of every CPPRE_PROG_* kind with an input of the right shape: (name, program, vars)."""
    o, m1, c, b = _programs_2d(), _mods()["1d"], synth.NS, synth.BURGERS
    C2 = _mods()["2d"].ConvOperator
    grid = (2, 6, 12, 16)
    one = lambda seed: cuda(synth.fields(seed, 1, grid))[:, 0]
    out = [("linear", programs.linear(_wave_op()), one(1)),
           ("linear_1d", programs.linear(m1.ConvOperator("x", 2)), one(2)[:, :, 0].contiguous()),   # the nd == 2 squeeze path
           ("burgers", programs.burgers(m1.ConvOperator("t", 1), m1.ConvOperator("x", 1), m1.ConvOperator("x", 2),
                                        b["dx"], b["dt"], b["nu"]), one(3)[:, :, 0].contiguous()),
           ("ns_continuity", programs.ns_continuity(o["D_x"], o["D_y"], c["dx"], c["dy"]), cuda(synth.fields(4, 3, grid))),
           ("ns_momentum", programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], c["dx"], c["dy"], c["dt"],
                                                c["nu"]), cuda(synth.fields(5, 3, grid)))]
    for eq in ("continuity", "momentum", "energy", "induction", "gauss"):
        out.append(("mhd_" + eq, programs.mhd(eq, o["D_t"], o["D_x"], o["D_y"], synth.MHD["gamma"]),
                    cuda(synth.fields(6, 6, grid, positive=(0, 3)))))
    gam = torch.tensor(5 / 3, dtype=torch.float32)
    D = [C2(domain="t", order=1, scale=1), C2(domain="x", order=1, scale=1), C2(domain="y", order=1, scale=1),
         C2(domain="x", order=2, scale=gam), C2(domain="y", order=2, scale=gam)]
    R = torch.linspace(1.5, 2.5, grid[-1])
    for eq, coef in (("continuity", 3.4), ("temperature", 2.25e-7)):
        out.append(("jorek_" + eq, programs.jorek(eq, *D, R, torch.tensor(coef), gam),
                    cuda(synth.fields(7, 3, grid, positive=(0, 2)))))
    return out


def test_autograd_forward_equals_the_fused_forward_for_every_program_kind():
    """the gradient path restates all programs in torch (programs._FORMULAS) beside csrc/programs.cuh.
    This pins the two copies together: for EVERY program kind (and the 1-D squeeze path) the residual computed on
    the autograd graph equals the fused kernel's on the same input, fp32 tolerance (2e-5 of max|ref|: the two sides
    order their fp32 sums differently)."""
    seen = set()
    for name, prog, v in _every_program():
        fused = prog.residual(v)
        assert not fused.requires_grad
        on_graph = prog.residual(v.clone().requires_grad_(True))
        assert on_graph.requires_grad, name
        assert_close_rel(on_graph.detach().cpu().numpy(), fused.cpu().numpy(), 2e-5, name)
        seen.add(prog.kind)
    assert seen == set(programs._FORMULAS), "a program kind has no autograd-vs-fused check"


def test_burgers_gradient_equals_torch_autograd_through_the_plain_expression():
    """Gradcheck-style comparison for a 1-D program: d(sum res^2)/du through the library's autograd path against
    torch.autograd through the plain PyTorch fp32 expression of the same residual (oracle/torch_port.py)."""
    from oracle import torch_port as TP
    b, m1 = synth.BURGERS, _mods()["1d"]
    prog = programs.burgers(m1.ConvOperator("t", 1), m1.ConvOperator("x", 1), m1.ConvOperator("x", 2),
                            b["dx"], b["dt"], b["nu"])
    u = cuda(synth.fields(11, 1, (3, 10, 4, 24)))[:, 0, :, 0].contiguous().requires_grad_(True)
    (prog.residual(u) ** 2).sum().backward()
    u2 = u.detach().cpu().clone().requires_grad_(True)
    (TP.burgers(u2, b["dx"], b["dt"], b["nu"]) ** 2).sum().backward()
    assert_close_rel(u.grad.cpu().numpy(), u2.grad.numpy(), 2e-5, "d loss / d u (Burgers)")


def test_scoring_call_on_a_grad_carrying_input_warns_once_about_the_unfused_path():
    """a scoring entry point handed a tensor that requires grad (a model output made outside no_grad)
    leaves the fused kernel for the differentiable path; that is announced once per entry point, not silent."""
    import warnings
    o, c = _programs_2d(), synth.NS
    prog = programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], c["dx"], c["dy"], c["dt"], c["nu"])
    v = cuda(synth.fields(5, 3, (2, 6, 12, 16)))
    programs._WARNED_UNFUSED.clear()
    with pytest.warns(RuntimeWarning, match="unfused differentiable path"):
        on_graph = prog.joint_scores(v.clone().requires_grad_(True))
    with warnings.catch_warnings():
        warnings.simplefilter("error")
        again = prog.joint_scores(v.clone().requires_grad_(True))       # announced once
        fused = prog.joint_scores(v)                                     # the fused path never warns
    assert torch.equal(on_graph, again) and torch.allclose(on_graph.detach(), fused, rtol=1e-4, atol=0)


class _MinMax:
    """Shape of the normaliser the scripts use: attributes a, b; decode(x) = (x - b) / a
    [ref: Tests/test_ns.py:111-134; class in the un-vendored Neural_PDE package]."""

    def __init__(self, a, b):
        self.a, self.b = torch.tensor(a), torch.tensor(b)


def test_normaliser_decode_fold_equals_the_residual_of_the_decoded_field():
    """SURVEY.md section 8 (f-1), second half: ``pred = out_normalizer.decode(pred)`` directly upstream of the
    residual call [ref: Joint/Wave_Residuals_CP.py:213-217] is folded into K1 -- no pass over the tensor.
    Parity: against the ORACLE evaluated on the explicitly decoded field (fp32), interior cells, 1e-5 of
    max|ref|; on all three kernels (TMA, t-innermost, generic), every epilogue, per-variable and scalar
    coefficients, a MinMax-style object; boundary=True takes the explicit decode and equals it bit for bit."""
    import bench
    o, n = _programs_2d(), synth.NS
    ns = programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], n["dx"], n["dy"], n["dt"], n["nu"])
    sc, of = [2.0, 0.5, 3.0], [0.3, -0.2, 1.0]
    for shape, path in (((3, 12, 16, 128), "tma"), ((2, 8, 9, 30), "generic")):
        enc = synth.fields(5, 3, shape)
        dec = (enc * np.asarray(sc, np.float32).reshape(1, 3, 1, 1, 1) + np.asarray(of, np.float32).reshape(1, 3, 1, 1, 1)).astype(np.float32)
        want = O.residual_ns_momentum(dec[:, 0], dec[:, 1], dec[:, 2], n["dx"], n["dy"], n["dt"], n["nu"])[:, 1:-1, 1:-1, 1:-1]
        got = ns.residual(cuda(enc), decode=(sc, of))
        assert L.last_path() == path
        assert_close_rel(got.cpu().numpy(), want, TOL)
        if path == "tma":
            vt = _tin(cuda(enc))
            got_t = ns.residual(vt, decode=(sc, of))
            assert L.last_path() == "tin"
            assert_close_rel(got_t.cpu().numpy(), want, TOL)
        # the explicit form (boundary=True) is what the fold replaces: identical to decoding by hand
        full = ns.residual(cuda(enc), boundary=True, decode=(sc, of))
        by_hand = ns.residual(cuda(enc) * torch.tensor(sc, device="cuda").view(1, 3, 1, 1, 1)
                              + torch.tensor(of, device="cuda").view(1, 3, 1, 1, 1), boundary=True)
        assert torch.equal(full, by_hand)
        assert_close_rel(full[:, 1:-1, 1:-1, 1:-1].cpu().numpy(), want, TOL)
        # fused epilogues with the fold against the same epilogues on the decoded tensor
        dec_t = cuda(dec)
        w = torch.rand(shape[1:], device="cuda") + 0.1
        a = ns.joint_scores(cuda(enc), inv_modulation=w, decode=(sc, of))
        b = ns.joint_scores(dec_t, inv_modulation=w)
        assert torch.allclose(a, b, rtol=1e-4, atol=0), (a, b)
        q = float(np.quantile(np.abs(want), 0.8))
        ca = ns.coverage_counts(cuda(enc), q, decode=(sc, of))
        cb = ns.coverage_counts(dec_t, q)
        assert int((ca - cb).abs().max()) <= max(2, int(1e-3 * want[0].size)), (ca, cb)
        sa = ns.abs_scores(cuda(enc), decode=(sc, of))
        assert_close_rel(sa.cpu().numpy(), np.abs(want), TOL)
    # MHD momentum (6 variables, division by the decoded density) on the TMA and t-innermost kernels
    mhd = bench.make_program("mhd", "momentum")
    enc = synth.fields(9, 6, (2, 8, 16, 128), positive=(0, 3))
    sc6, of6 = [1.5, 2.0, 0.5, 2.5, 1.0, 3.0], [0.25, 0.1, -0.3, 0.5, 0.0, 0.2]
    dec = (enc * np.asarray(sc6, np.float32).reshape(1, 6, 1, 1, 1) + np.asarray(of6, np.float32).reshape(1, 6, 1, 1, 1)).astype(np.float32)
    want = O.residual_mhd("momentum", *(dec[:, j] for j in range(6)), synth.MHD["gamma"])[:, 1:-1, 1:-1, 1:-1]
    got = mhd.residual(cuda(enc), decode=(sc6, of6))
    assert L.last_path() == "tma"
    assert_close_rel(got.cpu().numpy(), want, TOL)
    got = mhd.residual(_tin(cuda(enc)), decode=(sc6, of6))
    assert L.last_path() == "tin"
    assert_close_rel(got.cpu().numpy(), want, TOL)
    # single-field linear operator (wave), scalar coefficients and the MinMax-object form: (x - b) / a
    wave = bench.make_program("wave")
    u = synth.fields(3, 1, (2, 10, 16, 128))[:, 0]
    nrm = _MinMax(0.25, -0.5)
    dec = ((u - np.float32(-0.5)) / np.float32(0.25)).astype(np.float32)
    want = O.correlate(dec, O.build_kernel("2d", "t", 2) - np.float32(synth.WAVE["c2"]) * O.build_kernel("2d", ("x", "y"), 2))[:, 1:-1, 1:-1, 1:-1]
    got = wave.residual(cuda(u), decode=nrm)
    assert L.last_path() == "tma"
    assert_close_rel(got.cpu().numpy(), want, TOL)
    # Taylor-4 stencil (5 points per axis): from the first kept layer it reaches the padding, where the reference
    # pads the DECODED field, so the fold does not apply.  The library refuses the flag; the program object decodes
    # explicitly and still equals the oracle on the decoded field (an offset != 0 is what would expose a wrong fold)
    lap4 = programs.linear(_mods()["2d"].ConvOperator(("x", "y"), 2, taylor_order=4))
    assert not lap4._fold_ok() and wave._fold_ok()
    u4 = synth.fields(4, 1, (2, 8, 16, 128))[:, 0]
    dec4 = (u4 * np.float32(2.0) + np.float32(0.75)).astype(np.float32)
    want4 = O.correlate(dec4, O.build_kernel("2d", ("x", "y"), 2, taylor_order=4))[:, 1:-1, 1:-1, 1:-1]
    got4 = lap4.residual(cuda(u4), decode=(2.0, 0.75))
    assert_close_rel(got4.cpu().numpy(), want4, TOL)
    assert_close_rel(lap4.abs_scores(cuda(u4), decode=(2.0, 0.75)).cpu().numpy(), np.abs(want4), TOL)
    with pytest.raises(NotImplementedError, match="within one cell"):
        ops.run_program(L.PROG_LINEAR, [cuda(u4)], *lap4._slots(), [], decode=([2.0], [0.75]))
    # 1-D + time (batch-rows TMA mode): Burgers with a scalar decode
    import ConvOps_1d as C1
    b = synth.BURGERS
    bur = programs.burgers(C1.ConvOperator("t", 1), C1.ConvOperator("x", 1), C1.ConvOperator("x", 2), b["dx"], b["dt"], b["nu"])
    u1 = synth.fields(4, 1, (16, 12, 128))[:, 0]
    dec = (u1 * np.float32(1.7) + np.float32(0.4)).astype(np.float32)
    want = O.residual_burgers(dec, b["dx"], b["dt"], b["nu"])[:, 1:-1, 1:-1]
    got = bur.residual(cuda(u1), decode=(1.7, 0.4))
    assert L.last_path() == "tma"
    assert_close_rel(got.cpu().numpy(), want, TOL)
    with pytest.raises(ValueError):
        ns.residual(cuda(synth.fields(5, 3, (1, 6, 8, 8))), decode=([1.0, 2.0], [0.0, 0.0]))     # 2 scales, 3 variables
    with pytest.raises(NotImplementedError):
        ns.residual(cuda(synth.fields(5, 3, (1, 6, 8, 128))).bfloat16(), decode=(1.0, 0.0))


def test_flat_line_kernel_takes_ragged_contiguous_data_and_is_bit_equal_to_the_generic_kernel():
    """Contiguous star requests the TMA unit cannot stage -- N0 % 4 != 0 (254 x 254, 33 x 33, config 1's Nx = 197),
    unaligned bases -- run on the flat-line star kernel (stencil_tin_impl.cuh role 1: march = t, lines = x) instead
    of the one-thread-per-cell generic kernel.  Same arm order, same algebra: STORE fields, max-scores and coverage
    counts are BIT-EQUAL to the generic kernel's, for every program family, aligned (bulk copies, 128-bit epilogue)
    and unaligned rows (4-byte cp.async, scalar epilogue), odd and even lines, partial tiles, 1-D + time data.
    (MHD energy is the one program where nvcc contracts a packed product into the following packed sum -- e.g.
    g*p_gas + By^2 -- in the star kernels, while the generic kernel rounds them separately like NumPy: there the
    comparison is the usual 1e-5 of max|ref| and a handful of borderline cells in the counts.)"""
    import bench
    o, n = _programs_2d(), synth.NS
    progs = {"ns": (programs.ns_momentum(o["D_t"], o["D_x"], o["D_y"], o["D_xx_yy"], n["dx"], n["dy"], n["dt"], n["nu"]), 3),
             "cont": (programs.ns_continuity(o["D_x"], o["D_y"], n["dx"], n["dy"]), 3),
             "wave": (bench.make_program("wave"), 1)}
    for eq in ("continuity", "momentum", "energy", "induction", "gauss"):
        progs["mhd_" + eq] = (bench.make_program("mhd", eq), 6)
    g = torch.Generator(device="cuda").manual_seed(33)
    shapes = ((2, 5, 7, 197), (2, 4, 8, 254), (3, 6, 9, 254), (2, 3, 4, 6), (1, 7, 5, 33), (2, 5, 33, 33), (2, 4, 2, 2),
              (1, 3, 40, 510), (2, 6, 3, 255), (1, 2, 300, 10))
    for name, (prog, nf) in progs.items():
        for (bs, nt, nx, ny) in shapes:
            shape = (bs, nf, nt, nx, ny) if nf > 1 else (bs, nt, nx, ny)
            v = torch.randn(shape, device="cuda", generator=g)
            if name.startswith("mhd"):
                v = v.abs() + 0.5
            for misalign in (False, True):
                if misalign:                                            # base 4 bytes off a 16-byte boundary
                    buf = torch.empty(v.numel() + 1, device="cuda")
                    vv = buf[1:].view(v.shape)
                    vv.copy_(v)
                else:
                    vv = v
                prog.path = L.PATH_GENERIC
                want = prog.residual(vv, boundary=True)
                assert L.last_path() == "generic"
                prog.path = L.PATH_AUTO
                got = prog.residual(vv, boundary=True)
                assert L.last_path() == "flat", (name, shape, misalign)
                exact = name != "mhd_energy"
                if exact:
                    assert torch.equal(got, want), (name, shape, misalign, float((got - want).abs().max()))
                else:
                    assert_close_rel(got.cpu().numpy(), want.cpu().numpy(), TOL, f"{name} {shape}")
                ref = torch.randn(want.shape, device="cuda", generator=g)
                w = torch.rand(want.shape[1:], device="cuda", generator=g) + 0.1
                for interior in (True, False):
                    for red in ("max", "sum"):
                        prog.path = L.PATH_GENERIC
                        a = prog.joint_scores(vv, inv_modulation=w, ref=ref, interior=interior, reduce=red)
                        prog.path = L.PATH_AUTO
                        b = prog.joint_scores(vv, inv_modulation=w, ref=ref, interior=interior, reduce=red)
                        assert L.last_path() == "flat"
                        if red == "max" and exact:
                            assert torch.equal(a, b), (name, shape, interior)
                        else:
                            assert torch.allclose(a, b, rtol=2e-5, atol=1e-6), (name, shape, interior)
                    prog.path = L.PATH_GENERIC
                    a = prog.coverage_counts(vv, 0.7, half=w, centre=ref, interior=interior)
                    prog.path = L.PATH_AUTO
                    b = prog.coverage_counts(vv, 0.7, half=w, centre=ref, interior=interior)
                    assert L.last_path() == "flat"
                    assert torch.equal(a, b) if exact else int((a - b).abs().max()) <= 3, (name, shape, interior)
                prog.path = L.PATH_GENERIC
                a = prog.abs_scores(vv, ref=ref, boundary=True)
                prog.path = L.PATH_AUTO
                sc = torch.empty_like(want)
                b = prog.abs_scores(vv, ref=ref, boundary=True, out=sc)
                assert L.last_path() == "flat" and b.data_ptr() == sc.data_ptr()
                assert torch.equal(a, b) if exact else torch.allclose(a, b, rtol=0, atol=TOL * float(a.max()))
    # 1-D + time (ConvOps_1d): config 1's [n, 10, 197] advection operator and Burgers on ragged rows
    import ConvOps_1d as C1
    A = C1.ConvOperator()
    A.kernel = C1.ConvOperator("t", 1).kernel + 0.5 * C1.ConvOperator("x", 1).kernel
    bq = synth.BURGERS
    bur = programs.burgers(C1.ConvOperator("t", 1), C1.ConvOperator("x", 1), C1.ConvOperator("x", 2), bq["dx"], bq["dt"], bq["nu"])
    for prog in (programs.linear(A), bur):
        for shape in ((100, 10, 197), (7, 33, 50), (3, 5, 255), (3, 5, 510)):
            u = torch.randn(shape, device="cuda", generator=g)
            prog.path = L.PATH_GENERIC
            want = prog.residual(u, boundary=True)
            prog.path = L.PATH_AUTO
            got = prog.residual(u, boundary=True)
            assert L.last_path() == "flat", (prog.name, shape)
            assert torch.equal(got, want), (prog.name, shape)
            assert_close_rel(got.cpu().numpy(), (O.correlate(u.cpu().numpy(), A.kernel.numpy()) if prog.name == "linear"
                                                 else O.residual_burgers(u.cpu().numpy(), bq["dx"], bq["dt"], bq["nu"])), TOL)
    # what it declines still runs (generic) and PATH_FLAT then fails loudly
    W = bench.make_program("wave")
    u = torch.randn((2, 4, 6, 2050), device="cuda", generator=g)       # rows longer than a tile holds
    W.residual(u, boundary=True)
    assert L.last_path() == "generic"
    W.path = L.PATH_FLAT
    with pytest.raises(NotImplementedError):
        W.residual(u, boundary=True)
