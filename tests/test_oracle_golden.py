"""The NumPy oracle against outputs of the unmodified reference (tests/golden, written by
oracle/make_golden.py).  CPU only."""
import ast
import os

import numpy as np
import pytest

from conftest import assert_close_rel, load_golden
from oracle import cpu_oracle as O
from oracle import synth


def _parse_key(key):
    module, dom, order, ty, scale = key.split("|")
    return module, ast.literal_eval(dom), ast.literal_eval(order), int(ty), float(scale)


def test_kernel_tables_match_reference_constructors():
    g = load_golden("kernels")
    n = 0
    for key in g.files:
        if key.startswith("err|"):
            continue
        module, dom, order, ty, scale = _parse_key(key)
        want = g[key]
        got = O.build_kernel(module, dom, order, scale, ty)
        if want.size == 0:
            assert got is None, f"{key}: reference left no kernel, oracle built one"
        else:
            assert got is not None, f"{key}: oracle built no kernel"
            assert got.shape == want.shape, key
            np.testing.assert_array_equal(got, want, err_msg=key)   # bit-exact tables
        n += 1
    assert n > 400
    assert str(g["err|2d|conv"]) == "Unknown Convolution Method"


def test_reference_defects_are_reproduced():
    # 'y' == 't' in ConvOps_2d; 5x5x5 embedding at plane 1; Spatial x == y
    assert np.array_equal(O.build_kernel("2d", "y", 1), O.build_kernel("2d", "t", 1))
    k5 = O.build_kernel("2d", ("x", "y"), 2, 1.0, 4)
    assert k5.shape == (5, 5, 5) and np.count_nonzero(k5[1]) == 9 and np.count_nonzero(k5) == 9
    assert np.array_equal(O.build_kernel("spatial", "x", 1), O.build_kernel("spatial", "y", 1))
    assert O.build_kernel("1d", "x", 3) is None


@pytest.mark.parametrize("prefix,names", [
    ("2d", ["Dt", "Dx", "Dy", "Dtt", "Dxx", "I", "L3", "L5", "L7", "wave", "dense", "aniso"]),
    ("1d", ["Dt", "Dx", "Dxx", "I", "L3", "L5", "adv"]),
    ("0d", ["0_2", "1_2", "1_4", "2_2", "2_4", "2_6"]),
])
def test_direct_correlation(prefix, names):
    g = load_golden("conv")
    u = g[f"{prefix}|u"]
    for nm in names:
        assert_close_rel(O.correlate(u, g[f"{prefix}|{nm}|k"]), g[f"{prefix}|{nm}|y"],
                         what=f"{prefix}|{nm}")


def test_direct_correlation_permuted_view_and_valid():
    g = load_golden("conv")
    up = np.transpose(g["2d|perm|u"], (0, 3, 1, 2))
    assert_close_rel(O.correlate(up, g["2d|wave|k"]), g["2d|perm|y"], what="perm")
    for nm in ("Dx", "L3", "L5"):
        assert_close_rel(O.correlate(g["sp|u"], g[f"sp|{nm}|k"], pad=0), g[f"sp|{nm}|y"],
                         what=f"sp|{nm}")


def test_spectral_paths():
    g = load_golden("spectral")
    # Inverse paths on derivative stencils divide by ~1e-6 at the stencil's spectral zeros:
    # only the order of magnitude is reproducible (see make_golden.py); the well-conditioned
    # "wc" cases below pin the inverse arithmetic tightly.
    ILL = 0.5
    tol = 2e-5   # FFT round-off; the reference's own bound is 1e-4 abs (fft_conv_pytorch/utils.py:30-34)
    u, k = g["2d|u"], g["2d|k"]
    p3 = tuple(s // 2 for s in k.shape)
    assert_close_rel(O.fft_conv(u[:, None], k[None, None], padding=p3)[:, 0], g["2d|spec"], tol, "2d spec")
    assert_close_rel(O.fft_conv(u[:, None], k[None, None], padding=p3, inverse=True)[:, 0],
                     g["2d|spec_inv"], ILL, "2d spec inv")
    for corr in (0, 1):
        for sl in (0, 1):
            assert_close_rel(O.spectral_manual(u[:, None], k, "diff", bool(corr), bool(sl))[:, 0],
                             g[f"2d|diff|{corr}{sl}"], tol, f"2d diff {corr}{sl}")
            assert_close_rel(O.spectral_manual(u[:, None], k, "int", bool(corr), bool(sl))[:, 0],
                             g[f"2d|int|{corr}{sl}"], ILL, f"2d int {corr}{sl}")
    kc = g["2d|call_spec|k"]
    assert_close_rel(O.fft_conv(u[:, None], kc[None, None], padding=(1, 1, 1))[:, 0], g["2d|call_spec"], tol)
    # forward spectral == direct (relation R2)
    assert_close_rel(g["2d|spec"], O.correlate(u, k), tol, "R2")

    u1, k1 = g["1d|u"], g["1d|k"]
    assert_close_rel(O.fft_conv(u1[:, None], k1[None, None], padding=(1, 1))[:, 0], g["1d|spec"], tol)
    for corr in (0, 1):
        for sl in (0, 1):
            assert_close_rel(O.spectral_manual(u1[:, None], k1, "diff", bool(corr), bool(sl))[:, 0],
                             g[f"1d|diff|{corr}{sl}"], tol)
            assert_close_rel(O.spectral_manual(u1[:, None], k1, "int", bool(corr), bool(sl))[:, 0],
                             g[f"1d|int|{corr}{sl}"], ILL)
    u0, k0 = g["0d|u"], g["0d|k"]
    assert_close_rel(O.fft_conv(u0[:, None], k0[None, None], padding=2)[:, 0], g["0d|spec"], tol)
    for corr in (0, 1):
        for sl in (0, 1):
            assert_close_rel(O.spectral_manual(u0[:, None], k0, "diff", bool(corr), bool(sl))[:, 0],
                             g[f"0d|diff|{corr}{sl}"], tol)
            assert_close_rel(O.spectral_manual(u0[:, None], k0, "int", bool(corr), bool(sl),
                                               pad_field=False)[:, 0],
                             g[f"0d|int|{corr}{sl}"], ILL)
    us, ks = g["sp|u"], g["sp|k"]
    assert_close_rel(O.fft_conv(us, ks[None, None], padding=(1, 1)), g["sp|spec"], tol)
    assert_close_rel(O.fft_conv(us, ks[None, None], padding=(1, 1), inverse=True), g["sp|spec_inv"], ILL)
    assert_close_rel(O.spectral_manual(us, ks, "diff", True, True), g["sp|diff|11"], tol)
    assert_close_rel(O.spectral_manual(us, ks, "int", False, False), g["sp|int|00"], ILL)
    WC = 1e-4
    k = g["2d|wc|k"]
    assert_close_rel(O.fft_conv(u[:, None], k[None, None], padding=(1, 1, 1), inverse=True)[:, 0],
                     g["2d|wc|spec_inv"], WC, "2d wc spec_inv")
    assert_close_rel(O.spectral_manual(u[:, None], k, "int", False, False)[:, 0], g["2d|wc|int|00"], WC)
    assert_close_rel(O.spectral_manual(u[:, None], k, "int", True, True)[:, 0], g["2d|wc|int|11"], WC)
    d = O.spectral_manual(u[:, None], k, "diff", False, False)
    assert_close_rel(O.spectral_manual(d, k, "int", False, False)[:, 0], g["2d|wc|roundtrip"], WC)
    k = g["1d|wc|k"]
    assert_close_rel(O.spectral_manual(u1[:, None], k, "int", False, True)[:, 0], g["1d|wc|int|01"], WC)
    assert_close_rel(O.spectral_manual(u1[:, None], k, "int", True, False)[:, 0], g["1d|wc|int|10"], WC)
    k = g["0d|wc|k"]
    assert_close_rel(O.spectral_manual(u0[:, None], k, "int", False, True, pad_field=False)[:, 0],
                     g["0d|wc|int|01"], WC)
    assert_close_rel(O.spectral_manual(u0[:, None], k, "int", True, False, pad_field=False)[:, 0],
                     g["0d|wc|int|10"], WC)
    k = g["sp|wc|k"]
    assert_close_rel(O.fft_conv(us, k[None, None], padding=(1, 1), inverse=True), g["sp|wc|spec_inv"], WC)
    assert_close_rel(O.spectral_manual(us, k, "int", False, False), g["sp|wc|int|00"], WC)
    assert_close_rel(O.spectral_manual(us, k, "int", True, True), g["sp|wc|int|11"], WC)
    assert_close_rel(O.fft_conv(g["fc|sig"], g["fc|ker"], padding=(1.5,), bias=g["fc|bias"]), g["fc|same"], tol)
    WC = 1e-4
    k = g["2d|wc|k"]
    assert_close_rel(O.fft_conv(u[:, None], k[None, None], padding=(1, 1, 1), inverse=True)[:, 0],
                     g["2d|wc|spec_inv"], WC, "2d wc spec_inv")
    assert_close_rel(O.spectral_manual(u[:, None], k, "int", False, False)[:, 0], g["2d|wc|int|00"], WC)
    assert_close_rel(O.spectral_manual(u[:, None], k, "int", True, True)[:, 0], g["2d|wc|int|11"], WC)
    d = O.spectral_manual(u[:, None], k, "diff", False, False)
    assert_close_rel(O.spectral_manual(d, k, "int", False, False)[:, 0], g["2d|wc|roundtrip"], WC)
    k = g["1d|wc|k"]
    assert_close_rel(O.spectral_manual(u1[:, None], k, "int", False, True)[:, 0], g["1d|wc|int|01"], WC)
    assert_close_rel(O.spectral_manual(u1[:, None], k, "int", True, False)[:, 0], g["1d|wc|int|10"], WC)
    k = g["0d|wc|k"]
    assert_close_rel(O.spectral_manual(u0[:, None], k, "int", False, True, pad_field=False)[:, 0],
                     g["0d|wc|int|01"], WC)
    assert_close_rel(O.spectral_manual(u0[:, None], k, "int", True, False, pad_field=False)[:, 0],
                     g["0d|wc|int|10"], WC)
    k = g["sp|wc|k"]
    assert_close_rel(O.fft_conv(us, k[None, None], padding=(1, 1), inverse=True), g["sp|wc|spec_inv"], WC)
    assert_close_rel(O.spectral_manual(us, k, "int", False, False), g["sp|wc|int|00"], WC)
    assert_close_rel(O.spectral_manual(us, k, "int", True, True), g["sp|wc|int|11"], WC)
    assert_close_rel(O.fft_conv(g["fc|sig"], g["fc|ker"], padding=2, stride=2), g["fc|stride"], tol)


def test_boundary_manager_padding_bit_exact():
    g = load_golden("bc")
    types = [str(t) for t in g["types"]]
    x4, x2 = g["x4"], g["x2"]
    for ks in (3, 5, (3, 5)):
        for ty in types:
            t = {s: ty for s in ("left", "right", "top", "bottom")}
            v = {s: 1.25 for s in t}
            np.testing.assert_array_equal(O.bc_pad(x4, ks, t, v), g[f"all|{ks}|{ty}|x4"], err_msg=f"{ks} {ty}")
            np.testing.assert_array_equal(O.bc_pad(x2, ks, t, v), g[f"all|{ks}|{ty}|x2"])
    for a in types:
        for c in types:
            t = dict(left=a, right=c, top=c, bottom=a)
            v = dict(left=0.5, right=-0.5, top=2.0, bottom=-2.0)
            np.testing.assert_array_equal(O.bc_pad(x4, 5, t, v), g[f"mix|{a}|{c}"], err_msg=f"{a} {c}")
    sym = {s: "symmetric" for s in ("left", "right", "top", "bottom")}
    assert_close_rel(O.bc_apply_convolution(x2, g["conv|k"], sym), g["conv|sym|x2"])
    assert_close_rel(O.bc_apply_convolution(x2, g["conv|k"]), g["conv|per|x2"])


def test_vector_operators():
    g = load_golden("vector")
    a, b = g["a"], g["b"]
    gx = O.build_kernel("spatial", "x", 1, 1.7)
    lp = O.build_kernel("spatial", ("x", "y"), 2, 0.3)
    for bc in ("periodic", "neumann", "symmetric", "dirichlet", "free_slip"):
        S = lambda k, x: O.spatial_op(k, x, bc)
        assert_close_rel(S(gx, a) + S(gx, b), g[f"div|{bc}"], what=f"div {bc}")
        assert_close_rel(S(gx, b) - S(gx, a), g[f"curl|{bc}"], 1e-5, f"curl {bc}")
        assert_close_rel(S(lp, a), g[f"lap_s|{bc}"], what=f"lap {bc}")
        assert_close_rel(np.concatenate([S(lp, a), S(lp, b)], 1), g[f"lap_v|{bc}"])
    l5 = O.build_kernel("spatial", ("x", "y"), 2, 1.0, 4)
    assert_close_rel(O.spatial_op(l5, a, "periodic", 4), g["lap_s5|periodic"])
    g1 = O.build_kernel("spatial", "x", 1, 1.0)
    P = lambda x, bc="periodic": O.spatial_op(g1, x, bc)
    assert_close_rel(np.concatenate([P(a), P(a)], 1), g["grad|periodic"])
    assert_close_rel(np.concatenate([P(a, "neumann"), P(b, "neumann")], 1), g["grad2|neumann"])
    assert_close_rel(P(a) ** 2 + P(b) ** 2 + 2 * P(a) * P(b), g["vgrad|periodic"])
    for cls in ("Divergence", "Gradient", "Curl", "Laplace"):
        assert str(g[f"v2|{cls}|raises"]).startswith("ValueError: Unknown Convolution Method")
    p, q = g["v2|p"], g["v2|q"]
    kx, ky, kl = g["v2|gx"], g["v2|gy"], g["v2|lp"]
    assert_close_rel(O.vec2d_divergence(kx, ky, p, q), g["v2|div"])
    assert_close_rel(O.vec2d_gradient(kx, ky, p, q), g["v2|grad"])
    assert_close_rel(O.vec2d_curl(kx, ky, p, q), g["v2|curl"])
    assert_close_rel(O.vec2d_laplace(kl, p, q), g["v2|lap"])
    assert_close_rel(p * q + q * p, g["v2|dot"])
    assert_close_rel(p * q + q * q, g["v2|cross"])     # plus sign, as shipped


def test_residual_assemblies():
    from oracle import synth
    g = load_golden("residuals")
    c = synth.BURGERS
    assert_close_rel(O.residual_burgers(g["burgers|u"], c["dx"], c["dt"], c["nu"]), g["burgers|res"])
    uw = g["wave|u"][:, 0]
    np.testing.assert_array_equal(
        O.build_kernel("2d", "t", 2) - np.float32(0.25) * O.build_kernel("2d", ("x", "y"), 2), g["wave|k"])
    full = O.correlate(uw, g["wave|k"])
    assert_close_rel(full, g["wave|res_full"])
    assert_close_rel(O.interior(full, 3), g["wave|res"])
    v = g["ns|vars"]
    n = synth.NS
    assert_close_rel(O.residual_ns_continuity(v[:, 0], v[:, 1], n["dx"], n["dy"]), g["ns|cont"])
    mom = O.residual_ns_momentum(v[:, 0], v[:, 1], v[:, 2], n["dx"], n["dy"], n["dt"], n["nu"])
    assert_close_rel(mom, g["ns|mom"])
    assert_close_rel(mom, g["ns|mom_class"])
    m = g["mhd|vars"]
    f = [m[:, j] for j in range(6)]
    for eq in ("continuity", "momentum", "energy", "induction", "gauss"):
        assert_close_rel(O.residual_mhd(eq, *f), g[f"mhd|{eq}"], what=eq)
    assert_close_rel(O.residual_mhd("energy", *f), g["mhd|energy_class"])


def test_conformal_helpers_against_inline_twins():
    g = load_golden("cp")
    cal, val = g["cal"], g["val"]
    mod = O.modulation_func(cal, np.zeros_like(cal))
    np.testing.assert_array_equal(mod, g["modulation"])
    np.testing.assert_array_equal(O.modulation_func(np.zeros_like(cal), cal), g["modulation"])  # symmetric
    np.testing.assert_array_equal(O.ncf_metric_joint(cal, np.zeros_like(cal), mod), g["joint_scores"])
    for q in (1.5, 2.5, 3.5):
        sets = [-q * mod, q * mod]
        np.testing.assert_array_equal(O.filter_sims_joint(sets, val), g[f"joint_mask|{q}"])
        assert O.emp_cov_joint(sets, val) == float(g[f"joint_cov|{q}"])
    for within in (False, True):
        np.testing.assert_array_equal(O.filter_sims_within_bounds(-mod, mod, val, 0.5, within),
                                      g[f"filter_bounds|{int(within)}"])


def test_calibrate_rank_formula_matches_numpy():
    # 'higher' selects sorted[ceil((rows-1)*q)] with q = ceil((n+1)(1-alpha))/n in float64
    rng = np.random.default_rng(7)
    for n in (19, 20, 37, 100, 128, 500, 512, 1000, 2048, 4096):
        s = rng.standard_normal((n, 5)).astype(np.float32)
        for alpha in np.arange(0.05, 0.95 + 0.1, 0.1):
            try:
                want = O.calibrate(s, n, alpha)
            except ValueError:
                with pytest.raises(ValueError):
                    O.calibrate_rank(n, n, alpha)
                continue
            k = O.calibrate_rank(n, n, alpha)
            np.testing.assert_array_equal(np.sort(s, axis=0)[k], want)
            np.testing.assert_array_equal(O.select_rank(s, k), want)
    with pytest.raises(ValueError):
        O.calibrate(np.zeros(5, np.float32), 5, 0.05)     # q > 1 for tiny n
    s = rng.standard_normal((50, 3)).astype(np.float32)
    s[7, 1] = np.nan
    out = O.select_rank(s, 40)
    assert np.isnan(out[1]) and np.isnan(O.calibrate(s, 50, 0.2)[1]) and not np.isnan(out[0])


def test_float_order_key_is_monotone():
    x = np.array([-np.inf, -3.5, -0.0, 0.0, 1e-30, 2.0, np.inf], dtype=np.float32)
    k = O.float_order_key(x)
    assert np.all(np.diff(k.astype(np.int64)) >= 0)
    assert O.float_order_key(np.float32(np.nan)) > k[-1]


def test_torch_port_matches_reference():
    """The timed CPU baseline (oracle/torch_port.py) reproduces the reference's outputs."""
    import torch
    from oracle import synth, torch_port as TP
    g = load_golden("residuals")
    n, c = synth.NS, synth.BURGERS
    assert_close_rel(TP.ns_momentum(torch.from_numpy(g["ns|vars"]), n["dx"], n["dy"], n["dt"], n["nu"],
                                    boundary=True).numpy(), g["ns|mom"], 1e-6)
    assert_close_rel(TP.wave(torch.from_numpy(g["wave|u"][:, 0]), 0.25).numpy(), g["wave|res"], 1e-6)
    assert_close_rel(TP.burgers(torch.from_numpy(g["burgers|u"]), c["dx"], c["dt"], c["nu"], True).numpy(),
                     g["burgers|res"], 1e-6)
    assert_close_rel(TP.mhd_momentum(torch.from_numpy(g["mhd|vars"]), boundary=True).numpy(), g["mhd|momentum"], 1e-6)


def test_jorek_residuals_against_reference_golden():
    """tests/golden/jorek.npz = the JOREK reduced-MHD residuals composed from the UNMODIFIED reference's
    operators exactly as Marginal/JOREK_residuals_CP.py:203-242 does (oracle/make_golden.py gen_jorek)."""
    g = load_golden("jorek")
    o = O.ops_jorek()
    np.testing.assert_array_equal(o["D_RR"], g["k|D_RR"])          # scale = the re-bound gamma = 5/3
    np.testing.assert_array_equal(o["D_ZZ"], g["k|D_ZZ"])
    v = g["vars"]
    for eq, coef in (("continuity", 3.4), ("temperature", 2.25 * 1e-7)):
        got = O.residual_jorek(eq, v[:, 0], v[:, 1], v[:, 2], g["R"], coef, ops=o)
        assert_close_rel(got, g[eq], 1e-5, eq)


def test_reference_arm_runs_the_reference_own_code_and_equals_the_port():
    """bench.py's CPU arm (oracle/ref_arm.py): the operators come from the reference's own modules and the
    residual functions are the scripts' own ``def residual*`` statements, lifted from the syntax tree -- from the
    checkout in the build container, from the byte-for-byte snapshot (oracle/_ref/reference_hotpath.zip, packed by
    oracle/make_ref_snapshot.py at build time) elsewhere.  They must (a) not resolve to the drop-in mirror even when
    it is installed under the same module names, (b) equal oracle/torch_port.py bit for bit on every workload."""
    import torch
    from oracle import make_ref_snapshot, ref_arm, ref_shim, torch_port as TP
    if not ref_arm.available():
        pytest.skip("neither /root/reference nor oracle/_ref/reference_hotpath.zip is present")
    if os.path.exists(make_ref_snapshot.ARCHIVE):
        manifest = make_ref_snapshot.verify()
        assert set(manifest["files"]) == set(make_ref_snapshot.FILES)
    import cp_pre_b200
    cp_pre_b200.install()
    import Utils.ConvOps_2d as mirror
    cases = [("ns", "momentum", 3, (2, 6, 16, 16)), ("wave-joint", "", 1, (2, 6, 16, 16)), ("burgers", "", 1, (3, 8, 32))]
    cases += [("mhd", eq, 6, (2, 6, 16, 16)) for eq in ("continuity", "momentum", "energy", "induction", "gauss")]
    for wl, eq, nf, shape in cases:
        v = torch.from_numpy(synth.fields(3, nf, shape, positive=(0, 3) if wl == "mhd" else ()))
        v = v[:, 0] if nf == 1 else v
        got = ref_arm.residual_function(wl, eq)(v)
        if wl == "ns":
            want = TP.ns_momentum(v, synth.NS["dx"], synth.NS["dy"], synth.NS["dt"], synth.NS["nu"])
        elif wl == "mhd":
            want = TP.mhd(eq, v, synth.MHD["gamma"])
        elif wl == "burgers":
            want = TP.burgers(v, synth.BURGERS["dx"], synth.BURGERS["dt"], synth.BURGERS["nu"])
        else:
            want = TP.wave(v, synth.WAVE["c2"])
        assert torch.equal(got, want), (wl, eq)
    ns = ref_arm._ns
    assert ns.ConvOps_2d.ConvOperator is not mirror.ConvOperator
    assert ns.PRE_estimations.ConvOperator is ns.ConvOps_2d.ConvOperator
    root = os.path.abspath(ref_shim._root())
    assert os.path.abspath(ns.ConvOps_2d.__file__).startswith(root)
    import Utils.ConvOps_2d as again            # the mirror is back in place afterwards
    assert again is mirror
    adv = ref_arm.residual_function("advection")
    assert tuple(adv(torch.randn(4, 10, 197)).shape) == (4, 8, 195)


def test_periodic_boundary_residual_against_the_script_function():
    """[ref: Joint/NS_Residuals_CP.py:483-494] -- the oracle's restatement against the script's own function
    (lifted from its syntax tree like the residual functions of the CPU arm), all four walls, bit for bit."""
    import torch
    from oracle import ref_arm, ref_shim
    if not ref_arm.available():
        pytest.skip("no reference checkout / snapshot")
    env = {"torch": torch, "np": np, "dx": 0.01}
    ref_arm._lift("Joint/NS_Residuals_CP.py", env, {"periodic_bc_residual"}, operators=False)
    u = np.random.default_rng(3).standard_normal((3, 5, 6, 7)).astype(np.float32)
    for wall in ("top", "bottom", "left", "right"):
        want = env["periodic_bc_residual"](torch.from_numpy(u), wall).numpy()
        got = O.periodic_bc_residual(u, wall, 0.01)
        assert got.dtype == np.float32 and np.array_equal(got, want), wall
