"""Generate tests/golden/*.npz by RUNNING THE UNMODIFIED REFERENCE (build container only).

    python -m oracle.make_golden            # from the repo root

The reference ships no golden vectors or known-answer tests for this path,
so the fixtures are outputs of the reference itself on small seeded inputs.  Inputs are
stored next to the outputs, so nothing on the GPU box needs ``/root/reference``.
Provenance of every array: the ``ref.*`` call on the line that produces it.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_shim  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
NONE = np.zeros((0,), dtype=np.float32)  # marks "constructor left no .kernel"


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def N(t):
    return t.detach().cpu().numpy().copy()


def kernel_of(op):
    return N(op.kernel) if hasattr(op, "kernel") else NONE


def gen_kernels(ref):
    out = {}
    doms2 = ["t", "x", "y", ("x", "y"), ("x", "y", "t"), "xy", "z", None]
    for dom in doms2:
        for order in (0, 1, 2, 3):
            for ty in (2, 4, 6):
                for scale in (1.0, -0.37):
                    key = f"2d|{dom!r}|{order}|{ty}|{scale}"
                    out[key] = kernel_of(ref.ConvOps_2d.ConvOperator(dom, order, scale, ty))
    for dom in ["t", "x", ("x", "t"), ("x", "y"), "y", None]:
        for order in (0, 1, 2, 3):
            for ty in (2, 4, 6):
                for scale in (1.0, 2.5):
                    key = f"1d|{dom!r}|{order}|{ty}|{scale}"
                    out[key] = kernel_of(ref.ConvOps_1d.ConvOperator(dom, order, scale, ty))
    for order in (0, 1, 2, 3, None):
        for ty in (2, 4, 6):
            for scale in (1.0, 0.5):
                key = f"0d|None|{order}|{ty}|{scale}"
                out[key] = kernel_of(ref.ConvOps_0d.ConvOperator(order, scale, ty))
    for dom in ["x", "y", ("x", "y"), "t", None]:
        for order in (0, 1, 2):
            for ty in (2, 4, 6):
                for scale in (1.0, 3.0):
                    key = f"spatial|{dom!r}|{order}|{ty}|{scale}"
                    op = ref.ConvOps_Spatial.ConvOperator(dom, order, scale, ty, device="cpu")
                    out[key] = kernel_of(op)
    # constructor error behaviour
    try:
        ref.ConvOps_2d.ConvOperator("t", 1, conv="bogus")
        out["err|2d|conv"] = np.array(0)
    except ValueError as e:
        out["err|2d|conv"] = np.array(str(e))
    np.savez_compressed(os.path.join(OUT, "kernels.npz"), **out)
    return len(out)


def gen_conv(ref):
    rng = np.random.default_rng(101)
    out = {}
    C2 = ref.ConvOps_2d.ConvOperator
    u = rng.standard_normal((2, 6, 7, 8), dtype=np.float32)
    out["2d|u"] = u
    ops = {"Dt": C2("t", 1), "Dx": C2("x", 1), "Dy": C2("y", 1), "Dtt": C2("t", 2),
           "Dxx": C2("x", 2), "I": C2("x", 0), "L3": C2(("x", "y"), 2),
           "L5": C2(("x", "y"), 2, taylor_order=4), "L7": C2(("x", "y"), 2, taylor_order=6)}
    for name, op in ops.items():
        out[f"2d|{name}|k"] = N(op.kernel)
        out[f"2d|{name}|y"] = N(op(T(u)))
    W = C2()
    W.kernel = ops["Dtt"].kernel - 0.25 * ops["L3"].kernel      # additive composition
    out["2d|wave|k"] = N(W.kernel)
    out["2d|wave|y"] = N(W(T(u)))
    # a dense random 3x3x3 and an anisotropic (3,5,7) kernel through convolution(field, kernel)
    kd = rng.standard_normal((3, 3, 3), dtype=np.float32)
    ka = rng.standard_normal((3, 5, 7), dtype=np.float32)
    out["2d|dense|k"], out["2d|dense|y"] = kd, N(C2("t", 1).convolution(T(u), T(kd)))
    out["2d|aniso|k"], out["2d|aniso|y"] = ka, N(C2("t", 1).convolution(T(u), T(ka)))
    # permuted (t-innermost) input view, SURVEY fact 6
    up = rng.standard_normal((2, 7, 8, 6), dtype=np.float32)           # [BS,Nx,Ny,Nt]
    out["2d|perm|u"] = up
    out["2d|perm|y"] = N(W(T(up).permute(0, 3, 1, 2)))

    C1 = ref.ConvOps_1d.ConvOperator
    u1 = rng.standard_normal((3, 9, 10), dtype=np.float32)
    out["1d|u"] = u1
    ops1 = {"Dt": C1("t", 1), "Dx": C1("x", 1), "Dxx": C1("x", 2), "I": C1("t", 0),
            "L3": C1(("x", "t"), 2), "L5": C1(("x", "t"), 2, taylor_order=4)}
    for name, op in ops1.items():
        out[f"1d|{name}|k"] = N(op.kernel)
        out[f"1d|{name}|y"] = N(op(T(u1)))
    A = C1()
    A.kernel = ops1["Dt"].kernel + (1.0 * 2 * 0.005 / 0.01) * ops1["Dx"].kernel
    out["1d|adv|k"], out["1d|adv|y"] = N(A.kernel), N(A(T(u1)))

    C0 = ref.ConvOps_0d.ConvOperator
    u0 = rng.standard_normal((4, 12), dtype=np.float32)
    out["0d|u"] = u0
    for order, ty in ((0, 2), (1, 2), (1, 4), (2, 2), (2, 4), (2, 6)):
        op = C0(order, 1.5, ty)
        out[f"0d|{order}_{ty}|k"] = N(op.kernel)
        out[f"0d|{order}_{ty}|y"] = N(op(T(u0)))

    CS = ref.ConvOps_Spatial.ConvOperator
    us = rng.standard_normal((2, 1, 9, 10), dtype=np.float32)
    out["sp|u"] = us
    for name, op in {"Dx": CS("x", 1, 2.0, device="cpu"), "L3": CS(("x", "y"), 2, device="cpu"),
                     "L5": CS(("x", "y"), 2, 1.0, 4, device="cpu")}.items():
        out[f"sp|{name}|k"] = N(op.kernel)
        out[f"sp|{name}|y"] = N(op(T(us)))
    np.savez_compressed(os.path.join(OUT, "conv.npz"), **out)
    return len(out)


def gen_spectral(ref):
    rng = np.random.default_rng(202)
    out = {}
    C2, C1, C0 = (ref.ConvOps_2d.ConvOperator, ref.ConvOps_1d.ConvOperator,
                  ref.ConvOps_0d.ConvOperator)
    CS = ref.ConvOps_Spatial.ConvOperator
    u = rng.standard_normal((2, 6, 7, 9), dtype=np.float32)     # odd last dim -> even-pad path
    out["2d|u"] = u
    W = C2()
    W.kernel = C2("t", 2).kernel - 0.25 * C2(("x", "y"), 2).kernel
    out["2d|k"] = N(W.kernel)
    out["2d|spec"] = N(W.spectral_convolution(T(u)))
    out["2d|spec_inv"] = N(W.spectral_convolution(T(u), inverse=True))
    for corr in (False, True):
        for sl in (False, True):
            out[f"2d|diff|{int(corr)}{int(sl)}"] = N(W.differentiate(T(u), correlation=corr, slice_pad=sl))
            out[f"2d|int|{int(corr)}{int(sl)}"] = N(W.integrate(T(u), correlation=corr, slice_pad=sl))
    Ws = C2("t", 1, conv="spectral")
    out["2d|call_spec|k"] = N(Ws.kernel)
    out["2d|call_spec"] = N(Ws(T(u)))

    u1 = rng.standard_normal((3, 9, 10), dtype=np.float32)
    out["1d|u"] = u1
    A = C1()
    A.kernel = C1("t", 1).kernel + 0.7 * C1("x", 1).kernel
    out["1d|k"] = N(A.kernel)
    out["1d|spec"] = N(A.spectral_convolution(T(u1)))
    for corr in (False, True):
        for sl in (False, True):
            out[f"1d|diff|{int(corr)}{int(sl)}"] = N(A.differentiate(T(u1), correlation=corr, slice_pad=sl))
            out[f"1d|int|{int(corr)}{int(sl)}"] = N(A.integrate(T(u1), correlation=corr, slice_pad=sl))

    u0 = rng.standard_normal((4, 13), dtype=np.float32)
    out["0d|u"] = u0
    B = C0(2, 1.0, 4)
    out["0d|k"] = N(B.kernel)
    out["0d|spec"] = N(B.spectral_convolution(T(u0)))
    for corr in (False, True):
        for sl in (False, True):
            out[f"0d|diff|{int(corr)}{int(sl)}"] = N(B.differentiate(T(u0), correlation=corr, slice_pad=sl))
            out[f"0d|int|{int(corr)}{int(sl)}"] = N(B.integrate(T(u0), correlation=corr, slice_pad=sl))

    us = rng.standard_normal((2, 1, 8, 9), dtype=np.float32)
    out["sp|u"] = us
    S = CS(("x", "y"), 2, device="cpu")
    out["sp|k"] = N(S.kernel)
    out["sp|spec"] = N(S.spectral_convolution(T(us)))
    out["sp|spec_inv"] = N(S.spectral_convolution(T(us), inverse=True))
    out["sp|diff|11"] = N(S.differentiate(T(us), correlation=True, slice_pad=True))
    out["sp|int|00"] = N(S.integrate(T(us)))

    # Well-conditioned inverse cases.  Every derivative stencil has spectral zeros (DC at
    # least), so 1/(K^+1e-6) amplifies fp32 round-off by ~1e6 there and the reference's own
    # output is not reproducible to better than O(10 %) across FFT libraries.  Kernels of the
    # form 2*I + 0.1*Laplacian have |K^| in [1.2, 2]: the inverse path can be pinned tightly.
    Wc = C2()
    Wc.kernel = 2.0 * C2("x", 0).kernel + 0.1 * C2(("x", "y"), 2).kernel + 0.05 * C2("t", 2).kernel
    out["2d|wc|k"] = N(Wc.kernel)
    out["2d|wc|spec_inv"] = N(Wc.spectral_convolution(T(u), inverse=True))
    out["2d|wc|int|00"] = N(Wc.integrate(T(u)))
    out["2d|wc|int|11"] = N(Wc.integrate(T(u), correlation=True, slice_pad=True))
    out["2d|wc|roundtrip"] = N(Wc.integrate(Wc.differentiate(T(u), slice_pad=False), slice_pad=False))
    Ac = C1()
    Ac.kernel = 2.0 * C1("t", 0).kernel + 0.1 * C1(("x", "t"), 2).kernel
    out["1d|wc|k"] = N(Ac.kernel)
    out["1d|wc|int|01"] = N(Ac.integrate(T(u1)))
    out["1d|wc|int|10"] = N(Ac.integrate(T(u1), correlation=True, slice_pad=False))
    Bc = C0(0, 2.0)
    Bc.kernel = Bc.kernel + 0.1 * C0(2, 1.0, 2).kernel
    out["0d|wc|k"] = N(Bc.kernel)
    out["0d|wc|int|01"] = N(Bc.integrate(T(u0)))
    out["0d|wc|int|10"] = N(Bc.integrate(T(u0), correlation=True, slice_pad=False))
    Sc = CS("x", 0, 2.0, device="cpu")
    Sc.kernel = (Sc.kernel + 0.1 * S.kernel).detach()
    out["sp|wc|k"] = N(Sc.kernel)
    out["sp|wc|spec_inv"] = N(Sc.spectral_convolution(T(us), inverse=True))
    out["sp|wc|int|00"] = N(Sc.integrate(T(us)))
    out["sp|wc|int|11"] = N(Sc.integrate(T(us), correlation=True, slice_pad=True))

    # fft_conv proper: multi-channel broadcast, bias, stride, float padding
    sig = rng.standard_normal((2, 3, 11), dtype=np.float32)
    ker = rng.standard_normal((1, 3, 4), dtype=np.float32)       # even kernel
    bias = rng.standard_normal((3,), dtype=np.float32)
    out["fc|sig"], out["fc|ker"], out["fc|bias"] = sig, ker, bias
    out["fc|same"] = N(ref.fft_conv.fft_conv(T(sig), T(ker), bias=T(bias), padding="same"))
    out["fc|stride"] = N(ref.fft_conv.fft_conv(T(sig), T(ker), padding=2, stride=2))
    np.savez_compressed(os.path.join(OUT, "spectral.npz"), **out)
    return len(out)


def gen_bc(ref):
    rng = np.random.default_rng(303)
    out = {}
    BM = ref.boundary_conditions.BoundaryManager
    x4 = rng.standard_normal((2, 2, 6, 7), dtype=np.float32)
    x2 = rng.standard_normal((6, 7), dtype=np.float32)
    out["x4"], out["x2"] = x4, x2
    types = list(BM.SUPPORTED_TYPES)
    out["types"] = np.array(types)
    for ks in (3, 5, (3, 5)):
        for ty in types:
            b = BM(ks)
            b.set_all_boundaries(ty, 1.25)
            out[f"all|{ks}|{ty}|x4"] = N(b.pad_signal(T(x4)))
            out[f"all|{ks}|{ty}|x2"] = N(b.pad_signal(T(x2)))
    # every ordered (left,right) and (top,bottom) pair at kernel 5: exercises the
    # sequential-padding quirks (periodic right/bottom read the already padded tensor)
    for a in types:
        for c in types:
            b = BM(5)
            b.set_boundary_type("left", a, 0.5)
            b.set_boundary_type("right", c, -0.5)
            b.set_boundary_type("top", c, 2.0)
            b.set_boundary_type("bottom", a, -2.0)
            out[f"mix|{a}|{c}"] = N(b.pad_signal(T(x4)))
    k = rng.standard_normal((3, 3), dtype=np.float32)
    b = BM(3)
    b.set_all_boundaries("symmetric")
    out["conv|k"] = k
    out["conv|sym|x2"] = N(b.apply_convolution(T(x2), T(k)))
    b = BM(3)
    out["conv|per|x2"] = N(b.apply_convolution(T(x2), T(k)))
    try:
        BM(3).set_boundary_type("front", "periodic")
    except ValueError as e:
        out["err|side"] = np.array(str(e))
    try:
        BM(3).set_boundary_type("left", "robin")
    except ValueError as e:
        out["err|type"] = np.array(str(e))
    np.savez_compressed(os.path.join(OUT, "bc.npz"), **out)
    return len(out)


def gen_vector(ref):
    rng = np.random.default_rng(404)
    out = {}
    VS = ref.VectorConvOps_Spatial
    a = rng.standard_normal((2, 1, 8, 9), dtype=np.float32)
    b = rng.standard_normal((2, 1, 8, 9), dtype=np.float32)
    out["a"], out["b"] = a, b
    for bc in ("periodic", "neumann", "symmetric", "dirichlet", "free_slip"):
        out[f"div|{bc}"] = N(VS.Divergence(scale=1.7, boundary_cond=bc, device="cpu")(T(a), T(b)))
        out[f"curl|{bc}"] = N(VS.Curl(scale=1.7, boundary_cond=bc, device="cpu")(T(a), T(b)))
        out[f"lap_s|{bc}"] = N(VS.Laplace(scale=0.3, boundary_cond=bc, device="cpu")(T(a)))
        out[f"lap_v|{bc}"] = N(VS.Laplace(scale=0.3, boundary_cond=bc, scalar=False, device="cpu")(T(a), T(b)))
    out["lap_s5|periodic"] = N(VS.Laplace(taylor_order=4, device="cpu")(T(a)))
    # Gradient / Vector_Gradient hard-code cuda children (VectorConvOps_Spatial.py:38-39,147-148);
    # on this CPU box the swallowed .to() failure leaves an UNSCALED kernel, so scale=1 is the
    # only setting where CPU and GPU behaviour coincide.
    out["grad|periodic"] = N(VS.Gradient()(T(a)))
    out["grad2|neumann"] = N(VS.Gradient(boundary_cond="neumann")(T(a), T(b)))
    out["vgrad|periodic"] = N(VS.Vector_Gradient()(T(a), T(b)))
    v = np.concatenate([a, b], axis=1)
    w = rng.standard_normal((2, 2, 8, 9), dtype=np.float32)
    out["w"] = w
    out["dot"] = N(VS.dot(T(v), T(w)))
    out["cross"] = N(VS.cross(T(v), T(w)))
    out["vectorize"] = N(VS.vectorize(T(a), T(b)))
    # Utils/VectorConvOps.py: every constructor raises as shipped
    V2 = ref.VectorConvOps
    if isinstance(V2, Exception):
        out["v2|import_error"] = np.array(repr(V2))
    else:
        for cls in ("Divergence", "Gradient", "Curl", "Laplace"):
            try:
                getattr(V2, cls)()
                out[f"v2|{cls}|raises"] = np.array("")
            except Exception as e:
                out[f"v2|{cls}|raises"] = np.array(f"{type(e).__name__}: {e}")
        p = rng.standard_normal((2, 5, 6, 7), dtype=np.float32)
        q = rng.standard_normal((2, 5, 6, 7), dtype=np.float32)
        out["v2|p"], out["v2|q"] = p, q
        out["v2|dot"] = N(V2.dot((T(p), T(q)), (T(q), T(p))))
        out["v2|cross"] = N(V2.cross((T(p), T(q)), (T(q), T(q))))
        out["v2|vectorize"] = N(V2.vectorize(T(p), T(q)))
        # intended composition (Tests/NS_vector_convops.py:154-176): children by keyword
        C2 = ref.ConvOps_2d.ConvOperator
        gx, gy, lp = C2("x", 1, 1.3), C2("y", 1, 1.3), C2(("x", "y"), 2, 0.8)
        out["v2|gx"], out["v2|gy"], out["v2|lp"] = N(gx.kernel), N(gy.kernel), N(lp.kernel)
        out["v2|div"] = N(gx(T(p)) + gy(T(q)))
        out["v2|grad"] = N(torch.stack((gx(T(p)), gy(T(q)))))
        out["v2|curl"] = N(gx(T(q)) - gy(T(p)))
        out["v2|lap"] = N(torch.stack((lp(T(p)), lp(T(q)))))
    np.savez_compressed(os.path.join(OUT, "vector.npz"), **out)
    return len(out)


def gen_residuals(ref):
    """Residual formulas transcribed from the scripts, evaluated with the reference's own
    ConvOperator objects (the scripts themselves need missing data/checkpoints)."""
    from oracle import synth
    out = {}
    C2, C1 = ref.ConvOps_2d.ConvOperator, ref.ConvOps_1d.ConvOperator
    # --- Burgers [ref: Joint/Burgers_Residuals_CP.py:172-187]
    u1 = synth.fields(11, 1, (3, 10, 16))[:, 0]
    D_t, D_x, D_xx = C1("t", 1), C1("x", 1), C1("x", 2)
    dx, dt, nu = (torch.tensor(synth.BURGERS[k], dtype=torch.float32) for k in ("dx", "dt", "nu"))
    uu = T(u1)
    out["burgers|u"] = u1
    out["burgers|res"] = N(dx * D_t(uu) + dt * uu * D_x(uu) - nu * D_xx(uu) * (2 * dt / dx))
    # --- Wave through the class form [ref: Other_UQ/Evaluation/PRE_estimations.py:5-22]
    uw = synth.fields(12, 1, (2, 8, 12, 12))
    pw = ref.PRE_estimations.PRE_Wave(dt=0.5, dx=1.0, c=1.0)
    out["wave|u"] = uw
    out["wave|k"] = N(pw.D.kernel)
    out["wave|res_full"] = N(pw.residual(T(uw), boundary=True))
    out["wave|res"] = N(pw.residual(T(uw)))
    # --- NS [ref: Joint/NS_Residuals_CP.py:214-240]
    vn = synth.fields(13, 4, (2, 6, 12, 16))
    out["ns|vars"] = vn
    c = synth.NS
    dx, dy, dt, nu = np.asarray(c["dx"]), np.asarray(c["dy"]), c["dt"], c["nu"]
    D_t, D_x, D_y, D_xx_yy = C2("t", 1), C2("x", 1), C2("y", 1), C2(("x", "y"), 2)
    vars_ = T(vn)
    u, v, p = vars_[:, 0], vars_[:, 1], vars_[:, 2]
    out["ns|cont"] = N(D_x(u) + (dx / dy) * D_y(v))
    res_x = D_t(u)*dx*dy + u*D_x(u)*dt*dy + v*D_y(u)*dt*dx - nu*D_xx_yy(u)*dt + D_x(p)*dt*dy
    res_y = D_t(v)*dx*dy + u*D_x(v)*dt*dx + v*D_y(v)*dt*dy - nu*D_xx_yy(v)*dt + D_y(p)*dt*dx
    out["ns|mom"] = N(res_x + res_y)
    pn = ref.PRE_estimations.PRE_NS(dt, float(dx), float(dy))
    out["ns|mom_class"] = N(pn.residual(vars_, boundary=True))
    # --- MHD [ref: Joint/MHD_Residuals_CP.py:217-278]
    vm = synth.fields(14, 6, (2, 6, 12, 16), positive=(0, 3))
    out["mhd|vars"] = vm
    vars_ = T(vm)
    gamma = 5 / 3
    rho, u, v, p, Bx, By = (vars_[:, j] for j in range(6))
    out["mhd|continuity"] = N(D_t(rho) + u*D_x(rho) + rho*D_x(u) + v*D_y(rho) + rho*D_y(v))
    res_x = D_t(u) + u*D_x(u) + (1/rho)*D_x(p) - 2*(Bx/rho)*D_x(Bx) + v*D_y(u) - (By/rho)*D_y(Bx) - (Bx/rho)*D_y(By)
    res_y = D_t(v) + u*D_x(v) + (1/rho)*D_y(p) - 2*(By/rho)*D_y(By) + v*D_y(v) - (By/rho)*D_x(Bx) - (Bx/rho)*D_x(By)
    out["mhd|momentum"] = N(res_x + res_y)
    p_gas = p - 0.5*(Bx**2 + By**2)
    out["mhd|energy"] = N(D_t(rho) + u*D_x(p) + v*D_y(p) + (gamma-2)*(u*Bx+v*By)*(D_x(Bx) + D_y(By))
                          + (gamma*p_gas+By**2)*D_x(u) + (gamma*p_gas+Bx**2)*D_y(v) - Bx*By*(D_y(u) + D_x(v)))
    pm = ref.PRE_estimations.PRE_MHD(0.01, 1 / 128, 1 / 128)
    out["mhd|energy_class"] = N(pm.residual(vars_, boundary=True))
    res_x = D_t(Bx) - By*D_y(u) + Bx*D_y(v) - v*D_y(Bx) + u*D_y(By)
    res_y = D_t(By) + By*D_x(u) - Bx*D_x(v) - v*D_x(Bx) + u*D_x(By)
    out["mhd|induction"] = N(res_x + res_y)
    out["mhd|gauss"] = N(D_x(Bx) + D_y(By))
    np.savez_compressed(os.path.join(OUT, "residuals.npz"), **out)
    return len(out)


def gen_jorek(ref):
    """JOREK reduced-MHD residuals [ref: Marginal/JOREK_residuals_CP.py:184-242], transcribed and
    evaluated with the reference's own operators.  As in the script: alpha = beta = 1, but the
    second-order operators are built AFTER `gamma` was re-bound to tensor(5/3), so their kernels
    carry that factor; R = x_grid is a 1-D tensor and broadcasts along the LAST axis."""
    from oracle import synth
    out = {}
    C2 = ref.ConvOps_2d.ConvOperator
    vj = synth.fields(15, 3, (2, 6, 12, 16), positive=(0, 2))
    R = torch.linspace(1.5, 2.5, 16, dtype=torch.float32)
    D = torch.tensor(3.4, dtype=torch.float32)
    K = torch.tensor(2.25 * 1e-7, dtype=torch.float32)
    gamma = torch.tensor(5 / 3, dtype=torch.float32)
    D_t = C2(domain='t', order=1, scale=1)
    D_R = C2(domain='x', order=1, scale=1)
    D_Z = C2(domain='y', order=1, scale=1)
    D_RR = C2(domain='x', order=2, scale=gamma)
    D_ZZ = C2(domain='y', order=2, scale=gamma)
    vars_ = T(vj)
    rho, phi, Tt = vars_[:, 0], vars_[:, 1], vars_[:, 2]
    out["vars"], out["R"] = vj, N(R)
    out["k|D_RR"], out["k|D_ZZ"] = N(D_RR.kernel), N(D_ZZ.kernel)
    out["continuity"] = N(D_t(rho) - R*(D_R(rho)*D_Z(phi) - D_R(phi)*D_Z(rho)) - 2*rho*D_Z(phi)
                          - D*(D_RR(rho) + (1/R)*D_R(rho) + D_ZZ(rho)))
    out["temperature"] = N(Tt*D_t(rho) + rho*D_t(Tt) - rho*R*(D_R(Tt)*D_Z(phi) - D_R(phi)*D_Z(Tt))
                           + Tt*R*(D_R(rho)*D_Z(phi) - D_R(phi)*D_Z(rho)) + 2*gamma*rho*Tt*D_Z(phi)
                           + K * (D_RR(Tt) + (1/R)*D_R(Tt) + D_ZZ(Tt)))
    np.savez_compressed(os.path.join(OUT, "jorek.npz"), **out)
    return len(out)


def gen_cp(ref):
    """Conformal helpers: the module is absent from the reference; pins are its in-tree
    inline twins, evaluated verbatim with NumPy on seeded residual-like arrays."""
    rng = np.random.default_rng(505)
    out = {}
    cal = (0.05 * rng.standard_normal((40, 6, 9), dtype=np.float32)
           * (1 + rng.random((1, 6, 9), dtype=np.float32)))
    val = 0.05 * rng.standard_normal((25, 6, 9), dtype=np.float32)
    out["cal"], out["val"] = cal, val
    # [ref: Tests/test_advection_inv_sampling_marginal.py:428-431]
    modulation = np.std(cal, axis=0)
    out["modulation"] = modulation
    out["joint_scores"] = np.max(np.abs(cal) / modulation, axis=(1, 2))
    # [ref: Tests/test_advection_inv_sampling_marginal.py:465] joint coverage for given sets
    # [ref: Joint/Advection_Residuals_CP.py:327-329] filter_sims_joint
    for q in (1.5, 2.5, 3.5):
        sets = [-q * modulation, q * modulation]
        inside = ((val >= sets[0]).all(axis=(1, 2)) & (val <= sets[1]).all(axis=(1, 2)))
        out[f"joint_mask|{q}"] = inside
        out[f"joint_cov|{q}"] = np.array(inside.mean())
    # [ref: Active_Learning/Advection_AL_Marginal.py:169-197]
    lo, hi = -1.0 * modulation, 1.0 * modulation
    for within in (False, True):
        wb = ((val >= lo) & (val <= hi)) if within else ((val <= lo) | (val >= hi))
        out[f"filter_bounds|{int(within)}"] = wb.mean(axis=(1, 2)) >= 0.5
    np.savez_compressed(os.path.join(OUT, "cp.npz"), **out)
    return len(out)


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.manual_seed(0)
    ref = ref_shim.load()
    only = sys.argv[1:]            # e.g. `python -m oracle.make_golden gen_jorek` regenerates one file
    for fn in (gen_kernels, gen_conv, gen_spectral, gen_bc, gen_vector, gen_residuals, gen_jorek, gen_cp):
        if only and fn.__name__ not in only:
            continue
        n = fn(ref)
        print(f"{fn.__name__}: {n} arrays")


if __name__ == "__main__":
    main()
