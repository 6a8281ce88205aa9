"""NumPy restatement of the CP-PRE hot path (TEST INFRASTRUCTURE -- see oracle/__init__.py).

Citations ``[ref: path:line]`` are relative to the reference checkout
(gitvicky/CP-PRE @ 5b70067a).  Nothing here is imported by the product package.

All arithmetic on fields is done in float32, as the reference does (torch
default dtype float32); accumulation order is tap-by-tap in row-major tap order.
"""
from __future__ import annotations

import math
from fractions import Fraction as Fr

import numpy as np

F32 = np.float32

# --------------------------------------------------------------------------------------
# a1: stencil tables
# --------------------------------------------------------------------------------------
# 1-D central-difference coefficient rows, keyed (derivative order, Taylor order).
_LINE = {
    (0, None): [0, 1, 0],
    (1, 2): [-1, 0, 1],
    (1, 4): [Fr(1, 12), Fr(-2, 3), 0, Fr(2, 3), Fr(-1, 12)],
    (2, 2): [1, -2, 1],
    (2, 4): [Fr(-1, 12), Fr(4, 3), Fr(-5, 2), Fr(4, 3), Fr(-1, 12)],
    (2, 6): [Fr(1, 90), Fr(-3, 20), Fr(3, 2), Fr(-49, 18), Fr(3, 2), Fr(-3, 20), Fr(1, 90)],
}


def _f32(seq):
    # The reference writes Python float literals (e.g. -1/12) into a float32 tensor.
    return np.array([float(v) for v in seq], dtype=np.float64).astype(F32)


def _column_matrix(line):
    """k x k matrix with the 1-D stencil in the middle column
    [ref: Utils/ConvOps_2d.py:19-36 -- every dims==1 table has this shape]."""
    k = len(line)
    m = np.zeros((k, k), dtype=F32)
    m[:, k // 2] = _f32(line)
    return m


def _plus_matrix(line, centre):
    """Plus-shaped k x k matrix: the 1-D second-derivative row on the middle row and the
    middle column, with the centre entry as written in the reference
    [ref: Utils/ConvOps_2d.py:38-60]: -4 for Taylor 2 but the *1-D* centre (-5/2, -49/18)
    for Taylor 4/6."""
    k = len(line)
    m = np.zeros((k, k), dtype=F32)
    m[:, k // 2] = _f32(line)
    m[k // 2, :] = _f32(line)
    m[k // 2, k // 2] = F32(float(centre))
    return m


def get_stencil(module, dims, deriv_order, taylor_order=2):
    """Stencil lookup for module in {'0d','1d','2d','spatial'}.

    [ref: Utils/ConvOps_2d.py:17-63, Utils/ConvOps_1d.py:17-80, Utils/ConvOps_0d.py:21-47,
    Utils/ConvOps_Spatial.py:20-66].  Unknown combinations raise ValueError (TypeError for
    the malformed 3rd-order rows of the 1d module, ConvOps_1d.py:42,50).
    """
    if module == "0d":
        if deriv_order == 0:
            return _f32(_LINE[(0, None)])
        key = (deriv_order, taylor_order)
        if key in _LINE and key[0] in (1, 2):
            return _f32(_LINE[key])
        raise ValueError("Invalid stencil parameters")
    if dims == 1:
        if deriv_order == 0:
            return _column_matrix(_LINE[(0, None)])
        if deriv_order == 1 and taylor_order == 2:
            m = _column_matrix(_LINE[(1, 2)])
            if module == "spatial":  # [ref: Utils/ConvOps_Spatial.py:34-39] +-1/2
                m = (m * F32(0.5)).astype(F32)
            return m
        if deriv_order == 2 and taylor_order == 2:
            return _column_matrix(_LINE[(2, 2)])
        if module == "1d" and deriv_order == 3 and taylor_order in (2, 4):
            # [ref: Utils/ConvOps_1d.py:38-53] a missing comma makes the literal raise.
            raise TypeError("list indices must be integers or slices, not tuple")
    elif dims == 2:
        if deriv_order == 2 and taylor_order == 2:
            return _plus_matrix(_LINE[(2, 2)], -4)
        if deriv_order == 2 and taylor_order == 4:
            return _plus_matrix(_LINE[(2, 4)], Fr(-5, 2))
        if deriv_order == 2 and taylor_order == 6:
            return _plus_matrix(_LINE[(2, 6)], Fr(-49, 18))
    raise ValueError("Invalid stencil parameters")


def kernel_3d(stencil, axis):
    """[ref: Utils/ConvOps_2d.py:67-79] embed a k x k stencil at plane index **1** of a
    k^3 zero tensor along ``axis`` (index 1 even for k = 5, 7)."""
    k = stencil.shape[0]
    out = np.zeros((k, k, k), dtype=F32)
    if axis == 0:
        out[1, :, :] = stencil
    elif axis == 1:
        out[:, 1, :] = stencil
    elif axis == 2:
        out[:, :, 1] = stencil
    else:
        raise ValueError("Invalid axis. Must be either 0, 1 or 2")
    return out


_AXIS_2D = {"t": 2, "x": 0, "y": 1, ("x", "y"): 0, ("x", "y", "t"): 0}


def build_kernel(module, domain=None, order=None, scale=1.0, taylor_order=2):
    """What ``ConvOperator(domain, order, scale, taylor_order).kernel`` holds, or None
    when the reference constructor swallows a failure and leaves no ``.kernel``
    [ref: Utils/ConvOps_2d.py:96-125, ConvOps_1d.py:99-120, ConvOps_0d.py:64-76,
    ConvOps_Spatial.py:83-110]."""
    try:
        if module == "0d":
            st = get_stencil("0d", None, order, taylor_order)
            return (F32(scale) * st).astype(F32)
        dims = len(domain)
        st = get_stencil(module, dims, order, taylor_order)
        if module == "2d":
            if domain not in _AXIS_2D:
                raise ValueError("Invalid Domain")
            ker = kernel_3d(st, _AXIS_2D[domain])
        elif module == "1d":
            if domain == "t" or domain == ("x", "t"):
                ker = st
            elif domain == "x":
                ker = st.T.copy()
            else:
                raise ValueError("Invalid Domain")
        elif module == "spatial":
            if domain not in ("x", "y", ("x", "y")):
                raise ValueError("Invalid Domain")
            ker = st  # never transposed: 'x' and 'y' get the same kernel
        else:
            raise ValueError(module)
        return (F32(scale) * ker).astype(F32)
    except Exception:
        return None


# --------------------------------------------------------------------------------------
# a4: direct cross-correlation
# --------------------------------------------------------------------------------------
def correlate(field, kernel, pad=None):
    """Cross-correlation (no kernel flip), stride 1, zero padding ``pad`` per axis over the
    trailing ``kernel.ndim`` axes of ``field``; leading axes are batch.

    pad=None -> k//2 per axis ("same", [ref: Utils/ConvOps_2d.py:149, ConvOps_1d.py:150,
    ConvOps_0d.py:103]); pad=0 -> valid ([ref: Utils/ConvOps_Spatial.py:135]).
    out[..., i] = sum_a K[a] * in[..., i + a - pad]."""
    field = np.asarray(field, dtype=F32)
    kernel = np.asarray(kernel, dtype=F32)
    nd = kernel.ndim
    if pad is None:
        pad = tuple(k // 2 for k in kernel.shape)
    elif isinstance(pad, int):
        pad = (pad,) * nd
    lead = field.ndim - nd
    widths = [(0, 0)] * lead + [(p, p) for p in pad]
    fp = np.pad(field, widths)
    out_shape = field.shape[:lead] + tuple(
        field.shape[lead + d] + 2 * pad[d] - kernel.shape[d] + 1 for d in range(nd))
    out = np.zeros(out_shape, dtype=F32)
    for idx in np.ndindex(*kernel.shape):
        c = kernel[idx]
        if c == 0:
            continue
        sl = (slice(None),) * lead + tuple(
            slice(idx[d], idx[d] + out_shape[lead + d]) for d in range(nd))
        out += c * fp[sl]
    return out


# --------------------------------------------------------------------------------------
# a5/a6: spectral path
# --------------------------------------------------------------------------------------
def fft_conv(signal, kernel, padding=0, inverse=False, stride=1, bias=None):
    """[ref: Utils/fft_conv_pytorch/fft_conv.py:35-131] (dilation 1, constant padding).
    ``signal`` [B, C, *spatial], ``kernel`` [Co, Ci, *k] -- the reference never contracts
    channels (plain broadcasting product, :115), so neither does this."""
    signal = np.asarray(signal, dtype=F32)
    kernel = np.asarray(kernel, dtype=F32)
    n = signal.ndim - 2
    pads = (padding,) * n if isinstance(padding, (int, float)) else tuple(padding)
    strides = (stride,) * n if isinstance(stride, int) else tuple(stride)
    widths = [(0, 0), (0, 0)] + [(math.floor(p), math.ceil(p)) for p in pads]
    sig = np.pad(signal, widths)
    size = sig.shape
    if sig.shape[-1] % 2:
        sig = np.pad(sig, [(0, 0)] * (sig.ndim - 1) + [(0, 1)])
    kp = np.pad(kernel, [(0, 0), (0, 0)] + [(0, sig.shape[i] - kernel.shape[i])
                                           for i in range(2, sig.ndim)])
    axes = tuple(range(2, sig.ndim))
    sf = np.fft.rfftn(sig, axes=axes)
    kf = np.conj(np.fft.rfftn(kp, axes=axes))
    if inverse:
        kf = 1.0 / (kf + 1e-6)
    out = np.fft.irfftn(sf * kf, axes=axes)   # last axis is even by construction
    crop = (slice(None), slice(None)) + tuple(
        slice(0, size[i] - kernel.shape[i] + 1, strides[i - 2]) for i in range(2, sig.ndim))
    out = np.ascontiguousarray(out[crop]).astype(F32)
    if bias is not None:
        out = out + np.asarray(bias, dtype=F32).reshape((1, -1) + (1,) * n)
    return out


def spectral_manual(field, kernel, mode, correlation=False, slice_pad=True, eps=1e-6,
                    pad_field=True):
    """``differentiate`` (mode='diff') / ``integrate`` (mode='int') of the ConvOperators
    [ref: Utils/ConvOps_2d.py:179-284, ConvOps_1d.py:178-285, ConvOps_0d.py:133-237,
    ConvOps_Spatial.py:161-252].  ``field`` has a channel axis at 1.  pad_field=False
    reproduces ConvOps_0d.integrate, which discards its padding (0d:201-202)."""
    field = np.asarray(field, dtype=F32)
    kernel = np.asarray(kernel, dtype=F32)
    nd = kernel.ndim
    p = kernel.shape[-1] // 2
    if pad_field:
        pf = np.pad(field, [(0, 0), (0, 0)] + [(p, p)] * nd)
    else:
        pf = field
    axes = tuple(range(2, pf.ndim))
    kp = np.pad(kernel[None, None], [(0, 0), (0, 0)] + [(0, pf.shape[2 + d] - kernel.shape[d])
                                                        for d in range(nd)])
    ff = np.fft.rfftn(pf, axes=axes)
    kf = np.fft.rfftn(kp, axes=axes)
    if mode == "int":
        kf = 1.0 / (kf + eps)
    if correlation:
        kf = np.conj(kf)
    # no output size given, exactly like the reference: an odd padded last axis comes back
    # one sample shorter (2*(m-1)) -- reproduced, not corrected
    out = np.fft.irfftn(ff * kf, axes=axes)
    if slice_pad:
        crop = (slice(None), slice(None)) + tuple(
            slice(0, pf.shape[2 + d] - kernel.shape[d] + 1) for d in range(nd))
        out = np.ascontiguousarray(out[crop])
    return out.astype(F32)


# --------------------------------------------------------------------------------------
# a11: BoundaryManager padding
# --------------------------------------------------------------------------------------
_BC_TYPES = ("dirichlet", "neumann", "periodic", "symmetric", "free_slip", "outflow")


def _pad_one_side(x, axis, side, width, bc_type, value):
    """One sequential step of [ref: Utils/boundary_conditions.py:103-176]."""
    if width <= 0:
        return x
    n = x.shape[axis]
    take = lambda idx: np.take(x, idx, axis=axis)
    lo = side in ("left", "top")
    if bc_type == "dirichlet":
        shp = list(x.shape)
        shp[axis] = width
        ghost = np.full(shp, value, dtype=x.dtype)
    elif bc_type in ("neumann", "outflow"):
        ghost = take([0] * width if lo else [n - 1] * width)
    elif bc_type == "periodic":
        # low side: last `width` entries; high side: FIRST `width` entries of the tensor
        # as it is *now* (already low-padded) [ref: boundary_conditions.py:132-135,170-173]
        ghost = take(list(range(n - width, n)) if lo else list(range(width)))
    elif bc_type == "symmetric":
        if width >= n:
            raise RuntimeError("reflect padding must be smaller than the dimension")
        ghost = take(list(range(width, 0, -1)) if lo else [n - 2 - k for k in range(width)])
    else:  # free_slip: no branch in the reference -> nothing is padded
        return x
    return np.concatenate([ghost, x] if lo else [x, ghost], axis=axis)


def bc_pad(signal, kernel_size, types=None, values=None):
    """``BoundaryManager(kernel_size).pad_signal(signal)`` with per-side types/values.
    Sides are processed left, right (last axis) then top, bottom (second-to-last axis)."""
    kh, kw = (kernel_size, kernel_size) if isinstance(kernel_size, int) else kernel_size
    types = dict(left="periodic", right="periodic", top="periodic", bottom="periodic") \
        if types is None else types
    values = dict(left=0.0, right=0.0, top=0.0, bottom=0.0) if values is None else values
    x = np.asarray(signal, dtype=F32)
    for side, axis, width in (("left", -1, kw // 2), ("right", -1, kw // 2),
                              ("top", -2, kh // 2), ("bottom", -2, kh // 2)):
        ax = x.ndim + axis
        x = _pad_one_side(x, ax, side, width, types[side], F32(values[side]))
    return x


def bc_apply_convolution(signal, kernel, types=None, values=None):
    """[ref: Utils/boundary_conditions.py:184-211] pad + valid conv2d."""
    padded = bc_pad(signal, tuple(kernel.shape), types, values)
    return correlate(padded, kernel, pad=0)


# --------------------------------------------------------------------------------------
# a9/a10: vector operators
# --------------------------------------------------------------------------------------
def vec2d_divergence(kx, ky, a, b):
    """[ref: Utils/VectorConvOps.py:29-39] with children built by keyword (the shipped
    constructors raise -- SURVEY fact 8)."""
    return correlate(a, kx) + correlate(b, ky)


def vec2d_gradient(kx, ky, a, b=None):
    b = a if b is None else b
    return np.stack((correlate(a, kx), correlate(b, ky)))


def vec2d_curl(kx, ky, a, b):
    return correlate(b, kx) - correlate(a, ky)


def vec2d_laplace(kl, a, b=None):
    b = a if b is None else b
    return np.stack((correlate(a, kl), correlate(b, kl)))


def spatial_op(kernel, x, bc="periodic", taylor_order=2):
    """BC-pad with BoundaryManager((taylor+1, taylor+1)) then valid conv
    [ref: Utils/VectorConvOps_Spatial.py:41-55]."""
    t = {s: bc for s in ("left", "right", "top", "bottom")}
    return correlate(bc_pad(x, (taylor_order + 1, taylor_order + 1), t), kernel, pad=0)


# --------------------------------------------------------------------------------------
# a7: residual assemblies (expressed on kernels, so reference defects such as D_y == D_t
# propagate automatically)
# --------------------------------------------------------------------------------------
def ops_2d():
    """The operator set every 2-D script builds [ref: Joint/NS_Residuals_CP.py:214-219]."""
    return dict(D_t=build_kernel("2d", "t", 1), D_x=build_kernel("2d", "x", 1),
                D_y=build_kernel("2d", "y", 1), D_xx_yy=build_kernel("2d", ("x", "y"), 2))


def ops_1d():
    """[ref: Joint/Burgers_Residuals_CP.py:172-175]"""
    return dict(D_t=build_kernel("1d", "t", 1), D_x=build_kernel("1d", "x", 1),
                D_xx=build_kernel("1d", "x", 2))


def residual_burgers(u, dx, dt, nu, ops=None):
    """[ref: Joint/Burgers_Residuals_CP.py:182-187] (full field, boundary=True)."""
    o = ops or ops_1d()
    dx, dt, nu = F32(dx), F32(dt), F32(nu)
    u = np.asarray(u, dtype=F32)
    return dx * correlate(u, o["D_t"]) + dt * u * correlate(u, o["D_x"]) \
        - nu * correlate(u, o["D_xx"]) * (F32(2) * dt / dx)


def residual_ns_continuity(u, v, dx, dy, ops=None):
    """[ref: Joint/NS_Residuals_CP.py:222-228]"""
    o = ops or ops_2d()
    return correlate(u, o["D_x"]) + F32(dx / dy) * correlate(v, o["D_y"])


def residual_ns_momentum(u, v, p, dx, dy, dt, nu, ops=None):
    """[ref: Joint/NS_Residuals_CP.py:231-240; Other_UQ/Evaluation/PRE_estimations.py:42-52]"""
    o = ops or ops_2d()
    u, v, p = (np.asarray(a, dtype=F32) for a in (u, v, p))
    dx, dy, dt, nu = F32(dx), F32(dy), F32(dt), F32(nu)
    D = lambda k, f: correlate(f, o[k])
    rx = D("D_t", u) * dx * dy + u * D("D_x", u) * dt * dy + v * D("D_y", u) * dt * dx \
        - nu * D("D_xx_yy", u) * dt + D("D_x", p) * dt * dy
    ry = D("D_t", v) * dx * dy + u * D("D_x", v) * dt * dx + v * D("D_y", v) * dt * dy \
        - nu * D("D_xx_yy", v) * dt + D("D_y", p) * dt * dx
    return rx + ry


def residual_mhd(equation, rho, u, v, p, Bx, By, gamma=5.0 / 3.0, ops=None):
    """[ref: Joint/MHD_Residuals_CP.py:225-278]"""
    o = ops or ops_2d()
    rho, u, v, p, Bx, By = (np.asarray(a, dtype=F32) for a in (rho, u, v, p, Bx, By))
    Dt = lambda f: correlate(f, o["D_t"])
    Dx = lambda f: correlate(f, o["D_x"])
    Dy = lambda f: correlate(f, o["D_y"])
    one, two, half = F32(1), F32(2), F32(0.5)
    g = F32(gamma)
    if equation == "continuity":
        return Dt(rho) + u * Dx(rho) + rho * Dx(u) + v * Dy(rho) + rho * Dy(v)
    if equation == "momentum":
        rx = Dt(u) + u * Dx(u) + (one / rho) * Dx(p) - two * (Bx / rho) * Dx(Bx) + v * Dy(u) \
            - (By / rho) * Dy(Bx) - (Bx / rho) * Dy(By)
        ry = Dt(v) + u * Dx(v) + (one / rho) * Dy(p) - two * (By / rho) * Dy(By) + v * Dy(v) \
            - (By / rho) * Dx(Bx) - (Bx / rho) * Dx(By)
        return rx + ry
    if equation == "energy":
        pg = p - half * (Bx ** 2 + By ** 2)
        gm2 = F32(gamma - 2)
        return Dt(rho) + u * Dx(p) + v * Dy(p) + gm2 * (u * Bx + v * By) * (Dx(Bx) + Dy(By)) \
            + (g * pg + By ** 2) * Dx(u) + (g * pg + Bx ** 2) * Dy(v) - Bx * By * (Dy(u) + Dx(v))
    if equation == "induction":
        rx = Dt(Bx) - By * Dy(u) + Bx * Dy(v) - v * Dy(Bx) + u * Dy(By)
        ry = Dt(By) + By * Dx(u) - Bx * Dx(v) - v * Dx(Bx) + u * Dx(By)
        return rx + ry
    if equation in ("gauss", "divergence"):
        return Dx(Bx) + Dy(By)
    raise ValueError(equation)


def ops_jorek(scale2=5.0 / 3.0):
    """The operator set of the JOREK script [ref: Marginal/JOREK_residuals_CP.py:188-208]: first-order
    operators with scale 1 (alpha = beta = 1 after :192); D_RR / D_ZZ are built with ``scale=gamma``
    AFTER gamma was re-bound to tensor(5/3) at :201, so they carry that factor."""
    return dict(D_t=build_kernel("2d", "t", 1), D_R=build_kernel("2d", "x", 1), D_Z=build_kernel("2d", "y", 1),
                D_RR=build_kernel("2d", "x", 2, scale2), D_ZZ=build_kernel("2d", "y", 2, scale2))


def residual_jorek(equation, rho, phi, T, R, coef, gamma=5.0 / 3.0, ops=None):
    """Reduced-MHD continuity / temperature residuals [ref: Marginal/JOREK_residuals_CP.py:211-242].
    ``R`` is the 1-D major-radius grid: as a 1-D tensor it broadcasts along the LAST axis of the
    [BS,Nt,Nx,Ny] fields (the script passes x_grid, so this needs Nx == Ny there).  ``coef`` is the
    diffusivity D (continuity) or conductivity K (temperature)."""
    o = ops or ops_jorek()
    rho, phi, T, R = (np.asarray(a, dtype=F32) for a in (rho, phi, T, R))
    Dt = lambda f: correlate(f, o["D_t"])
    DR = lambda f: correlate(f, o["D_R"])
    DZ = lambda f: correlate(f, o["D_Z"])
    DRR = lambda f: correlate(f, o["D_RR"])
    DZZ = lambda f: correlate(f, o["D_ZZ"])
    one, two = F32(1), F32(2)
    c, g = F32(coef), F32(gamma)
    if equation == "continuity":
        return Dt(rho) - R * (DR(rho) * DZ(phi) - DR(phi) * DZ(rho)) - two * rho * DZ(phi) \
            - c * (DRR(rho) + (one / R) * DR(rho) + DZZ(rho))
    if equation == "temperature":
        return T * Dt(rho) + rho * Dt(T) - rho * R * (DR(T) * DZ(phi) - DR(phi) * DZ(T)) \
            + T * R * (DR(rho) * DZ(phi) - DR(phi) * DZ(rho)) + two * g * rho * T * DZ(phi) \
            + c * (DRR(T) + (one / R) * DR(T) + DZZ(T))
    raise ValueError(equation)


def periodic_bc_residual(u, wall="right", dx=1.0):
    """[ref: Joint/NS_Residuals_CP.py:483-494] edge-value residual of a periodic boundary."""
    u = np.asarray(u, dtype=F32)
    if wall == "top":
        res = u[..., 0, :] - u[..., -1, :]
    elif wall == "bottom":
        res = u[..., -1, :] - u[..., 0, :]
    elif wall == "left":
        res = u[..., :, 0] - u[..., :, -1]
    elif wall == "right":
        res = u[..., :, -1] - u[..., :, 0]
    else:
        raise UnboundLocalError(wall)
    return (res * F32(dx)).astype(F32)


def interior(res, nd):
    """[ref: Joint/Wave_Residuals_CP.py:184] strip one cell on each of the last nd axes."""
    return res[(Ellipsis,) + (slice(1, -1),) * nd]


# --------------------------------------------------------------------------------------
# a12-a16: conformal helpers (un-vendored Neural_PDE/UQ/inductive_cp.py @ fdf1dcc3)
# --------------------------------------------------------------------------------------
def calibrate_quantile_level(n, alpha):
    """Split-CP level ceil((n+1)(1-alpha))/n evaluated in float64."""
    return np.ceil((n + 1) * (1 - alpha)) / n


def calibrate_rank(n_rows, n, alpha, dtype=F32):
    """0-based order statistic NumPy's method='higher' selects: ceil((n_rows-1)*q) in the dtype of
    the score array (NumPy 2.x casts a scalar q to it), float32 for this path's score arrays."""
    q = calibrate_quantile_level(n, alpha)
    if not (0.0 <= q <= 1.0):
        raise ValueError("Quantiles must be in the range [0, 1]")
    return int(np.ceil((n_rows - 1) * np.dtype(dtype).type(q)))


def calibrate(scores, n, alpha):
    """PARITY UNPINNED: source absent from the reference tree; signature and axis-0
    behaviour from the call sites (e.g. [ref: Joint/NS_Residuals_CP.py:311,
    Other_UQ/Utils/training_utils.py:511]).  Restated as standard split conformal
    prediction."""
    return np.quantile(scores, calibrate_quantile_level(n, alpha), axis=0, method="higher")


def modulation_func(a, b):
    """[ref twin: Tests/test_advection_inv_sampling_marginal.py:428] population std (ddof=0)
    over the sample axis of the difference; symmetric in (a, b)."""
    return np.std(np.asarray(a) - np.asarray(b), axis=0)


def ncf_metric_joint(a, b, modulation):
    """[ref twin: Tests/test_advection_inv_sampling_marginal.py:430-431]"""
    a = np.asarray(a)
    return np.max(np.abs(a - np.asarray(b)) / modulation, axis=tuple(range(1, a.ndim)))


def emp_cov(prediction_sets, y_response):
    """PARITY UNPINNED (no in-tree twin); call sites [ref: Marginal/NS_Residuals_CP.py:312,337]."""
    y = np.asarray(y_response)
    return ((y >= prediction_sets[0]) & (y <= prediction_sets[1])).mean()


def emp_cov_joint(prediction_sets, y_response):
    """[ref twin: Tests/test_advection_inv_sampling_marginal.py:465]"""
    return filter_sims_joint(prediction_sets, y_response).mean()


def filter_sims_joint(prediction_sets, y_response):
    """[ref: Joint/Advection_Residuals_CP.py:327-329]"""
    y = np.asarray(y_response)
    axes = tuple(range(1, y.ndim))
    return (y >= prediction_sets[0]).all(axis=axes) & (y <= prediction_sets[1]).all(axis=axes)


def filter_sims_within_bounds(lower, upper, samples, threshold, within=False):
    """[ref: Active_Learning/Advection_AL_Marginal.py:169-197]"""
    s = np.asarray(samples)
    inside = ((s >= lower) & (s <= upper)) if within else ((s <= lower) | (s >= upper))
    return inside.mean(axis=tuple(range(1, s.ndim))) >= threshold


# --------------------------------------------------------------------------------------
# brute-force helpers used by tests of the select / histogram kernels
# --------------------------------------------------------------------------------------
def float_order_key(x):
    """uint32 key whose unsigned order equals the total order np.sort uses on float32
    (negatives reversed, then positives; NaNs with the sign bit clear sort last)."""
    b = np.asarray(x, dtype=F32).view(np.uint32)
    return np.where(b & np.uint32(0x80000000), ~b, b | np.uint32(0x80000000)).astype(np.uint32)


def select_rank(scores, rank):
    """k-th smallest along axis 0, NaN if the column holds a NaN (np.quantile behaviour)."""
    s = np.sort(np.asarray(scores, dtype=F32), axis=0)
    out = s[rank]
    return np.where(np.isnan(s[-1]), F32(np.nan), out)


# --------------------------------------------------------------------------------------
# multi-GPU marginal quantile: NumPy restatement of the three K3w kernels (include/cppre.h),
# used as their parity oracle and injected into cp_pre_b200/dist.py by the gloo CPU tests
# --------------------------------------------------------------------------------------
def ordered_keys(x):
    """int32 keys whose signed order is the float32 order (NaN with a clear sign bit on top)."""
    return (float_order_key(x) ^ np.uint32(0x80000000)).view(np.int32).reshape(np.shape(x))


def window_compact(scores, lo, hi, slots, slab):
    """(win [blocks, slots, slab] -inf padded, cnt [blocks*slab] int32, overflowing cell indices):
    per cell the first `slots` samples with lo <= key <= hi in row order; cnt = #{key > hi} + (1 << 24 if hi is a NaN key: some rank holds a NaN in this
    cell).  A NaN `hi` key leaves the cell's window empty."""
    s = np.asarray(scores, dtype=F32)
    n = s.shape[0]
    flat = s.reshape(n, -1)
    cells = flat.shape[1]
    keys = ordered_keys(flat).astype(np.int64)
    lo = np.asarray(lo).reshape(-1).astype(np.int64)
    hi = np.asarray(hi).reshape(-1).astype(np.int64)
    blocks = -(-cells // slab)
    win = np.full((blocks, slots, slab), -np.inf, dtype=F32)
    cnt = np.zeros(blocks * slab, dtype=np.int32)
    cnt[cells:] = 1 << 24                    # padding tail of the last slab: flagged like a NaN cell
    inf_key = int(ordered_keys(np.array([np.inf], dtype=F32))[0])
    overflow = []
    for c in range(cells):
        col = keys[:, c]
        is_open = bool(hi[c] > inf_key or lo[c] > hi[c])
        cnt[c] = int((col > hi[c]).sum()) + ((1 << 24) if is_open else 0)
        if is_open:
            continue
        inside = flat[(col >= lo[c]) & (col <= hi[c]), c]
        if len(inside) > slots:
            overflow.append(c)
            inside = inside[:slots]
        win[c // slab, :len(inside), c % slab] = inside
    return win, cnt, overflow


def select_from_top(cand, cnt, kt, network=0):
    """(out [cells], overflowing cell indices): the (kt - (cnt & 0xffffff))-th largest of each
    column; NaN where cnt >> 24 != 0; overflow where that rank is outside [1, rows] or, with
    network > 0, where the column holds more than `network` candidates other than -inf."""
    cand = np.asarray(cand, dtype=F32)
    rows, cells = cand.shape
    order = np.argsort(float_order_key(cand), axis=0, kind="stable")
    srt = np.take_along_axis(cand, order, axis=0)
    out = np.full(cells, np.nan, dtype=F32)
    overflow = []
    for c in range(cells):
        packed = int(cnt[c])
        if packed >> 24:
            continue
        want = kt - (packed & 0xFFFFFF)
        real = int((cand[:, c] != -np.inf).sum())
        if want < 1 or want > rows or (network and real > network):
            overflow.append(c)
            continue
        out[c] = srt[rows - want, c]
    return out, overflow
