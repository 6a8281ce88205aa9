"""Fused residual programs: one libcppre kernel per PDE residual, plus the calibration
pipelines built on them.

A program binds the reference's *live* operator objects (so whatever is in ``.kernel`` at call
time is what runs) to one of the library's residual kinds.  ``residual`` returns what the
reference's ``residual(vars, boundary)`` functions return; ``abs_scores`` / ``joint_scores`` run
the fused epilogues (|r - ref|, per-cell weight, per-sample max) so no intermediate tensor makes
an HBM round trip.

    prog = programs.ns_momentum(D_t, D_x, D_y, D_xx_yy, dx, dy, dt, nu)
    res  = prog.residual(vars)                       # [BS, Nt-2, Nx-2, Ny-2], like the scripts
    s    = prog.joint_scores(vars, inv_modulation)   # [BS], nothing but the scores is written
"""
from __future__ import annotations

from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib as L
from . import ops


_WARNED_UNFUSED = set()


def _warn_unfused(entry: str):
    """Once per entry point: a scoring call (calibration / UQ, not a training loss) whose input carries
    ``requires_grad`` -- typically a model output produced outside ``torch.no_grad()`` -- leaves the fused kernel
    for the differentiable path: one K1 launch per derivative slot plus torch elementwise ops, and a full-size
    score tensor.  Correct, but several times slower and heavier; say so instead of degrading silently."""
    if entry in _WARNED_UNFUSED:
        return
    _WARNED_UNFUSED.add(entry)
    import warnings
    warnings.warn(f"cp_pre_b200: {entry}() received a tensor that requires grad and is taking the unfused differentiable "
                  "path (one kernel launch per derivative, the full score field materialised). If no gradient is "
                  "wanted, call it under torch.no_grad() or pass .detach() to get the fused kernel.",
                  RuntimeWarning, stacklevel=3)


class ResidualProgram:
    """``kind`` evaluated with taps from ``operators`` on the fields ``field_index`` of vars."""

    def __init__(self, kind: int, operators: Sequence, scalars: Sequence[float],
                 field_index: Optional[Sequence[int]], ndim_space_time: int, name: str):
        self.kind, self.operators, self.scalars = kind, list(operators), [float(s) for s in scalars]
        self.field_index = None if field_index is None else list(field_index)
        self.nd = ndim_space_time          # 3 for [BS,Nt,Nx,Ny] data, 2 for [BS,Nt,Nx]
        self.name = name
        self.path = L.PATH_AUTO

    # -- plumbing ---------------------------------------------------------------------
    def _slots(self):
        ks, pads = [], []
        for op in self.operators:
            k = op._taps(op.kernel)         # AttributeError if the operator has no kernel
            k3 = op._fl.kernel3(k)
            ks.append(k3)
            pads.append(tuple(s // 2 for s in k3.shape))
        return ks, pads

    def _fields(self, vars_):
        # bf16-input mode: a bfloat16 CUDA tensor is passed through as it is (the TMA kernel stages
        # bf16 tiles and computes in fp32); everything else is brought to float32 like the reference
        bf16 = isinstance(vars_, torch.Tensor) and vars_.dtype == torch.bfloat16
        v = ops.to_cuda(vars_, torch.bfloat16 if bf16 else torch.float32)
        if self.field_index is None:        # single-field programs take the field itself
            fs = [v]
        else:
            fs = [v[:, j] for j in self.field_index]
        if self.nd == 2:
            fs = [f.unsqueeze(2) for f in fs]
        if getattr(self, "coef_field", None) is not None:
            # a 1-D coefficient grid that the reference broadcasts along the LAST axis (JOREK's R):
            # appended as a stride-0 view, never materialised
            cf = ops.to_cuda(self.coef_field, v.dtype).reshape(-1)
            if fs[0].dim() == 4 and cf.numel() != fs[0].shape[-1]:
                raise RuntimeError(f"The size of tensor a ({fs[0].shape[-1]}) must match the size of tensor b "
                                   f"({cf.numel()}) at non-singleton dimension {fs[0].dim() - 1}")
            # one [Nx, Ny] plane (Nx*Ny floats), viewed as the same plane for every instant and sample: the
            # flat-line star kernel stages it like a field; no [BS,Nt,Nx,Ny] copy is ever made
            plane = cf.view(1, -1).expand(fs[0].shape[2], -1).contiguous()
            fs.append(plane.view((1, 1) + tuple(plane.shape)).expand(fs[0].shape))
        if fs[0].dim() != 4:
            raise RuntimeError(f"{self.name}: expected {'[BS,F,Nt,Nx,Ny]' if self.field_index else '[BS,Nt,...]'} input")
        return fs

    def _interior_flags(self):
        return (L.FLAG_INTERIOR_0 | L.FLAG_INTERIOR_2 | (L.FLAG_INTERIOR_1 if self.nd == 3 else 0))

    def _unsq(self, t):
        """[..., Nt, Nx] <-> library layout [..., Nt, 1, Nx] for 1-D+time data."""
        return t.unsqueeze(-2) if (t is not None and self.nd == 2) else t

    def _crop(self, full):
        if self.nd == 2:
            return full[..., 1:-1, 1:-1]
        return full[..., 1:-1, 1:-1, 1:-1]

    def _run(self, vars_, decode=None, **kw):
        ks, pads = self._slots()
        fs = self._fields(vars_)
        return ops.run_program(self.kind, fs, ks, pads, self.scalars, path=self.path,
                               decode=self._decode_per_field(decode, vars_, len(fs)), **kw)

    # -- normaliser-decode fold ------------------------------------------------
    @staticmethod
    def _affine(decode):
        """(scale, offset) of ``decoded = scale * x + offset``.  Accepts the pair itself or a normaliser object
        of the MinMax kind the scripts use -- attributes ``a`` / ``b`` with ``decode(x) = (x - b) / a``
        [ref: Tests/test_ns.py:111-134; the class lives in the un-vendored Neural_PDE package]."""
        if hasattr(decode, "a") and hasattr(decode, "b"):
            a = torch.as_tensor(decode.a, dtype=torch.float64).reshape(-1)
            b = torch.as_tensor(decode.b, dtype=torch.float64).reshape(-1)
            return (1.0 / a).tolist(), (-b / a).tolist()
        sc, of = decode
        flat = lambda v: torch.as_tensor(v, dtype=torch.float64).reshape(-1).tolist()
        return flat(sc), flat(of)

    def _affine_per_variable(self, decode, vars_):
        """(scales, offsets), one value per variable (channel) of ``vars``; scalars apply to every variable."""
        sc, of = self._affine(decode)
        n_vars = 1 if self.field_index is None else int(vars_.shape[1])
        sc = sc * n_vars if len(sc) == 1 else sc
        of = of * n_vars if len(of) == 1 else of
        if len(sc) != n_vars or len(of) != n_vars:
            raise ValueError(f"decode needs one scale and one offset, or one per variable ({n_vars})")
        return sc, of

    def _decode_per_field(self, decode, vars_, n_fields):
        """(scales, offsets) per program field; a coefficient field (JOREK's R) is never decoded."""
        if decode is None:
            return None
        sc, of = self._affine_per_variable(decode, vars_)
        idx = [0] if self.field_index is None else self.field_index
        return ([sc[j] for j in idx] + [1.0] * (n_fields - len(idx)),
                [of[j] for j in idx] + [0.0] * (n_fields - len(idx)))

    def _fold_ok(self):
        """Whether the decode fold reproduces the reference on the kept interior ``[1:-1]``: only while no non-zero
        tap lies more than one cell from the centre.  A Taylor-4/6 stencil reaches the zero padding from the first
        kept layer, where the reference has padded the DECODED field; those operators take the explicit decode."""
        for k in self._slots()[0]:
            for ax, s in enumerate(k.shape):
                if s > 3:
                    outer = [i for i in range(s) if abs(i - s // 2) > 1]
                    if np.any(np.take(np.asarray(k), outer, axis=ax) != 0):
                        return False
        return True

    def _decoded(self, vars_, decode, keep_graph=False):
        """The explicit decode (one elementwise pass): what the fold replaces.  Used where the fold is not
        defined -- boundary=True results (in the reference the shell cells see the zero padding of the DECODED
        field) and gradient-carrying calls."""
        sc, of = self._affine_per_variable(decode, vars_)
        v = ops.to_cuda(vars_, keep_graph=keep_graph)
        if self.field_index is None:
            return v * sc[0] + of[0]
        shape = (1, -1) + (1,) * (v.dim() - 2)
        return (v * torch.tensor(sc, dtype=torch.float32, device=v.device).view(shape)
                + torch.tensor(of, dtype=torch.float32, device=v.device).view(shape))

    # -- autograd -----------------------------------------------------------
    def _needs_grad(self, vars_):
        return ops.needs_grad(vars_, *[getattr(op, "kernel", None) for op in self.operators])

    def _residual_autograd(self, vars_):
        """The FULL residual field on the autograd graph: every derivative slot is one differentiable
        K1 launch (``_LinearStencilFn``: backward = K1 with the adjoint taps, plus the per-tap reduction
        for ``kernel.requires_grad``) and the program's algebra is torch elementwise ops, so the product
        rule of the nonlinear programs is autograd's.  Used only when a gradient is asked for (a PRE loss,
        [ref: Physics_Informed/Wave_FNO_PI.py:202-228]); values agree with the fused kernel to fp32 rounding."""
        from ._operator import differentiable_apply
        v = ops.to_cuda(vars_, keep_graph=True)
        fs = [v] if self.field_index is None else [v[:, j] for j in self.field_index]
        if self.nd == 2:
            fs = [f.unsqueeze(2) for f in fs]
        if getattr(self, "coef_field", None) is not None:
            cf = ops.to_cuda(self.coef_field, keep_graph=True).reshape(-1)
            fs.append(cf.view(1, 1, 1, -1).expand(fs[0].shape))
        cache = {}

        def D(slot, f):
            if (slot, f) not in cache:
                cache[(slot, f)] = differentiable_apply(self.operators[slot], fs[f].contiguous())
            return cache[(slot, f)]

        full = _FORMULAS[self.kind](lambda f: fs[f], D, self.scalars)
        return full.squeeze(2) if self.nd == 2 else full

    # -- the reference-facing call ------------------------------------------------------
    def residual(self, vars_, boundary=False, decode=None):
        """The PDE residual; interior view (``[...,1:-1,1:-1(,1:-1)]``) unless boundary=True.
        ``decode`` = (scale, offset) -- scalars or one value per variable of ``vars`` -- or a MinMax-style
        normaliser object: the residual of ``scale * vars + offset`` (``out_normalizer.decode(pred)``,
        [ref: Joint/Wave_Residuals_CP.py:213-217]) with the decode folded into the kernel's loads."""
        staged = ops.Staged(vars_)
        if decode is not None and (boundary or self._needs_grad(vars_) or not self._fold_ok()):
            vars_, decode = self._decoded(vars_, decode, keep_graph=True), None
        if self._needs_grad(vars_):
            full = self._residual_autograd(vars_)
        else:
            full = self._run(vars_, decode=decode)
            if self.nd == 2:
                full = full.squeeze(2)
        return staged.back(full if boundary else self._crop(full))

    __call__ = residual

    # -- fused epilogues ----------------------------------------------------------------
    def abs_scores(self, vars_, ref=None, boundary=False, out=None, decode=None):
        """|residual - ref| (the marginal nonconformity score) in one pass.  ``ref`` is a FULL
        residual field (boundary=True form) or None; ``out`` an optional caller-owned FULL-size
        CUDA destination (e.g. a batch slice of a larger score tensor); ``decode`` as in ``residual``
        (with ``out`` / boundary=True the shell cells of the full field are not those of the decoded field)."""
        staged = ops.Staged(vars_)
        if decode is not None and ((boundary and out is None) or self._needs_grad(vars_) or ops.needs_grad(ref)
                                   or not self._fold_ok()):
            vars_, decode = self._decoded(vars_, decode, keep_graph=True), None
        if self._needs_grad(vars_) or ops.needs_grad(ref):
            _warn_unfused("abs_scores")
            full = self._residual_autograd(vars_)
            if ref is not None:
                full = full - ops.to_cuda(ref, keep_graph=True)
            full = full.abs()
            if out is not None:
                raise ValueError("out= cannot be combined with a gradient-carrying call")
            return staged.back(full if boundary else self._crop(full))
        r = None if ref is None else self._unsq(ops.to_cuda(ref))
        full = self._run(vars_, decode=decode, flags=L.FLAG_ABS, ref=r, out=self._unsq(out))
        if self.nd == 2:
            full = full.squeeze(2)
        return staged.back(full if boundary else self._crop(full))

    def joint_scores(self, vars_, inv_modulation=None, ref=None, interior=True, reduce="max", decode=None):
        """Per-sample reduce over cells of |residual - ref| * inv_modulation; only [BS] floats
        are written.  ``inv_modulation`` / ``ref`` are FULL-size ([Nt,Nx(,Ny)] / [BS,Nt,...]);
        ``decode`` as in ``residual`` (folded when ``interior``)."""
        staged = ops.Staged(vars_)
        if decode is not None and (not interior or self._needs_grad(vars_) or ops.needs_grad(ref, inv_modulation)
                                   or not self._fold_ok()):
            vars_, decode = self._decoded(vars_, decode, keep_graph=True), None
        if self._needs_grad(vars_) or ops.needs_grad(ref, inv_modulation):
            _warn_unfused("joint_scores")
            v = self.abs_scores(vars_, ref=ref, boundary=True)
            if inv_modulation is not None:
                v = v * ops.to_cuda(inv_modulation, keep_graph=True)
            v = (self._crop(v) if interior else v).flatten(1)
            return staged.back(v.amax(dim=1) if reduce == "max" else v.sum(dim=1))
        w = None if inv_modulation is None else self._unsq(ops.to_cuda(inv_modulation))
        r = None if ref is None else self._unsq(ops.to_cuda(ref))
        flags = L.FLAG_ABS | (self._interior_flags() if interior else 0)
        out = self._run(vars_, decode=decode, mode=L.EPI_SCORE, flags=flags, weight=w, ref=r,
                        reduce=L.REDUCE_MAX if reduce == "max" else L.REDUCE_SUM)
        return staged.back(out)


    def coverage_counts(self, vars_, q, half=None, centre=None, interior=True, decode=None):
        """Per-sample number of cells with centre - q*half <= residual(vars) <= centre + q*half,
        int32 [BS]; neither the residual of ``vars_`` nor the prediction sets are written.
        ``half`` ([Nt,Nx(,Ny)], e.g. qhat per cell with q=1, or the modulation with q=qhat) and
        ``centre`` ([BS,Nt,...], e.g. the predicted residual) are FULL-size; None = 1 / 0.
        With ``interior`` the shell cells the reference crops are not counted
        [ref: Marginal/NS_Residuals_CP.py:311-312,336-337; Joint/NS_Residuals_CP.py:312,336]."""
        staged = ops.Staged(vars_)
        if decode is not None and (not interior or not self._fold_ok()):
            vars_, decode = self._decoded(vars_, decode), None
        w = None if half is None else self._unsq(ops.to_cuda(half))
        r = None if centre is None else self._unsq(ops.to_cuda(centre))
        # q: Python float, or the one-element device tensor a device-side quantile returned (the
        # kernel reads it itself: no host round trip between calibrate and coverage)
        band = q if isinstance(q, torch.Tensor) and q.is_cuda else float(q)
        out = self._run(vars_, decode=decode, mode=L.EPI_COVER, flags=self._interior_flags() if interior else 0,
                        weight=w, ref=r, band_q=band)
        return staged.back(out)

    def interior_cells(self, vars_shape):
        """Number of cells the ``interior`` crop keeps for an input of this shape."""
        dims = vars_shape[-3:] if self.nd == 3 else vars_shape[-2:]
        n = 1
        for d in dims:
            n *= max(int(d) - 2, 0)
        return n


# ---------------------------------------------------------------------------------------
# the programs' algebra in torch ops (autograd path only; csrc/programs.cuh is the fused form).
# V(f): field f of the program; D(slot, f): derivative slot applied to field f; s: scalars.
# ---------------------------------------------------------------------------------------
def _f_ns_momentum(V, D, s):
    u, v = V(0), V(1)
    a, b, c, d = s[:4]
    rx = a * D(0, 0) + b * (u * D(1, 0)) + c * (v * D(2, 0)) - d * D(3, 0) + b * D(1, 2)
    ry = a * D(0, 1) + c * (u * D(1, 1)) + b * (v * D(2, 1)) - d * D(3, 1) + c * D(2, 2)
    return rx + ry


def _f_mhd_momentum(V, D, s):
    rho, u, v, Bx, By = V(0), V(1), V(2), V(4), V(5)
    ir = 1.0 / rho
    rx = D(0, 1) + u * D(1, 1) + ir * D(1, 3) - 2 * (Bx * ir) * D(1, 4) + v * D(2, 1) - (By * ir) * D(2, 4) - (Bx * ir) * D(2, 5)
    ry = D(0, 2) + u * D(1, 2) + ir * D(2, 3) - 2 * (By * ir) * D(2, 5) + v * D(2, 2) - (By * ir) * D(1, 4) - (Bx * ir) * D(1, 5)
    return rx + ry


def _f_mhd_energy(V, D, s):
    g = s[0]
    u, v, p, Bx, By = V(1), V(2), V(3), V(4), V(5)
    pg = p - 0.5 * (Bx ** 2 + By ** 2)
    return (D(0, 0) + u * D(1, 3) + v * D(2, 3) + (g - 2) * (u * Bx + v * By) * (D(1, 4) + D(2, 5))
            + (g * pg + By ** 2) * D(1, 1) + (g * pg + Bx ** 2) * D(2, 2) - Bx * By * (D(2, 1) + D(1, 2)))


def _f_mhd_induction(V, D, s):
    u, v, Bx, By = V(0), V(1), V(2), V(3)
    rx = D(0, 2) - By * D(2, 0) + Bx * D(2, 1) - v * D(2, 2) + u * D(2, 3)
    ry = D(0, 3) + By * D(1, 0) - Bx * D(1, 1) - v * D(1, 2) + u * D(1, 3)
    return rx + ry


def _f_jorek_continuity(V, D, s):
    rho, R = V(0), V(3)
    return (D(0, 0) - R * (D(1, 0) * D(2, 1) - D(1, 1) * D(2, 0)) - 2 * rho * D(2, 1)
            - s[0] * (D(3, 0) + (1.0 / R) * D(1, 0) + D(4, 0)))


def _f_jorek_temperature(V, D, s):
    rho, T, R = V(0), V(2), V(3)
    return (T * D(0, 0) + rho * D(0, 2) - rho * R * (D(1, 2) * D(2, 1) - D(1, 1) * D(2, 2))
            + T * R * (D(1, 0) * D(2, 1) - D(1, 1) * D(2, 0)) + 2 * s[1] * rho * T * D(2, 1)
            + s[0] * (D(3, 2) + (1.0 / R) * D(1, 2) + D(4, 2)))


_FORMULAS = {
    L.PROG_LINEAR: lambda V, D, s: D(0, 0),
    L.PROG_BURGERS: lambda V, D, s: s[0] * D(0, 0) + s[1] * (V(0) * D(1, 0)) - s[2] * D(2, 0),
    L.PROG_NS_CONTINUITY: lambda V, D, s: D(0, 0) + s[0] * D(1, 1),
    L.PROG_NS_MOMENTUM: _f_ns_momentum,
    L.PROG_MHD_CONTINUITY: lambda V, D, s: D(0, 0) + V(1) * D(1, 0) + V(0) * D(1, 1) + V(2) * D(2, 0) + V(0) * D(2, 2),
    L.PROG_MHD_MOMENTUM: _f_mhd_momentum,
    L.PROG_MHD_ENERGY: _f_mhd_energy,
    L.PROG_MHD_INDUCTION: _f_mhd_induction,
    L.PROG_MHD_GAUSS: lambda V, D, s: D(1, 0) + D(2, 1),
    L.PROG_JOREK_CONTINUITY: _f_jorek_continuity,
    L.PROG_JOREK_TEMPERATURE: _f_jorek_temperature,
}


# ---------------------------------------------------------------------------------------
# factories (argument meaning = the variables of the reference scripts)
# ---------------------------------------------------------------------------------------
def _f(x):
    """Python / NumPy / 0-d tensor scalar -> float (the scripts mix all three)."""
    return float(x.item() if hasattr(x, "item") else x)


def linear(D):
    """res = D(u) for an additive-kernel operator (advection, wave)
    [ref: Joint/Wave_Residuals_CP.py:173-184, Marginal/Advection_Residuals_CP.py:160-164]."""
    return ResidualProgram(L.PROG_LINEAR, [D], [], None, 3 if D._flavour == "2d" else 2, "linear")


def burgers(D_t, D_x, D_xx, dx, dt, nu):
    """dx*D_t(u) + dt*u*D_x(u) - nu*D_xx(u)*(2*dt/dx)  [ref: Joint/Burgers_Residuals_CP.py:182-187]"""
    dx, dt, nu = np.float32(_f(dx)), np.float32(_f(dt)), np.float32(_f(nu))
    return ResidualProgram(L.PROG_BURGERS, [D_t, D_x, D_xx], [dx, dt, nu * (np.float32(2) * dt / dx)], None, 2,
                           "burgers")


def ns_continuity(D_x, D_y, dx, dy):
    """D_x(u) + (dx/dy)*D_y(v)  [ref: Joint/NS_Residuals_CP.py:222-228]"""
    return ResidualProgram(L.PROG_NS_CONTINUITY, [D_x, D_y], [_f(dx) / _f(dy)], [0, 1], 3, "ns_continuity")


def ns_momentum(D_t, D_x, D_y, D_xx_yy, dx, dy, dt, nu):
    """res_x + res_y  [ref: Joint/NS_Residuals_CP.py:231-240]"""
    dx, dy, dt, nu = _f(dx), _f(dy), _f(dt), _f(nu)
    return ResidualProgram(L.PROG_NS_MOMENTUM, [D_t, D_x, D_y, D_xx_yy], [dx * dy, dt * dy, dt * dx, nu * dt],
                           [0, 1, 2], 3, "ns_momentum")


_MHD = {"continuity": (L.PROG_MHD_CONTINUITY, [0, 1, 2]), "momentum": (L.PROG_MHD_MOMENTUM, [0, 1, 2, 3, 4, 5]),
        "energy": (L.PROG_MHD_ENERGY, [0, 1, 2, 3, 4, 5]), "induction": (L.PROG_MHD_INDUCTION, [1, 2, 4, 5]),
        "gauss": (L.PROG_MHD_GAUSS, [4, 5]), "divergence": (L.PROG_MHD_GAUSS, [4, 5])}


def mhd(equation, D_t, D_x, D_y, gamma=5 / 3):
    """Ideal-MHD residuals on vars = [rho,u,v,p,Bx,By]  [ref: Joint/MHD_Residuals_CP.py:225-278]"""
    kind, idx = _MHD[equation]
    return ResidualProgram(kind, [D_t, D_x, D_y], [_f(gamma)], idx, 3, "mhd_" + equation)


def jorek(equation, D_t, D_R, D_Z, D_RR, D_ZZ, R, coef, gamma=5 / 3):
    """JOREK reduced-MHD residuals on vars = [rho, phi, T]; ``R`` the 1-D major-radius grid (broadcast
    along the last axis, as the script's ``R * (...)`` does), ``coef`` the diffusivity D
    (equation="continuity") or conductivity K ("temperature")  [ref: Marginal/JOREK_residuals_CP.py:211-242]."""
    kind = {"continuity": L.PROG_JOREK_CONTINUITY, "temperature": L.PROG_JOREK_TEMPERATURE}[equation]
    scalars = [_f(coef)] if equation == "continuity" else [_f(coef), _f(gamma)]
    prog = ResidualProgram(kind, [D_t, D_R, D_Z, D_RR, D_ZZ], scalars, [0, 1, 2], 3, "jorek_" + equation)
    prog.coef_field = R
    return prog


def periodic_bc_residual(u, wall="right", dx=1.0):
    """PRE over a periodic boundary: the edge-value difference ``(u[wall] - u[opposite wall]) * dx`` of a
    [..., Nx, Ny] field (e.g. the vorticity ``w`` as [BS, Nt, Nx, Ny]) -> [..., Ny] ('top' / 'bottom') or
    [..., Nx] ('left' / 'right'); the result goes through the same Joint CP helpers as any residual
    [ref: Joint/NS_Residuals_CP.py:483-512].  An unknown wall raises like the reference (its ``res`` is unbound)."""
    if wall not in ("top", "bottom", "left", "right"):
        raise UnboundLocalError("cannot access local variable 'res' where it is not associated with a value")
    staged = ops.Staged(u)
    return staged.back(ops.edge_difference(ops.to_cuda(u), wall, _f(dx)))


# ---------------------------------------------------------------------------------------
# calibration pipelines
# ---------------------------------------------------------------------------------------
def joint_calibration(program: ResidualProgram, cal_pred, cal_out=None, eps: float = 0.0, group=None,
                      out: Optional[torch.Tensor] = None):
    """Joint CP calibration [ref: Joint/NS_Residuals_CP.py:289-306, Joint/MHD_Residuals_CP.py:350-351]:
        modulation = std_samples(res(cal_out) - res(cal_pred)) (+ eps)
        scores[i]  = max_cells |res(cal_out)_i - res(cal_pred)_i| / modulation
    cal_out=None is the data-free PRE form (residual of the prediction alone).  With an initialised
    ``torch.distributed`` process group of more than one rank the inputs are this rank's shard of
    the batch: the modulation is the std over the GLOBAL batch (per-rank fp64 moments merged over
    NCCL, dist.global_cell_std) and the scores are rank-local.  ``out`` is an optional caller-owned
    FULL-size [n, Nt, Nx(, Ny)] CUDA buffer for the stored signed difference.
    Returns (modulation_full, inv_modulation_full, scores); the *_full arrays cover every cell
    ([Nt,Nx(,Ny)]); the reference's arrays are their ``[1:-1,...]`` interior.

    Bytes per calibration point: 4F + 4 (K1 STORE) + 4 (K2 moments) + 4 (K2 scores) = 4F + 12, against
    8F for the recompute form (modulation pass + score pass with nothing stored): storing wins from
    F = 4 fields up (MHD momentum / energy: 36 vs 48 B/point)."""
    from . import dist as D
    ref = None if cal_out is None else program.residual(ops.to_cuda(cal_out), boundary=True)
    # signed difference res_pred - res_out, stored once (full field, shell included)
    signed = program._run(ops.to_cuda(cal_pred), ref=None if ref is None else program._unsq(ref),
                          out=program._unsq(out))
    if program.nd == 2:
        signed = signed.squeeze(2)
    if D._world(group) > 1:
        mod, inv = D.global_cell_std(signed, None, eps, group)
    else:
        mod, inv = ops.cell_modulation(signed, None, eps)
    # the max runs over the interior only: shell cells get modulation = +inf, so |d|/inf = 0
    # never wins -- no cropped copy of the residual tensor is made
    masked = torch.full_like(mod, float("inf"))
    program._crop(masked).copy_(program._crop(mod))
    scores = ops.joint_score(signed, None, masked)
    return mod, inv, scores


def joint_quantile(scores: torch.Tensor, n: int, alpha: float, group=None) -> torch.Tensor:
    """calibrate(scores, n, alpha) for the 1-D joint scores, over all ranks of ``group``, as a
    one-element DEVICE tensor (K3d radix select; the 257-counter histograms are all-reduced over
    NCCL between passes).  Hand it to ``coverage_counts`` as ``q``.
    [ref: Joint/MHD_Residuals_CP.py:399, Joint/NS_Residuals_CP.py:311]"""
    from . import dist as D
    from .Neural_PDE.UQ.inductive_cp import calibrate_rank
    total = D.global_count(scores.numel(), group)
    return D.joint_quantile_device(scores, calibrate_rank(total, n, alpha), group)


def marginal_calibration(program: ResidualProgram, cal_pred, cal_out=None, n=None, alphas=(0.1,)):
    """Marginal CP calibration [ref: Marginal/NS_Residuals_CP.py:289-311]: per-cell quantiles of
    |res(cal_out) - res(cal_pred)| (or |res(cal_pred)| data-free) for every alpha in one pass."""
    from .Neural_PDE.UQ.inductive_cp import calibrate_rank
    ref = None if cal_out is None else program.residual(ops.to_cuda(cal_out), boundary=True)
    scores = program.abs_scores(ops.to_cuda(cal_pred), ref=ref)
    rows = scores.shape[0]
    ranks = [calibrate_rank(rows, rows if n is None else n, a) for a in alphas]
    return scores, ops.select_ranks(scores, ranks)
