"""The reference's CPU execution of the hot path, restated with the SAME library calls it makes
(``F.conv1d/2d/3d`` with zero "same" padding + elementwise torch ops, NumPy for the conformal
step) -- TEST INFRASTRUCTURE, see oracle/__init__.py.  This is the timed CPU baseline
(``cpu_baseline.kind = "port"`` / ``bench.py --impl reference``): multi-threaded through
torch's intra-op pool exactly like the reference, NumPy parts single-threaded by nature.

Validated against tests/golden (outputs of the unmodified reference) in
tests/test_oracle_golden.py::test_torch_port_matches_reference.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

from . import cpu_oracle as O


def conv_same(field: torch.Tensor, kernel: torch.Tensor) -> torch.Tensor:
    """[ref: Utils/ConvOps_2d.py:135-150, ConvOps_1d.py:130-150, ConvOps_0d.py:84-105]"""
    nd = kernel.dim()
    conv = {1: F.conv1d, 2: F.conv2d, 3: F.conv3d}[nd]
    pad = tuple(k // 2 for k in kernel.shape)
    return conv(field.unsqueeze(1), kernel[None, None], padding=pad if nd > 1 else pad[0]).squeeze(1)


class Ops2D:
    """[ref: Joint/NS_Residuals_CP.py:214-219] -- kernels from the oracle tables (D_y == D_t)."""

    def __init__(self):
        k = O.ops_2d()
        self.kt, self.kx, self.ky, self.kl = (torch.from_numpy(k[n]) for n in ("D_t", "D_x", "D_y", "D_xx_yy"))

    def D_t(self, f): return conv_same(f, self.kt)
    def D_x(self, f): return conv_same(f, self.kx)
    def D_y(self, f): return conv_same(f, self.ky)
    def D_xx_yy(self, f): return conv_same(f, self.kl)


def ns_momentum(vars_: torch.Tensor, dx, dy, dt, nu, ops: Ops2D | None = None, boundary=False):
    """[ref: Joint/NS_Residuals_CP.py:231-240]: 10 conv3d + ~30 elementwise temporaries."""
    o = ops or Ops2D()
    u, v, p = vars_[:, 0], vars_[:, 1], vars_[:, 2]
    res_x = o.D_t(u)*dx*dy + u*o.D_x(u)*dt*dy + v*o.D_y(u)*dt*dx - nu*o.D_xx_yy(u)*dt + o.D_x(p)*dt*dy
    res_y = o.D_t(v)*dx*dy + u*o.D_x(v)*dt*dx + v*o.D_y(v)*dt*dy - nu*o.D_xx_yy(v)*dt + o.D_y(p)*dt*dx
    r = res_x + res_y
    return r if boundary else r[..., 1:-1, 1:-1, 1:-1]


def wave(u: torch.Tensor, c2=0.25, boundary=False):
    """[ref: Joint/Wave_Residuals_CP.py:168-184]"""
    k = torch.from_numpy(O.build_kernel("2d", "t", 2)) - c2 * torch.from_numpy(O.build_kernel("2d", ("x", "y"), 2))
    r = conv_same(u, k)
    return r if boundary else r[..., 1:-1, 1:-1, 1:-1]


def burgers(u: torch.Tensor, dx, dt, nu, boundary=False):
    """[ref: Joint/Burgers_Residuals_CP.py:172-187]"""
    k = O.ops_1d()
    kt, kx, kxx = (torch.from_numpy(k[n]) for n in ("D_t", "D_x", "D_xx"))
    dx, dt, nu = (torch.tensor(v, dtype=torch.float32) for v in (dx, dt, nu))
    r = dx * conv_same(u, kt) + dt * u * conv_same(u, kx) - nu * conv_same(u, kxx) * (2 * dt / dx)
    return r if boundary else r[..., 1:-1, 1:-1]


def mhd_momentum(vars_: torch.Tensor, ops: Ops2D | None = None, boundary=False):
    """[ref: Joint/MHD_Residuals_CP.py:234-243]"""
    o = ops or Ops2D()
    rho, u, v, p, Bx, By = (vars_[:, j] for j in range(6))
    D_t, D_x, D_y = o.D_t, o.D_x, o.D_y
    res_x = D_t(u) + u*D_x(u) + (1/rho)*D_x(p) - 2*(Bx/rho)*D_x(Bx) + v*D_y(u) - (By/rho)*D_y(Bx) - (Bx/rho)*D_y(By)
    res_y = D_t(v) + u*D_x(v) + (1/rho)*D_y(p) - 2*(By/rho)*D_y(By) + v*D_y(v) - (By/rho)*D_x(Bx) - (Bx/rho)*D_x(By)
    r = res_x + res_y
    return r if boundary else r[..., 1:-1, 1:-1, 1:-1]


def mhd(equation: str, vars_: torch.Tensor, gamma=5.0 / 3.0, ops: Ops2D | None = None, boundary=False):
    """All five ideal-MHD residuals [ref: Joint/MHD_Residuals_CP.py:225-278]."""
    o = ops or Ops2D()
    rho, u, v, p, Bx, By = (vars_[:, j] for j in range(6))
    D_t, D_x, D_y = o.D_t, o.D_x, o.D_y
    if equation == "momentum":
        return mhd_momentum(vars_, o, boundary)
    if equation == "continuity":
        r = D_t(rho) + u*D_x(rho) + rho*D_x(u) + v*D_y(rho) + rho*D_y(v)
    elif equation == "energy":
        p_gas = p - 0.5*(Bx**2 + By**2)
        r = D_t(rho) + u*D_x(p) + v*D_y(p) + (gamma-2)*(u*Bx+v*By)*(D_x(Bx) + D_y(By)) + (gamma*p_gas+By**2)*D_x(u) \
            + (gamma*p_gas+Bx**2)*D_y(v) - Bx*By*(D_y(u) + D_x(v))
    elif equation == "induction":
        r = (D_t(Bx) - By*D_y(u) + Bx*D_y(v) - v*D_y(Bx) + u*D_y(By)) \
            + (D_t(By) + By*D_x(u) - Bx*D_x(v) - v*D_x(Bx) + u*D_x(By))
    elif equation in ("gauss", "divergence"):
        r = D_x(Bx) + D_y(By)
    else:
        raise ValueError(equation)
    return r if boundary else r[..., 1:-1, 1:-1, 1:-1]


def marginal_cp(cal_res: torch.Tensor, val_res: torch.Tensor, alpha: float):
    """[ref: Marginal/NS_Residuals_CP.py:327-337] data-free marginal CP on stored residuals."""
    scores = np.abs(cal_res.numpy())
    qhat = O.calibrate(scores, len(scores), alpha)
    return qhat, O.emp_cov([-qhat, qhat], val_res.numpy())


def joint_cp(cal_res: torch.Tensor, val_res: torch.Tensor, alpha: float, eps: float = 0.0):
    """[ref: Joint/NS_Residuals_CP.py:305-337] data-free joint CP on stored residuals."""
    a = cal_res.numpy()
    mod = O.modulation_func(a, np.zeros(a.shape)) + eps
    scores = O.ncf_metric_joint(a, np.zeros(a.shape), mod)
    qhat = O.calibrate(scores, len(scores), alpha)
    return qhat, mod, O.emp_cov_joint([-qhat * mod, qhat * mod], val_res.numpy())
