"""Drop-in for ``Other_UQ/Evaluation/PRE_estimations.py``: class-form PDE residuals
``PRE_Wave`` / ``PRE_NS`` / ``PRE_MHD`` with ``residual(vars, boundary=False)``.

Each residual is ONE fused libcppre kernel (``cp_pre_b200.programs``) instead of the reference's
10 ``F.conv3d`` calls plus ~30 elementwise temporaries; the derivative taps are read from the
live ``ConvOperator.kernel`` tensors at call time, so user edits to ``.kernel`` (and the shipped
``D_y == D_t`` defect) carry over unchanged.
"""
import torch

from cp_pre_b200 import programs
from cp_pre_b200.Utils.ConvOps_2d import ConvOperator


class PRE_Wave():
    def __init__(self, dt, dx, c=1.0):
        D_tt = ConvOperator('t', 2)
        D_xx_yy = ConvOperator(('x', 'y'), 2)
        self.D = ConvOperator()
        c = torch.tensor(c, dtype=torch.float32)
        self.D.kernel = D_tt.kernel - (c * dt / dx) ** 2 * D_xx_yy.kernel
        self.program = programs.linear(self.D)

    def residual(self, uu, boundary=False):
        return self.program.residual(uu[:, 0], boundary=boundary)


class PRE_NS():
    def __init__(self, dt, dx, dy):
        self.dt, self.dx, self.dy = dt, dx, dy
        self.D_t = ConvOperator(domain='t', order=1)
        self.D_x = ConvOperator(domain='x', order=1)
        self.D_y = ConvOperator(domain='y', order=1)
        self.D_x_y = ConvOperator(domain=('x', 'y'), order=1)      # no kernel, as in the reference
        self.D_xx_yy = ConvOperator(domain=('x', 'y'), order=2)
        self.program = programs.ns_momentum(self.D_t, self.D_x, self.D_y, self.D_xx_yy, dx, dy, dt, nu=0.001)

    def residual(self, vars, boundary=False):
        return self.program.residual(vars, boundary=boundary)


class PRE_MHD():
    def __init__(self, dt, dx, dy):
        self.dt, self.dx, self.dy = dt, dx, dy
        self.D_t = ConvOperator(domain='t', order=1)
        self.D_x = ConvOperator(domain='x', order=1)
        self.D_y = ConvOperator(domain='y', order=1)
        self.D_x_y = ConvOperator(domain=('x', 'y'), order=1)
        self.D_xx_yy = ConvOperator(domain=('x', 'y'), order=2)
        self.program = programs.mhd("energy", self.D_t, self.D_x, self.D_y, gamma=5 / 3)

    def residual(self, vars, boundary=False):
        return self.program.residual(vars, boundary=boundary)
