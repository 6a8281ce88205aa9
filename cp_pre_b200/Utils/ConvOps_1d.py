"""Drop-in for ``Utils/ConvOps_1d.py``: operators on [BS, Nt, Nx] fields (k x k kernels; the
'x' operators are the transposed 't' stencils).  Executed by libcppre's stencil kernel.
The third-order rows of the reference table are malformed there and therefore unreachable here
too: ``ConvOperator('x', 3)`` yields an object without ``.kernel``."""
import numpy as np  # noqa: F401
import torch
import torch.nn.functional as F  # noqa: F401

from cp_pre_b200._operator import StencilOperator, lookup_stencil
from cp_pre_b200.Utils.fft_conv_pytorch.fft_conv import *  # noqa: F401,F403


def get_stencil(dims, deriv_order, taylor_order=2):
    return lookup_stencil("1d", dims, deriv_order, taylor_order)


def pad_kernel(grid, kernel):
    k = kernel.shape[0]
    _, nt, nx = grid.shape[:3]
    return torch.nn.functional.pad(kernel, (0, nx - k, 0, nt - k), "constant", 0)


class ConvOperator(StencilOperator):
    """Derivative / integral operator on 1-D+time data as a convolution stencil."""

    _flavour = "1d"

    def __init__(self, domain=None, order=None, scale=1.0, taylor_order=2, conv='direct', device='cpu'):
        try:
            self.domain = domain
            self.dims = len(self.domain)
            self.order = order
            self.stencil = get_stencil(self.dims, self.order, taylor_order)
            if self.domain == 'x':
                self.stencil = self.stencil.T
            elif self.domain != 't' and self.domain != ('x', 't'):
                raise ValueError("Invalid Domain. Must be either x or t")
            self.kernel = self.stencil
            self.kernel = scale * self.kernel
            self.kernel = self.kernel.to(device)
        except Exception:
            pass
        self._select_conv(conv)

    def integrate(self, field, kernel=None, correlation=False, slice_pad=True, eps=1e-6):
        return self._spectral_manual(field, kernel, "int", correlation, slice_pad, eps)


def convection_solution(initial_condition, c, dt, nt):
    """u(x, t) = f(x - c t) sampled on nt steps [ref: Utils/ConvOps_1d.py:316-329]."""
    x = torch.arange(initial_condition.shape[2]).unsqueeze(0)
    return torch.cat([initial_condition(x - c * i * dt) for i in range(nt)], dim=1)
