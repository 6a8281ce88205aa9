"""Drop-in for ``Utils/ConvOps_2d.py``: finite-difference operators on [BS, Nt, Nx, Ny] fields.

``ConvOperator(domain, order, scale, taylor_order, conv, device, requires_grad)``; the operator
is its public ``.kernel`` (a k x k x k float32 tensor) and ``D(field)`` is the zero-padded
cross-correlation with it -- executed by libcppre's stencil kernel, not by ``F.conv3d``.
Reference behaviours kept on purpose: ``domain='y'`` builds the same kernel
as ``'t'``; 5^3 / 7^3 kernels sit at plane index 1; any constructor failure is swallowed and
leaves an object without ``.kernel``.
"""
import numpy as np  # noqa: F401
import torch
import torch.nn as nn  # noqa: F401
import torch.nn.functional as F  # noqa: F401

from cp_pre_b200._operator import StencilOperator, embed_3d, lookup_stencil
from cp_pre_b200.Utils.fft_conv_pytorch.fft_conv import *  # noqa: F401,F403  (re-export, like the reference)

_AXIS = {"t": 2, "x": 0, "y": 1, ("x", "y"): 0, ("x", "y", "t"): 0}


def get_stencil(dims, deriv_order, taylor_order=2):
    return lookup_stencil("2d", dims, deriv_order, taylor_order)


def kernel_3d(stencil, axis):
    return embed_3d(stencil, axis)


def pad_kernel(grid, kernel):
    """Zero-extends a cubic kernel to the grid's (Nt, Nx, Ny) extent -- argument order of the
    reference kept as is (it pairs nx with the last axis)."""
    k = kernel.shape[0]
    _, nt, nx, ny = grid.shape[:4]
    return torch.nn.functional.pad(kernel, (0, nx - k, 0, ny - k, 0, nt - k), "constant", 0)


class ConvOperator(StencilOperator):
    """Derivative / integral operator on 2-D+time data as a convolution stencil."""

    _flavour = "2d"

    def __init__(self, domain=None, order=None, scale=1.0, taylor_order=2, conv='direct', device='cpu',
                 requires_grad=False):
        try:
            self.domain = domain
            self.dims = len(self.domain)
            self.order = order
            self.stencil = get_stencil(self.dims, self.order, taylor_order)
            if self.domain not in _AXIS:
                raise ValueError("Invalid Domain. Must be either x,y or t")
            self.axis = _AXIS[self.domain]
            self.kernel = kernel_3d(self.stencil, self.axis)
            self.kernel = scale * self.kernel
            self.kernel = self.kernel.to(device)
            if requires_grad == True:  # noqa: E712
                self.kernel.requires_grad_ = True   # (sic) the reference sets an attribute, not the flag
        except Exception:
            pass
        self._select_conv(conv)

    def integrate(self, field, kernel=None, correlation=False, slice_pad=False, eps=1e-6):
        return self._spectral_manual(field, kernel, "int", correlation, slice_pad, eps)
