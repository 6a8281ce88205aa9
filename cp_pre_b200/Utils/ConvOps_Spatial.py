"""Drop-in for ``Utils/ConvOps_Spatial.py``: space-only operators on [BS, 1, Nx, Ny] fields.

Differences from the 2-D+time module that the reference has and this keeps: **valid**
correlation (the caller pads first, normally with ``BoundaryManager``), first derivative
weights +-1/2, default ``device=cuda``, kernel = (scale tensor with requires_grad) * stencil,
and -- since the stencil is never transposed by axis -- ``'x'`` and ``'y'`` operators carry the
same kernel."""
import numpy as np  # noqa: F401
import torch
import torch.nn as nn  # noqa: F401
import torch.nn.functional as F  # noqa: F401

from cp_pre_b200._operator import StencilOperator, lookup_stencil
from cp_pre_b200.Utils.fft_conv_pytorch.fft_conv import *  # noqa: F401,F403

_AXIS = {"x": 0, "y": 1, ("x", "y"): 0}


def get_stencil(dims, deriv_order, taylor_order=2):
    return lookup_stencil("spatial", dims, deriv_order, taylor_order)


def pad_kernel(grid, kernel):
    k = kernel.shape[0]
    _, nt, nx, ny = grid.shape[:4]
    return torch.nn.functional.pad(kernel, (0, nx - k, 0, ny - k, 0, nt - k), "constant", 0)


class ConvOperator(StencilOperator):
    """Spatial derivative operator (valid correlation on a pre-padded field)."""

    _flavour = "spatial"

    def __init__(self, domain=None, order=None, scale=1.0, taylor_order=2, conv='direct',
                 device=torch.device("cuda"), requires_grad=False):
        try:
            self.domain = domain
            self.dims = len(self.domain)
            self.order = order
            self.stencil = get_stencil(self.dims, self.order, taylor_order)
            if self.domain not in _AXIS:
                raise ValueError("Invalid Domain. Must be either x,y or their combination")
            self.axis = _AXIS[self.domain]
            self.kernel = self.stencil
            self.kernel = self.kernel.to(device)
            self.scale = torch.tensor(scale, dtype=torch.float32, device=device, requires_grad=True)
            self.kernel = self.scale * self.kernel
            if requires_grad == True:  # noqa: E712
                self.kernel.requires_grad_ = True   # (sic)
        except Exception:
            pass
        self._select_conv(conv)

    def integrate(self, field, kernel=None, correlation=False, slice_pad=False, eps=1e-6):
        return self._spectral_manual(field, kernel, "int", correlation, slice_pad, eps)
