"""Drop-in for ``Utils/ConvOps_0d.py``: temporal operators on [BS, Nt] ODE trajectories (true
1-D kernels; derivative orders 1-2 at Taylor orders 2/4/6).  Executed by libcppre's stencil
kernel.  ``integrate`` discards its zero padding exactly like the reference (0d:201-202)."""
import torch  # noqa: F401
import torch.nn.functional as F  # noqa: F401

from cp_pre_b200._operator import StencilOperator, lookup_stencil
from cp_pre_b200.Utils.fft_conv_pytorch.fft_conv import *  # noqa: F401,F403


def get_stencil(deriv_order, taylor_order=2):
    return lookup_stencil("0d", None, deriv_order, taylor_order)


class ConvOperator(StencilOperator):
    """Temporal derivative / integral operator for ODE data."""

    _flavour = "0d"

    def __init__(self, order=None, scale=1.0, taylor_order=2, conv='direct', device='cpu',
                 requires_grad=False):
        try:
            self.order = order
            self.stencil = get_stencil(self.order, taylor_order)
            self.kernel = (scale * self.stencil).to(device)
            if requires_grad:
                self.kernel.requires_grad_ = True   # (sic)
        except Exception:
            pass
        self._select_conv(conv)

    def spectral_convolution(self, field, kernel=None):
        return super().spectral_convolution(field, kernel)

    def integrate(self, field, kernel=None, correlation=False, slice_pad=True, eps=1e-6):
        return self._spectral_manual(field, kernel, "int", correlation, slice_pad, eps, pad_field=False)
