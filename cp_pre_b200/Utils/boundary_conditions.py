"""Drop-in for ``Utils/boundary_conditions.py``: ``BoundaryManager`` pads a [B, C, H, W] (or
[H, W]) signal side by side -- left, right, top, bottom -- with dirichlet / neumann / periodic /
symmetric / outflow ghosts ('free_slip' pads nothing, as in the reference).

The padding is an index map evaluated inside libcppre (``cppre_bc_pad``); ``apply_convolution``
never builds the padded tensor: the valid correlation reads through the same map
(``cppre_bc_conv2d``).  The reference's sequential-padding quirks (a periodic right/bottom
ghost repeats the left/top ghost) follow from the map's construction.
"""
import numpy as np
import torch

from cp_pre_b200 import _lib as L
from cp_pre_b200 import ops

_SIDES = ("left", "right", "top", "bottom")
_CODE = {"dirichlet": L.BC_DIRICHLET, "neumann": L.BC_NEUMANN, "outflow": L.BC_NEUMANN,
         "periodic": L.BC_PERIODIC, "symmetric": L.BC_SYMMETRIC, "free_slip": L.BC_NONE}


class BoundaryManager:
    """Per-side boundary conditions for convolution operators."""

    SUPPORTED_TYPES = ['dirichlet', 'neumann', 'periodic', 'symmetric', 'free_slip', 'outflow']

    def __init__(self, kernel_size):
        if isinstance(kernel_size, int):
            self.kernel_height = self.kernel_width = kernel_size
        else:
            self.kernel_height, self.kernel_width = kernel_size
        self.pad_left = self.pad_right = self.kernel_width // 2
        self.pad_top = self.pad_bottom = self.kernel_height // 2
        self.boundary_types = {s: 'periodic' for s in _SIDES}
        self.boundary_values = {s: 0.0 for s in _SIDES}

    def set_boundary_type(self, side, bc_type, value=0.0):
        if side not in _SIDES:
            raise ValueError(f"Unknown side: {side}. Use 'left', 'right', 'top', or 'bottom'")
        if bc_type.lower() not in self.SUPPORTED_TYPES:
            raise ValueError(f"Unsupported boundary type: {bc_type}")
        self.boundary_types[side] = bc_type.lower()
        self.boundary_values[side] = value

    def set_all_boundaries(self, bc_type, value=0.0):
        for side in _SIDES:
            self.set_boundary_type(side, bc_type, value)

    # -- native descriptor ------------------------------------------------------------
    def descriptor(self) -> L.BC:
        widths = (self.pad_left, self.pad_right, self.pad_top, self.pad_bottom)
        return ops.make_bc([_CODE[self.boundary_types[s]] for s in _SIDES], widths,
                           [float(self.boundary_values[s]) for s in _SIDES])

    def index_map(self, height, width):
        """(rows, row_values, cols, col_values): source indices of the padded image (-1 = Dirichlet
        constant, value alongside).  Host only."""
        return ops.bc_index_map(self.descriptor(), height, width)

    def pad_signal(self, signal):
        if signal.dim() < 2:
            raise IndexError("too many indices for tensor of dimension %d" % signal.dim())
        staged = ops.Staged(signal)
        x = ops.to_cuda(signal)
        return staged.back(ops.bc_pad(x, self.descriptor()))

    def apply_convolution(self, signal, kernel):
        """Valid correlation of the padded signal with ``kernel`` ([kh, kw] tensor)."""
        staged = ops.Staged(signal)
        x = ops.to_cuda(signal)
        out = ops.bc_conv2d(x, self.descriptor(), kernel.detach().cpu().numpy().astype(np.float32))
        if out.dim() == 4:               # the reference's squeeze(0).squeeze(0) of the result
            out = out.squeeze(0).squeeze(0)
        return staged.back(out)
