"""Shared machinery of the four drop-in ``ConvOperator`` classes (ODE, 1-D+time, 2-D+time,
space-only).  The public modules under ``cp_pre_b200/Utils`` only bind a flavour to it.

Behaviour contract (b): the operator *is* its public, mutable
``.kernel`` tensor; ``D(field)`` cross-correlates the field with whatever tensor is there,
zero "same" padding (space-only flavour: no padding).  Constructor failures are swallowed and
leave an object without ``.kernel``, exactly like the reference; an unknown ``conv`` raises
``ValueError("Unknown Convolution Method")``.

All arithmetic runs in libcppre's CUDA kernels (``ops.run_program``); there is no PyTorch
convolution anywhere on this path.
"""
from __future__ import annotations

import weakref
from fractions import Fraction

import numpy as np
import torch

from . import _lib as L
from . import ops

# ---------------------------------------------------------------------------------------
# stencil tables  [ref: Utils/ConvOps_2d.py:17-63, ConvOps_1d.py:17-80, ConvOps_0d.py:21-47,
# ConvOps_Spatial.py:20-66].  Stored once as exact rationals; the reference writes the same
# values as Python float literals into float32 tensors.
# ---------------------------------------------------------------------------------------
_Q = Fraction
_CENTRAL = {
    # (derivative order, Taylor order) -> central-difference weights
    (1, 2): (-1, 0, 1),
    (1, 4): (_Q(1, 12), _Q(-2, 3), 0, _Q(2, 3), _Q(-1, 12)),
    (2, 2): (1, -2, 1),
    (2, 4): (_Q(-1, 12), _Q(4, 3), _Q(-5, 2), _Q(4, 3), _Q(-1, 12)),
    (2, 6): (_Q(1, 90), _Q(-3, 20), _Q(3, 2), _Q(-49, 18), _Q(3, 2), _Q(-3, 20), _Q(1, 90)),
}
_IDENTITY = (0, 1, 0)


def _weights(seq) -> torch.Tensor:
    return torch.tensor([float(v) for v in seq], dtype=torch.float64).to(torch.float32)


def _upright(weights) -> torch.Tensor:
    """k x k matrix carrying the 1-D weights down its middle column."""
    w = _weights(weights)
    m = torch.zeros(len(w), len(w), dtype=torch.float32)
    m[:, len(w) // 2] = w
    return m


def _plus(weights, centre) -> torch.Tensor:
    """Plus-shaped matrix: the 1-D weights on the middle row and column, centre as given.
    The reference's Taylor-4/6 centres are the 1-D values (-5/2, -49/18), not doubled."""
    m = _upright(weights)
    m[len(m) // 2, :] = _weights(weights)
    m[len(m) // 2, len(m) // 2] = float(centre)
    return m


def lookup_stencil(flavour: str, dims, deriv_order, taylor_order=2) -> torch.Tensor:
    """``get_stencil`` of the four reference modules.  Raises ValueError for an unknown
    combination (TypeError for the 1-D module's malformed third-order rows)."""
    if flavour == "0d":
        if deriv_order == 0:
            return _weights(_IDENTITY)
        if deriv_order in (1, 2) and (deriv_order, taylor_order) in _CENTRAL:
            return _weights(_CENTRAL[(deriv_order, taylor_order)])
        raise ValueError("Invalid stencil parameters")
    if dims == 1:
        if deriv_order == 0:
            return _upright(_IDENTITY)
        if (deriv_order, taylor_order) in ((1, 2), (2, 2)):
            m = _upright(_CENTRAL[(deriv_order, taylor_order)])
            if flavour == "spatial" and deriv_order == 1:
                m = m * 0.5          # [ref: Utils/ConvOps_Spatial.py:34-39]
            return m
        if flavour == "1d" and deriv_order == 3 and taylor_order in (2, 4):
            # [ref: Utils/ConvOps_1d.py:38-53]: a missing comma makes the literal itself fail
            raise TypeError("list indices must be integers or slices, not tuple")
    elif dims == 2 and deriv_order == 2:
        centre = {2: -4, 4: _Q(-5, 2), 6: _Q(-49, 18)}.get(taylor_order)
        if centre is not None:
            return _plus(_CENTRAL[(2, taylor_order)], centre)
    raise ValueError("Invalid stencil parameters")


def embed_3d(stencil: torch.Tensor, axis: int) -> torch.Tensor:
    """``kernel_3d`` [ref: Utils/ConvOps_2d.py:67-79]: the k x k stencil goes to plane index 1
    (sic, also for k = 5 and 7) of a k^3 zero tensor, orthogonal to ``axis``."""
    k = stencil.shape[0]
    cube = torch.zeros(k, k, k)
    if axis not in (0, 1, 2):
        raise ValueError("Invalid axis. Must be either 0, 1 or 2")
    index = [slice(None)] * 3
    index[axis] = 1
    cube[tuple(index)] = stencil
    return cube


# ---------------------------------------------------------------------------------------
# flavours
# ---------------------------------------------------------------------------------------
class Flavour:
    """How one reference module lays its data out, mapped to the library's [B,N2,N1,N0]."""

    def __init__(self, name, field_ndim, kernel_ndim, same_padding, channel_dim):
        self.name = name
        self.field_ndim = field_ndim        # ndim of the field the user passes
        self.kernel_ndim = kernel_ndim
        self.same_padding = same_padding
        self.channel_dim = channel_dim      # True: data carries a size-1 channel axis at 1

    def kernel3(self, k: np.ndarray) -> np.ndarray:
        return k.reshape((1,) * (3 - k.ndim) + k.shape) if self.name != "1d" else k.reshape(
            k.shape[0], 1, k.shape[1])

    def field4(self, f: torch.Tensor) -> torch.Tensor:
        if self.name == "2d":
            return f
        if self.name == "1d":
            return f.unsqueeze(2)                      # [B,Nt,Nx] -> [B,Nt,1,Nx]
        if self.name == "0d":
            return f.unsqueeze(1).unsqueeze(1)         # [B,Nt] -> [B,1,1,Nt]
        return f                                       # spatial: [B,1,Nx,Ny]

    def unfield4(self, out: torch.Tensor) -> torch.Tensor:
        if self.name == "1d":
            return out.squeeze(2)
        if self.name == "0d":
            return out.squeeze(1).squeeze(1)
        return out


FLAVOURS = {
    "2d": Flavour("2d", 4, 3, True, False),
    "1d": Flavour("1d", 3, 2, True, False),
    "0d": Flavour("0d", 2, 1, True, False),
    "spatial": Flavour("spatial", 4, 2, False, True),
}


def taps_of(kernel: torch.Tensor) -> np.ndarray:
    """Host copy of a kernel tensor's coefficients (detached; float32 required like the
    reference, whose F.conv* raises on a dtype mismatch with the float32 field)."""
    if not isinstance(kernel, torch.Tensor):
        raise TypeError("kernel must be a torch.Tensor")
    if kernel.dtype != torch.float32:
        raise RuntimeError(f"expected scalar type Float but found {kernel.dtype} (kernel must be float32)")
    return kernel.detach().cpu().contiguous().numpy()


class _LinearStencilFn(torch.autograd.Function):
    """Autograd through ``D(field)`` (PRE as a training loss, [ref: Physics_Informed/Wave_FNO_PI.py:202-228]).
    d/d(field): the adjoint of a zero-padded cross-correlation is the cross-correlation with the flipped
    taps and padding k-1-p -- one more K1 launch.  d/d(kernel) (``D.kernel.requires_grad = True``, :206):
    grad_K[a,b,c] = sum over samples and cells of grad[o] * field[o + (a,b,c) - pad], one reduction per
    entry of the dense kernel tensor (training-time only; every entry, zeros included, gets its gradient
    because the optimiser sees the dense tensor)."""

    @staticmethod
    def forward(ctx, field4, kern, k3, pad):
        ctx.k3, ctx.pad = k3, pad
        ctx.kern_shape = None if kern is None else tuple(kern.shape)
        ctx.kern_device = None if kern is None else kern.device
        if kern is not None and kern.requires_grad:
            ctx.save_for_backward(field4)
        return ops.run_program(L.PROG_LINEAR, [field4.detach()], [k3], [pad])

    @staticmethod
    def backward(ctx, grad):
        import torch.nn.functional as F
        k3, pad = ctx.k3, ctx.pad
        grad = grad.contiguous()
        g_field = g_kern = None
        if ctx.needs_input_grad[0]:
            flipped = np.ascontiguousarray(k3[::-1, ::-1, ::-1])
            adj_pad = tuple(k - 1 - p for k, p in zip(k3.shape, pad))
            g_field = ops.run_program(L.PROG_LINEAR, [grad], [flipped], [adj_pad])
        if ctx.kern_shape is not None and ctx.needs_input_grad[1]:
            (x4,) = ctx.saved_tensors
            xp = F.pad(x4.detach(), (pad[2], pad[2], pad[1], pad[1], pad[0], pad[0]))
            O2, O1, O0 = grad.shape[1:]
            gk = torch.empty(k3.shape, dtype=torch.float32, device=grad.device)
            for a in range(k3.shape[0]):
                for b in range(k3.shape[1]):
                    for c in range(k3.shape[2]):
                        gk[a, b, c] = (xp[:, a:a + O2, b:b + O1, c:c + O0] * grad).sum()
            g_kern = gk.reshape(ctx.kern_shape).to(ctx.kern_device)
        return g_field, g_kern, None, None


def differentiable_apply(op, x4: torch.Tensor) -> torch.Tensor:
    """``op``'s kernel applied to a library-layout field [B,N2,N1,N0] on the autograd graph (w.r.t. the
    field and, if it requires grad, the kernel tensor)."""
    kern = op.kernel
    k3 = op._fl.kernel3(op._taps(kern))
    pad = tuple(s // 2 for s in k3.shape) if op._fl.same_padding else (0, 0, 0)
    return _LinearStencilFn.apply(x4, kern if kern.requires_grad else None, k3, pad)


class StencilOperator:
    """Base of the drop-in ``ConvOperator`` classes."""

    _flavour = "2d"

    # --- construction ---------------------------------------------------------------
    def _select_conv(self, conv):
        if conv == "direct":
            self.conv = self.convolution
        elif conv == "spectral":
            self.conv = self.spectral_convolution
        else:
            raise ValueError("Unknown Convolution Method")

    # --- helpers ----------------------------------------------------------------------
    @property
    def _fl(self) -> Flavour:
        return FLAVOURS[self._flavour]

    def _prep_field(self, field):
        """-> (CUDA fp32 field with the flavour's canonical ndim, Staged, squeeze_channel)."""
        if not isinstance(field, torch.Tensor):
            raise TypeError("conv: argument 'input' must be Tensor")
        fl = self._fl
        staged = ops.Staged(field)
        squeeze = False
        if fl.name in ("1d", "0d"):
            if field.dim() == fl.field_ndim + 1:
                # the reference passes an already channelled tensor straight to conv; only a
                # single channel is meaningful for a [1,1,...] weight
                if field.shape[1] != 1:
                    raise RuntimeError(
                        f"expected input to have 1 channels, but got {field.shape[1]} channels instead")
                field = field[:, 0]
                squeeze = True        # .squeeze(1) of the reference removes it again
            elif field.dim() != fl.field_ndim:
                raise RuntimeError(f"Expected {fl.field_ndim}D or {fl.field_ndim + 1}D input, got {field.dim()}D")
        elif fl.name == "2d":
            if field.dim() != 4:
                raise RuntimeError(f"Expected 4D (BS, Nt, Nx, Ny) input to conv3d, got {field.dim()}D")
        else:
            if field.dim() != 4:
                raise RuntimeError(f"Expected 4D (BS, 1, Nx, Ny) input to conv2d, got {field.dim()}D")
            if field.shape[1] != 1:
                raise RuntimeError(
                    f"expected input to have 1 channels, but got {field.shape[1]} channels instead")
        if field.dtype != torch.float32:
            raise RuntimeError(f"expected scalar type Float but found {field.dtype}")
        return field, staged, squeeze

    def _kernel(self, kernel=None):
        if kernel is not None:
            self.kernel = kernel            # sticky overwrite [ref: Utils/ConvOps_2d.py:146-147]
        return self.kernel                  # AttributeError when construction failed, like the reference

    def _taps(self, kern):
        """Host taps of ``kern``; re-read whenever the tensor object or its version changes
        (``.kernel`` is public and mutable), cached otherwise so a CUDA-resident kernel does not
        cost a device-to-host sync per call."""
        hit = getattr(self, "_tap_cache", None)
        if hit is not None and hit[0]() is kern and hit[1] == kern._version:
            return hit[2]
        k = taps_of(kern).copy()
        self._tap_cache = (weakref.ref(kern), kern._version, k)
        return k

    # --- a4: direct path ----------------------------------------------------------------
    def convolution(self, field, kernel=None):
        kern = self._kernel(kernel)
        fl = self._fl
        k3 = fl.kernel3(self._taps(kern))
        if kern.dim() != fl.kernel_ndim:
            raise RuntimeError(f"weight of ndim {kern.dim()} does not fit a {fl.name} operator")
        x, staged, _ = self._prep_field(field)
        pad = tuple(s // 2 for s in k3.shape) if fl.same_padding else (0, 0, 0)
        if ops.needs_grad(x, kern):         # differentiable w.r.t. the field and the kernel tensor
            x4 = fl.field4(ops.to_cuda(x, keep_graph=True))
            out = _LinearStencilFn.apply(x4, kern if kern.requires_grad else None, k3, pad)
        else:
            out = ops.run_program(L.PROG_LINEAR, [fl.field4(ops.to_cuda(x))], [k3], [pad])
        return staged.back(fl.unfield4(out))

    # --- a5: spectral path ----------------------------------------------------------------
    def spectral_convolution(self, field, kernel=None, inverse=False):
        from .Utils.fft_conv_pytorch.fft_conv import fft_conv
        kern = self._kernel(kernel)
        fl = self._fl
        x, staged, squeeze = self._prep_field(field)
        if fl.name != "spatial":
            x = x.unsqueeze(1)
        pad = tuple(s // 2 for s in kern.shape)
        out = fft_conv(x, kern.unsqueeze(0).unsqueeze(0), padding=pad if len(pad) > 1 else pad[0],
                       inverse=inverse)
        if fl.name != "spatial":
            out = out.squeeze(1)
        return out                          # fft_conv already returns on the field's device

    # --- a6: manual spectral differentiate / integrate -------------------------------------
    def _spectral_manual(self, field, kernel, mode, correlation, slice_pad, eps, pad_field=True):
        import torch.nn.functional as F
        kern = self._kernel(kernel)
        fl = self._fl
        x, staged, _ = self._prep_field(field)
        grad_path = ops.needs_grad(x, kern)
        xc = ops.to_cuda(x, keep_graph=grad_path)
        if fl.name != "spatial":
            xc = xc.unsqueeze(1)
        nd = fl.kernel_ndim
        p = kern.size(-1) // 2
        padded = F.pad(xc, (p, p) * nd) if pad_field else xc
        dims = tuple(range(2, padded.ndim))
        k = ops.to_cuda(kern, keep_graph=grad_path)
        kpad = []
        for i in reversed(range(nd)):
            kpad += [0, padded.size(2 + i) - k.size(i)]
        kp = F.pad(k, kpad)
        f_hat = torch.fft.rfftn(padded, dim=dims)
        k_hat = torch.fft.rfftn(kp, dim=tuple(range(kp.ndim)))
        flags = 0
        if mode == "int":
            flags |= L.SPEC_INVERSE
            if correlation:
                flags |= L.SPEC_CONJ_LAST
        elif correlation:
            flags |= L.SPEC_CONJ_FIRST
        if grad_path:       # same arithmetic in torch complex ops: differentiable w.r.t. field and kernel
            g = k_hat.conj() if flags & L.SPEC_CONJ_FIRST else k_hat
            if flags & L.SPEC_INVERSE:
                g = 1.0 / (g + eps)
            if flags & L.SPEC_CONJ_LAST:
                g = g.conj()
            prod = f_hat * g
        else:
            prod = ops.spectral_pointwise(f_hat, k_hat, flags, eps)
        out = torch.fft.irfftn(prod, dim=dims)          # no size given, like the reference
        if slice_pad:
            crop = (slice(None), slice(None)) + tuple(
                slice(0, padded.size(2 + i) - k.size(i) + 1) for i in range(nd))
            out = out[crop].contiguous()
        if fl.name != "spatial":
            out = out.squeeze(1)
        return staged.back(out)

    def differentiate(self, field, kernel=None, correlation=False, slice_pad=True):
        return self._spectral_manual(field, kernel, "diff", correlation, slice_pad, 0.0)

    def forward(self, field):
        return self.conv(field, self.kernel)

    def __call__(self, inputs):
        return self.forward(inputs)
