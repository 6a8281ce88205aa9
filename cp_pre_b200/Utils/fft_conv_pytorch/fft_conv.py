"""Drop-in for ``Utils/fft_conv_pytorch/fft_conv.py``.

Same signature and semantics as the reference's ``fft_conv`` (which is fkodom/fft-conv-pytorch
plus an ``inverse=`` regularised deconvolution): zero/other padding of the signal, even-length
fix-up of the last axis, kernel zero-padded to the signal size, **cross-correlation**
(conjugated kernel spectrum), optional ``1/(K+1e-6)``, crop with stride, optional bias.

Execution: the FFTs are cuFFT via ``torch.fft`` (a real FFT workload); the complex temporaries
between them (conj, reciprocal, product) are one libcppre kernel (``cppre_spectral_pointwise``).
Forward, un-strided, single-channel requests -- everything ``ConvOperator.spectral_convolution``
issues -- are mathematically the direct correlation and are served by the stencil kernel K1
unless ``force_fft=True``: "the spectral-derivative path uses an FFT only where it really is one".
"""
from functools import partial
from math import ceil, floor
from typing import Iterable, Tuple, Union

import numpy as np
import torch
import torch.nn.functional as f
from torch import Tensor, nn
from torch.fft import irfftn, rfftn  # noqa: F401  (re-exported like the reference module)

from cp_pre_b200 import _lib as L
from cp_pre_b200 import ops


def to_ntuple(val: Union[int, Iterable[int]], n: int) -> Tuple[int, ...]:
    """``val`` repeated n times, or checked to have n entries."""
    if isinstance(val, Iterable):
        out = tuple(val)
        if len(out) != n:
            raise ValueError(f"Cannot cast tuple of length {len(out)} to length {n}.")
        return out
    return n * (val,)


def _direct_eligible(signal, kernel, bias, padding_, padding_mode, stride_, dilation_, inverse):
    n = signal.ndim - 2
    return (not inverse and bias is None and padding_mode == "constant" and n in (1, 2, 3)
            and signal.shape[1] == 1 and tuple(kernel.shape[:2]) == (1, 1)
            and all(s == 1 for s in stride_) and all(d == 1 for d in dilation_)
            and all(float(p).is_integer() and 0 <= p <= 6 for p in padding_)
            and all(1 <= k <= 7 for k in kernel.shape[2:])
            and signal.dtype == torch.float32 and kernel.dtype == torch.float32)


def fft_conv(signal: Tensor, kernel: Tensor, bias: Tensor = None,
             padding: Union[int, Iterable[int], str] = 0, padding_mode: str = "constant",
             stride: Union[int, Iterable[int]] = 1, dilation: Union[int, Iterable[int]] = 1,
             groups: int = 1, inverse: bool = False, force_fft: bool = False) -> Tensor:
    """N-d cross-correlation through the convolution theorem.

    signal [B, C, *spatial], kernel [Co, Ci, *k].  Like the reference, channels are never
    contracted (the spectra are multiplied with plain broadcasting) and ``groups`` is accepted
    but unused.  Returns a tensor on ``signal``'s device.
    """
    n = signal.ndim - 2
    stride_ = to_ntuple(stride, n=n)
    dilation_ = to_ntuple(dilation, n=n)
    if isinstance(padding, str):
        if padding != "same":
            raise ValueError(f"Padding mode {padding} not supported.")
        if stride != 1 or dilation != 1:
            raise ValueError("stride must be 1 for padding='same'.")
        padding_ = [(k - 1) / 2 for k in kernel.shape[2:]]
    else:
        padding_ = to_ntuple(padding, n=n)

    home = signal.device
    # autograd: the fused kernels carry no graph, so a call whose signal / kernel / bias requires grad
    # (an _FFTConv layer being trained, a PRE loss through the spectral path) runs the same arithmetic
    # in torch ops (cuFFT + complex elementwise), which is differentiable in all three
    grad_path = ops.needs_grad(signal, kernel, bias)
    if not force_fft and not grad_path and _direct_eligible(signal, kernel, bias, padding_, padding_mode, stride_, dilation_,
                                          inverse):
        # forward spectral == direct correlation (relation R2): one pass of K1, no FFT
        x = ops.to_cuda(signal)[:, 0]
        k = kernel.detach().cpu().numpy()[0, 0].astype(np.float32)
        x4 = x.reshape((x.shape[0],) + (1,) * (3 - n) + tuple(x.shape[1:]))
        k3 = k.reshape((1,) * (3 - n) + k.shape)
        pad3 = (0,) * (3 - n) + tuple(int(p) for p in padding_)
        out = ops.run_program(L.PROG_LINEAR, [x4], [k3], [pad3])
        out = out.reshape((out.shape[0], 1) + tuple(out.shape[1 + (3 - n):]))
        return out if out.device == home else out.to(home)

    sig = ops.to_cuda(signal, keep_graph=grad_path)
    ker = ops.to_cuda(kernel, keep_graph=grad_path)
    # dilation: spread the taps with zeros (kron with a one-hot offset), trimming the tail
    if any(d != 1 for d in dilation_):
        offset = torch.zeros(1, 1, *dilation_, device=sig.device, dtype=sig.dtype)
        offset[(slice(None), slice(None), *((0,) * n))] = 1.0
        cutoff = tuple(slice(None, -d + 1 if d != 1 else None) for d in dilation_)
        ker = torch.kron(ker, offset)[(slice(None), slice(None)) + cutoff]
    sig_pad = [r(p) for p in padding_[::-1] for r in (floor, ceil)]
    sig = f.pad(sig, sig_pad, mode=padding_mode)
    size = sig.size()                       # before the even-length fix-up
    if sig.size(-1) % 2 != 0:
        sig = f.pad(sig, [0, 1])
    kpad = [v for i in reversed(range(2, sig.ndim)) for v in (0, sig.size(i) - ker.size(i))]
    dims = tuple(range(2, sig.ndim))
    sig_hat = rfftn(sig.float(), dim=dims)
    ker_hat = rfftn(f.pad(ker, kpad).float(), dim=dims)
    flags = L.SPEC_CONJ_FIRST | (L.SPEC_INVERSE if inverse else 0)
    if grad_path:
        gk = ker_hat.conj()
        if inverse:
            gk = 1.0 / (gk + 1e-6)
        prod = sig_hat * gk
    elif ker_hat.shape[0] == 1 and ker_hat.shape[1] == 1:
        prod = ops.spectral_pointwise(sig_hat, ker_hat[0, 0], flags, 1e-6)
    else:
        # multi-channel kernels broadcast against the signal spectrum like the reference's
        # `signal_fr * kernel_fr`; g(K) is evaluated once by the kernel, the broadcast by torch
        flat = ker_hat.reshape(-1)
        gk = ops.spectral_pointwise(torch.ones_like(flat).reshape(1, -1), flat, flags, 1e-6)
        prod = sig_hat * gk.reshape(ker_hat.shape)
    out = irfftn(prod, dim=dims)
    crop = (slice(None), slice(None)) + tuple(
        slice(0, size[i] - ker.size(i) + 1, stride_[i - 2]) for i in range(2, sig.ndim))
    out = out[crop].contiguous()
    if bias is not None:
        out = out + ops.to_cuda(bias, keep_graph=grad_path).view((1, -1) + (1,) * n)
    return out if out.device == home else out.to(home)


class _FFTConv(nn.Module):
    """Layer form of :func:`fft_conv` (random weights, like ``nn.ConvNd``)."""

    def __init__(self, in_channels: int, out_channels: int, kernel_size: Union[int, Iterable[int]],
                 padding: Union[int, Iterable[int]] = 0, padding_mode: str = "constant",
                 stride: Union[int, Iterable[int]] = 1, dilation: Union[int, Iterable[int]] = 1,
                 groups: int = 1, bias: bool = True, ndim: int = 1, inverse: bool = False):
        super().__init__()
        for name, ch in (("in_channels", in_channels), ("out_channels", out_channels)):
            if ch % groups != 0:
                raise ValueError(f"'{name}' must be divisible by 'groups'."
                                 f"Found: {name}={ch}, groups={groups}.")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.padding, self.padding_mode = kernel_size, padding, padding_mode
        self.stride, self.dilation, self.groups = stride, dilation, groups
        self.use_bias, self.inverse = bias, inverse
        self.weight = nn.Parameter(torch.randn(out_channels, in_channels // groups,
                                               *to_ntuple(kernel_size, ndim)))
        self.bias = nn.Parameter(torch.randn(out_channels)) if bias else None

    def forward(self, signal):
        return fft_conv(signal, self.weight, bias=self.bias, padding=self.padding,
                        padding_mode=self.padding_mode, stride=self.stride, dilation=self.dilation,
                        groups=self.groups, inverse=self.inverse)


FFTConv1d = partial(_FFTConv, ndim=1)
FFTConv2d = partial(_FFTConv, ndim=2)
FFTConv3d = partial(_FFTConv, ndim=3)
