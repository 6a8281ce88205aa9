"""Tensor-level wrappers over the libcppre C ABI.

PyTorch is plumbing here: it owns device memory and streams; every computation is one of the
library's CUDA kernels, enqueued on the caller's current stream.  Inputs living on the host
(CPU tensors, NumPy arrays -- the reference's default) are staged to the
current CUDA device and results are brought back, so the reference-facing API is a drop-in; a
host without CUDA raises -- there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib as L


# --------------------------------------------------------------------------------------
# device plumbing
# --------------------------------------------------------------------------------------
def require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("cp_pre_b200 needs a CUDA device (B200, sm_100a): it has no CPU fallback")


def _stream(device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _ptr(t: Optional[torch.Tensor]):
    return C.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else C.c_void_p(0)


class Staged:
    """Remembers where a user array came from so results can be returned the same way."""

    def __init__(self, x):
        self.numpy = isinstance(x, np.ndarray) or np.isscalar(x)
        self.device = None if self.numpy else (x.device if isinstance(x, torch.Tensor) else None)

    def back(self, t: torch.Tensor):
        if self.numpy:
            return t.detach().cpu().numpy()
        if self.device is not None and t.device != self.device:
            return t.to(self.device)
        return t


def needs_grad(*xs) -> bool:
    """True when autograd is recording and one of ``xs`` is a tensor that requires grad: such calls
    take the differentiable paths (the fused kernels have no graph), never a silent detach."""
    return torch.is_grad_enabled() and any(isinstance(x, torch.Tensor) and x.requires_grad for x in xs)


def to_cuda(x, dtype=torch.float32, keep_graph: bool = False) -> torch.Tensor:
    """Host array / CPU tensor -> current CUDA device (H2D); CUDA tensor -> itself.  Detached unless
    ``keep_graph`` (the differentiable paths), in which case dtype / device moves stay on the graph."""
    require_cuda()
    if isinstance(x, torch.Tensor):
        t = x if (keep_graph and needs_grad(x)) else x.detach()
    else:
        t = torch.from_numpy(np.ascontiguousarray(x))
    if t.dtype != dtype:
        t = t.to(dtype)
    if not t.is_cuda:
        t = t.cuda(non_blocking=False)
    return t


def _rows(t: torch.Tensor):
    """[n, ...] -> ([n, cells] with unit cell stride, row stride)."""
    n = t.shape[0]
    flat = t.reshape(n, -1)
    if flat.numel() and flat.shape[1] > 1 and flat.stride(1) != 1:
        flat = flat.contiguous()
    rs = flat.stride(0) if flat.shape[0] > 1 else (flat.shape[1] if flat.dim() > 1 else 1)
    return flat, int(rs)


# --------------------------------------------------------------------------------------
# K1: residual programs
# --------------------------------------------------------------------------------------
def _axis_order(t: torch.Tensor):
    """Permutation of the logical axes (1,2,3) that puts the largest stride first."""
    strides = [abs(t.stride(d)) if t.shape[d] > 1 else -1 for d in (1, 2, 3)]
    # extent-1 axes keep their place; the others are ordered by stride
    real = [d for d in range(3) if strides[d] >= 0]
    order = sorted(real, key=lambda d: -strides[d])
    perm = list(range(3))
    for slot, d in zip(real, order):
        perm[slot] = d
    return perm


def _to_logical_order(f: torch.Tensor, perm) -> torch.Tensor:
    """A contiguous [B,N2,N1,N0] copy of a permuted view.  The FNO layout -- memory order
    (B, N1, N0, N2), every sample one contiguous block -- goes through the library's tiled
    transpose (cppre_transpose_inner); anything else through torch's copy."""
    B, N2, N1, N0 = (int(v) for v in f.shape)
    M = N1 * N0
    if (perm == [1, 2, 0] and f.dtype == torch.float32 and N2 <= 256 and B < 65536 and f.stride(1) == 1
            and f.stride(3) == N2 and f.stride(2) == N0 * N2 and (B == 1 or f.stride(0) >= M * N2)):
        out = torch.empty((B, N2, N1, N0), dtype=torch.float32, device=f.device)
        with torch.cuda.device(f.device):
            L.check(L.load().cppre_transpose_inner(_ptr(f), int(f.stride(0)) if B > 1 else M * N2, B, M, N2, _ptr(out),
                                                   _stream(f.device)))
        return out
    return f.contiguous()


def _is_tin(t: torch.Tensor) -> bool:
    """A logical [B,Nt,Nx,Ny] view of [B,Nx,Ny,Nt] memory: what ``pred.permute(0,1,4,2,3)[:, j]`` of an FNO
    output is."""
    if t.dim() != 4 or t.numel() == 0:
        return False
    B, N2, N1, N0 = t.shape
    return (N2 > 1 and t.stride(1) == 1 and (N0 == 1 or t.stride(3) == N2) and (N1 == 1 or t.stride(2) == N0 * N2)
            and (B == 1 or t.stride(0) >= N1 * N0 * N2))


def _as_tin(t: torch.Tensor) -> torch.Tensor:
    """The same logical [B,Nt,Nx,Ny] values stored t-innermost (a copy unless it already is)."""
    return t if _is_tin(t) else t.permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)


def _residual_call(kind, fields, kernels, pads, scalars, mode, flags, reduce, ref, weight, weight_stride, path, out,
                   band_q, query_tin=False, decode=None):
    """Fills the C structs for one cppre_residual call and makes it (or, ``query_tin``, only asks whether the
    t-innermost kernel would take it).  ``out`` is the STORE destination or None for the reducing modes."""
    lib = L.load()
    f0 = fields[0]
    B, N2, N1, N0 = (int(v) for v in f0.shape)
    prog = L.Program()
    prog.kind, prog.n_fields, prog.n_slots = kind, len(fields), len(kernels)
    for s, (k, p) in enumerate(zip(kernels, pads)):
        for d in range(3):
            prog.ksize[s][d] = k.shape[d]
            prog.pad[s][d] = p[d]
        prog.h_taps[s] = k.ctypes.data
    for i, v in enumerate(scalars):
        prog.scalars[i] = float(v)
    fl = (L.Field * len(fields))()
    for i, f in enumerate(fields):
        fl[i].d_data = f.data_ptr() if f.numel() else 0
        for d in range(4):
            fl[i].stride[d] = f.stride(d)
        fl[i].scale, fl[i].offset = (1.0, 0.0) if decode is None else (float(decode[0][i]), float(decode[1][i]))
    epi = L.Epilogue()
    epi.mode, epi.flags, epi.reduce = mode, flags | (L.FLAG_DECODE if decode is not None else 0), reduce
    dev = f0.device
    if isinstance(band_q, torch.Tensor):
        if band_q.numel() != 1 or band_q.dtype != torch.float32 or band_q.device != dev:
            raise ValueError(f"band_q must be a Python float or a one-element float32 tensor on {dev}")
        epi.d_band_q = band_q.data_ptr()
    else:
        epi.band_q = float(band_q)
    if ref is not None:
        epi.d_ref = ref.data_ptr()
        for d in range(4):
            epi.ref_stride[d] = ref.stride(d)
    if weight is not None:
        epi.d_weight = weight.data_ptr()
        if weight_stride is not None:
            for d in range(3):
                epi.weight_stride[d] = int(weight_stride[d])
    res = None
    with torch.cuda.device(dev):
        if mode == L.EPI_STORE:
            epi.d_out = out.data_ptr() if out.numel() else 0
            for d in range(4):
                epi.out_stride[d] = out.stride(d)
            res = out
        else:
            res = torch.empty((B,), dtype=torch.float32 if mode == L.EPI_SCORE else torch.int32, device=dev)
            epi.d_scores = res.data_ptr() if B else 0
        shape = (C.c_int64 * 4)(B, N2, N1, N0)
        if query_tin:
            return bool(lib.cppre_residual_tin_eligible(C.byref(prog), shape, fl, C.byref(epi)))
        L.check(lib.cppre_residual(C.byref(prog), shape, fl, C.byref(epi), int(path), _stream(dev)))
    return res


def run_program(kind: int, fields: Sequence[torch.Tensor], kernels: Sequence[np.ndarray],
                pads: Sequence[Sequence[int]], scalars: Sequence[float] = (), *,
                mode: int = L.EPI_STORE, flags: int = 0, reduce: int = L.REDUCE_MAX,
                ref: Optional[torch.Tensor] = None, weight: Optional[torch.Tensor] = None,
                path: int = L.PATH_AUTO, canonicalise: bool = True,
                out: Optional[torch.Tensor] = None, band_q=0.0, decode=None) -> torch.Tensor:
    """Runs one fused residual program.

    fields   CUDA fp32 tensors, all of logical shape [B, N2, N1, N0], any strides (views welcome)
    kernels  one dense fp32 ndarray [k2, k1, k0] per derivative slot (cross-correlation taps)
    pads     per-slot zero padding (p2, p1, p0)
    Returns the [B, O2, O1, O0] field (STORE), the [B] fp32 scores (SCORE) or the [B] int32
    counts of cells with ref - band_q*weight <= residual <= ref + band_q*weight (COVER).
    ``band_q`` is a Python float or a one-element float32 CUDA tensor (e.g. the device scalar
    ``radix_select`` returns): the kernel then reads q itself and nothing crosses to the host.
    ``decode`` = (scales, offsets), one float per field: the program reads scale_f * x + offset_f (the
    normaliser-decode fold, CPPRE_FLAG_DECODE in include/cppre.h; exact on the interior cells only).
    """
    L.load()
    require_cuda()
    f0 = fields[0]
    if any(f.shape != f0.shape for f in fields):
        raise ValueError("all fields of a residual program must share one shape")
    bf16 = f0.dtype == torch.bfloat16          # bf16-input mode (include/cppre.h, CPPRE_FLAG_IN_BF16)
    if decode is not None:
        if bf16:
            raise NotImplementedError("the decode fold takes float32 fields only")
        if len(decode) != 2 or len(decode[0]) != len(fields) or len(decode[1]) != len(fields):
            raise ValueError("decode must be (scales, offsets) with one value per field")
    if any(f.dtype != f0.dtype for f in fields) or f0.dtype not in (torch.float32, torch.bfloat16):
        raise RuntimeError("expected scalar type Float (fields must all be float32, or all bfloat16)")
    kernels = [np.ascontiguousarray(k, dtype=np.float32) for k in kernels]
    pads = [tuple(int(v) for v in p) for p in pads]
    for k in kernels:
        if k.ndim != 3:
            raise ValueError("slot kernels must be 3-D [k2,k1,k0]")
    if f0.dim() != 4:
        raise ValueError("fields must be [B, N2, N1, N0]")
    B, N2, N1, N0 = (int(v) for v in f0.shape)
    O = [n + 2 * p - k + 1 for n, p, k in zip((N2, N1, N0), pads[0], kernels[0].shape)]
    if min(O) < 0:
        raise RuntimeError(f"kernel size {tuple(kernels[0].shape)} can't be greater than the padded input {(N2, N1, N0)}")
    if ref is not None and tuple(ref.shape) != (B, *O):
        raise ValueError(f"reference residual has shape {tuple(ref.shape)}, expected {(B, *O)}")
    if weight is not None and tuple(weight.shape) != tuple(O):
        raise ValueError(f"weight has shape {tuple(weight.shape)}, expected {tuple(O)}")
    if out is not None and mode == L.EPI_STORE and (tuple(out.shape) != (B, *O) or out.dtype != torch.float32
                                                    or out.device != f0.device):
        raise ValueError(f"out must be a float32 {(B, *O)} tensor on {f0.device}")
    eflags = flags | (L.FLAG_IN_BF16 if bf16 else 0)

    # t-innermost ingestion: [B,Nt,Nx,Ny] views of [B,Nx,Ny,Nt] memory -- FNO outputs
    # handed on as pred.permute(0,1,4,2,3) -- are processed IN PLACE by the t-innermost star kernel.  The
    # result, the reference residual and the per-cell weight use the same physical order; the caller sees
    # logical shapes throughout.
    if (canonicalise and path in (L.PATH_AUTO, L.PATH_TIN) and not bf16 and tuple(O) == (N2, N1, N0)
            and all(_is_tin(f) for f in fields) and (out is None or mode != L.EPI_STORE or _is_tin(out))):
        t_ref = None if ref is None else _as_tin(ref)
        t_w, t_ws = None, None
        if weight is not None:
            t_w = weight.permute(1, 2, 0).contiguous()                    # physical [Nx][Ny][Nt]
            t_ws = (1, N0 * N2, N2)
        t_out = out
        if mode == L.EPI_STORE and t_out is None:
            t_out = torch.empty((B, N1, N0, N2), dtype=torch.float32, device=f0.device).permute(0, 3, 1, 2)
        args = (kind, list(fields), kernels, pads, scalars, mode, eflags, reduce, t_ref, t_w, t_ws)
        if _residual_call(*args, L.PATH_TIN, t_out, band_q, query_tin=True, decode=decode):
            return _residual_call(*args, L.PATH_TIN, t_out, band_q, decode=decode)
        if path == L.PATH_TIN:
            raise NotImplementedError("request does not qualify for the t-innermost star kernel")

    # Physical-layout canonicalisation: a permuted view such
    # as pred.permute(0,1,4,2,3) is processed in its memory order -- fields, taps, padding,
    # weight and reference are permuted alike and the result is returned as the inverse
    # permuted view, so no .contiguous() copy of the input is ever made.
    perm = [0, 1, 2]
    if canonicalise and f0.numel() > 0 and all(_axis_order(f) == _axis_order(f0) for f in fields):
        perm = _axis_order(f0)
    if perm != [0, 1, 2]:
        # Memory order pays only when its innermost axis can fill the 128-column TMA tiles.  Views that the
        # t-innermost kernel did not take (odd Nt, Taylor-4/6 taps, bf16) are copied once: the FNO layout
        # through the library's tiled transpose, anything else through torch.
        n0_mem, n0_log = int(f0.shape[1 + perm[2]]), int(f0.shape[3])
        if (n0_mem % 4 != 0 or n0_mem < 128) and n0_log % 4 == 0 and n0_log >= n0_mem:
            fields = [_to_logical_order(f, perm) for f in fields]
            f0 = fields[0]
            perm = [0, 1, 2]
    if perm != [0, 1, 2]:
        tperm = [0] + [1 + d for d in perm]
        fields = [f.permute(tperm) for f in fields]
        kernels = [np.ascontiguousarray(np.transpose(k, perm)) for k in kernels]
        pads = [tuple(p[d] for d in perm) for p in pads]
        if ref is not None:
            ref = ref.permute(tperm)
        if weight is not None:
            weight = weight.permute(perm).contiguous()
        bit = {0: L.FLAG_INTERIOR_2, 1: L.FLAG_INTERIOR_1, 2: L.FLAG_INTERIOR_0}
        nf = eflags & (L.FLAG_ABS | L.FLAG_IN_BF16)
        for new_axis, old_axis in enumerate(perm):
            if eflags & bit[old_axis]:
                nf |= bit[new_axis]
        eflags = nf
        f0 = fields[0]
        if out is not None:
            out = out.permute(tperm)

    B, N2, N1, N0 = (int(v) for v in f0.shape)
    O = [n + 2 * p - k + 1 for n, p, k in zip((N2, N1, N0), pads[0], kernels[0].shape)]
    if weight is not None:
        weight = weight.contiguous()
    if mode == L.EPI_STORE and out is None:
        out = torch.empty((B, *O), dtype=torch.float32, device=f0.device)
    res = _residual_call(kind, list(fields), kernels, pads, scalars, mode, eflags, reduce, ref, weight, None, path,
                         out, band_q, decode=decode)
    if mode == L.EPI_STORE and perm != [0, 1, 2]:
        inv = [0] * 3
        for new_axis, old_axis in enumerate(perm):
            inv[old_axis] = new_axis
        res = res.permute([0] + [1 + d for d in inv])
    return res


# --------------------------------------------------------------------------------------
# K2 / K3 / K5: conformal calibration
# --------------------------------------------------------------------------------------
def cell_moments(a: torch.Tensor, b: Optional[torch.Tensor] = None):
    """Per-cell (mean, M2) in fp64 of a-b over axis 0; a, b CUDA fp32 [n, ...]."""
    lib = L.load()
    fa, sa = _rows(a)
    fb, sb = (None, 0) if b is None else _rows(b)
    n, cells = fa.shape
    mean = torch.empty(cells, dtype=torch.float64, device=a.device)
    m2 = torch.empty(cells, dtype=torch.float64, device=a.device)
    with torch.cuda.device(a.device):
        L.check(lib.cppre_cell_moments(_ptr(fa), sa, _ptr(fb), sb, n, cells, _ptr(mean), _ptr(m2),
                                       _stream(a.device)))
    return mean, m2


def cell_std_from_m2(m2: torch.Tensor, count: float, eps: float = 0.0, want_inv: bool = False):
    lib = L.load()
    cells = m2.numel()
    sd = torch.empty(cells, dtype=torch.float32, device=m2.device)
    inv = torch.empty(cells, dtype=torch.float32, device=m2.device) if want_inv else None
    with torch.cuda.device(m2.device):
        L.check(lib.cppre_cell_std(_ptr(m2), float(count), float(eps), cells, _ptr(sd), _ptr(inv),
                                   _stream(m2.device)))
    return (sd, inv) if want_inv else sd


def cell_std(a: torch.Tensor, b: Optional[torch.Tensor] = None, eps: float = 0.0, want_inv: bool = False):
    """Population std (ddof = 0) over axis 0 of a-b, shaped like a[0]."""
    if a.shape[0] == 0:
        raise ValueError("std over an empty sample axis")
    _, m2 = cell_moments(a, b)
    res = cell_std_from_m2(m2, a.shape[0], eps, want_inv)
    if want_inv:
        return res[0].reshape(a.shape[1:]), res[1].reshape(a.shape[1:])
    return res.reshape(a.shape[1:])


def cell_modulation(a: torch.Tensor, b: Optional[torch.Tensor] = None, eps: float = 0.0):
    """(std + eps, 1/(std + eps)) of a-b over axis 0 in ONE launch (single-GPU modulation_func;
    same arithmetic as cell_moments + cell_std_from_m2), shaped like a[0]."""
    if a.shape[0] == 0:
        raise ValueError("std over an empty sample axis")
    lib = L.load()
    fa, sa = _rows(a)
    fb, sb = (None, 0) if b is None else _rows(b)
    n, cells = fa.shape
    sd = torch.empty(cells, dtype=torch.float32, device=a.device)
    inv = torch.empty(cells, dtype=torch.float32, device=a.device)
    with torch.cuda.device(a.device):
        L.check(lib.cppre_cell_modulation(_ptr(fa), sa, _ptr(fb), sb, n, cells, float(eps), _ptr(sd), _ptr(inv),
                                          _stream(a.device)))
    return sd.reshape(a.shape[1:]), inv.reshape(a.shape[1:])


def joint_score(a: torch.Tensor, b: Optional[torch.Tensor], mod: torch.Tensor) -> torch.Tensor:
    lib = L.load()
    fa, sa = _rows(a)
    fb, sb = (None, 0) if b is None else _rows(b)
    n, cells = fa.shape
    m = mod.reshape(-1).contiguous()
    if m.numel() != cells:
        raise ValueError(f"modulation has {m.numel()} cells, residuals have {cells}")
    out = torch.empty(n, dtype=torch.float32, device=a.device)
    with torch.cuda.device(a.device):
        L.check(lib.cppre_joint_score(_ptr(fa), sa, _ptr(fb), sb, _ptr(m), n, cells, _ptr(out),
                                      _stream(a.device)))
    return out


SEGMENT_SLOTS = 60          # candidate slots per segment and cell of the segmented select (as in the multi-GPU exchange)


def _segmented_select(fs: torch.Tensor, rank: int):
    """One order statistic near the TOP of columns longer than the register networks hold (n > 2048), on one
    GPU, with the kernels of the multi-GPU candidate-window exchange (dist._window_quantile) and the row
    segments of the column in the role of the ranks: the J-th largest of every segment (K3t, J = ceil(kt/S) <= 64)
    brackets the kt-th largest of the whole column; one more pass collects the few samples inside the bracket
    (K3w) and the answer is picked among them (select_from_top).  Two passes over the scores instead of the nine
    of the radix select.  Returns None when the request does not fit (the caller then takes the radix kernel);
    cells whose bracket overflowed the candidate slots are redone exactly with the radix kernel."""
    n, cells = fs.shape
    kt = n - int(rank)                           # the kt-th largest
    if kt < 1 or kt > 1024:
        return None
    S = max(2, -(-kt // 64))                     # segments: J = ceil(kt / S) <= 64
    if S > 16 or n // S <= 256:                  # owner network 64 * S <= 1024; segments long enough for the top-K kernel
        return None
    J = -(-kt // S)
    bounds = [(i * n) // S for i in range(S + 1)]
    segs = [fs[bounds[i]:bounds[i + 1]] for i in range(S)]
    keys = torch.stack([ordered_keys(select_ranks(seg, [seg.shape[0] - J])[0]) for seg in segs])
    lo, hi = keys.min(0).values, keys.max(0).values
    dev = fs.device
    over = torch.zeros(1, dtype=torch.int32, device=dev)
    over_cells = torch.full((8192,), -1, dtype=torch.int32, device=dev)
    wins, cnt = [], None
    for seg in segs:
        w, c = window_compact(seg, lo, hi, SEGMENT_SLOTS, cells, over, over_cells)      # one block: slab = all cells
        wins.append(w[0])
        cnt = c if cnt is None else cnt + c
    net = 32
    while net < 64 * S and net < S * SEGMENT_SLOTS:
        net *= 2
    q = select_from_top(torch.cat(wins), cnt.to(torch.int32), kt, over, over_cells, 0, min(net, 1024))
    n_over = int(over.item())
    if n_over:
        if n_over > over_cells.numel():
            return None
        ids = torch.unique(over_cells[over_cells >= 0]).to(torch.int64)
        q[ids] = _select_ranks_native(fs[:, ids].contiguous(), [int(rank)])[0]
    return q


def select_ranks(scores: torch.Tensor, ranks: Sequence[int]) -> torch.Tensor:
    """[len(ranks), *scores.shape[1:]]: the ranks[r]-th smallest along axis 0 (NaN-aware)."""
    fs, rs = _rows(scores)
    n, cells = fs.shape
    ranks = [int(r) for r in ranks]
    # long columns: above 1024 samples the register sort runs at < 200 Gscores/s and above 2048 only the radix
    # select is left; a calibration rank (near the top) is served by the segmented window select instead
    if n > 1024 and cells >= 32 and len(ranks) <= 4 and os.environ.get("CPPRE_SEGMENTED_SELECT", "1") != "0":
        parts = [_segmented_select(fs, r) for r in ranks]
        if all(p is not None for p in parts):
            return torch.stack(parts).reshape((len(ranks),) + tuple(scores.shape[1:]))
    return _select_ranks_native(scores, ranks)


def _select_ranks_native(scores: torch.Tensor, ranks: Sequence[int]) -> torch.Tensor:
    """cppre_select_ranks: top-K streaming / register sort / radix select, chosen by the library."""
    lib = L.load()
    fs, rs = _rows(scores)
    n, cells = fs.shape
    ranks = [int(r) for r in ranks]
    out = torch.empty((len(ranks), cells), dtype=torch.float32, device=scores.device)
    arr = (C.c_int32 * max(len(ranks), 1))(*ranks)
    with torch.cuda.device(scores.device):
        L.check(lib.cppre_select_ranks(_ptr(fs), rs, n, cells, arr, len(ranks), _ptr(out), None,
                                       _stream(scores.device)))
    return out.reshape((len(ranks),) + tuple(scores.shape[1:]))


def ordered_keys(x: torch.Tensor) -> torch.Tensor:
    """float32 -> int32 keys whose signed order is the float order (NaN with a clear sign bit on top)."""
    lib = L.load()
    xc = x.contiguous()
    out = torch.empty(xc.shape, dtype=torch.int32, device=x.device)
    with torch.cuda.device(x.device):
        L.check(lib.cppre_ordered_keys(_ptr(xc), xc.numel(), _ptr(out), _stream(x.device)))
    return out


def window_compact(scores: torch.Tensor, lo: torch.Tensor, hi: torch.Tensor, slots: int, slab: int,
                   overflow: torch.Tensor, overflow_cells: Optional[torch.Tensor] = None,
                   peer_ptrs: Optional[Sequence[int]] = None, my_rank: int = 0):
    """Second pass of the multi-GPU marginal quantile (K3w).  scores [n, cells...]; lo / hi int32
    ordered keys per cell.  Returns (win [blocks, slots, slab] float32, cnt [blocks*slab] int32);
    ``overflow`` (int32 [1], caller-zeroed) is incremented per cell with more than ``slots``
    window samples; their indices are written to ``overflow_cells`` (int32 [cap]) at the counter's
    old value while that is below cap.  With ``peer_ptrs`` (every rank's receive buffer
    [world, slots, slab], mapped here) block r is stored straight into rank r's buffer at
    [my_rank] -- the exchange is fused into the kernel -- and ``win`` is returned as None."""
    lib = L.load()
    fs, rs = _rows(scores)
    n, cells = fs.shape
    cap = 0 if overflow_cells is None else overflow_cells.numel()
    blocks = (cells + slab - 1) // slab
    fused = peer_ptrs is not None
    win = None if fused else torch.empty((blocks, slots, slab), dtype=torch.float32, device=scores.device)
    cnt = torch.zeros(blocks * slab, dtype=torch.int32, device=scores.device)
    if blocks * slab != cells:      # padding tail of the last slab: no candidates; flagged like a NaN cell
        if not fused:               # (fused: the owner never reads a flagged cell's candidates)
            win[-1, :, cells - (blocks - 1) * slab:] = float("-inf")
        cnt[cells:] = 1 << 24
    peers = (C.c_void_p * len(peer_ptrs))(*peer_ptrs) if fused else None
    with torch.cuda.device(scores.device):
        L.check(lib.cppre_window_compact(_ptr(fs), rs, n, cells, _ptr(lo.reshape(-1).contiguous()),
                                         _ptr(hi.reshape(-1).contiguous()), int(slots), int(slab), _ptr(win),
                                         _ptr(cnt), _ptr(overflow), _ptr(overflow_cells), cap, peers,
                                         len(peer_ptrs) if fused else 0, int(my_rank), _stream(scores.device)))
    return win, cnt


def select_from_top(cand: torch.Tensor, cnt: torch.Tensor, kt: int, overflow: torch.Tensor,
                    overflow_cells: Optional[torch.Tensor] = None, cell_offset: int = 0,
                    network: int = 0) -> torch.Tensor:
    """out[c] = the (kt - (cnt[c] & 0xffffff))-th largest of cand[:, c]; NaN where cnt[c] >> 24.
    ``network`` > 0: pack the non-padding candidates of each cell and sort that many keys only.
    Cells that cannot be answered are counted in ``overflow`` and listed (cell_offset + c)."""
    lib = L.load()
    fc, rs = _rows(cand)
    rows, cells = fc.shape
    cap = 0 if overflow_cells is None else overflow_cells.numel()
    out = torch.empty(cells, dtype=torch.float32, device=cand.device)
    with torch.cuda.device(cand.device):
        L.check(lib.cppre_select_from_top(_ptr(fc), rs, rows, cells, _ptr(cnt.contiguous()), int(kt), int(network),
                                          _ptr(out), _ptr(overflow), _ptr(overflow_cells), cap, int(cell_offset),
                                          _stream(cand.device)))
    return out


def _ptr_array(ptrs: Sequence[int]):
    return (C.c_void_p * len(ptrs))(*[int(p) for p in ptrs])


def peer_copy(src: torch.Tensor, dst_offset_words: int, peer_dst: Sequence[int]):
    """src (contiguous, 4-byte elements) -> every listed peer buffer at word offset dst_offset_words."""
    if not peer_dst or src.numel() == 0:
        return
    lib = L.load()
    assert src.is_contiguous() and src.element_size() == 4
    with torch.cuda.device(src.device):
        L.check(lib.cppre_peer_copy(_ptr(src), src.numel(), int(dst_offset_words), _ptr_array(peer_dst), len(peer_dst),
                                    _stream(src.device)))


def peer_bcast_keys(vals: torch.Tensor, cells_pad: int, peer_keys: Sequence[int], my_rank: int):
    """K3x-1: ordered uint32 key of vals[c] -> every peer's key buffer [world, cells_pad] at [my_rank, c]."""
    lib = L.load()
    v = vals.reshape(-1).contiguous()
    with torch.cuda.device(v.device):
        L.check(lib.cppre_peer_bcast_keys(_ptr(v), v.numel(), int(cells_pad), _ptr_array(peer_keys), len(peer_keys),
                                          int(my_rank), _stream(v.device)))


def window_exchange(scores: torch.Tensor, keys: torch.Tensor, slots: int, slab: int, peer_recv: Sequence[int],
                    my_rank: int, overflow: torch.Tensor, overflow_cells: Optional[torch.Tensor] = None,
                    key_mode: int = 0):
    """K3x-2: scores [n, cells...]; keys = this rank's [world, cells_pad] key buffer (int32 storage of uint32
    keys; key_mode 0) or the MAX-all-reduced [2, cells_pad] pair [~min, max] of signed keys (key_mode 1).
    Headers + candidates go straight into the owners' receive buffers [world, slab/32, 1+slots, 32]
    (slab a multiple of 32: dist.fused_slab)."""
    lib = L.load()
    fs, rs = _rows(scores)
    n, cells = fs.shape
    cap = 0 if overflow_cells is None else overflow_cells.numel()
    with torch.cuda.device(scores.device):
        L.check(lib.cppre_window_exchange(_ptr(fs), rs, n, cells, _ptr(keys), int(keys.shape[1]), int(key_mode), int(slots), int(slab),
                                          _ptr_array(peer_recv), len(peer_recv), int(my_rank), _ptr(overflow),
                                          _ptr(overflow_cells), cap, _stream(scores.device)))


def owner_select(recv: torch.Tensor, cells: int, kt: int, network: int, cell_offset: int, peer_q: Sequence[int],
                 overflow: torch.Tensor, overflow_cells: Optional[torch.Tensor] = None):
    """K3x-3: recv = this rank's [world, slab/32, 1+slots, 32] receive buffer (per sender and 32-cell tile:
    header row + dense candidate rows); the answers of its `cells` live cells are stored into every peer's
    q buffer at [cell_offset + c]."""
    lib = L.load()
    w, tiles, rows, lanes = recv.shape
    assert lanes == 32
    slab = tiles * 32
    cap = 0 if overflow_cells is None else overflow_cells.numel()
    with torch.cuda.device(recv.device):
        L.check(lib.cppre_owner_select(_ptr(recv), int(w), int(rows - 1), int(slab), int(cells), int(kt), int(network),
                                       int(cell_offset), _ptr_array(peer_q), len(peer_q), _ptr(overflow),
                                       _ptr(overflow_cells), cap, _stream(recv.device)))


def radix_hist(x: torch.Tensor, prefix: int, mask: int, shift: int, hist: torch.Tensor):
    lib = L.load()
    flat = x.reshape(-1).contiguous()
    with torch.cuda.device(x.device):
        L.check(lib.cppre_radix_hist(_ptr(flat), flat.numel(), prefix, mask, shift, _ptr(hist),
                                     _stream(x.device)))
    return hist


def radix_select(scores, k: int, allreduce=None) -> torch.Tensor:
    """The k-th smallest (0-based) of a flat score array -- or of the union of a LIST of arrays (scores
    that arrive in chunks) and, when ``allreduce`` (a callable that sums an int32 CUDA tensor over the
    ranks in place) is given, of ALL ranks' scores -- as a one-element float32 CUDA tensor.  Four 8-bit
    radix passes whose state lives on the device: nothing crosses to the host, so the result can be
    handed to the coverage launch as ``band_q`` right away.  NaN when a score is NaN (np.quantile) or
    k is out of range."""
    lib = L.load()
    parts = [t.reshape(-1).contiguous() for t in (scores if isinstance(scores, (list, tuple)) else [scores])]
    dev = parts[0].device
    state = torch.empty(4, dtype=torch.int32, device=dev)
    hist = torch.empty(257, dtype=torch.int32, device=dev)
    out = torch.empty(1, dtype=torch.float32, device=dev)
    if allreduce is None and len(parts) == 1 and parts[0].numel() <= (1 << 20):
        with torch.cuda.device(dev):      # everything on this GPU: one launch, one CTA
            L.check(lib.cppre_radix_select(_ptr(parts[0]), parts[0].numel(), int(k), _ptr(out), _stream(dev)))
        return out
    with torch.cuda.device(dev):
        st = _stream(dev)
        L.check(lib.cppre_radix_select_begin(_ptr(state), _ptr(hist), int(k), st))
        for p in range(4):
            for flat in parts:
                L.check(lib.cppre_radix_select_hist(_ptr(flat), flat.numel(), _ptr(state), p, _ptr(hist), st))
            if allreduce is not None:
                allreduce(hist)
            L.check(lib.cppre_radix_select_pick(_ptr(hist), _ptr(state), p, _ptr(out), st))
    return out


def inside_count(y: torch.Tensor, lo: torch.Tensor, hi: torch.Tensor) -> torch.Tensor:
    """uint32-valued int32 tensor [n]: cells of sample i with lo <= y <= hi.  lo/hi are
    [cells...] (broadcast over samples) or [n, cells...]."""
    lib = L.load()
    fy, sy = _rows(y)
    n, cells = fy.shape

    def bound(t):
        t = t.to(torch.float32)
        if t.numel() == 1:
            t = t.reshape(1).expand(cells)
        if t.numel() == cells:
            return t.reshape(-1).contiguous(), 0
        if t.numel() == n * cells:
            f, s = _rows(t.reshape(y.shape))
            return f, s
        f, s = _rows(torch.broadcast_to(t, y.shape).contiguous())
        return f, s

    flo, slo = bound(lo)
    fhi, shi = bound(hi)
    out = torch.empty(n, dtype=torch.int32, device=y.device)
    with torch.cuda.device(y.device):
        L.check(lib.cppre_inside_count(_ptr(fy), sy, _ptr(flo), slo, _ptr(fhi), shi, n, cells, _ptr(out),
                                       _stream(y.device)))
    return out


def inside_count_band(y: torch.Tensor, centre: Optional[torch.Tensor], half: Optional[torch.Tensor],
                      q: float) -> torch.Tensor:
    """Cells with centre - q*half <= y <= centre + q*half (sets never materialised)."""
    lib = L.load()
    fy, sy = _rows(y)
    n, cells = fy.shape
    fc, sc = (None, 0) if centre is None else _rows(centre)
    fh = None if half is None else half.reshape(-1).contiguous()
    out = torch.empty(n, dtype=torch.int32, device=y.device)
    with torch.cuda.device(y.device):
        L.check(lib.cppre_inside_count_band(_ptr(fy), sy, _ptr(fc), sc, _ptr(fh), float(q), n, cells,
                                            _ptr(out), _stream(y.device)))
    return out


# --------------------------------------------------------------------------------------
# boundary conditions, spectral pointwise
# --------------------------------------------------------------------------------------
def make_bc(types, widths, values) -> L.BC:
    bc = L.BC()
    for i in range(4):
        bc.type[i], bc.width[i], bc.value[i] = int(types[i]), int(widths[i]), float(values[i])
    return bc


def bc_padded_shape(bc: L.BC, h: int, w: int):
    hp, wp = C.c_int64(), C.c_int64()
    L.check(L.load().cppre_bc_padded_shape(C.byref(bc), h, w, C.byref(hp), C.byref(wp)))
    return hp.value, wp.value


def bc_index_map(bc: L.BC, h: int, w: int):
    """(rows, row_values, cols, col_values): source indices (-1 = Dirichlet constant) and the
    constants, straight from the library's host-side resolver."""
    hp, wp = bc_padded_shape(bc, h, w)
    rows, cols = (C.c_int32 * max(hp, 1))(), (C.c_int32 * max(wp, 1))()
    rv, cv = (C.c_float * max(hp, 1))(), (C.c_float * max(wp, 1))()
    L.check(L.load().cppre_bc_index_map(C.byref(bc), h, w, rows, rv, cols, cv))
    return (np.array(rows[:hp], dtype=np.int32), np.array(rv[:hp], dtype=np.float32),
            np.array(cols[:wp], dtype=np.int32), np.array(cv[:wp], dtype=np.float32))


def bc_pad(x: torch.Tensor, bc: L.BC) -> torch.Tensor:
    """x CUDA fp32 [..., H, W] -> [..., Hp, Wp]."""
    lib = L.load()
    h, w = x.shape[-2:]
    hp, wp = bc_padded_shape(bc, h, w)
    xc = x.contiguous()
    n_img = xc.numel() // max(h * w, 1) if h * w else int(np.prod(x.shape[:-2]))
    out = torch.empty(tuple(x.shape[:-2]) + (hp, wp), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        L.check(lib.cppre_bc_pad(_ptr(xc), n_img, h, w, C.byref(bc), _ptr(out), _stream(x.device)))
    return out


def bc_conv2d(x: torch.Tensor, bc: L.BC, kernel: np.ndarray) -> torch.Tensor:
    lib = L.load()
    h, w = x.shape[-2:]
    hp, wp = bc_padded_shape(bc, h, w)
    k = np.ascontiguousarray(kernel, dtype=np.float32)
    kh, kw = k.shape
    if hp - kh + 1 < 1 or wp - kw + 1 < 1:
        raise RuntimeError(f"Kernel size {(kh, kw)} can't be greater than actual input size {(hp, wp)}")
    xc = x.contiguous()
    n_img = int(np.prod(x.shape[:-2]))
    out = torch.empty(tuple(x.shape[:-2]) + (hp - kh + 1, wp - kw + 1), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        L.check(lib.cppre_bc_conv2d(_ptr(xc), n_img, h, w, C.byref(bc),
                                    k.ctypes.data_as(C.POINTER(C.c_float)), kh, kw, _ptr(out),
                                    _stream(x.device)))
    return out


def edge_difference(u: torch.Tensor, wall: str, scale: float) -> torch.Tensor:
    """(u[..., wall line] - u[..., opposite line]) * scale over the last two axes of a CUDA fp32 tensor (any view
    whose leading axes flatten to one stride); [..., W] for 'top' / 'bottom', [..., H] for 'left' / 'right'."""
    walls = {"top": L.WALL_TOP, "bottom": L.WALL_BOTTOM, "left": L.WALL_LEFT, "right": L.WALL_RIGHT}
    if wall not in walls:
        raise ValueError(f"unknown wall {wall!r}")
    if u.dim() < 2:
        raise IndexError("too many indices for tensor of dimension %d" % u.dim())
    h, w = (int(v) for v in u.shape[-2:])
    lead = tuple(u.shape[:-2])
    n_img = int(np.prod(lead)) if lead else 1
    flat = u.reshape(n_img, h, w)                      # a view whenever the leading axes flatten to one stride
    out = torch.empty(lead + ((w,) if wall in ("top", "bottom") else (h,)), dtype=torch.float32, device=u.device)
    with torch.cuda.device(u.device):
        L.check(L.load().cppre_edge_difference(_ptr(flat), n_img, h, w, flat.stride(0), flat.stride(1), flat.stride(2),
                                               walls[wall], float(scale), _ptr(out), _stream(u.device)))
    return out


def spectral_pointwise(sig_hat: torch.Tensor, ker_hat: torch.Tensor, flags: int, eps: float = 1e-6):
    """sig_hat complex64 [B, ...], ker_hat complex64 [...]; returns sig_hat * g(ker_hat)."""
    lib = L.load()
    s = sig_hat.contiguous()
    k = ker_hat.contiguous()
    m = k.numel()
    batch = s.numel() // max(m, 1)
    out = torch.empty_like(s)
    sr, kr, orr = (torch.view_as_real(t) for t in (s, k, out))
    with torch.cuda.device(s.device):
        L.check(lib.cppre_spectral_pointwise(_ptr(sr), _ptr(kr), batch, m, flags, float(eps), _ptr(orr),
                                             _stream(s.device)))
    return out
