"""Drop-in for ``Neural_PDE/UQ/inductive_cp.py`` -- the five conformal helpers every script
star-imports (``from Neural_PDE.UQ.inductive_cp import *``).

The module itself is NOT in the reference tree (un-vendored submodule Neural_PDE @ fdf1dcc3);
signatures and semantics are taken from its 150+ call sites and from the in-tree inline twins
(SURVEY.md section 8 a12-a16, 8c):

    calibrate(scores, n, alpha)          split-CP quantile along axis 0          **parity unpinned**
    emp_cov(prediction_sets, y)          mean over all elements of lo<=y<=hi     **parity unpinned**
    emp_cov_joint(prediction_sets, y)    mean over samples of all-cells-inside   (twin: Tests/test_advection_inv_sampling_marginal.py:465)
    modulation_func(a, b)                std over samples of a-b                 (twin: ...:428)
    ncf_metric_joint(a, b, modulation)   per-sample max |a-b|/modulation         (twin: ...:430-431)

Inputs may be NumPy arrays (what the scripts pass), CPU tensors or CUDA tensors; results come
back in kind.  All arithmetic is libcppre CUDA kernels: K2 (cell statistics), K3 (order
statistics), K5 (coverage counts).  Score arrays are processed in float32.
"""
import numpy as np
import torch

from cp_pre_b200 import ops

__all__ = ["calibrate", "calibrate_many", "calibrate_level", "calibrate_rank", "emp_cov", "emp_cov_joint",
           "modulation_func", "ncf_metric_joint", "filter_sims_joint", "filter_sims_within_bounds"]


def calibrate_level(n, alpha):
    """ceil((n+1)(1-alpha))/n, evaluated in float64 like NumPy would."""
    return np.ceil((n + 1) * (1 - alpha)) / n


def calibrate_rank(n_rows, n, alpha, dtype=np.float32):
    """0-based order statistic that ``np.quantile(scores, q, method='higher')`` returns for the level
    above: ceil((n_rows-1)*q) evaluated in the dtype of the SCORE array, as NumPy 2.x does (a Python /
    NumPy float q is cast to the array's dtype first, ``numpy/lib/_function_base_impl.py: quantile``):
    float32 for the float32 score arrays this library processes, float64 when the caller hands over
    float64 scores.  The two differ when n != n_rows (rows=14, n=13, alpha=0.5: index 8 in float32
    arithmetic, 7 in float64).  q outside [0,1] raises like NumPy.

    This follows NumPy >= 2 (the version this image and the golden vectors use).  NumPy 1.x kept a Python-float
    q in float64 whatever the array's dtype, so on such an install ``np.quantile`` on float32 scores would pick
    the float64 index; the library does not branch on ``np.__version__`` -- its contract is the NumPy 2 result."""
    q = calibrate_level(n, alpha)
    if not (0.0 <= q <= 1.0):
        raise ValueError("Quantiles must be in the range [0, 1]")
    ft = np.dtype(dtype).type
    return int(np.ceil((n_rows - 1) * ft(q)))


def _rank_dtype(scores):
    """float64 score arrays keep NumPy's float64 index arithmetic; everything else is float32."""
    dt = scores.dtype if isinstance(scores, np.ndarray) else getattr(scores, "dtype", None)
    return np.float64 if dt in (np.float64, torch.float64) else np.float32


def _out(staged, t, scalar):
    if staged.numpy:
        a = t.detach().cpu().numpy()
        return a.reshape(()).astype(np.float32)[()] if scalar else a
    t = staged.back(t)
    return t.reshape(()) if scalar else t


def calibrate_many(scores, n, alphas):
    """Quantiles for a whole alpha grid from ONE pass over the scores (the scripts re-run
    ``calibrate`` 9-10 times on the same array).  Returns [len(alphas), *scores.shape[1:]]."""
    staged = ops.Staged(scores)
    s = ops.to_cuda(scores)
    if s.dim() == 0 or s.shape[0] == 0:
        raise ValueError("calibrate needs at least one score along axis 0")
    ranks = [calibrate_rank(s.shape[0], n, a, _rank_dtype(scores)) for a in alphas]
    out = ops.select_ranks(s, ranks)
    return out.detach().cpu().numpy() if staged.numpy else staged.back(out)


def calibrate(scores, n, alpha):
    """qhat = the ceil((n+1)(1-alpha))/n empirical quantile of ``scores`` along axis 0
    (method 'higher'); [cells...] for N-d scores, a scalar for 1-D scores."""
    staged = ops.Staged(scores)
    s = ops.to_cuda(scores)
    if s.dim() == 0 or s.shape[0] == 0:
        raise ValueError("calibrate needs at least one score along axis 0")
    k = calibrate_rank(s.shape[0], n, alpha, _rank_dtype(scores))
    out = ops.select_ranks(s, [k])[0]
    return _out(staged, out, scalar=(s.dim() == 1))


def _bounds(prediction_sets, y):
    lo, hi = prediction_sets[0], prediction_sets[1]
    return (ops.to_cuda(np.asarray(lo) if not isinstance(lo, torch.Tensor) else lo),
            ops.to_cuda(np.asarray(hi) if not isinstance(hi, torch.Tensor) else hi))


def _inside(prediction_sets, y_response):
    y = ops.to_cuda(y_response)
    if y.dim() == 1:
        y = y.reshape(-1, 1)
    lo, hi = _bounds(prediction_sets, y)
    cells = y[0].numel() if y.shape[0] else 0
    return ops.inside_count(y, lo, hi), cells, y.shape[0]


def emp_cov(prediction_sets, y_response):
    """Marginal empirical coverage: fraction of ALL elements with lo <= y <= hi."""
    inside, cells, n = _inside(prediction_sets, y_response)
    total = n * cells
    return np.float64(int(inside.to(torch.int64).sum().item())) / total if total else np.float64("nan")


def filter_sims_joint(prediction_sets, y_response):
    """bool[n]: samples whose every cell lies inside the set
    [ref: Joint/Advection_Residuals_CP.py:327-329]."""
    staged = ops.Staged(y_response)
    inside, cells, _ = _inside(prediction_sets, y_response)
    mask = inside == cells
    return mask.cpu().numpy() if staged.numpy else staged.back(mask)


def emp_cov_joint(prediction_sets, y_response):
    """Joint empirical coverage: fraction of samples entirely inside the set."""
    inside, cells, n = _inside(prediction_sets, y_response)
    return np.float64(int((inside == cells).sum().item())) / n if n else np.float64("nan")


def filter_sims_within_bounds(lower_bound, upper_bound, samples, threshold, within=False):
    """bool[n]: fraction of cells inside (within=True) or on/outside (within=False) the bounds
    is >= threshold [ref: Active_Learning/Advection_AL_Marginal.py:169-197]."""
    staged = ops.Staged(samples)
    y = ops.to_cuda(samples)
    lo, hi = _bounds((lower_bound, upper_bound), y)
    cells = y[0].numel()
    if within:
        cnt = ops.inside_count(y, lo, hi).to(torch.float64)
    else:
        # (y <= lo) | (y >= hi) by inclusion-exclusion on three closed-interval counts
        inf = torch.tensor(float("inf"), device=y.device)
        le_lo = ops.inside_count(y, -inf, lo)
        ge_hi = ops.inside_count(y, hi, inf)
        both = ops.inside_count(y, hi, lo)                  # hi <= y <= lo: counted twice above
        cnt = (le_lo + ge_hi - both).to(torch.float64)
    mask = (cnt / cells) >= threshold
    return mask.cpu().numpy() if staged.numpy else staged.back(mask)


def modulation_func(a, b):
    """Per-cell population std over the sample axis of a-b (symmetric in a, b)."""
    staged = ops.Staged(a)
    zero_b = (not isinstance(b, torch.Tensor)) and not np.any(b)     # np.zeros(...) "no data" form
    ta = ops.to_cuda(a)
    out = ops.cell_std(ta, None if zero_b else ops.to_cuda(b))
    return out.detach().cpu().numpy() if staged.numpy else staged.back(out)


def ncf_metric_joint(a, b, modulation):
    """Per-sample max over cells of |a-b| / modulation."""
    staged = ops.Staged(a)
    zero_b = (not isinstance(b, torch.Tensor)) and not np.any(b)
    ta, tm = ops.to_cuda(a), ops.to_cuda(modulation)
    out = ops.joint_score(ta, None if zero_b else ops.to_cuda(b), tm)
    return out.detach().cpu().numpy() if staged.numpy else staged.back(out)
