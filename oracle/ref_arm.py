"""The reference's OWN code for the hot path, as the timed CPU arm of bench.py -- TEST / MEASUREMENT
INFRASTRUCTURE (see oracle/__init__.py); never imported by the product package.

What runs here is unmodified reference code, taken from the checkout (build container) or from the
byte-for-byte snapshot ``oracle/_ref/reference_hotpath.zip`` (GPU box; oracle/make_ref_snapshot.py):

* the operators: ``Utils/ConvOps_1d.py`` / ``Utils/ConvOps_2d.py`` ``ConvOperator`` (``F.conv2d`` / ``F.conv3d`` on
  the CPU, torch intra-op threads), imported as modules;
* the residual assembly: the reference keeps it in notebook-style scripts that cannot be imported (they load
  data and checkpoints top to bottom), so the statements that BUILD THE OPERATORS and the ``def residual*``
  functions are lifted out of the script's syntax tree and executed as they stand -- e.g.
  ``Joint/NS_Residuals_CP.py:211-240`` -- with the physical constants of the synthetic config bound to the
  names the script uses (``dx``, ``dt``, ``nu`` ...).  No line of the residual algebra is restated.

The five conformal helpers (``calibrate``, ``emp_cov``, ``emp_cov_joint``, ``modulation_func``,
``ncf_metric_joint``) live in a submodule that is absent from the reference tree; for those the
oracle's NumPy restatement is called, exactly where the scripts call the originals.
"""
from __future__ import annotations

import ast
import os

import numpy as np
import torch

from . import cpu_oracle as O
from . import ref_shim, synth

_OPERATOR_NAMES = {"D_t", "D_x", "D_y", "D_x_y", "D_xx_yy", "D_tt", "D_xx", "D", "disc"}
# workload -> (script, operator module attribute of the shim namespace, residual function, field slice)
_SCRIPTS = {
    "ns": ("Joint/NS_Residuals_CP.py", "ConvOps_2d", "residual_momentum"),
    "wave": ("Joint/Wave_Residuals_CP.py", "ConvOps_2d", "residual"),
    "wave-joint": ("Joint/Wave_Residuals_CP.py", "ConvOps_2d", "residual"),
    "burgers": ("Joint/Burgers_Residuals_CP.py", "ConvOps_1d", "residual"),
    "mhd": ("Joint/MHD_Residuals_CP.py", "ConvOps_2d", None),          # residual_<equation>
    "advection": ("Marginal/Advection_Residuals_CP.py", "ConvOps_1d", None),   # the operator D itself
}
_ns = None


def available() -> bool:
    root = ref_shim._root()
    return root is not None and all(os.path.exists(os.path.join(root, s[0])) for s in _SCRIPTS.values())


def source() -> str:
    return ref_shim.source()


def _lift(script: str, namespace: dict, functions: set, operators: bool = True) -> dict:
    """Executes, in ``namespace``, the module-level statements of ``script`` that build the operators
    (assignments to the names in _OPERATOR_NAMES, incl. ``D.kernel = ...``) and the listed function
    definitions -- in source order, untouched.  ``operators=False`` lifts the functions only."""
    path = os.path.join(ref_shim._root(), script)
    tree = ast.parse(open(path).read(), filename=path)
    keep = []
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in functions:
            keep.append(node)
        elif operators and isinstance(node, ast.Assign) and len(node.targets) == 1:
            t = node.targets[0]
            name = t.id if isinstance(t, ast.Name) else (t.value.id if isinstance(t, ast.Attribute) and isinstance(t.value, ast.Name) else None)
            if name in _OPERATOR_NAMES:
                keep.append(node)
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), namespace)
    return namespace


def residual_function(workload: str, equation: str = "momentum"):
    """``residual(vars) -> torch.Tensor`` (interior crop, like the scripts call it) built from the reference's
    own statements.  ``vars`` is [BS, F, Nt, Nx, Ny] (multi-field), [BS, Nt, Nx(, Ny)] (single field)."""
    global _ns
    if _ns is None:
        _ns = ref_shim.load()
    script, opmod, fname = _SCRIPTS[workload]
    env = {"torch": torch, "np": np, "ConvOperator": getattr(_ns, opmod).ConvOperator}
    f32 = lambda v: torch.tensor(v, dtype=torch.float32)
    if workload == "ns":
        c = synth.NS                       # scripts: dx = np.asarray(x[-1]-x[-2]); dy = dx; dt, nu floats
        env.update(dx=np.asarray(c["dx"], dtype=np.float32), dy=np.asarray(c["dy"], dtype=np.float32), dt=c["dt"], nu=c["nu"])
    elif workload in ("wave", "wave-joint"):
        env.update(c=1.0, dt=np.sqrt(synth.WAVE["c2"]), dx=1.0)      # (c*dt/dx)**2 = 0.25
    elif workload == "burgers":
        c = synth.BURGERS                  # Joint/Burgers_Residuals_CP.py:177-179 makes them fp32 tensors
        env.update(dx=f32(c["dx"]), dt=f32(c["dt"]), nu=f32(c["nu"]))
    elif workload == "mhd":
        env.update(gamma=synth.MHD["gamma"])
        fname = "residual_" + ("gauss" if equation == "divergence" else equation)
    elif workload == "advection":
        a = synth.ADVECTION                # Marginal/Advection_Residuals_CP.py:86-93,160-164
        env.update(v=a["v"], dt=a["t_end"] / a["nt"], dx=(a["x_max"] - a["x_min"]) / a["nx"])
    _lift(script, env, {fname} if fname else set())
    if workload == "advection":
        D = env["D"]
        return lambda u: D(u)[..., 1:-1, 1:-1]
    return env[fname]


# the conformal step on stored residuals, where the scripts call the un-vendored helpers
def marginal_cp(cal_res: torch.Tensor, val_res: torch.Tensor, alpha: float):
    """[ref: Marginal/NS_Residuals_CP.py:327-337]"""
    scores = np.abs(cal_res.numpy())
    qhat = O.calibrate(scores, len(scores), alpha)
    return qhat, O.emp_cov([-qhat, qhat], val_res.numpy())


def joint_cp(cal_res: torch.Tensor, val_res: torch.Tensor, alpha: float, eps: float = 0.0):
    """[ref: Joint/NS_Residuals_CP.py:305-337, Joint/MHD_Residuals_CP.py:350-351]"""
    a = cal_res.numpy()
    mod = O.modulation_func(a, np.zeros(a.shape)) + eps
    scores = O.ncf_metric_joint(a, np.zeros(a.shape), mod)
    qhat = O.calibrate(scores, len(scores), alpha)
    return qhat, mod, O.emp_cov_joint([-qhat * mod, qhat * mod], val_res.numpy())
