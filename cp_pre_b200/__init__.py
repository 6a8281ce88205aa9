"""cp_pre_b200 -- B200-native (sm_100a) implementation of CP-PRE's hot path: PDE physics-residual
(PRE) evaluation with finite-difference stencil operators plus marginal / joint conformal
calibration, behind the reference's own Python API.

Layout (mirrors the reference checkout so its import statements keep working after
``cp_pre_b200.install()``):

    Utils/ConvOps_0d|1d|2d|Spatial.py, VectorConvOps(_Spatial).py, boundary_conditions.py,
    Utils/fft_conv_pytorch/fft_conv.py        the operator API
    Neural_PDE/UQ/inductive_cp.py             calibrate / emp_cov / ... (a12-a16)
    Other_UQ/Evaluation/PRE_estimations.py    PRE_Wave / PRE_NS / PRE_MHD class forms (a7)
    programs.py                               fused residual programs + joint/marginal pipelines
    dist.py                                   batch-sharded multi-GPU calibration (NCCL)
    csrc/ + libcppre.so                       the CUDA kernels behind the C ABI (include/cppre.h)

There is no CPU fallback anywhere in this package.
"""
import os
import sys

__version__ = "0.1.0"
PACKAGE_DIR = os.path.dirname(os.path.abspath(__file__))


def install():
    """Puts the mirror of the reference's directory layout on ``sys.path`` so that
    ``from Utils.ConvOps_2d import ConvOperator``, ``from ConvOps_1d import ConvOperator`` and
    ``from Neural_PDE.UQ.inductive_cp import *`` resolve to this package."""
    root = os.path.dirname(PACKAGE_DIR)
    for p in (os.path.join(PACKAGE_DIR, "Utils"), PACKAGE_DIR, root):
        if p not in sys.path:
            sys.path.insert(0, p)
    for name in ("Utils", "Neural_PDE", "Other_UQ"):
        mod = sys.modules.get(name)
        if mod is not None and not str(getattr(mod, "__path__", [""])[0]).startswith(PACKAGE_DIR):
            # another checkout's namespace package of the same name is already imported
            for key in [k for k in sys.modules if k == name or k.startswith(name + ".")]:
                del sys.modules[key]
