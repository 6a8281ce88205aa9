"""Import the UNMODIFIED reference (TEST INFRASTRUCTURE; build container only).

``/root/reference`` exists only where the repo is built, never on the GPU box.  There the
only source is ``oracle/_ref/reference_hotpath.zip`` -- the byte-for-byte snapshot of the hot-path
modules that ``oracle/make_ref_snapshot.py`` packs at build time (git-ignored, it travels with
the working tree) -- which is unpacked into a temporary directory and imported the same way.
Used by ``oracle/make_golden.py`` to pin the NumPy restatement against the reference's own
outputs, and by ``oracle/ref_arm.py`` (the timed CPU arm of ``bench.py``).  GPU tests and
``smoke()`` never call it.

Recipe: matplotlib is not installed, and two reference modules import
it at module level (Utils/ConvOps_1d.py:314, Utils/boundary_conditions.py:4), so stub
modules are inserted; ``/root/reference`` and ``/root/reference/Utils`` go on ``sys.path``
exactly as the reference's scripts arrange it.
"""
from __future__ import annotations

import importlib
import importlib.util
import os
import sys
import types
import warnings

REFERENCE_ROOT = os.environ.get("CPPRE_REFERENCE_ROOT", "/root/reference")
_ARCHIVE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "reference_hotpath.zip")
_unpacked = None


def _root() -> str | None:
    """The reference checkout, else the unpacked build-time snapshot, else None."""
    global _unpacked
    if os.path.isdir(os.path.join(REFERENCE_ROOT, "Utils")):
        return REFERENCE_ROOT
    if _unpacked is None and os.path.exists(_ARCHIVE):
        import atexit
        import shutil
        import tempfile
        import zipfile
        from .make_ref_snapshot import verify
        verify(_ARCHIVE)
        d = tempfile.mkdtemp(prefix="cppre_ref_")
        atexit.register(shutil.rmtree, d, ignore_errors=True)
        with zipfile.ZipFile(_ARCHIVE) as z:
            z.extractall(d)
        _unpacked = d
    return _unpacked


def source() -> str:
    """Where ``load()`` takes the reference from: "checkout", "snapshot" or "absent"."""
    r = _root()
    return "absent" if r is None else ("checkout" if r == REFERENCE_ROOT else "snapshot")


def available() -> bool:
    return _root() is not None


def load():
    """Returns a namespace with the reference modules.  Our own drop-in mirrors use the same
    module names, so any of those already imported are evicted first."""
    root = _root()
    if root is None:
        raise RuntimeError(f"no reference checkout at {REFERENCE_ROOT} and no snapshot at {_ARCHIVE}")
    for name in ("matplotlib", "matplotlib.pyplot"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    names = ["Utils", "Utils.ConvOps_2d", "Utils.ConvOps_1d", "Utils.VectorConvOps",
             "ConvOps_2d", "ConvOps_1d", "ConvOps_0d", "ConvOps_Spatial", "boundary_conditions",
             "VectorConvOps_Spatial", "fft_conv_pytorch", "fft_conv_pytorch.fft_conv",
             "fft_conv_pytorch.utils", "Other_UQ", "Other_UQ.Evaluation",
             "Other_UQ.Evaluation.PRE_estimations"]
    saved = {n: sys.modules.pop(n) for n in names if n in sys.modules}
    old_path = list(sys.path)
    # The drop-in mirror (cp_pre_b200.install()) uses the SAME module names and its ``Utils`` is a regular package,
    # which would win over the reference's namespace package wherever it sits on sys.path: take the mirror's
    # directories off the path while the reference is imported.
    mirror = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "cp_pre_b200")
    sys.path[:] = [root, os.path.join(root, "Utils")] + [
        p for p in old_path if not os.path.abspath(p or ".").startswith(mirror)]
    importlib.invalidate_caches()
    ns = types.SimpleNamespace()
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            ns.ConvOps_2d = importlib.import_module("Utils.ConvOps_2d")
            ns.ConvOps_1d = importlib.import_module("ConvOps_1d")
            ns.ConvOps_0d = importlib.import_module("ConvOps_0d")
            ns.ConvOps_Spatial = importlib.import_module("ConvOps_Spatial")
            ns.boundary_conditions = importlib.import_module("boundary_conditions")
            ns.VectorConvOps_Spatial = importlib.import_module("VectorConvOps_Spatial")
            ns.fft_conv = importlib.import_module("fft_conv_pytorch.fft_conv")
            try:
                ns.VectorConvOps = importlib.import_module("Utils.VectorConvOps")
            except Exception as exc:  # pragma: no cover
                ns.VectorConvOps = exc
            # PRE_estimations lives in a directory without __init__.py: load by path.
            spec = importlib.util.spec_from_file_location(
                "_ref_PRE_estimations",
                os.path.join(root, "Other_UQ", "Evaluation", "PRE_estimations.py"))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            ns.PRE_estimations = mod
        for m in (ns.ConvOps_2d, ns.ConvOps_1d, ns.ConvOps_0d, ns.ConvOps_Spatial, ns.boundary_conditions, ns.fft_conv):
            if not os.path.abspath(m.__file__).startswith(os.path.abspath(root)):
                raise RuntimeError(f"{m.__name__} resolved to {m.__file__}, not to the reference under {root}")
    finally:
        sys.path[:] = old_path
        for n in names:
            sys.modules.pop(n, None)
        sys.modules.update(saved)
    return ns
