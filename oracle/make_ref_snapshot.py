"""Snapshot of the UNMODIFIED reference hot-path modules for the CPU arm of bench.py -- TEST / MEASUREMENT
INFRASTRUCTURE (see oracle/__init__.py); never imported by the product package.

The reference is a directory of Python scripts (no setup.py / pyproject: nothing to pip-install, nothing to
compile), and ``/root/reference`` does not exist on the GPU box.  So that ``bench.py --impl reference`` and the
``cpu_baseline`` leg can time the reference's OWN code there (``cpu_baseline.kind = "reference"``) instead of
the oracle's port, ``__graft_entry__.build()`` calls ``snapshot()`` in the build container: the files listed
in ``FILES`` are packed verbatim, byte for byte, into ``oracle/_ref/reference_hotpath.zip`` with a manifest of
their SHA-256 digests.  ``oracle/_ref/`` is git-ignored (reference sources never enter this repository's
history) but travels to the GPU box with the working tree, like the built ``libcppre.so``.

``oracle/ref_shim.py`` unpacks the archive into a temporary directory at run time and imports it exactly as it
imports ``/root/reference`` in the build container.  Without the archive the bench falls back to
``oracle/torch_port.py`` (``kind = "port"``), which is pinned to the same reference outputs.
"""
from __future__ import annotations

import hashlib
import json
import os
import zipfile

HERE = os.path.dirname(os.path.abspath(__file__))
ARCHIVE = os.path.join(HERE, "_ref", "reference_hotpath.zip")
# the hot path of SURVEY.md section 8(a): operator API, spectral backend, boundary manager, vector operators and
# the class forms of the residuals; nothing else of the reference is needed to run it
FILES = [
    "Utils/ConvOps_0d.py", "Utils/ConvOps_1d.py", "Utils/ConvOps_2d.py", "Utils/ConvOps_Spatial.py",
    "Utils/VectorConvOps.py", "Utils/VectorConvOps_Spatial.py", "Utils/boundary_conditions.py",
    "Utils/fft_conv_pytorch/__init__.py", "Utils/fft_conv_pytorch/fft_conv.py", "Utils/fft_conv_pytorch/utils.py",
    "Other_UQ/Evaluation/PRE_estimations.py",
    # notebook-style scripts: only their operator-building statements and `def residual*` functions are executed
    # (lifted from the syntax tree by oracle/ref_arm.py); the rest of each file is never run
    "Joint/NS_Residuals_CP.py", "Joint/MHD_Residuals_CP.py", "Joint/Burgers_Residuals_CP.py",
    "Joint/Wave_Residuals_CP.py", "Marginal/Advection_Residuals_CP.py",
]


def snapshot(reference_root: str = "/root/reference") -> str | None:
    """Packs FILES from ``reference_root`` into ARCHIVE (fixed timestamps: the archive is reproducible).
    Returns the archive path, or None when the reference checkout is not there (GPU box: the archive that
    travelled with the tree is used as it is)."""
    if not os.path.isdir(os.path.join(reference_root, "Utils")):
        return ARCHIVE if os.path.exists(ARCHIVE) else None
    os.makedirs(os.path.dirname(ARCHIVE), exist_ok=True)
    manifest = {"source": reference_root, "files": {}}
    sub = os.path.join(reference_root, ".SUBMODULES.json")
    if os.path.exists(sub):
        try:
            manifest["reference_meta"] = json.load(open(sub))
        except ValueError:
            pass
    tmp = ARCHIVE + ".tmp"
    with zipfile.ZipFile(tmp, "w", zipfile.ZIP_DEFLATED) as z:
        for rel in FILES:
            data = open(os.path.join(reference_root, rel), "rb").read()
            manifest["files"][rel] = hashlib.sha256(data).hexdigest()
            z.writestr(zipfile.ZipInfo(rel, date_time=(2020, 1, 1, 0, 0, 0)), data)
        z.writestr(zipfile.ZipInfo("MANIFEST.json", date_time=(2020, 1, 1, 0, 0, 0)),
                   json.dumps(manifest, indent=1, sort_keys=True))
    os.replace(tmp, ARCHIVE)
    return ARCHIVE


def verify(archive: str = ARCHIVE) -> dict:
    """Re-hashes every member against the manifest (raises on a mismatch); returns the manifest."""
    with zipfile.ZipFile(archive) as z:
        manifest = json.loads(z.read("MANIFEST.json"))
        for rel, want in manifest["files"].items():
            got = hashlib.sha256(z.read(rel)).hexdigest()
            if got != want:
                raise RuntimeError(f"{rel}: digest {got} differs from the manifest's {want}")
    return manifest


if __name__ == "__main__":
    path = snapshot()
    print(path, "" if path is None else json.dumps(verify(path)["files"], indent=1))
