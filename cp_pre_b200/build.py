"""Builds libcppre.so in-tree:  python -m cp_pre_b200.build  (nvcc, sm_100a; no GPU needed)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def build(jobs: int = 0, verbose: bool = False) -> str:
    jobs = jobs or max(4, min(16, os.cpu_count() or 8))
    cmd = ["make", "-C", os.path.join(HERE, "csrc"), f"-j{jobs}"]
    res = subprocess.run(cmd, capture_output=not verbose, text=True)
    if res.returncode != 0:
        sys.stderr.write((res.stdout or "") + (res.stderr or ""))
        raise RuntimeError("building libcppre.so failed")
    return os.path.join(HERE, "libcppre.so")


if __name__ == "__main__":
    print(build(verbose=True))
