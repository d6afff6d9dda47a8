"""Build the C-ABI CUDA library (libgsr_b200.so) in-tree with nvcc for sm_100a.

The .so is git-ignored but travels to the GPU box with the repo snapshot.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libgsr_b200.so")
SOURCES = ["gsr_api.cu", "gsr_preprocess.cu", "gsr_binning.cu", "gsr_depthsort.cu", "gsr_blend.cu", "gsr_train.cu", "gsr_peer.cu"]
HEADERS = ["gsr_common.cuh", "gsr_math.cuh", os.path.join("..", "..", "include", "gsr_b200.h")]

NVCC_FLAGS = [
    "-O3", "-std=c++17",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
    "--shared",
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; the CUDA library cannot be built")


def is_stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/*.cu into libgsr_b200.so (no-op when up to date).  Returns the library path.

    Safe under torchrun: an exclusive file lock serialises the ranks (the first one in builds, the others find the
    library fresh when they get the lock), and every build writes to its own temporary file before the atomic rename."""
    if not force and not is_stale():
        return LIB_PATH
    import fcntl
    import tempfile
    with open(LIB_PATH + ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not is_stale():        # another process built it while we waited
                return LIB_PATH
            fd, tmp = tempfile.mkstemp(prefix="libgsr_b200.", suffix=".so.tmp", dir=PKG_DIR)
            os.close(fd)
            cmd = [find_nvcc(), *NVCC_FLAGS, "-o", tmp, *[os.path.join(CSRC, f) for f in SOURCES]]
            if verbose:
                print(" ".join(cmd), file=sys.stderr)
            res = subprocess.run(cmd, capture_output=True, text=True)
            if res.returncode != 0:
                try:
                    os.unlink(tmp)
                except OSError:
                    pass
                raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
            os.chmod(tmp, 0o755)
            os.replace(tmp, LIB_PATH)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
