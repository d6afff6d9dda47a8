#!/usr/bin/env python
"""baseline/build_ref.py -- install the UNMODIFIED reference into the git-ignored baseline/_ref/ (reference arm of bench.py).

What it does (only where /root/reference exists, i.e. in the build container; the GPU box uses the prebuilt tree, which
travels with the repository snapshot because baseline/_ref is git-ignored but not gpurun-ignored):

1. the reference's own torch extensions, built by the reference's OWN setup.py, untouched:
       submodules/diff-gaussian-rasterization      -> baseline/_ref/diff_gaussian_rasterization/{__init__.py,_C*.so}
       submodules/diff-gaussian-rasterization-amr  -> baseline/_ref/diff_gaussian_rasterization_amr/{__init__.py,_C*.so}
   `python -m pip install --no-index --no-build-isolation --no-deps --target baseline/_ref <copy of the submodule>`
   from a copy under a temporary directory (/root/reference is read-only and setup.py builds in-tree).  The only
   accommodations are environment variables, no source edit:
       TORCH_CUDA_ARCH_LIST=10.0a            sm_100a cubins (torch.utils.cpp_extension reads it)
       NVCC_APPEND_FLAGS="-include cstdint"  gcc 13 needs <cstdint> for cuda_rasterizer/rasterizer_impl.h:24,40
   These are the binaries `bench.py --impl reference` times through the reference's stock Python packages
   (GaussianRasterizer -> _RasterizeGaussians -> _C.rasterize_gaussians, the path
   /root/reference/gaussian_renderer/__init__.py:86-106 times).

2. the reference's callers, copied verbatim (never committed) for the drop-in acceptance test:
       gaussian_renderer/__init__.py, gaussian_renderer_amr/__init__.py, utils/sh_utils.py  -> baseline/_ref/callers/...

Run:  python baseline/build_ref.py [--force] [--reference /root/reference]
"""
from __future__ import annotations

import argparse
import glob
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")

EXTENSIONS = {   # package name -> submodule directory under <reference>/submodules
    "diff_gaussian_rasterization": "diff-gaussian-rasterization",
    "diff_gaussian_rasterization_amr": "diff-gaussian-rasterization-amr",
}
CALLERS = ["gaussian_renderer/__init__.py", "gaussian_renderer_amr/__init__.py", "utils/sh_utils.py"]


def ext_path(pkg: str):
    hits = glob.glob(os.path.join(DEST, pkg, "_C*.so"))
    return hits[0] if hits else None


def available(pkg: str = "diff_gaussian_rasterization") -> bool:
    return ext_path(pkg) is not None and os.path.exists(os.path.join(DEST, pkg, "__init__.py"))


def callers_available() -> bool:
    return all(os.path.exists(os.path.join(DEST, "callers", c)) for c in CALLERS)


def _newest_source(src_dir: str) -> float:
    newest = 0.0
    for root, _, files in os.walk(src_dir):
        if "third_party" in root:
            continue
        for f in files:
            if f.endswith((".cu", ".cpp", ".h", ".py")):
                newest = max(newest, os.path.getmtime(os.path.join(root, f)))
    return newest


def build_extension(pkg: str, reference: str, force: bool = False, verbose: bool = True) -> str:
    src = os.path.join(reference, "submodules", EXTENSIONS[pkg])
    if not os.path.isdir(src):
        raise RuntimeError(f"{src} not found")
    have = ext_path(pkg)
    if have and not force and os.path.getmtime(have) >= _newest_source(src):
        return have
    os.makedirs(DEST, exist_ok=True)
    with tempfile.TemporaryDirectory(prefix="gsr_refbuild_") as tmp:
        work = os.path.join(tmp, EXTENSIONS[pkg])
        shutil.copytree(src, work, ignore=shutil.ignore_patterns("build", "*.egg-info", "__pycache__"))
        env = dict(os.environ)
        env["TORCH_CUDA_ARCH_LIST"] = "10.0a"
        env["NVCC_APPEND_FLAGS"] = (env.get("NVCC_APPEND_FLAGS", "") + " -include cstdint").strip()
        env.setdefault("MAX_JOBS", str(os.cpu_count() or 4))
        cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps", "--upgrade",
               "--find-links", "/opt/wheelhouse", "--target", DEST, work]
        if verbose:
            print("+", " ".join(cmd), file=sys.stderr)
        res = subprocess.run(cmd, env=env, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"pip install of the reference {pkg} failed:\n{res.stdout[-3000:]}\n{res.stderr[-3000:]}")
    out = ext_path(pkg)
    if not out:
        raise RuntimeError(f"pip reported success but {pkg}/_C*.so is missing under {DEST}")
    return out


def stage_callers(reference: str) -> None:
    for rel in CALLERS:
        src = os.path.join(reference, rel)
        dst = os.path.join(DEST, "callers", rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)       # verbatim; baseline/_ref/ is git-ignored, nothing of it enters the history


def build_all(reference: str = "/root/reference", force: bool = False, verbose: bool = True) -> dict:
    out = {}
    stage_callers(reference)
    for pkg in EXTENSIONS:
        out[pkg] = build_extension(pkg, reference, force, verbose)
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default=os.environ.get("GSR_REFERENCE", "/root/reference"))
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--only", default=None, choices=list(EXTENSIONS))
    a = ap.parse_args()
    if a.only:
        stage_callers(a.reference)
        print(a.only, "->", build_extension(a.only, a.reference, a.force))
    else:
        for k, v in build_all(a.reference, a.force).items():
            print(k, "->", v)
