"""Build libpillarfront.so in-tree with plain nvcc for sm_100a (no JIT cache, no torch extension machinery).

    python -m 3d_object_detection_b200.build        # (via importlib; the directory name is not an identifier)

The library lands in 3d_object_detection_b200/lib/ -- git-ignored, but it travels to the GPU box with the
repo snapshot.  nvcc cross-compiles here without a GPU.
"""
from __future__ import annotations

import os
import shutil
import subprocess
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
LIB_DIR = PKG / "lib"
LIB = LIB_DIR / "libpillarfront.so"
SOURCES = ["pf_api.cu", "pf_voxelize.cu", "pf_gather.cu", "pf_scatter.cu", "pf_anchor.cu", "pf_nms.cu", "pf_post.cu", "pf_runner.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "--shared", "-Xcompiler", "-fPIC,-Wall,-Wextra,-fvisibility=hidden,-fopenmp", "-lgomp",
              "-Xptxas", "-v", "--fmad=false"]


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def needs_build() -> bool:
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    deps = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + [ROOT / "include" / "pillarfront.h", Path(__file__)]
    return any(d.stat().st_mtime > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and not needs_build():
        return LIB
    LIB_DIR.mkdir(exist_ok=True)
    tmp = LIB.with_suffix(".so.tmp")
    extra = os.environ.get("PF_NVCC_EXTRA", "").split()      # experiments only, e.g. PF_NVCC_EXTRA=-DPF_TR_MINB=4
    cmd = [nvcc_path(), *NVCC_FLAGS, *extra, "-I", str(ROOT / "include"), "-I", str(CSRC),
           "-o", str(tmp)] + [str(CSRC / s) for s in SOURCES]
    res = subprocess.run(cmd, capture_output=True, text=True)
    (LIB_DIR / "build.log").write_text("$ " + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stdout + res.stderr)
    os.replace(tmp, LIB)
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose=True))
