"""TEST INFRASTRUCTURE ONLY.  g++ build of the host emulation of the post-processing kernels (post_emul.cpp)."""
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parents[1]
CSRC = ROOT / "3d_object_detection_b200" / "csrc"
LIB = HERE / "_build" / "libpost_emul.so"
VOX_LIB = HERE / "_build" / "libvoxel_emul.so"


def build(force=False):
    deps = [HERE / "post_emul.cpp", HERE / "cuda_emul.h", CSRC / "pf_post_kernels.cuh", CSRC / "pf_post_desc.h"]
    if not force and LIB.exists() and all(d.stat().st_mtime <= LIB.stat().st_mtime for d in deps):
        return LIB
    LIB.parent.mkdir(exist_ok=True)
    cmd = ["g++", "-O1", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", "-Wall", "-Wno-unknown-pragmas",
           "-I", str(CSRC), "-I", str(HERE), "-o", str(LIB), str(HERE / "post_emul.cpp")]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("g++ failed:\n" + res.stdout + res.stderr)
    return LIB


def build_voxel(force=False):
    """g++ build of the host emulation of the voxelisation kernels (voxel_emul.cpp)."""
    deps = [HERE / "voxel_emul.cpp", HERE / "cuda_emul.h", HERE / "warp_emul.h", CSRC / "pf_voxel_kernels.cuh",
            CSRC / "pf_internal.cuh", ROOT / "include" / "pillarfront.h"]
    if not force and VOX_LIB.exists() and all(d.stat().st_mtime <= VOX_LIB.stat().st_mtime for d in deps):
        return VOX_LIB
    VOX_LIB.parent.mkdir(exist_ok=True)
    cmd = ["g++", "-O1", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", "-Wall", "-Wno-unknown-pragmas",
           "-Wno-unused-function", "-I", str(CSRC), "-I", str(HERE), "-I", str(ROOT / "include"), "-o", str(VOX_LIB),
           str(HERE / "voxel_emul.cpp")]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("g++ failed:\n" + res.stdout + res.stderr)
    return VOX_LIB
