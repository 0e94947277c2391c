"""Compile the C oracle (oracle/pillar_oracle.c) into oracle/_build/libpillar_oracle.so with plain gcc.

ORACLE -- TEST INFRASTRUCTURE ONLY.  Building the checker is not using it: __graft_entry__.build() calls
this so the prebuilt library travels to the GPU box; the product package never imports it.

There is no ``oracle/_ref``: the reference is pure Python (numba JIT + torch), it has no C/C++ sources to
compile, and it cannot travel to the GPU box.  Its outputs are pinned as fixtures in tests/golden/ by
oracle/make_golden.py instead.
"""
from __future__ import annotations

import os
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
SRC = HERE / "pillar_oracle.c"
OUT_DIR = HERE / "_build"
LIB = OUT_DIR / "libpillar_oracle.so"


def build(force: bool = False) -> Path:
    OUT_DIR.mkdir(exist_ok=True)
    if not force and LIB.exists() and LIB.stat().st_mtime >= SRC.stat().st_mtime:
        return LIB
    cmd = ["gcc", "-O2", "-ffp-contract=off", "-fopenmp", "-shared", "-fPIC", "-Wall", "-Wextra",
           "-o", str(LIB) + ".tmp", str(SRC), "-lm"]
    subprocess.run(cmd, check=True)
    os.replace(str(LIB) + ".tmp", LIB)
    return LIB


if __name__ == "__main__":
    print(build(force=True))
