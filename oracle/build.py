"""Build the oracle's C restatements (test infrastructure) into oracle/_build/liboracle.so.

The reference is pure Python (no C/C++ sources), so there is
nothing to compile into oracle/_ref/; the Python reference is instead imported in the
build container by tools/make_golden.py to produce tests/golden/*.npz.
"""
import os
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
OUT = HERE / "_build" / "liboracle.so"
SRCS = [HERE / "csrc" / "dft.c", HERE / "csrc" / "wgrid.c", HERE / "csrc" / "wgrid_fast.c"]


def build(force: bool = False) -> Path:
    deps = SRCS + [HERE / "csrc" / "wgrid_impl.h"]
    if not force and OUT.exists() and all(OUT.stat().st_mtime >= d.stat().st_mtime for d in deps):
        return OUT
    OUT.parent.mkdir(exist_ok=True)
    cmd = ["gcc", "-O3", "-march=x86-64-v3", "-fopenmp", "-fPIC", "-shared", "-std=gnu11", "-fno-math-errno",
           "-o", str(OUT)] + [str(s) for s in SRCS] + ["-lm"]
    subprocess.run(cmd, check=True)
    return OUT


if __name__ == "__main__":
    print(build(force=True))
