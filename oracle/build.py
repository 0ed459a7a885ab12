"""Builds the C oracles with gcc (TEST INFRASTRUCTURE ONLY).

    python -m oracle.build

Outputs go to oracle/_build/ (git-ignored, but shipped to the GPU box by gpurun).
The reference itself has no compilable C/C++ for this path (the
reference is pure Python over un-vendored aubio / CTranslate2), so there is no
oracle/_ref to build.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_build")
LIB = os.path.join(OUT, "libyin_oracle.so")
SRC = os.path.join(HERE, "yin_aubio.c")


def build(force: bool = False) -> str:
    os.makedirs(OUT, exist_ok=True)
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    # -ffp-contract=off: no FMA contraction, i.e. what a baseline x86-64 aubio wheel executes.
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-std=c99", "-ffp-contract=off", "-fno-fast-math",
           "-o", LIB, SRC, "-lm"]
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
