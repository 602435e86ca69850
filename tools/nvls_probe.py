#!/usr/bin/env python
"""Build and run tools/nvls_probe.cu (in-switch all_reduce with multimem.ld_reduce / multimem.st,
one process driving all visible GPUs).  Round-2 groundwork; see the header of the .cu file.

    python tools/nvls_probe.py [N]        # builds tools/_nvls_probe if needed, runs it on N GPUs
    python tools/nvls_probe.py --build-only
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
EXE = os.path.join(HERE, "_nvls_probe")
SRC = os.path.join(HERE, "nvls_probe.cu")


def build():
    if os.path.exists(EXE) and os.path.getmtime(EXE) >= os.path.getmtime(SRC):
        return
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-o", EXE, SRC,
                    f"-L{cuda}/lib64/stubs", "-lcuda"], check=True)


if __name__ == "__main__":
    build()
    if "--build-only" not in sys.argv:
        args = [a for a in sys.argv[1:] if not a.startswith("-")]
        sys.exit(subprocess.run([EXE, *args]).returncode)
