"""Build the oracle's native pieces.

  oracle/liboracle.so          gcc, from oracle/*.c (our CPU restatements)
  oracle/_ref/libgpu_flow_ref.so  nvcc, from the UNMODIFIED reference source
                               /root/reference/lib/flow_c/gpu_flow_kernel.cu, compiled where it
                               lies (only when /root/reference is present; nothing is copied
                               into the repo).  It needs a GPU to run, so it is used only by
                               `-m gpu` tests and bench legs as the device-side reference.

TEST INFRASTRUCTURE ONLY -- see oracle/__init__.py.
"""
from __future__ import annotations

import glob
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")
REF_DIR = os.path.join(HERE, "_ref")
REF_LIB = os.path.join(REF_DIR, "libgpu_flow_ref.so")
REFERENCE_FLOW_DIR = "/root/reference/lib/flow_c"


def _newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def build_oracle(force=False):
    srcs = sorted(glob.glob(os.path.join(HERE, "*.c")))
    if not force and _newer(LIB, srcs + [__file__]):
        return LIB
    # -ffp-contract=off: every fused operation in the sources is an explicit fmaf()
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off", "-fno-fast-math",
           "-std=c11", "-o", LIB, *srcs, "-lm"]
    subprocess.check_call(cmd)
    return LIB


def build_ref(force=False):
    """Compile the reference's own CUDA flow kernel for sm_100a (no source is copied)."""
    src = os.path.join(REFERENCE_FLOW_DIR, "gpu_flow_kernel.cu")
    shim = os.path.join(HERE, "ref_flow_shim.cu")  # #includes `src` in place + a device-pointer launch
    if not os.path.exists(src):
        return REF_LIB if os.path.exists(REF_LIB) else None
    if not force and _newer(REF_LIB, [shim, src]):
        return REF_LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    os.makedirs(REF_DIR, exist_ok=True)
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-Xcompiler",
           "-fPIC", "-shared", "-I", REFERENCE_FLOW_DIR, "-o", REF_LIB, shim]
    subprocess.check_call(cmd)
    return REF_LIB


if __name__ == "__main__":
    print(build_oracle(force=True))
    print(build_ref(force=True))
