"""Builds libgs_b200.so (hand-written sm_100a kernels + C ABI) in-tree with nvcc.

Two arithmetic profiles, chosen per translation unit ("Arithmetic profiles"):
  literal fp32  -- project.cu, blend_fwd.cu: -fmad=false, so only explicit fmaf() fuses and the
                   results are bit-exact with oracle/gs_oracle.c
  fast fp32     -- blend_bwd.cu, point_bwd.cu, loss.cu (and the integer kernels): default contraction
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
# GS_B200_LIB_NAME: another file name for the library (tuning experiments build variants side by side)
LIB = os.path.join(HERE, os.environ.get("GS_B200_LIB_NAME", "libgs_b200.so"))

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "-I", INCLUDE, "-I", CSRC]
UNITS = {
    "capi.cu": [],
    "project.cu": ["-fmad=false"],
    "keys.cu": [],
    "sort.cu": [],
    "bin.cu": [],
    "blend_fwd.cu": ["-fmad=false"],
    "blend_bwd.cu": [],
    "point_bwd.cu": [],
    "point_bwd_batch.cu": [],
    "loss.cu": [],
    "train_ops.cu": [],
}


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libgs_b200.so cannot be built")


def _stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(INCLUDE, "gs_b200.h"),
                                                               os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not _stale():
        return LIB
    nvcc = nvcc_path()
    objdir = os.path.join(HERE, "build", os.path.basename(LIB).replace(".so", "") if "GS_B200_LIB_NAME" in os.environ else "")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    for unit, extra in UNITS.items():
        obj = os.path.join(objdir, unit.replace(".cu", ".o"))
        tune = os.environ.get("GS_EXTRA_NVCC_FLAGS", "").split()  # tuning experiments only
        cmd = [nvcc, *ARCH, *COMMON, *extra, *tune, "-c", os.path.join(CSRC, unit), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd), flush=True)
        procs.append((unit, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    objs = []
    for unit, obj, proc in procs:
        out, _ = proc.communicate()
        if proc.returncode != 0:
            raise RuntimeError(f"nvcc failed on {unit}:\n{out.decode()}")
        if verbose:
            print(out.decode())
        objs.append(obj)
    cmd = [nvcc, *ARCH, "-shared", "-o", LIB, *objs, "-lcudart"]
    subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
