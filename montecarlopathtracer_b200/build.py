"""Builds libmcpt.so (hand-written sm_100a CUDA + C++ host code) in-tree with nvcc/g++.

    python -m montecarlopathtracer_b200.build [--force]

nvcc cross-compiles for sm_100a without a GPU.  Flags that matter:
  -gencode arch=compute_100a,code=sm_100a   B200 only, no PTX fallback for other parts
  no -use_fast_math; -fmad=false for parity.cu (the kernels compared bit for bit with the CPU oracle);
  -fmad=true for launch.cu                  the render kernels' triangle test is explicit round-to-nearest
                                            intrinsics (bit-exact whatever the flag); shading and the walk's
                                            steering may contract (+2.6 % at K2, profiles/r2_tuning.md)
  -ffp-contract=off (host)                  the KD builder and the triangle pre-computation must
                                            round like the reference's x86 build
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
# MCPT_BUILD_NAME=x builds lib/libmcpt_x.so next to the shipped library (A/B builds for the GPU box: tracer.py loads
# the library named by MCPT_LIB); extra nvcc flags for launch.cu in MCPT_NVCC_EXTRA.
_NAME = os.environ.get("MCPT_BUILD_NAME", "")
OBJDIR = os.path.join(HERE, "build" + ("_" + _NAME if _NAME else ""))
LIB = os.path.join(LIBDIR, "libmcpt" + ("_" + _NAME if _NAME else "") + ".so")
CLI = os.path.join(LIBDIR, "mcpt_cli")

NVCC = os.environ.get("MCPT_NVCC", "/usr/local/cuda/bin/nvcc")
CXX = "/usr/bin/g++"
CUDA_INC = "/usr/local/cuda/include"

CU_SOURCES = ["launch.cu", "parity.cu", "kd_build_gpu.cu", "noalloc.cu"]
CPP_SOURCES = ["kd_build.cpp", "kd_build_sah.cpp", "scene_io.cpp", "device_scene.cpp", "kd_ropes.cpp", "obj_loader.cpp", "image_io.cpp", "abi_geometry.cpp", "abi_render.cpp", "abi_dist.cpp"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-lineinfo", "-diag-suppress=549",
              "-Xcompiler", "-fPIC", "-ccbin", CXX]
# parity.cu (kernels compared bit for bit with the CPU oracle) is never contracted; launch.cu (render kernels: the
# triangle test is explicit _rn intrinsics, shading uses the hardware units) takes MCPT_FMAD (profiles/r2_tuning.md).
FMAD = {"launch.cu": os.environ.get("MCPT_FMAD", "true"), "parity.cu": "false", "kd_build_gpu.cu": "false", "noalloc.cu": "false"}
CXX_FLAGS = ["-std=c++17", "-O2", "-fPIC", "-fopenmp", "-ffp-contract=off", "-Wall", "-I" + CUDA_INC]


def _newer(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _headers() -> list[str]:
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    hs.append(os.path.join(HERE, "..", "include", "mcpt.h"))
    hs.append(os.path.abspath(__file__))
    return hs


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(OBJDIR, exist_ok=True)
    hdrs = _headers()
    objs = []
    for src in CU_SOURCES:
        o = os.path.join(OBJDIR, src + ".o")
        objs.append(o)
        if force or _newer(o, [os.path.join(CSRC, src)] + hdrs):
            cmd = [NVCC] + NVCC_FLAGS + ["-fmad=" + FMAD[src], "-c", os.path.join(CSRC, src), "-o", o]
            if src == "launch.cu":
                cmd[1:1] = os.environ.get("MCPT_NVCC_EXTRA", "").split()
            if os.environ.get("MCPT_TUNE"):       # tuning build: every render_kernel_sm variant (launch.cu)
                cmd.insert(1, "-DMCPT_TUNE")
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            subprocess.check_call(cmd)
    for src in CPP_SOURCES:
        o = os.path.join(OBJDIR, src + ".o")
        objs.append(o)
        if force or _newer(o, [os.path.join(CSRC, src)] + hdrs):
            subprocess.check_call([CXX] + CXX_FLAGS + ["-c", os.path.join(CSRC, src), "-o", o])
    if force or _newer(LIB, objs):
        subprocess.check_call([NVCC, "-shared", "-ccbin", CXX, "-o", LIB] + objs +
                              ["-cudart", "static", "-Xcompiler", "-fopenmp", "-lgomp", "-ldl", "-Wno-deprecated-gpu-targets"])
    cli_src = os.path.join(CSRC, "cli.cpp")
    if not _NAME and (force or _newer(CLI, [cli_src, LIB] + hdrs)):   # headless driver, links only against the C ABI
        subprocess.check_call([CXX, "-std=c++17", "-O2", cli_src, "-o", CLI, "-L" + LIBDIR, "-lmcpt", "-Wl,-rpath,$ORIGIN"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
