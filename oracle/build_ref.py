#!/usr/bin/env python
"""Compile the REFERENCE's own CUDA kernels, unmodified, into oracle/_ref/  --  TEST INFRASTRUCTURE.

The reference's preprocess / reproject stages exist only as CUDA kernels:
    RAFTStereo/src/stereo_preprocess.cu, CREStereo/src/CREStereo_preprocess.cu,
    FastACVNet_plus/src/FastACVNet_plus_preprocess.cu, HitNet/src/HitNet_preprocess.cu
Each compiles from its single source file plus its own header; the only external include is the
OpenCV umbrella header, which those files do not use (oracle/shim/opencv2/opencv.hpp is an empty
stand-in).  The sources are compiled FROM WHERE THEY LIE under /root/reference -- nothing is copied
into this repository -- and the outputs go only to oracle/_ref/ (git-ignored, but shipped to the
GPU box with the snapshot).  The rest of the reference (the *Algorithm.cpp / TRT*.cpp files) needs
OpenCV C++ and TensorRT, which this image does not have: unbuildable here.

The GPU parity tests use these libraries to pin the oracle's restatement of stages a4/a5 against
the reference's real kernels, and bench.py can time them as the "reference kernels on B200" row.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("SA_REFERENCE_ROOT", "/root/reference")
OUT_DIR = os.path.join(HERE, "_ref")

SOURCES = {  # model -> (source, include dir)
    "RAFTStereo": ("RAFTStereo/src/stereo_preprocess.cu", "RAFTStereo/include"),
    "CREStereo": ("CREStereo/src/CREStereo_preprocess.cu", "CREStereo/include"),
    "FastACVNet_plus": ("FastACVNet_plus/src/FastACVNet_plus_preprocess.cu", "FastACVNet_plus/include"),
    "HitNet": ("HitNet/src/HitNet_preprocess.cu", "HitNet/include"),
}


def lib_path(model: str) -> str:
    return os.path.join(OUT_DIR, f"libref_{model}.so")


def build(force: bool = False):
    """Returns {model: path} for what exists afterwards.  A missing reference tree is not an error
    (the GPU box has none): previously built files are used as they are."""
    out = {}
    have_ref = os.path.isdir(REF)
    os.makedirs(OUT_DIR, exist_ok=True)
    for model, (src, inc) in SOURCES.items():
        dst = lib_path(model)
        s = os.path.join(REF, src)
        if have_ref and os.path.exists(s) and (force or not os.path.exists(dst)):
            cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-shared",
                   "-Xcompiler", "-fPIC", "-I", os.path.join(HERE, "shim"), "-I", os.path.join(REF, inc),
                   "-o", dst, s]
            subprocess.check_call(cmd)
        if os.path.exists(dst):
            out[model] = dst
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
