"""Build libstereo_geom.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo)."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
OBJ = os.path.join(PKG, "_obj")
LIB = os.path.join(PKG, "libstereo_geom.so")

SOURCES = ["api.cu", "maps.cu", "rectify.cu", "reproject.cu", "post.cu", "multi.cpp", "jpeg.cpp", "calib_yaml.cpp", "stereo_rectify.cpp", "compat.cpp"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC,-fvisibility=hidden",
              "-Xptxas", "-v", "--expt-relaxed-constexpr"]


def _deps_mtime() -> float:
    t = 0.0
    for root in (CSRC, os.path.join(os.path.dirname(PKG), "include")):
        for f in os.listdir(root):
            t = max(t, os.path.getmtime(os.path.join(root, f)))
    return t


def _compile(src: str, verbose: bool) -> str:
    obj = os.path.join(OBJ, src + ".o")
    extra = os.environ.get("SA_EXTRA_NVCC_FLAGS", "").split()      # tuning experiments, e.g. -DSA_BATCH_MINB=3
    cmd = ["nvcc", *NVCC_FLAGS, *extra, "-x", "cu", "-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    with open(obj + ".ptxas.log", "w") as f:
        f.write(r.stderr)
    if verbose:
        sys.stderr.write(r.stderr)
    return obj


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= _deps_mtime():
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), SOURCES))
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB, *objs, "-lcudart_static", "-lpthread", "-ldl", "-lrt"]
    subprocess.check_call(cmd)
    return LIB


def build_variant(name: str, flags: list[str], sources=("rectify.cu",)) -> str:
    """Tuning experiments: recompile `sources` with extra nvcc flags (e.g. -DSA_BATCH16_MINB_FIXED=4) and link them with the
    regular objects into tools/exp/_bin/libstereo_geom_<name>.so; run with SA_GEOM_LIB=<that path>."""
    build()
    out_dir = os.path.join(os.path.dirname(PKG), "tools", "exp", "_bin")
    os.makedirs(out_dir, exist_ok=True)
    objs = []
    for src in SOURCES:
        if src in sources:
            obj = os.path.join(out_dir, f"{name}_{src}.o")
            r = subprocess.run(["nvcc", *NVCC_FLAGS, *flags, "-x", "cu", "-c", os.path.join(CSRC, src), "-o", obj], capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(r.stderr)
            with open(obj + ".ptxas.log", "w") as f:
                f.write(r.stderr)
            objs.append(obj)
        else:
            objs.append(os.path.join(OBJ, src + ".o"))
    lib = os.path.join(out_dir, f"libstereo_geom_{name}.so")
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", lib, *objs, "-lcudart_static", "-lpthread", "-ldl", "-lrt"])
    return lib


if __name__ == "__main__":
    if "--variant" in sys.argv:                               # python -m stereoalgorithms_b200.build --variant NAME FLAG...
        i = sys.argv.index("--variant")
        print(build_variant(sys.argv[i + 1], sys.argv[i + 2:]))
        sys.exit(0)
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
