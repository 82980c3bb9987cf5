"""The CMake route (CMakeLists.txt at the repo root): `cmake -B build && cmake --build build` must produce a
libstereo_geom.so for sm_100a that exports exactly what include/*.h declares, plus - when an OpenCV is on hand - the four
per-model libraries under the reference's own target names (FastACVNet_plus/src/CMakeLists.txt:3,28 and siblings) that
export the reference's unmangled entry points.  CPU only: nvcc cross-compiles; the image has no OpenCV C++ headers, so
the stand-in cv::Mat of tests/fake_opencv plays that part."""
import ctypes as C
import os
import shutil
import subprocess

import pytest

from conftest import ROOT
from test_host_cpu import _declared_symbols
from test_shims_cpu import EXPECTED

TARGETS = {"fastacvnet_plus": "FastACVNet_plus_Algorithm", "raft_stereo": "RAFTStereoAlgorithm", "HitNet": "HitNetAlgorithm",
           "crestereo": "CREStereoAlgorithm"}


@pytest.fixture(scope="module")
def build_dir(tmp_path_factory):
    if not shutil.which("cmake") or not shutil.which("nvcc"):
        pytest.skip("cmake / nvcc not available")
    b = str(tmp_path_factory.mktemp("cmake_build"))
    gen = ["-G", "Ninja"] if shutil.which("ninja") else []
    cfg = subprocess.run(["cmake", "-S", ROOT, "-B", b, *gen, "-DSA_OPENCV_INCLUDE_DIR=" + os.path.join(ROOT, "tests", "fake_opencv")],
                         capture_output=True, text=True)
    assert cfg.returncode == 0, cfg.stdout + cfg.stderr
    bld = subprocess.run(["cmake", "--build", b, "-j", str(os.cpu_count() or 4)], capture_output=True, text=True)
    assert bld.returncode == 0, bld.stdout[-4000:] + bld.stderr[-4000:]
    return b


def _exports(so):
    out = subprocess.check_output(["nm", "-D", "--defined-only", so], text=True)
    return {l.split()[-1] for l in out.splitlines() if " T " in l}


def test_cmake_builds_the_c_abi_library_for_sm_100a(build_dir):
    so = os.path.join(build_dir, "libstereo_geom.so")
    assert os.path.exists(so)
    assert {s for s in _exports(so) if s.startswith("sa_")} == set(_declared_symbols())
    lib = C.CDLL(so)                                                   # loads without a GPU (static CUDA runtime)
    lib.sa_version.restype = C.c_char_p
    assert lib.sa_version().startswith(b"stereo_geom_b200")
    # device code for sm_100a and nothing else
    arch = subprocess.check_output(["cuobjdump", "-lelf", so], text=True)
    assert "sm_100a" in arch and not [l for l in arch.splitlines() if "sm_" in l and "sm_100a" not in l]


@pytest.mark.parametrize("target", list(TARGETS))
def test_cmake_builds_the_reference_named_libraries(build_dir, target):
    so = os.path.join(build_dir, f"lib{target}.so")
    assert os.path.exists(so)
    assert _exports(so) == EXPECTED[TARGETS[target]]
    needed = subprocess.check_output(["readelf", "-d", so], text=True)
    assert "libstereo_geom.so" in needed
