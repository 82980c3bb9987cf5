"""The C++ drop-in shims (shims/*.cpp) compile against a cv::Mat and export exactly the reference's unmangled names.
CPU only: built with the stand-in cv::Mat of tests/fake_opencv (the image has no OpenCV C++ headers)."""
import os
import subprocess

import pytest

from conftest import ROOT

EXPECTED = {   # reference headers: */include/*Algorithm.h
    "FastACVNet_plus_Algorithm": {"Initialize", "RunFastACVNet_plus", "RunFastACVNet_plus_RectifyImage", "Version", "Release"},
    "CREStereoAlgorithm": {"Initialize", "RunCREStereo", "RunCREStereo_RectifyImage", "Version", "Release"},
    "RAFTStereoAlgorithm": {"Initialize", "RunRAFTStereo", "Version", "Release"},
    "HitNetAlgorithm": {"Initialize", "RunHitNet", "Version", "Release"},
}


def build_shim(name, out_dir):
    lib = os.path.join(ROOT, "stereoalgorithms_b200", "libstereo_geom.so")
    so = os.path.join(str(out_dir), f"lib{name}.so")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-fPIC", "-shared", "-I", os.path.join(ROOT, "tests", "fake_opencv"), "-o", so,
                           os.path.join(ROOT, "shims", name + ".cpp"), lib, "-Wl,-rpath," + os.path.dirname(lib)])
    return so


@pytest.mark.parametrize("name", list(EXPECTED))
def test_shim_exports_reference_names(name, tmp_path):
    so = build_shim(name, tmp_path)
    out = subprocess.check_output(["nm", "-D", "--defined-only", so], text=True)
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert exported == EXPECTED[name]
