"""Pins the oracle's restatement of stages a4/a5 against the REFERENCE'S OWN CUDA kernels, compiled unmodified from
/root/reference into oracle/_ref/ by oracle/build_ref.py (the .so files travel to the GPU box; the sources do not).
Also checks the product against them directly.  Needs a B200: -m gpu."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import build_ref
from oracle import oracle as O

pytestmark = pytest.mark.gpu

# mangled names of the reference's host launchers (e.g. RAFTStereo/include/stereo_preprocess.h:15-17)
SYMS = {
    "RAFTStereo": ("_Z21RAFTStereo_preprocessPhPfiiP11CUstream_st", "_Z23cuda_reprojectImageTo3DPhPfS0_S0_ii", O.RAFTSTEREO),
    "CREStereo": ("_Z20CREStereo_preprocessPhPfiiP11CUstream_st", "_Z28CREStereo_reprojectImageTo3DPhPfS0_S0_ii", O.CRESTEREO),
    "FastACVNet_plus": ("_Z26FastACVNet_plus_preprocessPhPfiiP11CUstream_st", "_Z34FastACVNet_plus_reprojectImageTo3DPhPfS0_S0_ii", O.FASTACVNET_PLUS),
    "HitNet": ("_Z17HitNet_preprocessPhS_iiPfP11CUstream_st", "_Z25HitNet_reprojectImageTo3DPhPfS0_S0_ii", O.HITNET),
}


def _lib(model):
    path = build_ref.lib_path(model)
    if not os.path.exists(path):
        pytest.skip(f"{path} not built (oracle/build_ref.py needs the reference tree)")
    return C.CDLL(path)


@pytest.mark.parametrize("model", list(SYMS))
def test_reference_kernels_vs_oracle_and_product(model, demo_pair, calibs):
    import torch
    from stereoalgorithms_b200.geometry import DENSE, StereoGeometry
    pre, rep, mid = SYMS[model]
    lib = _lib(model)
    W, H = 640, 480
    c = calibs["calib_a"]
    rl, rr = O.rectify_pair(c, *demo_pair)
    d = O.synthetic_disparity(W, H, 77)
    if mid == O.RAFTSTEREO:
        d = -d
    vp, ci = C.c_void_p, C.c_int
    dl, dr = torch.from_numpy(rl).cuda(), torch.from_numpy(rr).cuda()
    # ---- preprocess ----
    if model == "HitNet":
        t = torch.empty((6, H, W), dtype=torch.float32, device="cuda")
        f = getattr(lib, pre)
        f.argtypes = [vp, vp, ci, ci, vp, vp]
        f(dl.data_ptr(), dr.data_ptr(), W, H, t.data_ptr(), None)
        torch.cuda.synchronize()
        ref_t = [t.cpu().numpy()]
        want_t = [O.preprocess(mid, rl, rr)]
    else:
        tl = torch.empty((3, H, W), dtype=torch.float32, device="cuda")
        tr = torch.empty((3, H, W), dtype=torch.float32, device="cuda")
        f = getattr(lib, pre)
        f.argtypes = [vp, vp, ci, ci, vp]
        f(dl.data_ptr(), tl.data_ptr(), W, H, None)
        f(dr.data_ptr(), tr.data_ptr(), W, H, None)
        torch.cuda.synchronize()
        ref_t = [tl.cpu().numpy(), tr.cpu().numpy()]
        want_t = list(O.preprocess(mid, rl, rr))
    for a, b in zip(ref_t, want_t):
        assert np.array_equal(a, b), f"{model}: oracle preprocess differs from the reference kernel"
    # ---- reproject ----
    q = torch.from_numpy(O.q_to_f32(c["Q"])).cuda()
    dd = torch.from_numpy(d.copy()).cuda()
    pts = torch.empty((H * W, 6), dtype=torch.float32, device="cuda")
    f = getattr(lib, rep)
    f.argtypes = [vp, vp, vp, vp, ci, ci]
    f(dl.data_ptr(), dd.data_ptr(), pts.data_ptr(), q.data_ptr(), H, W)
    torch.cuda.synchronize()
    ref_p, ref_d = pts.cpu().numpy(), dd.cpu().numpy()
    want_p, want_d = O.reproject_dense(mid, d, rl, c["Q"])
    assert np.array_equal(ref_d, want_d)
    assert np.array_equal(np.isnan(ref_p), np.isnan(want_p)) and np.array_equal(np.isinf(ref_p), np.isinf(want_p))
    fin = np.isfinite(want_p)
    assert np.all(np.abs(ref_p[fin] - want_p[fin]) <= 1e-5 * np.maximum(np.abs(ref_p[fin]), np.abs(want_p[fin])))
    # ---- product vs the reference kernels, no oracle in between ----
    with StereoGeometry(W, H, model=model) as g:
        g.set_calibration(c)
        g.rectify_preprocess(demo_pair[0][None], demo_pair[1][None])
        g.set_disparity(np.stack([d, d])[None] if mid == O.CRESTEREO else d[None, None])
        got_p, _ = g.reproject(1, DENSE)
        got_t = [g.tensor(1)[0]] if model == "HitNet" else [g.tensor(1, 0)[0], g.tensor(1, 1)[0]]
        for a, b in zip(got_t, ref_t):
            assert np.array_equal(a, b)
        assert np.array_equal(np.isnan(got_p[0]), np.isnan(ref_p)) and np.array_equal(np.isinf(got_p[0]), np.isinf(ref_p))
        assert np.all(np.abs(got_p[0][fin] - ref_p[fin]) <= 1e-5 * np.maximum(np.abs(got_p[0][fin]), np.abs(ref_p[fin])))
        assert np.array_equal(g.disparity(1)[0], ref_d)
