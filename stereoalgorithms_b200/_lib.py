"""ctypes loader for libstereo_geom.so.  There is no Python or CPU fallback: if the library is missing or does
not load, importing the bindings raises."""
from __future__ import annotations

import ctypes as C
import os

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SA_GEOM_LIB") or os.path.join(PKG, "libstereo_geom.so")   # the override is for tuning experiments (tools/)

# enum mirrors of include/stereo_geom.h
SA_OK = 0
SA_ERR_INVALID_ARG, SA_ERR_NOT_READY, SA_ERR_CUDA, SA_ERR_IO, SA_ERR_PARSE, SA_ERR_TOO_LARGE, SA_ERR_NOMEM = -1, -2, -3, -4, -5, -6, -7
MODEL_RAFTSTEREO, MODEL_HITNET, MODEL_CRESTEREO, MODEL_FASTACVNET_PLUS = 0, 1, 2, 3
F32, F16 = 0, 1
HOST, DEVICE = 0, 1
LEFT, RIGHT = 0, 1
DENSE, COMPACT = 0, 1
STAGE_RECTIFY, STAGE_RESIZE, STAGE_REPROJECT, STAGE_HEATMAP, STAGE_STUB, STAGE_MAPS = range(6)
FLAG_TIMING, FLAG_NO_HEATMAP_BUF, FLAG_NO_POINT_BUF = 1, 2, 4


class sa_config(C.Structure):
    _fields_ = [("struct_size", C.c_int), ("gpu_id", C.c_int), ("width", C.c_int), ("height", C.c_int),
                ("net_width", C.c_int), ("net_height", C.c_int), ("model", C.c_int), ("tensor_dtype", C.c_int),
                ("max_batch", C.c_int), ("flags", C.c_uint)]


class sa_calibration(C.Structure):
    _fields_ = [(k, C.c_double * n) for k, n in (("K_l", 9), ("D_l", 14), ("R_L", 9), ("P1", 12), ("K_r", 9), ("D_r", 14), ("R_R", 9),
                                                 ("P2", 12), ("Q", 16), ("R", 9), ("T", 3))] + [("nD_l", C.c_int), ("nD_r", C.c_int)]


class sa_image(C.Structure):
    _fields_ = [("data", C.c_void_p), ("rows", C.c_int), ("cols", C.c_int), ("step", C.c_size_t)]


ENGINE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p)

_vp, _i, _u, _sz, _cp = C.c_void_p, C.c_int, C.c_uint, C.c_size_t, C.c_char_p

# name -> (restype, argtypes); every symbol include/*.h declares
SIGNATURES = {
    "sa_create": (_i, [C.POINTER(sa_config), C.POINTER(_vp)]),
    "sa_destroy": (_i, [_vp]),
    "sa_version": (_cp, []),
    "sa_last_error": (_cp, [_vp]),
    "sa_load_calibration_yaml": (_i, [_vp, _cp]),
    "sa_set_calibration": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp]),
    "sa_set_Q": (_i, [_vp, _vp]),
    "sa_get_Q": (_i, [_vp, _vp]),
    "sa_set_maps": (_i, [_vp, _i, _vp, _vp]),
    "sa_get_maps": (_i, [_vp, _i, _vp, _vp]),
    "sa_rectify_preprocess": (_i, [_vp, _i, _vp, _vp, _sz, _i, _i, _vp]),
    "sa_input_tensor": (_vp, [_vp, _i]),
    "sa_disparity_buffer": (_vp, [_vp]),
    "sa_disparity_planes": (_i, [_vp]),
    "sa_stub_disparity": (_i, [_vp, _i, _u, _vp]),
    "sa_set_disparity": (_i, [_vp, _i, _vp, _i, _vp]),
    "sa_reproject": (_i, [_vp, _i, _i, _vp, _vp, _i, _vp]),
    "sa_pointcloud_buffer": (_vp, [_vp]),
    "sa_heatmap": (_i, [_vp, _i, _vp, _i, _vp]),
    "sa_disparity_u8": (_i, [_vp, _i, _vp, _i, _vp]),
    "sa_copy_rectified": (_i, [_vp, _i, _i, _vp, _i, _vp]),
    "sa_decode_capture_jpeg": (_i, [_vp, _i, C.POINTER(_vp), C.POINTER(_sz), _vp]),
    "sa_capture_buffer": (_vp, [_vp]),
    "sa_copy_capture": (_i, [_vp, _i, _vp, _i, _vp]),
    "sa_rectify_preprocess_jpeg": (_i, [_vp, _i, C.POINTER(_vp), C.POINTER(_sz), _i, _vp]),
    "sa_encode_jpeg": (_i, [_vp, _vp, _i, _i, _i, _i, _sz, _i, _vp, _sz, C.POINTER(_sz), _vp]),
    "sa_write_jpeg": (_i, [_vp, _cp, _vp, _i, _i, _i, _i, _sz, _i]),
    "sa_decode_jpeg": (_i, [_vp, _vp, _sz, _vp, _i, _sz, C.POINTER(_i), C.POINTER(_i), _vp]),
    "sa_copy_net_image": (_i, [_vp, _i, _i, _vp, _i, _vp]),
    "sa_copy_tensor": (_i, [_vp, _i, _i, _vp, _i, _vp]),
    "sa_copy_disparity": (_i, [_vp, _i, _vp, _i, _vp]),
    "sa_synchronize": (_i, [_vp]),
    "sa_download": (_i, [_vp, _i, _vp, _vp, _vp, _sz, _vp, _sz, _vp, _vp]),
    "sa_get_stream": (_vp, [_vp]),
    "sa_set_engine": (_i, [_vp, ENGINE_FN, _vp]),
    "sa_run_batch": (_i, [_vp, _i, _vp, _vp, _i, _u, _vp, _vp, _vp, _vp, _vp]),
    "sa_run_batch_multi": (_i, [C.POINTER(_vp), _i, _i, _vp, _vp, _i, _u, _vp, _vp, _vp, _vp, _vp, C.POINTER(_i)]),
    "sa_shard_range": (_i, [_i, _i, _i, C.POINTER(_i), C.POINTER(_i)]),
    "sa_parse_calibration_yaml": (_i, [_cp, C.POINTER(sa_calibration), C.c_char_p, _sz]),
    "sa_write_calibration_yaml": (_i, [_cp, C.POINTER(sa_calibration)]),
    "sa_stereo_rectify": (_i, [_vp, _vp, _i, _vp, _vp, _i, _i, _i, _vp, _vp, _i, C.c_double, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "sa_rectify_calibration": (_i, [C.POINTER(sa_calibration), _i, _i, _i, C.c_double, _i, _i, _vp, _vp]),
    "sa_scale_calibration": (_i, [C.POINTER(sa_calibration), C.c_double, C.c_double, C.c_double, C.POINTER(sa_calibration)]),
    "sa_norm_lut": (_i, [_i, _vp]),
    "sa_resize_table": (_i, [_i, _i, _vp]),
    "sa_reproject_compact_noalloc": (_i, [_vp, _sz, _vp, _sz, _sz, _vp, _i, _i, _i, _i, _vp, _vp, _vp, C.c_char_p, _sz]),
    "sa_host_alloc": (_i, [C.POINTER(_vp), _sz]),
    "sa_host_free": (_i, [_vp]),
    "sa_host_register": (_i, [_vp, _sz]),
    "sa_host_unregister": (_i, [_vp]),
    "sa_stage_time": (_i, [_vp, _i, C.POINTER(C.c_double), C.POINTER(C.c_long)]),
    "sa_reset_timing": (_i, [_vp]),
    "sa_launch_count": (C.c_long, [_vp]),
    "sa_set_option": (_i, [_vp, _cp, C.c_long]),
    # stereo_algorithms_compat.h
    "sa_ref_Initialize": (_vp, [_i, _cp, _i, _cp]),
    "sa_ref_Run": (_i, [_vp, _i, C.POINTER(sa_image), C.POINTER(sa_image), _vp, _vp]),
    "sa_ref_Version": (_cp, [_vp]),
    "sa_ref_Release": (_i, [_vp]),
    "sa_ref_SetEngine": (_i, [_vp, ENGINE_FN, _vp]),
    "sa_ref_SetStubSeed": (_i, [_vp, _u]),
    "sa_ref_SetHostRegister": (_i, [_vp, _i]),
    "sa_ref_Geometry": (_vp, [_vp]),
    "sa_ref_LastError": (_cp, [_vp]),
    "sa_ref_Heatmap": (_i, [_vp, _vp]),
    "sa_write_pointcloud_txt": (_i, [_cp, _vp, _sz]),
    "sa_write_pointcloud_ply": (_i, [_cp, _vp, _sz, _i, _i]),
    "sa_write_pointcloud_pcd": (_i, [_cp, _vp, _sz, _i]),
    "sa_write_pgm": (_i, [_cp, _vp, _i, _i, _sz]),
    "sa_write_ppm_bgr": (_i, [_cp, _vp, _i, _i, _sz]),
}

_lib = None


def load() -> C.CDLL:
    """Load the shared library (building is __graft_entry__.build()'s / stereoalgorithms_b200/build.py's job)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python -m stereoalgorithms_b200.build` (needs nvcc). "
                              "There is no fallback implementation.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)          # AttributeError if the library does not export a declared symbol
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib
