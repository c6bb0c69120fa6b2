"""ctypes binding of libdspslam_sm100.so (the C ABI declared in include/dspslam.h).

The library is built in-tree by `__graft_entry__.build()` / `make -C dsp-slam-rgbd_b200/csrc`.
There is no fallback: a missing library or a failing call raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdspslam_sm100.so")

MODE_FP32, MODE_TC, MODE_TC_FAST, MODE_TC_EXACT = 0, 1, 4, 5
ST_OK, ST_FEW_POINTS, ST_NAN = 0, 1, 2
GRID_REFERENCE_SHEAR, GRID_LATTICE = 0, 1
DBG_STRIDE = 5120
ABI_VERSION = 2

EXPORTS = [
    "dspslam_abi_version", "dspslam_last_error", "dspslam_launch_count", "dspslam_mode_available", "dspslam_create",
    "dspslam_destroy", "dspslam_sdf", "dspslam_sdf_jac", "dspslam_mean_sdf", "dspslam_grid_sdf",
    "dspslam_gn_joint", "dspslam_gn_pose_only", "dspslam_profile_enable", "dspslam_profile_read",
    "dspslam_marching_cubes", "dspslam_frame_build_kitti", "dspslam_frame_build_mono",
    "dspslam_pg_from_matrices", "dspslam_pg_linearize", "dspslam_pg_oplus", "dspslam_mean_sdf_noalloc", "dspslam_query",
]
Q_WSPLIT_MASK, Q_JAC_EARLY, Q_DEVICE = 0, 1, 2
NOALLOC_MAX_ROWS = 262144
FRAME_INFO = 8
FRAME_MAX_DETS = 256
FRAME_MAX_MASKS = 256


class DecoderDesc(C.Structure):
    _fields_ = [("code_len", C.c_int32), ("hidden", C.c_int32), ("n_hidden", C.c_int32),
                ("latent_in", C.c_int32)]


class JointParams(C.Structure):
    _fields_ = [("k1", C.c_float), ("k2", C.c_float), ("k3", C.c_float), ("k4", C.c_float),
                ("b1", C.c_float), ("b2", C.c_float), ("lr", C.c_float), ("s_damp", C.c_float),
                ("cut_off", C.c_float), ("num_iterations", C.c_int32),
                ("num_depth_samples", C.c_int32), ("ray_phases", C.c_int32), ("refine_window", C.c_float)]


class Profile(C.Structure):
    _fields_ = [(n, C.c_double) for n in ("ms_gn_fwd", "ms_gn_jac", "ms_rows_fwd", "ms_rows_jac")] + \
               [(n, C.c_int64) for n in ("rows_gn_fwd", "rows_gn_jac", "rows_fwd", "rows_jac",
                                         "n_gn_fwd", "n_gn_jac", "n_rows_fwd", "n_rows_jac", "rows_gn_inball",
                                         "rows_gn_refine", "n_gn_refine")] + [("ms_gn_refine", C.c_double), ("ms_gn_jac_early", C.c_double)]


class FrameParams(C.Structure):
    _fields_ = [("img_w", C.c_int32), ("img_h", C.c_int32), ("num_lidar_max", C.c_int32),
                ("min_mask_area", C.c_int32), ("downsample_ratio", C.c_int32), ("max_background", C.c_int32),
                ("K", C.c_float * 9), ("invK", C.c_float * 9), ("T_cam_velo", C.c_float * 12),
                ("k1", C.c_double), ("k2", C.c_double), ("K64", C.c_double * 9), ("invK64", C.c_double * 9)]


class FrameBox(C.Structure):
    _fields_ = [("T_obj_velo", C.c_float * 12), ("lo", C.c_float * 3), ("hi", C.c_float * 3),
                ("half", C.c_float * 3), ("is_front", C.c_int32)]


class NativeError(RuntimeError):
    pass


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise NativeError(
            "libdspslam_sm100.so is not built (%s). Run `python -c 'import __graft_entry__ as g; "
            "g.build()'` or `make -C dsp-slam-rgbd_b200/csrc`. There is no CPU fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, fp = C.c_void_p, C.c_int32, C.c_int64, C.c_void_p
    L.dspslam_abi_version.restype = C.c_int
    L.dspslam_last_error.restype = C.c_char_p
    L.dspslam_launch_count.restype = i64
    L.dspslam_launch_count.argtypes = [vp]
    L.dspslam_query.restype = C.c_int
    L.dspslam_query.argtypes = [vp, C.c_int]
    L.dspslam_mode_available.restype = C.c_int
    L.dspslam_mode_available.argtypes = [vp, C.c_int]
    L.dspslam_create.restype = C.c_int
    L.dspslam_create.argtypes = [C.POINTER(DecoderDesc), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                 C.c_int, C.POINTER(C.c_void_p)]
    L.dspslam_destroy.restype = None
    L.dspslam_destroy.argtypes = [vp]
    L.dspslam_sdf.restype = C.c_int
    L.dspslam_sdf.argtypes = [vp, vp, C.c_int, fp, fp, i64, fp]
    L.dspslam_sdf_jac.restype = C.c_int
    L.dspslam_sdf_jac.argtypes = [vp, vp, C.c_int, fp, fp, i64, fp, fp]
    L.dspslam_mean_sdf.restype = C.c_int
    L.dspslam_mean_sdf.argtypes = [vp, vp, C.c_int, fp, fp, i64, fp]
    L.dspslam_mean_sdf_noalloc.restype = C.c_int
    L.dspslam_mean_sdf_noalloc.argtypes = [vp, vp, C.c_int, fp, fp, i64, fp]
    L.dspslam_grid_sdf.restype = C.c_int
    L.dspslam_grid_sdf.argtypes = [vp, vp, C.c_int, fp, C.c_int, i64, i64, C.c_int, fp]
    L.dspslam_marching_cubes.restype = C.c_int
    L.dspslam_marching_cubes.argtypes = [vp, vp, fp, C.c_int, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int,
                                         fp, fp, fp]
    L.dspslam_gn_joint.restype = C.c_int
    L.dspslam_gn_joint.argtypes = [vp, vp, C.c_int, C.POINTER(JointParams), C.c_int, fp, fp, fp, fp, fp, fp,
                                   i64, i64, C.c_int, C.c_int, fp, fp, fp, fp, fp]
    L.dspslam_gn_pose_only.restype = C.c_int
    L.dspslam_gn_pose_only.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, fp, fp, i64, C.c_int, fp, fp, fp, fp]
    L.dspslam_profile_enable.restype = C.c_int
    L.dspslam_profile_enable.argtypes = [vp, C.c_int]
    L.dspslam_profile_read.restype = C.c_int
    L.dspslam_profile_read.argtypes = [vp, C.POINTER(Profile)]
    L.dspslam_frame_build_kitti.restype = C.c_int
    L.dspslam_frame_build_kitti.argtypes = [vp, vp, C.POINTER(FrameParams), fp, i64, C.POINTER(FrameBox), C.c_int,
                                            fp, C.c_int, fp, fp, fp, fp, fp, fp, fp, fp]
    L.dspslam_frame_build_mono.restype = C.c_int
    L.dspslam_frame_build_mono.argtypes = [vp, vp, C.POINTER(FrameParams), fp, C.c_int, fp, fp, fp]
    L.dspslam_pg_from_matrices.restype = C.c_int
    L.dspslam_pg_from_matrices.argtypes = [vp, vp, fp, C.c_int, fp]
    L.dspslam_pg_linearize.restype = C.c_int
    L.dspslam_pg_linearize.argtypes = [vp, vp, fp, C.c_int, fp, fp, fp, C.c_int, C.c_double, C.c_double,
                                       fp, fp, fp, fp, fp, fp, fp, fp, fp]
    L.dspslam_pg_oplus.restype = C.c_int
    L.dspslam_pg_oplus.argtypes = [vp, vp, fp, fp, C.c_int]
    if L.dspslam_abi_version() != ABI_VERSION:
        raise NativeError("libdspslam_sm100.so ABI version mismatch")
    _lib = L
    return L


def check(rc, what):
    if rc != 0:
        msg = lib().dspslam_last_error()
        raise NativeError("%s failed (code %d): %s" % (what, rc, msg.decode() if msg else "?"))
