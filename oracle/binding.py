"""ctypes binding of the CPU oracle (oracle/libag2_oracle.so).

TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module.  It reuses the
struct definitions of agile_grasp2_b200.capi (the oracle mirrors include/ag2.h
under the prefix ag2o_); the product never imports it.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from agile_grasp2_b200 import capi
from agile_grasp2_b200.scenes import WORKSPACE, reference_camera_poses

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libag2_oracle.so")
_api = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "ag2_oracle.cpp")
    stale = (not os.path.exists(LIB_PATH)) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src)
    if force or stale:
        subprocess.run(["make", "-C", _HERE] + (["-B"] if force else []), check=True, capture_output=True)
    return LIB_PATH


def load() -> capi.Api:
    global _api
    if _api is None:
        build()
        _api = capi.Api(LIB_PATH, "ag2o_")
    return _api


def reference_default_params() -> capi.Params:
    """The reference's compiled-in defaults, restated independently of
    ag2_default_params (grasp_detector.cpp:17-80, :111-127; workspace of
    launch/file_detect_grasps.launch:16)."""
    p = capi.Params()
    p.nn_radius_taubin = 0.01
    p.nn_radius_hands = 0.1
    p.num_orientations = 8
    p.num_threads = 1
    p.num_samples = 1000
    p.finger_width = 0.01
    p.hand_outer_diameter = 0.09
    p.hand_depth = 0.06
    p.hand_height = 0.02
    p.init_bite = 0.015
    left, right = reference_camera_poses()
    for i in range(16):
        p.cam_tf_left[i] = float(left.reshape(-1)[i])
        p.cam_tf_right[i] = float(right.reshape(-1)[i])
    for i in range(6):
        p.workspace[i] = WORKSPACE[i]
    p.antipodal_mode = capi.MODE_PREDICTION
    p.filter_half_grasps = 1
    p.prune = 1
    p.num_selected = 50
    p.min_score_diff = 500.0
    p.min_aperture = 0.03
    p.max_aperture = 0.07
    p.normals_radius = 0.01
    p.rng_seed = 0
    return p


def context(params: capi.Params, threads: int | None = None) -> capi.Context:
    p = params.copy()
    if threads is not None:
        p.num_threads = threads
    return capi.Context(p, api=load())


def point_normals(ctx: capi.Context, idx) -> np.ndarray:
    """Oracle normals of selected cloud points only (the per-point code of compute_normals)."""
    idx = np.ascontiguousarray(idx, dtype=np.int32)
    out = np.zeros((len(idx), 3), np.float32)
    rc = load().lib.ag2o_point_normals(ctx._h, idx.ctypes.data_as(C.c_void_p), C.c_size_t(len(idx)),
                                       out.ctypes.data_as(C.c_void_p))
    if rc:
        raise capi.Ag2Error(rc, "ag2o_point_normals")
    return out


def detmath(u1: np.ndarray, u2: np.ndarray):
    """(log(u1), cos(2 pi u2), Box-Muller normal) from include/ag2_detmath.h as compiled into the oracle."""
    u1 = np.ascontiguousarray(u1, dtype=np.float64)
    u2 = np.ascontiguousarray(u2, dtype=np.float64)
    out = [np.zeros(len(u1)) for _ in range(3)]
    load().lib.ag2o_detmath(u1.ctypes.data_as(C.c_void_p), u2.ctypes.data_as(C.c_void_p), C.c_size_t(len(u1)),
                            *[o.ctypes.data_as(C.c_void_p) for o in out])
    return out


def eigen33(cov: np.ndarray):
    cov = np.ascontiguousarray(cov, dtype=np.float32).reshape(9)
    val = C.c_float()
    vec = np.zeros(3, np.float32)
    load().lib.ag2o_eigen33(cov.ctypes.data_as(C.c_void_p), C.byref(val), vec.ctypes.data_as(C.c_void_p))
    return float(val.value), vec


def average_normal_axis(normals: np.ndarray, sample, cam_origin):
    normals = np.ascontiguousarray(normals, dtype=np.float64).reshape(-1, 3)
    sample = np.ascontiguousarray(sample, dtype=np.float64)
    cam_origin = np.ascontiguousarray(cam_origin, dtype=np.float64)
    out = [np.zeros(3) for _ in range(3)]
    load().lib.ag2o_average_normal_axis(
        normals.ctypes.data_as(C.c_void_p), C.c_int(len(normals)), sample.ctypes.data_as(C.c_void_p),
        cam_origin.ctypes.data_as(C.c_void_p), *[o.ctypes.data_as(C.c_void_p) for o in out])
    return out


def image_postprocess(img: np.ndarray) -> np.ndarray:
    img = np.ascontiguousarray(img, dtype=np.float32).reshape(60, 60, 3)
    out = np.zeros((60, 60, 3), np.uint8)
    load().lib.ag2o_image_postprocess(img.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
    return out


def finger_spacing(fw: float, od: float) -> np.ndarray:
    out = np.zeros(20)
    load().lib.ag2o_finger_spacing(C.c_double(fw), C.c_double(od), out.ctypes.data_as(C.c_void_p))
    return out


def deepen_depths(init_bite: float, hand_depth: float) -> np.ndarray:
    out = np.zeros(64)
    n = C.c_int()
    load().lib.ag2o_deepen_depths(C.c_double(init_bite), C.c_double(hand_depth), out.ctypes.data_as(C.c_void_p),
                                  C.c_int(64), C.byref(n))
    return out[: n.value]


def orientation_angles(R: int) -> np.ndarray:
    out = np.zeros(R)
    load().lib.ag2o_orientation_angles(C.c_int(R), out.ctypes.data_as(C.c_void_p))
    return out


def xavier_weights(seed: int, bias_scale: float = 0.0) -> dict:
    """Caffe 'xavier' filler (uniform(-s, s), s = sqrt(3 / fan_in)) from numpy's PCG64.
    bias_scale > 0 adds small random biases so the bias path is exercised in tests."""
    rng = np.random.default_rng(seed)
    out = {}
    for name, shape in capi.LENET_SHAPES:
        if name.endswith("_w"):
            fan_in = int(np.prod(shape[1:]))
            s = np.sqrt(3.0 / fan_in)
            out[name] = rng.uniform(-s, s, size=shape).astype(np.float32)
        else:
            out[name] = (rng.uniform(-1, 1, size=shape) * bias_scale).astype(np.float32)
    return out
