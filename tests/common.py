"""Shared helpers of the parity tests: drive libag2.so and the oracle with the same inputs."""
import numpy as np

from agile_grasp2_b200 import capi
from oracle import binding as ob


def make_params(mode=capi.MODE_GEOMETRIC, **kw):
    p = ob.reference_default_params()
    p.init_bite = 0.01  # launch files (file_detect_grasps.launch)
    p.antipodal_mode = mode
    p.num_threads = 8
    p.min_score_diff = 0.0
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def load_scene(ctx, sc, normals=None, sample_points=None):
    ctx.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"], normals=normals)
    if sample_points is not None:
        ctx.set_sample_points(sample_points)
    else:
        ctx.set_sample_indices(sc["samples"])


def both(params, sc, weights=None, normals=None, sample_points=None):
    g = capi.Context(params)
    o = ob.context(params)
    for c in (g, o):
        if weights is not None:
            c.set_lenet_weights(weights)
        load_scene(c, sc, normals, sample_points)
    return g, o


def angle(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    cosv = np.sum(a * b, -1) / (np.linalg.norm(a, axis=-1) * np.linalg.norm(b, axis=-1))
    return np.arccos(np.clip(cosv, -1, 1))


def hyp_int_fields_equal(a, b):
    for f in ["sample_id", "orientation_id", "half_antipodal", "full_antipodal", "finger_slot", "depth_step",
              "num_in_box"]:
        if not np.array_equal(a[f], b[f]):
            return f
    return None
