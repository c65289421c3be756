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


# ---- known-answer scenes: analytic surfaces with analytic normals, sample ON the surface --------------
def cylinder_scene(r=0.03, half_len=0.03, arc=0.00025, dz=0.002, center=(0.7, 0.0, 0.0)):
    """Upright cylinder of radius r (axis = world z) through `center`; points every `arc` m of arc and `dz` m of
    height, outward radial normals.  Returns xyz (float32), normals (float32), sample (float64, on the surface,
    facing the default cameras near the origin)."""
    n_ring = int(round(2 * np.pi * r / arc))
    th = np.arange(n_ring) * (2 * np.pi / n_ring)
    zs = np.arange(-half_len, half_len + 1e-9, dz)
    T, Z = np.meshgrid(th, zs)
    nrm = np.stack([np.cos(T).ravel(), np.sin(T).ravel(), np.zeros(T.size)], 1)
    xyz = (np.asarray(center)[None, :] + r * nrm + np.stack([np.zeros(T.size), np.zeros(T.size), Z.ravel()], 1))
    xyz = xyz.astype(np.float32)
    k = int(np.argmin(np.abs(th - np.pi)) + n_ring * (len(zs) // 2))  # theta = pi: the side facing -x (the cameras)
    return xyz, nrm.astype(np.float32), xyz[k].astype(np.float64)


def thin_box_scene(width=0.02, height=0.05, half_len=0.04, spacing=0.002, center=(0.7, 0.0, 0.0)):
    """A box `width` wide (two faces with normals -+x), `height` deep along y, closed by a front face at y = -height/2
    (normal -y), 2*half_len long in z.  Narrower than the hand aperture (0.07).  Sample on the -x face."""
    c = np.asarray(center)
    ys = np.arange(-height / 2, height / 2 + 1e-9, spacing)
    zs = np.arange(-half_len, half_len + 1e-9, spacing)
    Y, Z = np.meshgrid(ys, zs)
    n = Y.size
    left = np.stack([np.full(n, -width / 2), Y.ravel(), Z.ravel()], 1)
    right = np.stack([np.full(n, width / 2), Y.ravel(), Z.ravel()], 1)
    xs = np.arange(-width / 2 + spacing, width / 2 - spacing + 1e-9, spacing)
    X, Z2 = np.meshgrid(xs, zs)
    front = np.stack([X.ravel(), np.full(X.size, -height / 2), Z2.ravel()], 1)
    xyz = (np.concatenate([left, right, front]) + c[None, :]).astype(np.float32)
    nrm = np.concatenate([np.tile([-1.0, 0, 0], (n, 1)), np.tile([1.0, 0, 0], (n, 1)),
                          np.tile([0, -1.0, 0], (X.size, 1))]).astype(np.float32)
    k = int(np.argmin(np.abs(left[:, 1]) + np.abs(left[:, 2])))
    return xyz, nrm, xyz[k].astype(np.float64)


def shell_around(sample, radius=0.056, spacing=0.0015):
    """Points on a sphere of `radius` around `sample` (normals radial): whatever the hand frame, one of them lies
    behind the hand base (y < bite - hand_depth = -0.05), finger_hand.cpp:33-34."""
    n_lat = int(round(np.pi * radius / spacing))
    pts = []
    for i in range(n_lat + 1):
        phi = np.pi * i / n_lat
        n_lon = max(1, int(round(2 * np.pi * radius * np.sin(phi) / spacing)))
        lon = np.arange(n_lon) * (2 * np.pi / n_lon)
        pts.append(np.stack([np.sin(phi) * np.cos(lon), np.sin(phi) * np.sin(lon), np.full(n_lon, np.cos(phi))], 1))
    d = np.concatenate(pts)
    return (np.asarray(sample)[None, :] + radius * d).astype(np.float32), d.astype(np.float32)
