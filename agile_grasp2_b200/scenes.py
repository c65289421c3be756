"""Seeded synthetic tabletop / clutter clouds for the parity tests and bench.py.

Recipe: points are produced by pin-hole ray casting (640x480,
f = 525 px, c = (319.5, 239.5)) from the reference's camera pose(s)
(grasp_detector.cpp:111-127) against analytic primitives (table plane, boxes,
upright cylinders, spheres), Gaussian range noise sigma = 1 mm along the ray,
transformed to the base frame, stored float32.  The table top is the floor of an
axis-aligned room (back / side walls and ceiling ~1-2 m away), so every ray
returns within Kinect range and one view is exactly 307 200 points.
Sample indices are drawn (seeded, without replacement, ascending) from the points
that lie inside the reference's workspace (launch/file_detect_grasps.launch:16),
which is what preprocessPointCloud's filterWorkspace + subsampleUniformly select from.

This module only generates INPUTS; it is not on the product path.
"""
from __future__ import annotations

import numpy as np

WORKSPACE = (0.48, 0.92, -0.6, 0.4, -0.25, 1.0)
TABLE_Z = -0.20
IMG_W, IMG_H, FOCAL = 640, 480, 525.0
CX, CY = 319.5, 239.5

# grasp_detector.cpp:113-121
BASE_TF = np.array(
    [[0, 0.445417, 0.895323, 0.215], [1, 0, 0, -0.015], [0, 0.895323, -0.445417, 0.23], [0, 0, 0, 1]], dtype=np.float64
)
SQRT_TF = np.array(
    [[0.9366, -0.0162, 0.3500, -0.2863], [0.0151, 0.9999, 0.0058, 0.0058], [-0.3501, -0.0002, 0.9367, 0.0554],
     [0, 0, 0, 1]], dtype=np.float64
)


def reference_camera_poses():
    """cam_tf_left = base_tf * sqrt_tf^-1, cam_tf_right = base_tf * sqrt_tf."""
    return BASE_TF @ np.linalg.inv(SQRT_TF), BASE_TF @ SQRT_TF


def _rays(cam_tf: np.ndarray):
    u, v = np.meshgrid(np.arange(IMG_W, dtype=np.float64), np.arange(IMG_H, dtype=np.float64))
    d_cam = np.stack([(u - CX) / FOCAL, (v - CY) / FOCAL, np.ones_like(u)], axis=-1).reshape(-1, 3)
    d_cam /= np.linalg.norm(d_cam, axis=1, keepdims=True)
    d = d_cam @ cam_tf[:3, :3].T
    o = cam_tf[:3, 3]
    return o, d


ROOM_MIN = np.array([-0.5, -1.5, TABLE_Z])
ROOM_MAX = np.array([1.6, 1.5, 1.5])


def _hit_room(o, d):
    """First hit of the inside of the room box (the camera is inside)."""
    with np.errstate(divide="ignore", invalid="ignore"):
        t1 = (ROOM_MIN - o) / d
        t2 = (ROOM_MAX - o) / d
    t = np.maximum(t1, t2)  # exit distance per axis
    t[~np.isfinite(t)] = np.inf
    return np.min(t, axis=1)


def _hit_box(o, d, center, size, yaw):
    c, s = np.cos(yaw), np.sin(yaw)
    rot = np.array([[c, s, 0], [-s, c, 0], [0, 0, 1]])  # world -> box
    ob = rot @ (o - center)
    db = d @ rot.T
    half = np.asarray(size) / 2
    with np.errstate(divide="ignore", invalid="ignore"):
        t1 = (-half - ob) / db
        t2 = (half - ob) / db
    tmin = np.nanmax(np.minimum(t1, t2), axis=1)
    tmax = np.nanmin(np.maximum(t1, t2), axis=1)
    t = np.where((tmax >= tmin) & (tmin > 0), tmin, np.inf)
    return t


def _hit_cylinder(o, d, center_xy, z0, z1, radius):
    ox, oy = o[0] - center_xy[0], o[1] - center_xy[1]
    a = d[:, 0] ** 2 + d[:, 1] ** 2
    b = 2 * (ox * d[:, 0] + oy * d[:, 1])
    cc = ox * ox + oy * oy - radius * radius
    disc = b * b - 4 * a * cc
    with np.errstate(divide="ignore", invalid="ignore"):
        t = (-b - np.sqrt(np.maximum(disc, 0))) / (2 * a)
    z = o[2] + t * d[:, 2]
    t_side = np.where((disc > 0) & (t > 0) & (z >= z0) & (z <= z1), t, np.inf)
    with np.errstate(divide="ignore", invalid="ignore"):
        tc = (z1 - o[2]) / d[:, 2]
    px, py = ox + tc * d[:, 0], oy + tc * d[:, 1]
    t_cap = np.where((tc > 0) & (px * px + py * py <= radius * radius), tc, np.inf)
    return np.minimum(t_side, t_cap)


def _hit_sphere(o, d, center, radius):
    oc = o - center
    b = 2 * (d @ oc)
    cc = oc @ oc - radius * radius
    disc = b * b - 4 * cc
    t = (-b - np.sqrt(np.maximum(disc, 0))) / 2
    return np.where((disc > 0) & (t > 0), t, np.inf)


def _make_objects(rng: np.random.Generator, n_obj: int, kinds=("box", "cyl", "sphere")):
    """Non-interpenetrating objects resting on the table inside the workspace."""
    objs, placed = [], []
    tries = 0
    while len(objs) < n_obj and tries < 10000:
        tries += 1
        kind = kinds[int(rng.integers(len(kinds)))]
        x = rng.uniform(0.56, 0.84)
        y = rng.uniform(-0.42, 0.28)
        if kind == "box":
            size = rng.uniform(0.03, 0.08, size=3)
            size[2] = rng.uniform(0.05, 0.14)
            rad = 0.5 * float(np.hypot(size[0], size[1]))
            o = ("box", np.array([x, y, TABLE_Z + size[2] / 2]), size, rng.uniform(0, np.pi))
        elif kind == "cyl":
            r = rng.uniform(0.02, 0.04)
            h = rng.uniform(0.08, 0.16)
            rad = r
            o = ("cyl", np.array([x, y]), TABLE_Z, TABLE_Z + h, r)
        else:
            r = rng.uniform(0.03, 0.05)
            rad = r
            o = ("sphere", np.array([x, y, TABLE_Z + r]), r)
        if all(np.hypot(x - px, y - py) > rad + pr + 0.015 for px, py, pr in placed):
            placed.append((x, y, rad))
            objs.append(o)
    return objs


def _cast(cam_tf, objs, rng, noise=0.001):
    o, d = _rays(cam_tf)
    t = _hit_room(o, d)
    for ob in objs:
        if ob[0] == "box":
            t = np.minimum(t, _hit_box(o, d, ob[1], ob[2], ob[3]))
        elif ob[0] == "cyl":
            t = np.minimum(t, _hit_cylinder(o, d, ob[1], ob[2], ob[3], ob[4]))
        else:
            t = np.minimum(t, _hit_sphere(o, d, ob[1], ob[2]))
    ok = np.isfinite(t)
    t = t + rng.normal(0.0, noise, size=t.shape)
    pts = o[None, :] + t[:, None] * d
    return pts[ok].astype(np.float32)


def _in_workspace(xyz, ws=WORKSPACE):
    return (
        (xyz[:, 0] > ws[0]) & (xyz[:, 0] < ws[1]) & (xyz[:, 1] > ws[2]) & (xyz[:, 1] < ws[3]) & (xyz[:, 2] > ws[4])
        & (xyz[:, 2] < ws[5])
    )


def _draw_samples(rng, xyz, n):
    cand = np.flatnonzero(_in_workspace(xyz))
    n = min(n, cand.size)
    return np.sort(rng.choice(cand, size=n, replace=False)).astype(np.int32)


def ring_pose(center, radius, height, ang):
    """Camera on a ring around `center`, looking at it (z = optical axis)."""
    pos = np.array([center[0] + radius * np.cos(ang), center[1] + radius * np.sin(ang), center[2] + height])
    z = center - pos
    z /= np.linalg.norm(z)
    x = np.cross(z, np.array([0, 0, 1.0]))
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    T = np.eye(4)
    T[:3, 0], T[:3, 1], T[:3, 2], T[:3, 3] = x, y, z, pos
    return T


def make_scene(config: str, num_samples: int | None = None, seed: int | None = None, num_points: int | None = None):
    """Returns dict(xyz [N,3] f32, cam_mask [N] u8, num_cams, samples [S] i32, name).

    config: 'c1' tabletop 50k/100, 'c2' single view 307k/1000, 'c3' two-view clutter
    ~614k/10k, 'c4' 16-view fused ~4.9M/100k, 'tiny' (fast unit-test scene).
    """
    left, right = reference_camera_poses()
    if config == "c1":
        rng = np.random.default_rng(1 if seed is None else seed)
        objs = [
            ("box", np.array([0.66, -0.12, TABLE_Z + 0.06]), np.array([0.06, 0.10, 0.12]), 0.3),
            ("cyl", np.array([0.72, 0.10]), TABLE_Z, TABLE_Z + 0.15, 0.03),
        ]
        xyz = _cast(left, objs, rng)
        # keep the 50 000 points nearest the scene centre so the cloud is the dense tabletop region
        d = np.hypot(xyz[:, 0] - 0.70, xyz[:, 1] + 0.02)
        keep = np.sort(np.argsort(d, kind="stable")[: (num_points or 50000)])
        xyz = xyz[keep]
        cam = np.ones(len(xyz), np.uint8)
        S = 100 if num_samples is None else num_samples
        return dict(name="c1", xyz=xyz, cam_mask=cam, num_cams=1, samples=_draw_samples(rng, xyz, S))
    if config == "c2":
        rng = np.random.default_rng(2 if seed is None else seed)
        objs = _make_objects(rng, 5)
        xyz = _cast(left, objs, rng)
        cam = np.ones(len(xyz), np.uint8)
        S = 1000 if num_samples is None else num_samples
        return dict(name="c2", xyz=xyz, cam_mask=cam, num_cams=1, samples=_draw_samples(rng, xyz, S))
    if config == "c3":
        rng = np.random.default_rng(3 if seed is None else seed)
        objs = _make_objects(rng, 10)
        a = _cast(left, objs, rng)
        b = _cast(right, objs, rng)
        xyz = np.concatenate([a, b])
        cam = np.concatenate([np.full(len(a), 1, np.uint8), np.full(len(b), 2, np.uint8)])
        S = 10000 if num_samples is None else num_samples
        return dict(name="c3", xyz=xyz, cam_mask=cam, num_cams=2, samples=_draw_samples(rng, xyz, S))
    if config == "c4":
        rng = np.random.default_rng(4 if seed is None else seed)
        objs = _make_objects(rng, 10)
        center = np.array([0.70, -0.07, TABLE_Z + 0.05])
        parts, cams = [], []
        lo, ro = left[:3, 3], right[:3, 3]
        for k in range(16):
            T = ring_pose(center, 0.8, 0.45, 2 * np.pi * k / 16)
            p = _cast(T, objs, rng)
            src = 0 if np.linalg.norm(T[:3, 3] - lo) <= np.linalg.norm(T[:3, 3] - ro) else 1
            parts.append(p)
            cams.append(np.full(len(p), 1 << src, np.uint8))
        xyz = np.concatenate(parts)
        cam = np.concatenate(cams)
        S = 100000 if num_samples is None else num_samples
        return dict(name="c4", xyz=xyz, cam_mask=cam, num_cams=2, samples=_draw_samples(rng, xyz, S))
    if config == "tiny":
        # a coarse two-view clutter scene: every 4th pixel in each direction -> ~38k points
        rng = np.random.default_rng(7 if seed is None else seed)
        objs = _make_objects(rng, 6)
        a = _cast(left, objs, rng)
        b = _cast(right, objs, rng)
        sel = (np.arange(IMG_H)[:, None] % 2 == 0) & (np.arange(IMG_W)[None, :] % 2 == 0)
        sel = sel.reshape(-1)
        a, b = a[sel[: len(a)]], b[sel[: len(b)]]
        xyz = np.concatenate([a, b])
        # restrict to the workspace neighbourhood so the oracle finishes in seconds
        m = (xyz[:, 0] > 0.40) & (xyz[:, 0] < 1.0) & (xyz[:, 1] > -0.7) & (xyz[:, 1] < 0.5)
        cam = np.concatenate([np.full(len(a), 1, np.uint8), np.full(len(b), 2, np.uint8)])[m]
        xyz = xyz[m]
        S = 64 if num_samples is None else num_samples
        return dict(name="tiny", xyz=xyz, cam_mask=cam, num_cams=2, samples=_draw_samples(rng, xyz, S))
    raise ValueError(config)


def plates_scene(gap=0.05, extent=0.04, spacing=0.002, thickness_pts=1):
    """Known-answer scene: two parallel plates `gap` apart, normals along +-x of a hand
    frame aligned with the world axes.  Returns xyz, normals (pointing outward).  The plates
    are shorter than the hand depth so the hand base clears them; sample ON a plate
    (e.g. [-gap/2, 0, 0]) so the local frame has neighbours."""
    g = np.arange(-extent, extent + 1e-9, spacing)
    yy, zz = np.meshgrid(g, g)
    n = yy.size
    left = np.stack([np.full(n, -gap / 2), yy.ravel(), zz.ravel()], axis=1)
    right = np.stack([np.full(n, gap / 2), yy.ravel(), zz.ravel()], axis=1)
    xyz = np.concatenate([left, right]).astype(np.float32)
    nrm = np.concatenate([np.tile([-1.0, 0, 0], (n, 1)), np.tile([1.0, 0, 0], (n, 1))]).astype(np.float32)
    return xyz, nrm
