"""CPU tests: pin the oracle's third-party restatements against REAL libraries available here
(scipy cKDTree, numpy eigh, OpenCV 4.13 via cv2, torch CPU) and against known answers."""
import os

import numpy as np
import pytest

from agile_grasp2_b200 import capi, scenes
from common import make_params
from oracle import binding as ob

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def octx(tiny_scene):
    o = ob.context(make_params())
    o.set_cloud(tiny_scene["xyz"], tiny_scene["cam_mask"], tiny_scene["num_cams"])
    return o


def test_radius_search_vs_ckdtree(octx, tiny_scene):
    from scipy.spatial import cKDTree
    xyz = tiny_scene["xyz"]
    tree = cKDTree(xyz.astype(np.float64))
    rng = np.random.default_rng(1)
    for r in (0.01, 0.03, 0.1):
        r2f = np.float32(r * r)
        for i in rng.integers(0, len(xyz), 8):
            q = xyz[i]
            idx_b, d2_b = octx.radius_search(q, r, brute=True)
            idx_t, d2_t = octx.radius_search(q, r, brute=False)
            assert np.array_equal(idx_b, idx_t) and np.array_equal(d2_b, d2_t)  # kd-tree == brute force
            assert np.all(np.diff(d2_b) >= 0) and np.all(d2_b < r2f)
            ref = np.array(sorted(tree.query_ball_point(q.astype(np.float64), r)))
            d2_64 = np.sum((xyz.astype(np.float64) - q.astype(np.float64)) ** 2, axis=1)
            tie = np.abs(d2_64 - float(r2f)) <= 4 * np.spacing(np.float32(r2f))  # documented float-threshold ties
            sym = set(ref.tolist()) ^ set(idx_b.tolist())
            assert all(tie[j] for j in sym)


def test_eigen33_vs_numpy():
    rng = np.random.default_rng(2)
    for _ in range(200):
        a = rng.normal(size=(40, 3)) * np.array([1.0, 0.6, 0.05]) * 0.004
        rot, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        c = np.cov((a @ rot.T).T).astype(np.float32)
        val, vec = ob.eigen33(c)
        w, v = np.linalg.eigh(c.astype(np.float64))
        assert abs(val - w[0]) <= 2e-3 * w[2]
        ang = np.arccos(min(1.0, abs(float(vec @ v[:, 0]))))
        assert ang < 5e-3  # float closed-form solver on a well separated spectrum


def test_average_normal_axis_vs_numpy():
    rng = np.random.default_rng(3)
    for _ in range(50):
        base = rng.normal(size=3); base /= np.linalg.norm(base)
        t = rng.normal(size=3); t -= base * (t @ base); t /= np.linalg.norm(t)
        ang = rng.normal(0, 0.5, size=37)
        N = np.cos(ang)[:, None] * base + np.sin(ang)[:, None] * t + rng.normal(0, 0.02, size=(37, 3))
        N /= np.linalg.norm(N, axis=1, keepdims=True)
        sample, cam = rng.normal(size=3), rng.normal(size=3) + 2
        nrm, bino, axis = ob.average_normal_axis(N, sample, cam)
        w, v = np.linalg.eigh(N.T @ N)
        c = v[:, 0]
        assert abs(abs(axis @ c) - 1) < 1e-9
        j = np.argmax(np.sum((N @ N.T) ** 6, axis=0))
        n_ref = (np.eye(3) - np.outer(c, c)) @ N[j]; n_ref /= np.linalg.norm(n_ref)
        assert abs(abs(nrm @ n_ref) - 1) < 1e-9
        v2 = sample - cam
        assert nrm @ v2 <= 0 and bino @ v2 <= 0
        assert np.allclose(np.cross(nrm, bino), axis, atol=1e-12)


def test_image_tail_vs_opencv():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(4)
    for k in range(6):
        img = np.zeros((60, 60, 3), np.float32)
        m = rng.random((60, 60)) < (0.05 + 0.1 * k)
        v = np.abs(rng.normal(size=(60, 60, 3))); v /= np.linalg.norm(v, axis=2, keepdims=True)
        img[m] = v[m].astype(np.float32)
        ref = cv2.dilate(img, cv2.getStructuringElement(cv2.MORPH_RECT, (3, 3)))
        ref = cv2.cvtColor(ref, cv2.COLOR_BGR2RGB)
        # cv::Mat::convertTo(CV_8UC3, 255.0) == saturate_cast<uchar>(cvRound(v * 255)); values are >= 0 so
        # cv2.convertScaleAbs (same rounding, plus |.|) is the binding-reachable form of it
        ref8 = cv2.convertScaleAbs(ref, alpha=255.0)
        out = ob.image_postprocess(img)
        assert np.array_equal(out, ref8)
    gold = np.load(os.path.join(GOLD, "image_tail_cv2.npz"))
    for i in range(len(gold["inp"])):
        assert np.array_equal(ob.image_postprocess(gold["inp"][i]), gold["out"][i])


def test_lenet_vs_torch():
    torch = pytest.importorskip("torch")
    import torch.nn.functional as F
    w = ob.xavier_weights(7, bias_scale=0.3)
    o = ob.context(make_params(capi.MODE_PREDICTION))
    o.set_lenet_weights(w)
    rng = np.random.default_rng(5)
    imgs = rng.integers(0, 256, size=(9, 60, 60, 3), dtype=np.uint8)
    got = o.classify_images(imgs)
    x = torch.from_numpy(imgs).permute(0, 3, 1, 2).double()
    tw = {k: torch.from_numpy(v).double() for k, v in w.items()}
    y = F.max_pool2d(F.conv2d(x, tw["conv1_w"], tw["conv1_b"]), 2, 2)
    y = F.max_pool2d(F.conv2d(y, tw["conv2_w"], tw["conv2_b"]), 2, 2)
    y = F.relu(F.linear(y.reshape(len(imgs), -1), tw["ip1_w"], tw["ip1_b"]))
    y = F.linear(y, tw["ip2_w"], tw["ip2_b"]).numpy()
    assert np.all(np.abs(got - y) <= np.maximum(1e-3, 1e-5 * np.abs(y)))


def test_host_tables_match_reference_quirks():
    fs = ob.finger_spacing(0.01, 0.09)
    assert fs[9] == 5.204170427930421e-18 and fs[9] != 0.0  # last left slot is 5.2e-18, not 0
    assert fs[10] == 0.0 and abs(fs[19] - 0.08) < 1e-15
    assert np.all(np.diff(fs[:10]) > 0) and np.all(np.diff(fs[10:]) > 0)
    d = ob.deepen_depths(0.01, 0.06)
    assert len(d) == 10 and d[-1] == 0.05999999999999999
    assert len(ob.deepen_depths(0.015, 0.06)) == 9
    a = ob.orientation_angles(8)
    assert a[0] == -np.pi / 2 and np.allclose(np.diff(a), np.pi / 8)


def _run_points(xyz, nrm, pts, **kw):
    p = make_params(capi.MODE_GEOMETRIC, prune=0, **kw)
    o = ob.context(p)
    o.set_cloud(xyz, np.ones(len(xyz), np.uint8), 1, normals=nrm)
    o.set_sample_points(pts)
    o.detect()
    return o


def test_known_answer_plates_full_antipodal():
    gap = 0.05
    xyz, nrm = scenes.plates_scene(gap=gap)
    o = _run_points(xyz, nrm, np.array([[0.0, 0.0, 0.0]]), nn_radius_taubin=0.03)  # frame sees both plates
    h = o.hypotheses()
    assert len(h) >= 1
    full = h[h["full_antipodal"] == 1]
    assert len(full) >= 1                       # the orientation aligned with the plate normal
    assert np.allclose(full["width"], gap, atol=1e-6)
    assert np.all(np.abs(np.abs(full["binormal"][:, 0]) - 1) < 1e-6)  # closing direction = plate normal


def test_known_answer_single_plane_never_full():
    g = np.arange(-0.08, 0.08 + 1e-9, 0.002)
    yy, zz = np.meshgrid(g, g)
    xyz = np.stack([np.zeros(yy.size), yy.ravel(), zz.ravel()], 1).astype(np.float32)
    nrm = np.tile([1.0, 0, 0], (len(xyz), 1)).astype(np.float32)
    o = _run_points(xyz, nrm, np.array([[0.0, 0.0, 0.0]]))
    assert np.all(o.hypotheses()["full_antipodal"] == 0)


def test_sample_without_neighbours_is_dropped(tiny_scene):
    p = make_params(capi.MODE_GEOMETRIC)
    o = ob.context(p)
    o.set_cloud(tiny_scene["xyz"], tiny_scene["cam_mask"], tiny_scene["num_cams"])
    o.set_sample_points(np.array([[5.0, 5.0, 5.0], list(tiny_scene["xyz"][tiny_scene["samples"][0]])]))
    o.detect()
    f = o.frames()
    assert f["valid"][0] == 0 and f["valid"][1] == 1
    assert np.all(o.sweep_records()["num_in_box"][0] == -1)


def test_fewer_than_three_neighbours_gives_nan_normal():
    xyz = np.array([[0, 0, 0], [0.001, 0, 0], [1, 1, 1]], np.float32)
    o = ob.context(make_params())
    o.set_cloud(xyz, None, 1)
    o.compute_normals()
    assert np.all(np.isnan(o.normals()))


def test_score_exactly_at_threshold_is_accepted(tiny_scene):
    w = ob.xavier_weights(11, bias_scale=0.5)
    p = make_params(capi.MODE_PREDICTION, min_score_diff=0.0)
    o = ob.context(p); o.set_lenet_weights(w)
    o.set_cloud(tiny_scene["xyz"], tiny_scene["cam_mask"], tiny_scene["num_cams"])
    o.set_sample_indices(tiny_scene["samples"]); o.detect()
    l = o.logits(); s = (l[:, 1] - l[:, 0]).astype(np.float64)
    thr = float(np.sort(s)[len(s) // 2])
    p.min_score_diff = thr
    o.set_params(p); o.generate_hypotheses(); o.prune(); o.render_images(); o.classify(); o.select()
    assert o.num(capi.LIST_ANTIPODAL) == int(np.sum(s >= thr))  # >= : grasp_detector.cpp:202


def test_oracle_golden_regression(tiny_scene):
    """Golden fixture generated by tests/golden/make_golden.py from the oracle itself at the
    commit that first matched every live-library pin above (regression pin, not a reference pin)."""
    gold = np.load(os.path.join(GOLD, "oracle_tiny.npz"))
    o = ob.context(make_params(capi.MODE_GEOMETRIC))
    o.set_cloud(tiny_scene["xyz"], tiny_scene["cam_mask"], tiny_scene["num_cams"])
    o.set_sample_indices(tiny_scene["samples"]); o.detect()
    sw = o.sweep_records()
    assert np.array_equal(sw["fingers"], gold["fingers"]) and np.array_equal(sw["antipodal"], gold["antipodal"])
    assert np.array_equal(sw["num_in_box"], gold["num_in_box"])
    assert np.array_equal(o.normals().view(np.uint32), gold["normals_bits"])
    h = o.hypotheses()
    assert np.allclose(h["bottom"], gold["bottom"], atol=1e-12) and np.allclose(h["width"], gold["width"], atol=1e-12)


def test_preprocess_vs_numpy(tiny_scene):
    """f1: workspace filter + voxelisation against an independent numpy formulation
    (np.unique over lexicographic voxel ids).  Camera source as the reference computes it
    (cloud_camera.cpp:139-153): the i-th voxel of the sorted set takes the source of idx_cam_source[i], the
    i-th point that OPENED a voxel in scan order -- not of the voxel's own first point."""
    p = make_params()
    o = ob.context(p)
    xyz, cam = tiny_scene["xyz"].copy(), tiny_scene["cam_mask"]
    xyz[5] = np.nan
    o.preprocess_cloud(xyz, cam, tiny_scene["num_cams"], voxelize=True, voxel_size=0.003)
    oxyz, ocam = o.cloud()
    ws = np.array(list(p.workspace))
    x64 = xyz.astype(np.float64)
    keep = ((x64[:, 0] > ws[0]) & (x64[:, 0] < ws[1]) & (x64[:, 1] > ws[2]) & (x64[:, 1] < ws[3]) &
            (x64[:, 2] > ws[4]) & (x64[:, 2] < ws[5]))
    f, fc = xyz[keep], cam[keep]
    mn = f.min(0)
    cell = np.float32(0.003)
    v = np.floor((f - mn) / cell).astype(np.int64)
    key = (v[:, 0] << 42) | (v[:, 1] << 21) | v[:, 2]
    uk, first = np.unique(key, return_index=True)
    vv = v[first]
    exp = (vv.astype(np.float32) * cell + mn).astype(np.float32)
    assert oxyz.shape == exp.shape
    assert np.array_equal(oxyz.view(np.uint32), exp.view(np.uint32))
    assert np.array_equal(ocam, fc[np.sort(first)])  # openers in scan order, assigned to voxels in sorted order
    assert not np.array_equal(fc[np.sort(first)], fc[first])  # the two-camera scene tells the two readings apart
    idx = o.subsample_uniformly(100, seed=3)
    assert len(idx) == 100 and np.all(np.diff(idx) > 0) and idx[-1] < len(oxyz)


# ---- known-answer scenes with hand-derived expectations (the reference holds no golden vectors) ---------
from common import cylinder_scene, shell_around, thin_box_scene  # noqa: E402


def test_known_answer_cylinder_r3cm():
    """Cylinder r = 3 cm (diameter 6 cm < aperture 7 cm).  Derivation: all surface normals are perpendicular to the
    cylinder axis, so sum n n^T has a zero eigenvalue along it and the hand axis IS the cylinder axis; every
    cross-section the hand sees is a circle of radius r, whose extent along ANY closing direction perpendicular to
    the axis is 2r as soon as both tangent points are inside the closing region, and at the tangent points the
    normals are +-(closing direction): inside the 30 degree friction cones on both sides => FULL antipodal
    (antipodal.cpp:8-84), width = 2r up to the arc sampling (r(1 - cos(arc / 2r)) < 1e-8).  A closing region cut
    below the tangent points can still be FULL only while the boundary normals stay inside the cone:
    width >= 2 r cos(30 deg)."""
    r = 0.03
    xyz, nrm, sample = cylinder_scene(r=r)
    o = _run_points(xyz, nrm, sample[None, :], nn_radius_taubin=0.01)
    f = o.frames()
    assert f["valid"][0] == 1 and abs(abs(f["curvature_axis"][0][2]) - 1) < 1e-9  # hand axis = cylinder axis
    h = o.hypotheses()
    full = h[h["full_antipodal"] == 1]
    assert len(full) >= 1
    assert abs(full["width"].max() - 2 * r) < 1e-6
    assert np.all(full["width"] >= 2 * r * np.cos(np.pi / 6) - 1e-6) and np.all(h["width"] <= 2 * r + 1e-6)
    assert np.all(np.abs(np.abs(h["axis"][:, 2]) - 1) < 1e-9)


def test_known_answer_thin_box_inside_aperture():
    """A 2 cm wide box between the fingers (aperture 7 cm): for the orientation whose closing direction is the box's
    x axis the left / right extremal sets are the two side faces (normals -+x: inside both friction cones, same y and
    z ranges) => FULL antipodal with width exactly the box width; the front face (normal along -y) never enters a
    friction cone (its normal is perpendicular to every closing direction that is perpendicular to y ... and at most
    cos = sin(angle) otherwise), so no orientation reports a width above the box diagonal in the closing plane."""
    w = 0.02
    xyz, nrm, sample = thin_box_scene(width=w)
    o = _run_points(xyz, nrm, sample[None, :], nn_radius_taubin=0.03)
    h = o.hypotheses()
    full = h[h["full_antipodal"] == 1]
    aligned = full[np.abs(np.abs(full["binormal"][:, 0]) - 1) < 1e-6]
    assert len(aligned) >= 1
    assert np.allclose(aligned["width"], w, atol=1e-6)  # float32 coordinates: 0.69 and 0.71 are 2e-8 off
    assert np.all(h["width"] <= np.hypot(w, 0.05) + 1e-6)


def test_known_answer_point_below_hand_base_blocks_every_slot():
    """finger_hand.cpp:33-34: a point with y < bottom = bite - hand_depth makes evaluateFingers return all-false.  A
    shell of points 5.6 cm around the sample (> hand_depth - init_bite = 5 cm) puts such a point behind the hand base
    for EVERY frame and orientation (it needs |z| < hand_height = 2 cm: the shell's equator band): all 20 slots
    false for all orientations, no hypothesis.  Without the shell the same scene yields hands (control)."""
    xyz, nrm, sample = thin_box_scene()
    o = _run_points(xyz, nrm, sample[None, :], nn_radius_taubin=0.03)
    assert len(o.hypotheses()) >= 1 and np.any(o.sweep_records()["fingers"] != 0)
    sx, sn = shell_around(sample)
    o2 = _run_points(np.concatenate([xyz, sx]), np.concatenate([nrm, sn]), sample[None, :], nn_radius_taubin=0.03)
    sw = o2.sweep_records()
    assert np.array_equal(o2.frames()["normal"], o.frames()["normal"])  # the shell is outside the frame's neighbourhood
    assert np.all(sw["fingers"] == 0) and np.all(sw["emitted"] == 0) and len(o2.hypotheses()) == 0


@pytest.mark.parametrize("init_bite,hand_depth", [(0.0, 0.005), (0.0, 0.015), (0.0, 0.03), (0.01, 0.06), (0.015, 0.06),
                                                  (0.005, 0.01), (0.02, 0.02)])
def test_deepen_list_includes_an_exactly_reachable_hand_depth(init_bite, hand_depth):
    """finger_hand.cpp:122: `for (depth = init_bite + 0.005; depth <= hand_depth; depth += 0.005)` in IEEE double.
    Python floats ARE IEEE doubles: the same loop here is an independent restatement, including the cases where the
    repeated addition lands exactly on hand_depth (0 + 0.005 <= 0.005; 0.01 + 0.005 == 0.015) and the one where it
    lands one ulp above it and the last step is lost (0.005 * 3 additions against 0.03)."""
    want = []  # the deepen steps proper; step 0 of a sweep record is init_bite itself
    d = init_bite + 0.005
    while d <= hand_depth:
        want.append(d)
        d += 0.005
    got = ob.deepen_depths(init_bite, hand_depth)
    assert list(got) == want
    if (init_bite, hand_depth) == (0.0, 0.005):
        assert want == [0.005]             # the boundary value itself is a step
    if (init_bite, hand_depth) == (0.01, 0.06):
        assert want[-1] == 0.05999999999999999 and len(want) == 10


def deepen_scene():
    """A 2 cm wide, 8 mm deep box and a sample on the near edge of its -x face: in the hand frame whose closing
    direction is the box's x axis and whose approach is +y every point has 0 <= y <= 0.008."""
    xyz, nrm, _ = thin_box_scene(width=0.02, height=0.008)
    left = xyz[np.abs(xyz[:, 0] - np.float32(0.69)) < 1e-6]
    return xyz, nrm, left[np.argmin(left[:, 1] + 10 * np.abs(left[:, 2]))].astype(np.float64)


def test_known_answer_deepen_reaches_the_last_depth_in_free_space():
    """deepenHand with hand_depth exactly reachable: init_bite 0.01, hand_depth 0.025 -> the loop of
    finger_hand.cpp:122 yields 0.015, 0.02, 0.025 (0.02 + 0.005 == 0.025 in IEEE double: the `<=` admits the third
    step).  In the aligned frame of `deepen_scene` the box lies between the fingers of every hand that straddles it
    and no point is ever behind the hand base -- at the last step bottom = 0.025 - 0.025 = 0 and all y >= 0 -- so the
    loop runs to its end: depth_step = 3, top = 0.025 and bottom = 0.0 exactly.  One orientation further the box
    corner at y < 0 blocks the third step."""
    xyz, nrm, sample = deepen_scene()
    assert list(ob.deepen_depths(0.01, 0.025)) == [0.015, 0.02, 0.025]
    o = _run_points(xyz, nrm, sample[None, :], nn_radius_taubin=0.03, init_bite=0.01, hand_depth=0.025)
    h = o.hypotheses()
    aligned = h[(np.abs(np.abs(h["binormal"][:, 0]) - 1) < 1e-6) & (h["approach"][:, 1] > 0.999999)]
    assert len(aligned) == 1 and aligned["depth_step"][0] == 3
    assert aligned["hand_top"][0] == 0.025 and aligned["hand_bottom"][0] == 0.0
    assert h["depth_step"].min() < 3


def _expected_grasp_image(pts, nrm, aligns):
    """learning.cpp:143-209 restated with Python floats (IEEE double) and cv2 for the tail."""
    import math
    cv2 = pytest.importorskip("cv2")
    y = pts[:, 1] - (pts[:, 1].min() if aligns else 0.0)
    cellsize = 1.0 / float(60)
    acc = {}
    for j in range(len(pts)):
        hcell = math.floor(float(pts[j, 0]) / cellsize)
        vcell = math.floor(float(y[j]) / cellsize)
        acc.setdefault(hcell + vcell * 60, []).append(j)
    img = np.zeros((60, 60, 3), np.float32)
    for i, js in acc.items():
        if not (0 <= i < 3600):
            continue  # only cells 0..3599 are ever matched (:166-179)
        s = np.zeros(3)
        for j in js:
            s = s + nrm[j]
        a = np.abs((1.0 / np.sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2])) * s)
        img[59 - i // 60, i % 60] = a.astype(np.float32)
    img = cv2.dilate(img, cv2.getStructuringElement(cv2.MORPH_RECT, (3, 3)))
    img = cv2.cvtColor(img, cv2.COLOR_BGR2RGB)
    return cv2.convertScaleAbs(img, alpha=255.0)


@pytest.mark.parametrize("aligns", [False, True])
def test_image_cells_exactly_on_a_sixtieth_boundary(aligns):
    """learning.cpp:152-156: cell = floor(x / (1.0 / 60)).  For x = k / 60 the quotient is k or k - 1ulp depending on
    k (e.g. floor((29/60) / (1/60)) = 28, floor((30/60) / (1/60)) = 30): points on every k/60 in x and in y, one
    ulp below and above some of them, and cells aliasing across the row end (x' >= 1), against the double-precision
    restatement above + OpenCV's dilate / cvtColor / convertTo."""
    import math
    ks = np.arange(60)
    # in IEEE double floor((k / 60) / (1 / 60)) == k for every k < 60, but one ulp below k / 60 the quotient can round
    # back up to k: the cell then differs from floor(x * 60)'s -- which is why the division is kept as written
    assert all(math.floor((k / 60.0) / (1.0 / 60)) == k for k in ks)
    below = [np.nextafter(k / 60.0, 0.0) for k in ks[1:]]
    odd = [k for k, x in zip(ks[1:], below) if math.floor(x / (1.0 / 60)) != math.floor(x * 60.0)]
    assert odd == [3, 6, 12, 24, 33, 48]  # one ulp below k/60, yet cell k: these points are in the set below
    rng = np.random.default_rng(12)
    pts = []
    for k in ks:
        pts.append([k / 60.0, ((7 * k) % 60) / 60.0, 0.5])
        pts.append([np.nextafter(k / 60.0, 0.0) if k else 0.0, ((11 * k + 3) % 60) / 60.0, 0.25])
        pts.append([np.nextafter(k / 60.0, 1.0), ((13 * k + 5) % 60) / 60.0, 0.75])
    pts += [[1.0, 0.2, 0.5], [1.0 + 1.0 / 60, 0.3, 0.5], [0.5, 1.0 - 1e-12, 0.5], [0.5, 1.0, 0.5], [-1.0 / 60, 0.5, 0.5]]
    pts = np.array(pts, np.float64)
    nrm = rng.normal(size=pts.shape)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    want = _expected_grasp_image(pts, nrm, aligns)
    o = ob.context(make_params())
    got = o.render_point_sets(pts, nrm, [0, len(pts)], aligns=aligns)[0]
    assert np.array_equal(got, want), int(np.sum(got != want))


def test_eigen33_vs_numpy_at_the_normals_tolerance():
    """How far pcl::eigen33's float closed form (trigonometric roots, eigenvector from row cross products) is from
    the exact smallest eigenvector of its own float input, on covariances of the kind the path produces (1 cm
    neighbourhoods, thin in one direction): the eigenvalue agrees with numpy.linalg.eigh to 1e-6 of the largest one,
    the eigenvector only to ~5e-4 rad (measured worst case 4.8e-4 over 2 000 matrices, flat in the spectral gap from
    0.05 to 1).  That error belongs to the reference's algorithm -- PCL computes its normals with exactly this closed
    form -- so it cannot be held to the 1e-4 rad of the GPU-vs-oracle normals test, which compares two evaluations
    of the SAME closed form in the same operation order (bit-equal for > 99.9 % of the points).  What this pin shows is
    that the restatement is a correct eigen-solver to the accuracy of the algorithm: 1e-3 rad, five times tighter
    than round 1's bound and twice the measured worst case."""
    rng = np.random.default_rng(22)
    worst = 0.0
    for _ in range(600):
        a = rng.normal(size=(60, 3)) * np.array([1.0, rng.uniform(0.3, 1.0), rng.uniform(0.01, 0.15)]) * 0.004
        rot, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        c = np.cov((a @ rot.T).T).astype(np.float32)
        val, vec = ob.eigen33(c)
        w, v = np.linalg.eigh(c.astype(np.float64))
        sep = (w[1] - w[0]) / w[2]
        if sep < 0.05:
            continue  # near-degenerate neighbourhoods are excluded from the normals tolerance as well
        ang = float(np.arccos(min(1.0, abs(float(vec @ v[:, 0])))))
        worst = max(worst, ang)
        assert abs(val - w[0]) <= 1e-6 * w[2]
        assert ang <= 1e-3, (ang, sep)
    assert 1e-5 < worst  # the closed form is NOT exact: documents why the bound is not 1e-4


def test_reproducible_log_and_cos_of_the_sampler_vs_numpy():
    """include/ag2_detmath.h (deviation D5): log and cos(2 pi u) as fixed sequences of correctly rounded double
    operations, compiled into the oracle and into the device sampler.  Pinned here against numpy's libm: a few ulp,
    which is all Box-Muller needs; the bit-equality of GPU and oracle draws is tested on the GPU."""
    rng = np.random.default_rng(5)
    u1 = np.concatenate([rng.random(200000), 2.0 ** -rng.integers(1, 53, 2000), [1.0 - 2.0 ** -53, 0.5, 2.0 ** -53]])
    u1 = u1[u1 > 0]
    u2 = np.concatenate([rng.random(len(u1) - 9), [0.0, 0.125, 0.25, 0.375, 0.5, 0.625, 0.75, 0.875, 1.0 - 2.0 ** -53]])
    lg, cs, nrm = ob.detmath(u1, u2)
    ref_log = np.log(u1)
    assert np.max(np.abs(lg - ref_log) / np.maximum(np.abs(ref_log), 1e-300)) < 2e-15
    # cos(2 pi u) with the argument reduced exactly, as the header does (numpy's cos of a rounded 2 pi u is the worse one)
    k = np.floor(4 * u2 + 0.5)
    r = u2 - 0.25 * k
    z = 2 * np.pi * r
    ref_cos = np.where(k % 4 == 0, np.cos(z), np.where(k % 4 == 1, -np.sin(z), np.where(k % 4 == 2, -np.cos(z), np.sin(z))))
    assert np.max(np.abs(cs - ref_cos)) < 2e-15
    ref_n = np.sqrt(-2.0 * ref_log) * ref_cos
    assert np.max(np.abs(nrm - ref_n)) < 1e-13
    assert cs[len(u1) - 9] == 1.0 and abs(cs[len(u1) - 7]) < 1e-15 and cs[len(u1) - 5] == -1.0  # u = 0, 1/4, 1/2


def test_find_clusters_vs_numpy_restatement():
    """HandleSearch::findClusters (handle_search.cpp:4-80, remove_inliers = false) restated in numpy from the reference
    text, independently of the oracle's loops: same hands kept, same moved positions and mean scores."""
    from common import load_scene
    sc = scenes.make_scene("tiny", num_samples=700)
    min_inliers = 2
    o = ob.context(make_params(capi.MODE_NONE, min_inliers=min_inliers))
    load_scene(o, sc)
    o.detect()
    hands = o.hypotheses(capi.LIST_HYPOTHESES)
    assert len(hands) > 50
    o.find_clusters(capi.LIST_HYPOTHESES)
    got = o.hypotheses(capi.LIST_CLUSTERS)
    axis, bottom, score = hands["axis"], hands["bottom"], hands["score"]
    n = len(hands)
    keep, moved, mean = [], [], []
    cos_thresh = np.cos(15.0 * np.pi / 180.0)
    for i in range(n):
        aligned = np.abs(axis @ axis[i]) > cos_thresh
        delta = bottom[i] - bottom                       # delta_pos = pos_i - pos_j
        near = np.linalg.norm(delta, axis=1) <= 0.05
        proj = delta - np.outer(delta @ axis[i], axis[i])  # (I - a a^T) delta
        on_axis = np.linalg.norm(proj, axis=1) <= 0.005
        inl = aligned & near & on_axis
        inl[i] = False
        k = int(inl.sum())
        if k >= min_inliers:
            keep.append(i)
            moved.append(bottom[inl].sum(0) / k - bottom[i])
            mean.append(score[inl].sum() / k)
    assert 0 < len(keep) < n
    assert len(got) == len(keep)
    assert np.array_equal(got["sample_id"], hands["sample_id"][keep])
    assert np.array_equal(got["orientation_id"], hands["orientation_id"][keep])
    moved = np.array(moved)
    for f in ("bottom", "surface", "top"):
        assert np.allclose(got[f], hands[f][keep] + moved, rtol=0, atol=1e-12), f
    assert np.allclose(got["score"], np.array(mean), rtol=0, atol=1e-12)
    assert np.array_equal(got["axis"], hands["axis"][keep])


def test_prune_vs_numpy_restatement():
    """GraspDetector::pruneGraspsOnHandParameters (grasp_detector.cpp:363-395, called at :150-156 with the workspace as
    FLOAT arguments and the cloud's minimum z) restated in numpy from the reference text."""
    from common import load_scene
    sc = scenes.make_scene("tiny", num_samples=700)
    xyz = sc["xyz"]
    lo, hi = xyz.min(0), xyz.max(0)
    mid = 0.5 * (lo + hi)
    p = make_params(capi.MODE_NONE)
    ws = [float(lo[0] - 1.0), float(mid[0] + 0.02), float(mid[1] - 0.06), float(hi[1] + 1.0), -10.0, 10.0]  # cuts in x and y
    for k in range(6):
        p.workspace[k] = ws[k]
    for filt in (1, 0):
        p.filter_half_grasps = filt
        o = ob.context(p)
        load_scene(o, sc)
        o.detect()
        hands, got = o.hypotheses(capi.LIST_HYPOTHESES), o.hypotheses(capi.LIST_PRUNED)
        half_width = 0.5 * p.hand_outer_diameter
        corners = np.stack([hands["bottom"] + half_width * hands["binormal"], hands["bottom"] - half_width * hands["binormal"],
                            hands["top"] + half_width * hands["binormal"], hands["top"] - half_width * hands["binormal"],
                            hands["bottom"] - 0.10 * hands["approach"]], axis=1)  # hands x 5 x 3
        mn, mx = corners.min(1), corners.max(1)
        f32 = lambda v: float(np.float32(v))  # noqa: E731  (float parameters of the reference function)
        min_z = float(xyz[:, 2].min())        # pcl::getMinMax3D of the processed cloud
        keep = ((hands["width"] >= p.min_aperture) & (hands["width"] <= p.max_aperture) & (mn[:, 2] >= min_z) &
                (mn[:, 1] >= f32(ws[2])) & (mx[:, 1] <= f32(ws[3])) & (mn[:, 0] >= f32(ws[0])) & (mx[:, 0] <= f32(ws[1])))
        if filt:
            keep &= hands["half_antipodal"] != 0
        assert 0 < keep.sum() < len(hands), (filt, keep.sum(), len(hands))
        assert len(got) == keep.sum()
        assert np.array_equal(got["sample_id"], hands["sample_id"][keep])
        assert np.array_equal(got["orientation_id"], hands["orientation_id"][keep])


def test_antipodal_flags_vs_numpy_restatement():
    """Antipodal::evaluateGrasp (antipodal.cpp:8-84) on the closing region of every hypothesis of a scene, restated in
    numpy from the reference text on the oracle's own in-box sets and normals: extremal points within 0.003 of the
    scaled x range (x is scaled by 1 / 0.1, hand_search.cpp:399-409), normals within the 30 degree friction cone of the
    closing direction, HALF if one side has such points, FULL if both sides do and their y and z ranges overlap."""
    from common import load_scene
    sc = scenes.make_scene("tiny", num_samples=300)
    o = ob.context(make_params(capi.MODE_NONE))
    load_scene(o, sc)
    o.detect()
    hands = o.hypotheses(capi.LIST_HYPOTHESES)
    normals = o.normals().astype(np.float64)
    xyz = sc["xyz"]
    cos_f = np.cos(30.0 * np.pi / 180.0)
    n_half = n_full = 0
    for h, hy in enumerate(hands):
        ids = o.in_box(capi.LIST_HYPOTHESES, h)
        assert len(ids) == hy["num_in_box"] > 0
        sample = xyz[sc["samples"][hy["sample_id"]]]
        d = (xyz[ids] - sample).astype(np.float64)        # centred in float, then double (hand_search.cpp:209)
        x, y, z = d @ hy["binormal"], d @ hy["approach"], d @ hy["axis"]
        nx = normals[ids] @ hy["binormal"]
        xs = x * 10.0                                       # only x needs its scale: the threshold is absolute
        left = (xs < xs.min() + 0.003) & (-nx > cos_f)
        right = (xs > xs.max() - 0.003) & (nx > cos_f)
        half = bool(left.any() or right.any())
        full = False
        if left.any() and right.any():
            full = bool(min(y[left].max(), y[right].max()) > max(y[left].min(), y[right].min()) and
                        min(z[left].max(), z[right].max()) > max(z[left].min(), z[right].min()))
        assert half == bool(hy["half_antipodal"]) and full == bool(hy["full_antipodal"]), (h, half, full)
        n_half += half
        n_full += full
    assert len(hands) > 100 and 0 < n_full < n_half < len(hands)


def test_grasp_images_vs_numpy_restatement():
    """Learning::createGraspImages -> createInstance -> convertToImageRGB (learning.cpp:4-33, 221-249, 143-209) with the
    unit-box scaling of hand_search.cpp:399-409, restated in numpy from the reference text on the oracle's in-box sets
    and normals: cell of every point, per-cell double sums in list order, normalise, abs, float image, 3x3 dilation,
    BGR2RGB, x255 with round-half-even.  (The OpenCV tail alone is pinned against cv2 in test_image_tail_vs_opencv.)"""
    from common import load_scene
    sc = scenes.make_scene("tiny", num_samples=200)
    p = make_params(capi.MODE_NONE)
    o = ob.context(p)
    load_scene(o, sc)
    o.detect()
    hands = o.hypotheses(capi.LIST_PRUNED)
    imgs = o.render_images()
    assert len(hands) == len(imgs) > 30
    normals = o.normals().astype(np.float64)
    xyz = sc["xyz"]
    fs = ob.finger_spacing(p.finger_width, p.hand_outer_diameter)
    cellsize = 1.0 / 60.0
    differing = total = 0
    for h, hy in enumerate(hands):
        ids = o.in_box(capi.LIST_PRUNED, h)
        sample = xyz[sc["samples"][hy["sample_id"]]]
        d = (xyz[ids] - sample).astype(np.float64)
        frame = np.stack([hy["binormal"], hy["approach"], hy["axis"]], axis=1)  # frame_rot columns
        pts, nrm = d @ frame, normals[ids] @ frame
        e = int(hy["finger_slot"])
        left, right = fs[e] + p.finger_width, fs[10 + e]
        left_const = left - 0.5 * (0.1 - (right - left))
        xs = (1.0 / 0.1) * (pts[:, 0] - left_const)
        ys = (1.0 / (hy["hand_top"] - hy["hand_bottom"])) * (pts[:, 1] - hy["hand_bottom"])
        ys = ys - ys.min()  # aligns = true (default argument)
        cell = np.floor(xs / cellsize).astype(np.int64) + np.floor(ys / cellsize).astype(np.int64) * 60
        ok = (cell >= 0) & (cell < 3600)
        acc = np.zeros((3600, 3))
        np.add.at(acc, cell[ok], nrm[ok])  # unbuffered: the points of a cell are added in list order
        cnt = np.bincount(cell[ok], minlength=3600)
        img = np.zeros((60, 60, 3), np.float32)
        occ = np.nonzero(cnt)[0]
        v = np.abs(acc[occ] * (1.0 / np.linalg.norm(acc[occ], axis=1))[:, None])
        img[59 - occ // 60, occ % 60] = v.astype(np.float32)
        pad = np.zeros((62, 62, 3), np.float32)
        pad[1:61, 1:61] = img
        dil = np.max([pad[a:a + 60, b:b + 60] for a in range(3) for b in range(3)], axis=0)
        rgb = dil[:, :, ::-1]
        u8 = np.clip(np.rint(rgb * np.float32(255.0)), 0, 255).astype(np.uint8)  # rint = round half to even
        differing += int(np.count_nonzero(u8 != imgs[h]))
        total += u8.size
        assert np.max(np.abs(u8.astype(int) - imgs[h].astype(int))) <= 1, h
    assert differing <= 1e-5 * total, (differing, total)


@pytest.mark.parametrize("geom", [
    {},  # the launch files' hand (finger 1 cm, aperture 9 cm, depth 6 cm, height 2 cm, bite 1 cm), 8 orientations
    {"num_orientations": 16, "nn_radius_hands": 0.06},  # c5's corner: more orientations, smaller neighbourhood
    {"num_orientations": 5, "finger_width": 0.015, "hand_outer_diameter": 0.12, "hand_depth": 0.08, "hand_height": 0.03,
     "init_bite": 0.015, "nn_radius_hands": 0.12},  # a larger hand
    {"num_orientations": 3, "finger_width": 0.005, "hand_outer_diameter": 0.06, "hand_depth": 0.03, "hand_height": 0.01,
     "init_bite": 0.0},  # a small hand that starts at the surface
], ids=["default", "r16_small_radius", "large_hand", "small_hand"])
def test_finger_sweep_vs_numpy_restatement(geom):
    """HandSearch::calculateHand (hand_search.cpp:319-426) with FingerHand::evaluateFingers / evaluateHand / deepenHand /
    computePointsInClosingRegion (finger_hand.cpp:17-180, 313-325), restated in numpy from the reference text for every
    (sample, orientation) of a scene: crop at the hand height, rotate, finger placements free of points below the bite,
    hands = matching finger pairs, middle hand, deepen in 5 mm steps while both of its fingers stay free, closing region.
    The oracle's sweep records (finger mask, hand mask, chosen slot, deepen steps, in-box count, emitted) must equal it."""
    from common import load_scene
    sc = scenes.make_scene("tiny", num_samples=150 if not geom else 60)
    p = make_params(capi.MODE_NONE, **geom)
    o = ob.context(p)
    load_scene(o, sc)
    o.detect()
    frames, sweeps = o.frames(), o.sweep_records()
    xyz = sc["xyz"]
    fw, od, hd, hh, bite0 = p.finger_width, p.hand_outer_diameter, p.hand_depth, p.hand_height, p.init_bite
    fs = ob.finger_spacing(fw, od)
    angles = ob.orientation_angles(p.num_orientations)
    depths = ob.deepen_depths(bite0, hd)  # init_bite + 0.005, ... while <= hand_depth (accumulated in double)

    def fingers_at(pts, bite, only=None):
        out = np.zeros(20, bool)
        crop = pts[pts[:, 1] < bite]
        if len(crop) == 0 or np.any(crop[:, 1] < bite - hd):  # nothing between the fingers / the hand base would collide
            return out
        for i in (range(20) if only is None else (only, 10 + only)):
            out[i] = not np.any((crop[:, 0] > fs[i]) & (crop[:, 0] < fs[i] + fw))
        return out

    checked = emitted = 0
    for s, fr in enumerate(frames):
        if not fr["valid"]:
            continue
        q = xyz[sc["samples"][s]]
        ids, _ = o.radius_search(q, p.nn_radius_hands)
        pts = (xyz[ids] - q).astype(np.float64)  # centred in float, then double (hand_search.cpp:209)
        frame = np.stack([fr["normal"], fr["binormal"], fr["curvature_axis"]], axis=1)
        z = pts @ frame[:, 2]
        cropped = pts[(z > -hh) & (z < hh)]
        for r, ang in enumerate(angles):
            rot = np.array([[np.cos(ang), -np.sin(ang), 0.0], [np.sin(ang), np.cos(ang), 0.0], [0.0, 0.0, 1.0]])
            prot = cropped @ (frame @ rot)  # points_rot = frame_rot^T * points, one row per point
            rec = sweeps[s, r]
            f = fingers_at(prot, bite0)
            assert int(rec["fingers"]) == int(sum(1 << i for i in range(20) if f[i])), (s, r)
            hand = f[:10] & f[10:] if f.sum() > 2 else np.zeros(10, bool)
            assert int(rec["hand"]) == int(sum(1 << i for i in range(10) if hand[i])), (s, r)
            checked += 1
            if not hand.any():
                assert rec["slot"] == -1 and rec["emitted"] == 0
                continue
            idx = np.nonzero(hand)[0]
            e = int(idx[int(np.ceil(len(idx) / 2.0)) - 1])  # the middle hand
            top, steps = bite0, 0
            for dpt in depths:
                if fingers_at(prot, dpt, only=e).sum() < 2:
                    break
                top, steps = dpt, steps + 1
            left, right = fs[e] + fw, fs[10 + e]
            inbox = (prot[:, 1] < top) & (prot[:, 0] > left) & (prot[:, 0] < right)
            assert rec["slot"] == e and rec["depth_step"] == steps, (s, r, e, steps)
            assert rec["num_in_box"] == inbox.sum(), (s, r)
            assert rec["emitted"] == (1 if inbox.any() else 0)
            emitted += int(inbox.any())
    print("finger sweep pin:", geom, "pairs", checked, "emitted", emitted)
    assert checked > (900 if not geom else 100) and emitted > (60 if not geom else 10)


@pytest.mark.parametrize("geom", [
    {},
    {"num_orientations": 16, "nn_radius_hands": 0.06},
    {"num_orientations": 5, "finger_width": 0.015, "hand_outer_diameter": 0.12, "hand_depth": 0.08, "hand_height": 0.03,
     "init_bite": 0.015, "nn_radius_hands": 0.12},
], ids=["default", "r16_small_radius", "large_hand"])
def test_grasp_poses_vs_numpy_restatement(geom):
    """FingerHand::calculateGraspParameters (finger_hand.cpp:183-214) and the grasp width (hand_search.cpp:397) from the
    oracle's own frames, sweep records and in-box sets: surface / bottom / top = frame_rot * (center, y, 0) + sample with
    center = (left + right) / 2, y = min y of the rotated points / bite - hand_depth / bite; axes = columns of frame_rot."""
    from common import load_scene
    sc = scenes.make_scene("tiny", num_samples=150)
    p = make_params(capi.MODE_NONE, **geom)
    o = ob.context(p)
    load_scene(o, sc)
    o.detect()
    frames, hands = o.frames(), o.hypotheses(capi.LIST_HYPOTHESES)
    xyz = sc["xyz"]
    fs = ob.finger_spacing(p.finger_width, p.hand_outer_diameter)
    angles = ob.orientation_angles(p.num_orientations)
    depths = np.concatenate([[p.init_bite], ob.deepen_depths(p.init_bite, p.hand_depth)])
    assert len(hands) > (60 if not geom else 20)
    for h, hy in enumerate(hands):
        fr = frames[hy["sample_id"]]
        ang = angles[hy["orientation_id"]]
        frame = np.stack([fr["normal"], fr["binormal"], fr["curvature_axis"]], axis=1)
        rot = np.array([[np.cos(ang), -np.sin(ang), 0.0], [np.sin(ang), np.cos(ang), 0.0], [0.0, 0.0, 1.0]])
        frame_rot = frame @ rot
        for k, name in enumerate(("binormal", "approach", "axis")):
            assert np.allclose(hy[name], frame_rot[:, k], rtol=0, atol=1e-12), (h, name)
        q = xyz[sc["samples"][hy["sample_id"]]]
        ids, _ = o.radius_search(q, p.nn_radius_hands)
        pts = (xyz[ids] - q).astype(np.float64)
        z = pts @ frame[:, 2]
        prot = pts[(z > -p.hand_height) & (z < p.hand_height)] @ frame_rot
        e = int(hy["finger_slot"])
        left, right = fs[e] + p.finger_width, fs[10 + e]
        top = depths[hy["depth_step"]]
        assert abs(hy["hand_top"] - top) < 1e-15 and abs(hy["hand_bottom"] - (top - p.hand_depth)) < 1e-15
        center, surface = 0.5 * (left + right), prot[:, 1].min()
        sample = fr["sample"]
        for name, y in (("surface", surface), ("bottom", top - p.hand_depth), ("top", top)):
            assert np.allclose(hy[name], frame_rot @ np.array([center, y, 0.0]) + sample, rtol=0, atol=1e-12), (h, name)
        inbox = prot[(prot[:, 1] < top) & (prot[:, 0] > left) & (prot[:, 0] < right)]
        assert abs(hy["width"] - (inbox[:, 0].max() - inbox[:, 0].min())) < 1e-12


def test_normals_vs_numpy_restatement_of_pcl():
    """pcl::NormalEstimation::computePointNormal as the path uses it (hand_search.cpp:83-94, r = 0.01): neighbours in
    radiusSearch order, computeMeanAndCovarianceMatrix's single pass over nine FLOAT sums without centring, covariance =
    second moments - centroid products, normal = eigenvector of the smallest eigenvalue, flipped towards the viewpoint
    (0, 0, 0); fewer than three neighbours -> NaN.  The moments are restated in float32 in that order.  Fed to the
    float32 numpy restatement of PCL's eigen33 (`_pcl_eigen33`) they must give the oracle's normal BIT FOR BIT -- the order of the
    sums is part of the semantics; numpy's double-precision eigenvector of the same matrix agrees to 3e-3 rad where the
    two smallest eigenvalues are well separated (the float closed form is that far off on thin neighbourhoods)."""
    sc = scenes.make_scene("tiny")
    o = ob.context(make_params(capi.MODE_NONE))
    o.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"])
    o.compute_normals()
    got = o.normals()
    xyz = sc["xyz"]
    rng = np.random.default_rng(9)
    checked = 0
    for i in rng.choice(len(xyz), 400, replace=False):
        ids, d2 = o.radius_search(xyz[i], 0.01)
        if len(ids) < 3:
            assert np.all(np.isnan(got[i]))
            continue
        order = np.lexsort((ids, d2))  # ascending (d2, index), FLANN's order
        pts = xyz[ids[order]]
        acc = np.zeros(9, np.float32)
        for x, y, z in pts:  # sequential float32 sums, as PCL's loop
            acc += np.array([x * x, x * y, x * z, y * y, y * z, z * z, x, y, z], np.float32)
        acc = acc * (np.float32(1.0) / np.float32(len(pts)))  # Eigen 3.2's operator/=(scalar) multiplies by the reciprocal
        c = acc[6:9]
        cov = np.array([[acc[0] - c[0] * c[0], acc[1] - c[0] * c[1], acc[2] - c[0] * c[2]],
                        [acc[1] - c[0] * c[1], acc[3] - c[1] * c[1], acc[4] - c[1] * c[2]],
                        [acc[2] - c[0] * c[2], acc[4] - c[1] * c[2], acc[5] - c[2] * c[2]]], np.float32)
        _, ev = _pcl_eigen33(cov, False)  # the float32 numpy restatement further down, D2 trigonometry
        flip = np.float32(-1.0) if np.dot(-xyz[i], ev) < 0 else np.float32(1.0)  # flipNormalTowardsViewpoint, viewpoint (0, 0, 0)
        mine = (ev * flip).astype(np.float32)
        assert np.array_equal(np.abs(mine).view(np.uint32), np.abs(got[i]).view(np.uint32)), (i, mine, got[i])
        assert np.dot(mine.astype(np.float64), got[i].astype(np.float64)) > 0, i
        checked += 1
        w, v = np.linalg.eigh(cov.astype(np.float64))
        if w[1] - w[0] > 0.2 * max(abs(w[2]), 1e-30):
            cosang = abs(np.dot(v[:, 0], got[i].astype(np.float64))) / np.linalg.norm(got[i])
            assert np.arccos(np.clip(cosang, -1, 1)) < 3e-3, (i, w)
    assert checked > 150


def _splitmix64(x: int) -> int:
    """Published splitmix64 finaliser (Steele, Lea, Flood 2014), in Python integers."""
    m = (1 << 64) - 1
    x = (x + 0x9E3779B97F4A7C15) & m
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & m
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & m
    return x ^ (x >> 31)


@pytest.mark.parametrize("method", [1, 2])
def test_importance_samples_vs_numpy_restatement(tiny_scene, method):
    """ImportanceSampling's batch of one round (importance_sampling.cpp:47-86 with drawSamplesFromSumOfGaussians :115-125
    and drawSamplesFromMaxOfGaussians :128-157), restated in Python from the reference's text with deviation D5's
    counter-based stream in place of rand() / boost::mt19937 (one 64-bit hash per (seed, round,
    candidate, slot); slot 6 picks the hand, slots 2a / 2a+1 are the two uniforms of component a, Box-Muller; the random
    cloud points use candidate 0xFFFFFFFF + k, slot 7).  Semantics pinned: num_rand = int(prob * num_samples) (:47), the
    Gaussians first and the cloud points after them (:79), sigma used as the standard deviation of the draw AND,
    unsquared, in the exponent -1/(2 sigma) (:146, :152), `term` (:53), acceptance `p >= maxp` (:153), a rejected
    candidate consumes its draws.  libm's log / cos stand in for the oracle's reproducible ones: 1e-12 m."""
    import math

    m64 = (1 << 64) - 1

    def is_hash(seed, rnd, cand, j):
        k = _splitmix64(seed ^ 0x1234ABCD5678EF01)
        k = _splitmix64((k + rnd) & m64)
        k = _splitmix64((k + cand) & m64)
        return _splitmix64((k + j) & m64)

    def u01(h):
        return ((h >> 11) + 0.5) / 9007199254740992.0

    o = ob.context(make_params(capi.MODE_NONE))
    o.set_cloud(tiny_scene["xyz"], tiny_scene["cam_mask"], tiny_scene["num_cams"])
    xyz = tiny_scene["xyz"]
    rng = np.random.default_rng(21 + method)
    H = 37
    hands = np.zeros(H, capi.HYPOTHESIS_DTYPE)
    centres = xyz[rng.choice(len(xyz), 6, replace=False)].astype(np.float64)
    hands["surface"] = centres[rng.integers(0, 6, H)] + rng.normal(0, 0.01, (H, 3))  # overlapping Gaussians: rejections happen
    ip = o.default_is_params()
    ip.method, ip.num_samples, ip.seed, ip.radius, ip.prob_rand_samples = method, 90, 0xC0FFEE + method, 0.02, 0.3
    surf = hands["surface"]
    rejected = 0
    for rnd in (0, 3):
        got = o.draw_importance_samples(ip, rnd, hands)
        ns = int(ip.num_samples)
        num_rand = int(ip.prob_rand_samples * ns)
        num_gauss = ns - num_rand
        sigma = float(ip.radius)
        term = 1.0 / math.sqrt(math.pow(2.0 * math.pi, 3.0) * math.pow(sigma, 3.0))
        want = np.zeros((ns, 3))
        cand = j = 0
        while j < num_gauss:
            idx = is_hash(ip.seed, rnd, cand, 6) % H
            x = np.array([surf[idx][a] + math.sqrt(-2.0 * math.log(u01(is_hash(ip.seed, rnd, cand, 2 * a)))) *
                          math.cos(2.0 * math.pi * u01(is_hash(ip.seed, rnd, cand, 2 * a + 1))) * sigma for a in range(3)])
            cand += 1
            take = True
            if method == 2:
                maxp = 0.0
                for k in range(H):
                    d = x - surf[k]
                    maxp = max(maxp, term * math.exp((-1.0 / (2.0 * sigma)) * float(d @ d)))
                d = x - surf[idx]
                take = term * math.exp((-1.0 / (2.0 * sigma)) * float(d @ d)) >= maxp
            if take:
                want[j] = x
                j += 1
            else:
                rejected += 1
        for jj in range(num_gauss, ns):
            r = is_hash(ip.seed, rnd, 0xFFFFFFFF + (jj - num_gauss), 7) % len(xyz)
            want[jj] = xyz[r].astype(np.float64)
        assert np.abs(got - want).max() < 1e-12, (rnd, np.abs(got - want).max())
        assert np.array_equal(got[num_gauss:], want[num_gauss:])
    assert (rejected > 10) == (method == 2)


_LIBM = None


def _libm():
    """glibc's float trigonometry, what PCL's computeRoots calls on this platform."""
    global _LIBM
    if _LIBM is None:
        import ctypes

        _LIBM = ctypes.CDLL("libm.so.6")
        for f, k in (("atan2f", 2), ("cosf", 1), ("sinf", 1)):
            getattr(_LIBM, f).restype = ctypes.c_float
            getattr(_LIBM, f).argtypes = [ctypes.c_float] * k
    return _LIBM


# PCL 1.7 common/eigen.hpp computeRoots2 / computeRoots / eigen33 (smallest eigenpair of a symmetric float 3x3) in numpy
# float32 scalars, one rounding per C operation, left to right.  float_trig = True calls libm's atan2f / cosf / sinf as
# PCL does; False evaluates them in double and rounds (deviation D2).
def _pcl_roots2(b, c):
    d = np.float32(float(b) * float(b) - 4.0 * float(c))
    if d < 0: d = np.float32(0)
    sd = np.sqrt(d)
    return [np.float32(0), np.float32(0.5) * (b - sd), np.float32(0.5) * (b + sd)]
def _pcl_roots3(m, float_trig):
    m00, m01, m02, m11, m12, m22 = m[0,0], m[0,1], m[0,2], m[1,1], m[1,2], m[2,2]
    c0 = m00*m11*m22 + np.float32(2)*m01*m02*m12 - m00*m12*m12 - m11*m02*m02 - m22*m01*m01
    c1 = m00*m11 - m01*m01 + m00*m22 - m02*m02 + m11*m22 - m12*m12
    c2 = m00 + m11 + m22
    if abs(c0) < np.finfo(np.float32).eps: return _pcl_roots2(c2, c1)
    inv3 = np.float32(1.0/3.0); sqrt3 = np.sqrt(np.float32(3))
    c2o3 = c2 * inv3
    ao3 = (c1 - c2 * c2o3) * inv3
    if ao3 > 0: ao3 = np.float32(0)
    hb = np.float32(0.5) * (c0 + c2o3 * (np.float32(2) * c2o3 * c2o3 - c1))
    q = hb*hb + ao3*ao3*ao3
    if q > 0: q = np.float32(0)
    rho = np.sqrt(-ao3)
    if float_trig:
        th = np.float32(_libm().atan2f(float(np.sqrt(-q)), float(hb))) * inv3
        ct, st = np.float32(_libm().cosf(float(th))), np.float32(_libm().sinf(float(th)))
    else:
        th = np.float32(np.arctan2(float(np.sqrt(-q)), float(hb))) * inv3
        ct, st = np.float32(np.cos(float(th))), np.float32(np.sin(float(th)))
    r = [c2o3 + np.float32(2)*rho*ct, c2o3 - rho*(ct + sqrt3*st), c2o3 - rho*(ct - sqrt3*st)]
    if r[0] >= r[1]: r[0], r[1] = r[1], r[0]
    if r[1] >= r[2]:
        r[1], r[2] = r[2], r[1]
        if r[0] >= r[1]: r[0], r[1] = r[1], r[0]
    if r[0] <= 0: return _pcl_roots2(c2, c1)
    return r
def _pcl_eigen33(cov, float_trig):
    scale = np.abs(cov).max()
    if scale <= np.finfo(np.float32).tiny: scale = np.float32(1)
    m = (cov / scale).astype(np.float32)
    r = _pcl_roots3(m, float_trig)
    m = m.copy(); m[0,0] -= r[0]; m[1,1] -= r[0]; m[2,2] -= r[0]
    def cross(a, b): return np.array([a[1]*b[2]-a[2]*b[1], a[2]*b[0]-a[0]*b[2], a[0]*b[1]-a[1]*b[0]], np.float32)
    vs = [cross(m[0], m[1]), cross(m[0], m[2]), cross(m[1], m[2])]
    ls = [v[0]*v[0] + (v[1]*v[1] + v[2]*v[2]) for v in vs]
    l1, l2, l3 = ls
    k = 0 if (l1 >= l2 and l1 >= l3) else (1 if (l2 >= l1 and l2 >= l3) else 2)
    return r[0] * scale, (vs[k] / np.sqrt(ls[k])).astype(np.float32)


def test_pcl_eigen33_numpy_float32_restatement_and_size_of_deviation_d2(tiny_scene):
    """(1) The float32 restatement above, with D2's trigonometry, must reproduce the oracle's eigen33 BIT FOR BIT on the
    covariances of 1 500 real neighbourhoods.  (2) The same restatement with this platform's atan2f / cosf / sinf measures
    what D2 changes against PCL linked with this libm (glibc 2.39): 50 of the 1 500 normals differ, all in the last
    bits, the worst by 1.2e-7 rad -- three orders of magnitude below the 1e-4 rad the normals are held to.  Bounded here
    at 1e-6 rad and quoted in DESIGN.md section 2."""
    rng = np.random.default_rng(5)
    o = ob.context(make_params(capi.MODE_NONE))
    o.set_cloud(tiny_scene["xyz"], tiny_scene["cam_mask"], tiny_scene["num_cams"])
    xyz = tiny_scene["xyz"]
    n = differ = 0
    worst = 0.0
    for i in rng.choice(len(xyz), 1500, replace=False):
        ids, d2 = o.radius_search(xyz[i], 0.01)
        if len(ids) < 3:
            continue
        cov = np.cov(xyz[ids].astype(np.float64).T, bias=True).astype(np.float32)
        lam_o, ev_o = ob.eigen33(cov)
        lam, ev = _pcl_eigen33(cov, False)
        assert np.array_equal(ev.view(np.uint32), np.asarray(ev_o, np.float32).view(np.uint32)), (i, ev, ev_o)
        assert np.float32(lam) == np.float32(lam_o), (i, lam, lam_o)
        _, ev_f = _pcl_eigen33(cov, True)
        n += 1
        if not np.array_equal(ev, ev_f):
            differ += 1
            worst = max(worst, float(np.linalg.norm(np.cross(ev.astype(np.float64), ev_f.astype(np.float64)))))
    assert n > 1000
    assert worst < 1e-6, (differ, n, worst)
    print(f"D2: {differ} of {n} normals differ from libm float trigonometry, worst {worst:.2e} rad")


def test_local_frames_vs_numpy_restatement(tiny_scene):
    """HandSearch::calculateLocalFrames (hand_search.cpp:97-170) + LocalFrame (local_frame.cpp:4-15, 26-59) restated in
    numpy: radiusSearch(sample, nn_radius_taubin); no neighbour -> no frame (:122); min(50, n) normals drawn WITH
    replacement from the neighbour list (:126-137; rand() % n replaced by deviation D1's counter-based draw
    splitmix64(seed ^ splitmix64(sample_id << 8 ^ j)) % n); the camera counts of the DRAWN points, first maximum wins
    (:139-147, Eigen's maxCoeff); columns normalised (:149-150); curvature axis = eigenvector of the smallest eigenvalue
    of N N^T (:29-38, numpy.linalg.eigh here); normal = the drawn normal with the largest column sum of (N^T N)^6
    projected off the axis (:41-44); binormal = axis x normal; both turned against (sample - camera origin of the majority
    camera) (:50-54); axis = normal x binormal (:57).  Counts and camera exact, axes to 1e-9."""
    p = make_params(capi.MODE_NONE)
    o = ob.context(p)
    o.set_cloud(tiny_scene["xyz"], tiny_scene["cam_mask"], tiny_scene["num_cams"])
    samples = np.asarray(tiny_scene["samples"], np.int64)
    o.set_sample_indices(samples)
    o.compute_normals()
    o.generate_hypotheses()
    fr = o.frames()
    nrm = o.normals().astype(np.float64)
    xyz, cam = tiny_scene["xyz"], tiny_scene["cam_mask"]
    tfs = [np.array(p.cam_tf_left).reshape(4, 4), np.array(p.cam_tf_right).reshape(4, 4)]
    m64 = (1 << 64) - 1
    checked = 0
    for s, i in enumerate(samples):
        ids, d2 = o.radius_search(xyz[i], p.nn_radius_taubin)
        assert fr[s]["num_neighbors"] == len(ids)
        if len(ids) == 0:
            assert fr[s]["valid"] == 0
            continue
        ids = ids[np.lexsort((ids, d2))]
        m = min(50, len(ids))
        drawn = [ids[_splitmix64(int(p.rng_seed) ^ _splitmix64(((s << 8) ^ j) & m64)) % len(ids)] for j in range(m)]
        N = nrm[drawn].T  # 3 x m
        if not np.all(np.isfinite(N)):
            continue
        counts = [sum((cam[d] >> c) & 1 for d in drawn) for c in range(tiny_scene["num_cams"])]
        majority = int(np.argmax(counts))
        N = N / np.sqrt((N * N).sum(axis=0))
        w, v = np.linalg.eigh(N @ N.T)
        if min(w[1] - w[0], w[2] - w[1]) < 1e-6 * w[2]:
            continue
        axis = v[:, 0]
        k = int(np.argmax(((N.T @ N) ** 6).sum(axis=0)))
        npart = (np.eye(3) - np.outer(axis, axis)) @ N[:, k]
        normal = npart / np.linalg.norm(npart)
        binormal = np.cross(axis, normal)
        src = xyz[i].astype(np.float64) - tfs[majority][:3, 3]
        if normal @ src > 0:
            normal = -normal
        if binormal @ src > 0:
            binormal = -binormal
        axis = np.cross(normal, binormal)
        assert fr[s]["valid"] == 1 and fr[s]["majority_cam"] == majority, s
        assert np.array_equal(fr[s]["sample"], xyz[i].astype(np.float64))
        for name, want in (("normal", normal), ("binormal", binormal), ("curvature_axis", axis)):
            assert np.abs(fr[s][name] - want).max() < 1e-9, (s, name, fr[s][name], want)
        checked += 1
    assert checked >= 0.8 * len(samples)


@pytest.mark.parametrize("num_selected,min_score_diff", [(7, -1e30), (50, 0.0), (0, -10.0), (3, 5.0)])
def test_score_and_select_vs_numpy_restatement(tiny_scene, num_selected, min_score_diff):
    """GraspDetector::detectGraspPoses, PREDICTION branch and selection (grasp_detector.cpp:186-221, 238-252), in numpy:
    score = logits[1] - logits[0] as a FLOAT difference widened to double (:203); a hand is kept iff score >=
    min_score_diff (:205), marked full antipodal with that score (:208-209), in list order; then the num_selected best
    by score, best first (isScoreGreater), all of them when there are no more than num_selected (:241-252).  Equal
    scores keep list order (deviation D4: std::partial_sort leaves it unspecified)."""
    p = make_params(capi.MODE_PREDICTION, num_selected=num_selected, min_score_diff=min_score_diff)
    o = ob.context(p)
    o.set_lenet_weights(ob.xavier_weights(5))
    o.set_cloud(tiny_scene["xyz"], tiny_scene["cam_mask"], tiny_scene["num_cams"])
    o.set_sample_indices(tiny_scene["samples"])
    o.detect()
    pruned = o.hypotheses(capi.LIST_PRUNED)
    logits = o.logits()
    assert logits.dtype == np.float32 and logits.shape == (len(pruned), 2) and len(pruned) > 15
    score = (logits[:, 1] - logits[:, 0]).astype(np.float64)
    keep = np.flatnonzero(score >= min_score_diff)
    anti = o.hypotheses(capi.LIST_ANTIPODAL)
    assert len(anti) == len(keep) and (0 < len(keep) < len(pruned) or min_score_diff < -1e29)
    for name in ("sample_id", "orientation_id"):
        assert np.array_equal(anti[name], pruned[name][keep])
    assert np.array_equal(anti["score"], score[keep]) and np.all(anti["full_antipodal"] == 1)
    order = keep[np.argsort(-score[keep], kind="stable")]
    if 0 < num_selected < len(order):
        order = order[:num_selected]
    sel = o.hypotheses(capi.LIST_SELECTED)
    assert len(sel) == len(order)
    for name in ("sample_id", "orientation_id"):
        assert np.array_equal(sel[name], pruned[name][order]), name
    assert np.array_equal(sel["score"], score[order])
    assert np.array_equal(sel["surface"], pruned["surface"][order])
