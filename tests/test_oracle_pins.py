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
    (np.unique over lexicographic voxel ids; first point of a voxel gives the camera source)."""
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
    assert np.array_equal(ocam, fc[first])
    idx = o.subsample_uniformly(100, seed=3)
    assert len(idx) == 100 and np.all(np.diff(idx) > 0) and idx[-1] < len(oxyz)
