"""The header-only C++ shims (include/agile_grasp2/*.h) driven by tests/cpp/test_shim.cpp — the reference's
detect_grasps_file / test_cnn flows — and its result compared with the oracle run through the same steps."""
import os
import struct
import subprocess

import numpy as np
import pytest

from agile_grasp2_b200 import capi
from common import hyp_int_fields_equal, make_params
from oracle import binding as ob

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "test_shim")


def test_shim_headers_compile_standalone(tmp_path):
    """Every shim header is self-contained (CPU: compile only)."""
    for h in sorted(os.listdir(os.path.join(ROOT, "include", "agile_grasp2"))):
        src = tmp_path / (h + ".cpp")
        src.write_text('#include <agile_grasp2/%s>\nint main() { return 0; }\n' % h)
        r = subprocess.run(["/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++", "-std=c++17", "-fsyntax-only",
                            "-I" + os.path.join(ROOT, "include"), str(src)], capture_output=True, text=True)
        assert r.returncode == 0, (h, r.stderr[-2000:])


def test_shims_compile_against_the_eigen_interface():
    """types.h aliases its value types to Eigen's where <Eigen/Dense> is found.  Eigen is not in this image: a mock with
    Eigen's signatures and defaults (tests/cpp/mock_eigen) stands in for a compile-only check of that branch."""
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    for src in ("test_shim.cpp", "test_group.cpp", "test_selection.cpp"):
        r = subprocess.run([gxx, "-std=c++17", "-Wall", "-fsyntax-only", "-I" + os.path.join(ROOT, "include"),
                            "-I" + os.path.join(ROOT, "tests", "cpp", "mock_eigen"), os.path.join(ROOT, "tests", "cpp", src)],
                           capture_output=True, text=True)
        assert r.returncode == 0, (src, r.stderr[-2000:])
    probe = subprocess.run([gxx, "-std=c++17", "-E", "-dM", "-I" + os.path.join(ROOT, "include"),
                            "-I" + os.path.join(ROOT, "tests", "cpp", "mock_eigen"), "-x", "c++",
                            os.path.join(ROOT, "include", "agile_grasp2", "types.h")], capture_output=True, text=True)
    assert "AG2SHIM_HAVE_EIGEN" in probe.stdout


def _splitmix64(x):
    m = (1 << 64) - 1
    x = (x + 0x9E3779B97F4A7C15) & m
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & m
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & m
    return x ^ (x >> 31)


def _selection_draw(seed, what, k):
    m = (1 << 64) - 1
    return _splitmix64((_splitmix64(seed ^ ((0xA5A5F00DBAADC0DE + what) & m)) + k) & m)


@pytest.mark.parametrize("n,k,seed,cloud", [(10, 4, 5, 100), (40, 40, 1, 100), (40, 60, 1, 100), (200, 17, 99, 5000),
                                            (30, 12, 7, 10), (1, 1, 0, 1), (0, 5, 3, 10)])
def test_sample_selection_of_the_shims_vs_numpy_restatement(tmp_path, n, k, seed, cloud):
    """include/agile_grasp2/sample_selection.h + CloudCamera::subsampleSamples on the CPU (tests/cpp/test_selection.cpp,
    no ctx): CloudCamera::subsampleSamples (cloud_camera.cpp:181-206: all incoming samples when there are no more than
    num_samples, else the first num_samples of a shuffled sequence), the strict workspace filter on sample points
    (grasp_detector.cpp:305-313) and the draw with replacement from caller-supplied indices (grasp_detector.cpp:334-347),
    against the same rules written in Python with the documented counter-based stream."""
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    exe = tmp_path / "test_selection"
    r = subprocess.run([gxx, "-O1", "-std=c++17", "-Wall", "-I" + os.path.join(ROOT, "include"), "-o", str(exe),
                        os.path.join(ROOT, "tests", "cpp", "test_selection.cpp")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    nfile = tmp_path / "normals.csv"
    hfile = tmp_path / "hands.csv"
    out = subprocess.run([str(exe), str(n), str(k), str(seed), str(cloud), str(nfile), str(hfile)], capture_output=True, text=True,
                         check=True).stdout
    lines = {ln.split()[0]: ln.split()[1:] for ln in out.strip().splitlines()}

    def positions(count):
        seq = list(range(count))
        if k >= count:
            return seq
        for i in range(count - 1, 0, -1):
            j = _selection_draw(seed, 1, i) % (i + 1)
            seq[i], seq[j] = seq[j], seq[i]
        return seq[:k]

    assert [int(v) for v in lines["positions"]] == positions(n)
    if n > k:
        assert len(set(lines["positions"])) == k  # a subset without repetition
    idx = [7 * i + 3 for i in range(n)]
    want = [idx[_selection_draw(seed, 2, i) % n] for i in range(k)] if (n and k != n and k < cloud) else idx
    assert [int(v) for v in lines["indices"]] == want
    pts = np.array([[0.25 * (i % 7) - 0.25, 0.5 * (i % 6) - 0.5, -0.25 * (i % 5)] for i in range(n)]).reshape(-1, 3)
    inside = pts[(pts[:, 0] > 0) & (pts[:, 0] < 1) & (pts[:, 1] > 0) & (pts[:, 1] < 2) & (pts[:, 2] > -1) & (pts[:, 2] < 0)]
    assert int(lines["inside"][0]) == len(inside)
    got = np.array([float(v) for v in lines["samples"]]).reshape(-1, 3)
    assert np.array_equal(got, inside[positions(len(inside))].reshape(-1, 3))
    # CloudCamera::setSamples(msg) (cloud_camera.cpp:243-251) and writeNormalsToFile (:209-220)
    assert int(lines["all"][0]) == len(inside)
    text = nfile.read_text()
    rows = np.array([[float(v) for v in ln.split(",")] for ln in text.splitlines()]).reshape(-1, 3)
    assert np.array_equal(rows, inside)
    if len(inside):
        assert text.splitlines()[0] == ",".join("%.17g" % v for v in inside[0])
    # GraspHypothesis::writeHandsToFile (grasp_hypothesis.cpp:4-23) and convertToGraspMsg (:37-50)
    hl = hfile.read_text().splitlines()
    assert len(hl) == 3
    for h, ln in enumerate(hl):
        bottom = [0.1 * (h + 1) + a / 3.0 for a in range(3)]
        surface = [-0.2 * (h + 1) + a / 7.0 for a in range(3)]
        want = "".join("%.17g," % v for v in bottom + surface + [1, 0, 0, 0, 1, 0, 0, 0, 1]) + "%g" % (0.05 + h / 300.0)
        assert ln == want
    m = [float(v) for v in lines["msg"]]
    assert m[0] == 0.1 * 2 and m[1] == -0.2 * 2 + 2 / 7.0
    assert np.float32(m[2]) == np.float32(0.05 + 1 / 300.0) and np.float32(m[3]) == np.float32(1.0 / 3.0 + 1)
    assert m[2] != 0.05 + 1 / 300.0  # narrowed, not the double


@pytest.mark.gpu
def test_cpp_shim_flow_matches_oracle(tiny_scene, tmp_path):
    assert os.path.exists(BIN), "build tests/cpp first (__graft_entry__.build())"
    pcd, out = str(tmp_path / "cloud.pcd"), str(tmp_path / "out.bin")
    capi.pcd_write(pcd, tiny_scene["xyz"], binary=True)
    r = subprocess.run([BIN, pcd, out], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    raw = open(out, "rb").read()
    n_proc, n_smp = struct.unpack_from("<QQ", raw, 0)
    samples = np.frombuffer(raw, np.int32, n_smp, 16)
    (n_h,) = struct.unpack_from("<Q", raw, 16 + 4 * n_smp)
    hands = np.frombuffer(raw, capi.HYPOTHESIS_DTYPE, n_h, 24 + 4 * n_smp)
    # the same steps on the oracle
    p = make_params(capi.MODE_PREDICTION, min_score_diff=-1e9, num_selected=20, num_samples=64, num_threads=4, rng_seed=5)
    o = ob.context(p)
    with capi.Context(p) as gw:  # the shim's "no trained_file" weights: ag2_init_lenet_xavier(seed 1)
        gw.init_lenet_xavier(1)
        o.set_lenet_weights(gw.get_lenet_weights())
    o.preprocess_cloud(tiny_scene["xyz"], None, 1, voxelize=True, voxel_size=0.003)
    assert len(o.cloud()[0]) == n_proc
    assert np.array_equal(o.subsample_uniformly(64, seed=5), samples)
    o.detect()
    oh = o.hypotheses(capi.LIST_SELECTED)
    assert len(oh) == n_h
    # ranking by score: equal up to score ties within the logits tolerance
    tol = 1e-3 + 2e-5 * float(np.abs(o.logits()).max())
    assert np.all(np.abs(np.sort(hands["score"]) - np.sort(oh["score"])) <= tol)
    key = lambda h: sorted(zip(h["sample_id"].tolist(), h["orientation_id"].tolist()))
    assert key(hands) == key(oh)
    order_g = np.lexsort((hands["orientation_id"], hands["sample_id"]))
    order_o = np.lexsort((oh["orientation_id"], oh["sample_id"]))
    assert hyp_int_fields_equal(hands[order_g], oh[order_o]) is None
    assert np.allclose(hands[order_g]["bottom"], oh[order_o]["bottom"], rtol=0, atol=1e-9)
    # ClassificationNode flow: the same densities through a numpy restatement of the point generator
    # (classification_node.cpp:77-133), the oracle's image and LeNet
    pos = 24 + 4 * n_smp + 192 * n_h
    (n_pairs,) = struct.unpack_from("<Q", raw, pos)
    pair_scores = np.frombuffer(raw, np.float64, n_pairs, pos + 8)
    two = hands[:2]
    dens = np.array([0.0 if k % 5 == 3 else np.float32(0.35) + np.float32(0.11) * np.float32(k % 7)
                     for k in range(2 * len(two) * 6)], np.float32)
    want = []
    for i in range(2):
        sc = []
        for j, h in enumerate(two):
            pts, nrm = _fill_unit_box(h, dens, i * len(two) * 6 + j * 6)
            img = o.render_point_sets(pts, nrm, [0, len(pts)], aligns=False)
            lg = o.classify_images(img)[0]
            sc.append(np.float32(lg[1] - lg[0]))
        want.append(float(np.float32(sum(sc, np.float32(0.0)))) / len(two))
    assert np.allclose(pair_scores, want, rtol=2e-5, atol=2e-3), (pair_scores, want)


def _fill_unit_box(h, dens, start):
    min_sp, length = np.float32([0.04, 0.02]), np.float32([0.9, 0.9, 0.9, 0.9, 0.5, 0.5])
    x_start, y_start, var = np.float32([0.25, 0.25, 0.75, 0.75, 0.25, 0.25]), np.float32(0.1), [1, 1, 1, 1, 0, 0]
    frame = np.stack([h["binormal"], h["approach"], h["axis"]], axis=1)  # columns
    faces = np.stack([h["binormal"], -1.0 * h["binormal"], -1.0 * h["approach"]], axis=1)
    ncol = np.zeros((3, 3))
    for k in range(3):
        for r in range(3):
            s = 0.0
            for a in range(3):
                s += frame[a, r] * faces[a, k]
            ncol[r, k] = s
    pts, nrm = [], []
    for i in range(6):
        d = dens[start + i]
        if not d > 0:
            continue
        spacing = float(np.float32(min_sp[i // 4] / d))
        n = int(np.floor(float(length[i]) / spacing)) + 1
        p = [float(x_start[i]), float(y_start), 0.0]
        for _ in range(n):
            pts.append(list(p)); nrm.append(ncol[:, i // 2].copy())
            p[var[i]] += spacing
    return np.array(pts).reshape(-1, 3), np.array(nrm).reshape(-1, 3)


@pytest.mark.gpu
def test_detect_grasps_file_example_equals_the_c_abi_steps(tiny_scene, tmp_path):
    """examples/detect_grasps_file (the headless stand-in for src/nodes/detect_grasps_file.cpp +
    grasp_detection_node.cpp:98-120) on a .pcd of the tiny scene: its grasp file holds what the same steps give
    through the C ABI (preprocess, subsample, detect), and its GraspListMsg bytes have the wire size."""
    exe = os.path.join(ROOT, "examples", "detect_grasps_file")
    assert os.path.exists(exe), "build examples first (__graft_entry__.build())"
    pcd, hands_file, msg_file = str(tmp_path / "scene.pcd"), str(tmp_path / "hands.csv"), str(tmp_path / "grasps.bin")
    capi.pcd_write(pcd, tiny_scene["xyz"], binary=True)
    r = subprocess.run([exe, "cloud_file_name:=" + pcd, "num_samples:=64", "antipodal_mode:=1", "min_inliers:=0",
                        "min_score_diff:=-1e9", "rng_seed:=5", "_num_selected:=30", "hands_file:=" + hands_file, "msg_file:=" + msg_file],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    p = make_params(capi.MODE_PREDICTION, num_selected=30, num_samples=64, rng_seed=5, min_inliers=0, num_threads=4,
                    filter_half_grasps=0, min_aperture=0.03, max_aperture=0.08, min_score_diff=-1e9)
    with capi.Context(p) as g:
        g.init_lenet_xavier(1)  # the shim's weights when trained_file is empty
        g.preprocess_cloud(tiny_scene["xyz"], None, 1, voxelize=True, voxel_size=0.003)
        g.subsample_uniformly(64, seed=5)
        g.detect()
        want = g.hypotheses(capi.LIST_SELECTED)
    assert ("# grasps: %d;" % len(want)) in r.stdout and len(want) > 0
    rows = np.array([[float(v) for v in ln.split(",")] for ln in open(hands_file).read().splitlines()]).reshape(-1, 16)
    assert len(rows) == len(want)
    for k, name in enumerate(("bottom", "surface", "axis", "approach", "binormal")):
        assert np.array_equal(rows[:, 3 * k:3 * k + 3], want[name]), name
    assert np.allclose(rows[:, 15], want["width"], rtol=1e-5)
    assert os.path.getsize(msg_file) == 16 + 4 + 4 + len(want) * 152


@pytest.mark.gpu
@pytest.mark.skipif("__import__('torch').cuda.device_count() < 2")
def test_cpp_group_two_gpus_equals_one(tiny_scene, tmp_path):
    """tests/cpp/test_group.cpp: GraspDetector / ImportanceSampling with Params::num_gpus = 2 (ag2_group_create:
    host threads inside libag2, NVLink peer memory) against the same calls on one GPU, byte for byte."""
    binp = os.path.join(ROOT, "tests", "cpp", "test_group")
    assert os.path.exists(binp), "build tests/cpp first (__graft_entry__.build())"
    pcd = str(tmp_path / "cloud.pcd")
    capi.pcd_write(pcd, tiny_scene["xyz"], binary=True)
    r = subprocess.run([binp, pcd, "120", "2"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "TEST_GROUP OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    # the two-stage input distribution of three ranks and more (slices from their holders), forced on two
    r = subprocess.run([binp, pcd, "120", "2"], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, AG2_GROUP_TWO_STAGE="1"))
    assert r.returncode == 0 and "TEST_GROUP OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
