"""GPU parity: libag2.so (CUDA, through the C ABI) against the CPU oracle on the same seeded inputs."""
import numpy as np
import pytest

from agile_grasp2_b200 import capi, scenes
from common import angle, both, hyp_int_fields_equal, load_scene, make_params
from oracle import binding as ob

pytestmark = pytest.mark.gpu
POSE_FIELDS = ["surface", "bottom", "top", "axis", "approach", "binormal", "width", "hand_top", "hand_bottom"]


@pytest.fixture(scope="module")
def run_tiny(tiny_scene):
    p = make_params(capi.MODE_GEOMETRIC)
    g, o = both(p, tiny_scene)
    g.detect()
    o.detect()
    return g, o, tiny_scene


def test_radius_search_sets_bit_exact(run_tiny):
    g, o, sc = run_tiny
    rng = np.random.default_rng(0)
    xyz = sc["xyz"]
    for r in (0.01, 0.03, 0.1):
        for i in rng.integers(0, len(xyz), 6):
            q = xyz[i] + (rng.normal(0, 0.002, 3).astype(np.float32) if i % 2 else 0)
            gi, gd = g.radius_search(q, r)
            oi, od = o.radius_search(q, r, brute=True)
            assert np.array_equal(gi, oi)
            assert np.array_equal(gd.view(np.uint32), od.view(np.uint32))


def test_normals(run_tiny):
    g, o, _ = run_tiny
    gn, on = g.normals(), o.normals()
    assert np.array_equal(np.isnan(gn), np.isnan(on))
    ok = ~np.isnan(on).any(1)
    ang = angle(gn[ok], on[ok])
    bit_equal = np.mean(np.all(gn[ok].view(np.uint32) == on[ok].view(np.uint32), axis=1))
    print("normals: max angle %.3g rad, bit-equal fraction %.6f" % (ang.max(), bit_equal))
    assert ang.max() <= 1e-4  # tolerance of north_star
    assert bit_equal > 0.999


def test_frames(run_tiny):
    g, o, _ = run_tiny
    gf, of = g.frames(), o.frames()
    assert np.array_equal(gf["valid"], of["valid"])
    assert np.array_equal(gf["num_neighbors"], of["num_neighbors"])
    assert np.array_equal(gf["majority_cam"], of["majority_cam"])
    v = of["valid"] == 1
    for f in ("normal", "binormal", "curvature_axis"):
        assert angle(gf[f][v], of[f][v]).max() <= 1e-6
    assert np.array_equal(gf["sample"], of["sample"])


def test_sweep_records_bit_exact(run_tiny):
    g, o, _ = run_tiny
    gs, os_ = g.sweep_records(), o.sweep_records()
    for f in ("fingers", "hand", "slot", "depth_step", "num_in_box", "antipodal", "emitted"):
        bad = np.argwhere(gs[f] != os_[f])
        assert len(bad) == 0, (f, bad[:5], gs[f][tuple(bad[0])], os_[f][tuple(bad[0])])


@pytest.mark.parametrize("which", [capi.LIST_HYPOTHESES, capi.LIST_PRUNED, capi.LIST_ANTIPODAL, capi.LIST_SELECTED])
def test_hypothesis_lists(run_tiny, which):
    g, o, _ = run_tiny
    gh, oh = g.hypotheses(which), o.hypotheses(which)
    assert len(gh) == len(oh)
    assert hyp_int_fields_equal(gh, oh) is None
    for f in POSE_FIELDS:
        assert np.allclose(gh[f], oh[f], rtol=0, atol=1e-9), f
    if which == capi.LIST_HYPOTHESES:
        assert len(gh) > 10


def test_in_box_sets(run_tiny):
    g, o, _ = run_tiny
    n = g.num(capi.LIST_HYPOTHESES)
    for h in range(0, n, max(1, n // 12)):
        assert np.array_equal(g.in_box(capi.LIST_HYPOTHESES, h), o.in_box(capi.LIST_HYPOTHESES, h))


def test_stats_match(run_tiny):
    g, o, _ = run_tiny
    gs, os_ = g.stats(), o.stats()
    for k in ("num_frames", "num_hypotheses", "num_pruned", "num_antipodal", "num_selected", "sum_neighbors_hands",
              "sum_cropped", "sum_in_box", "sum_neighbors_normals"):
        assert gs[k] == os_[k], k
    assert gs["gpu_launches"] > 0


def test_prediction_mode_images_and_scores(tiny_scene):
    w = ob.xavier_weights(11, bias_scale=0.5)
    p = make_params(capi.MODE_PREDICTION, min_score_diff=0.0, num_selected=20)
    g, o = both(p, tiny_scene, weights=w)
    g.generate_hypotheses(); o.generate_hypotheses()
    g.prune(); o.prune()
    gi, oi = g.render_images(), o.render_images()
    assert gi.shape == oi.shape and gi.shape[0] > 5
    diff = gi.astype(int) - oi.astype(int)
    print("images: %d, differing bytes %d" % (gi.shape[0], np.count_nonzero(diff)))
    assert np.count_nonzero(diff) == 0
    g.classify(); o.classify()
    gl, ol = g.logits(), o.logits()
    tol = np.maximum(1e-3, 1e-5 * np.abs(ol))
    print("logits: max abs diff %.3g (scale %.3g)" % (np.abs(gl - ol).max(), np.abs(ol).max()))
    assert np.all(np.abs(gl - ol) <= tol)
    g.select(); o.select()
    gs, os_ = g.hypotheses(capi.LIST_SELECTED), o.hypotheses(capi.LIST_SELECTED)
    # accept sets identical except scores within tolerance of the threshold
    ga, oa = g.hypotheses(capi.LIST_ANTIPODAL), o.hypotheses(capi.LIST_ANTIPODAL)
    key = lambda h: set(zip(h["sample_id"].tolist(), h["orientation_id"].tolist()))
    sym = key(ga) ^ key(oa)
    near = {(s, r) for s, r, sc in zip(oh_all(o)["sample_id"], oh_all(o)["orientation_id"], scores_all(o))
            if abs(sc - p.min_score_diff) <= 1e-3}
    assert sym <= near
    assert len(gs) == len(os_)


def oh_all(o):
    return o.hypotheses(capi.LIST_PRUNED)


def scores_all(o):
    l = o.logits()
    return (l[:, 1] - l[:, 0]).astype(np.float64)


def test_classify_images_direct(tiny_scene):
    w = ob.xavier_weights(5, bias_scale=0.2)
    p = make_params(capi.MODE_PREDICTION)
    g = capi.Context(p); o = ob.context(p)
    g.set_lenet_weights(w); o.set_lenet_weights(w)
    rng = np.random.default_rng(3)
    imgs = rng.integers(0, 256, size=(37, 60, 60, 3), dtype=np.uint8)
    imgs[0] = 0; imgs[1] = 255
    gl, ol = g.classify_images(imgs), o.classify_images(imgs)
    tol = np.maximum(1e-3, 1e-5 * np.abs(ol))
    assert np.all(np.abs(gl - ol) <= tol), np.abs(gl - ol).max()


@pytest.mark.parametrize("s1,s2", [(64.0, 1.0 / 64.0), (1.0 / 512.0, 512.0), (4096.0, 1.0 / 4096.0), (3.0e-5, 1.0 / 3.0e-5)])
def test_classify_images_with_rescaled_layers(s1, s2):
    """conv2 runs on fp16 split operands behind power-of-two scales derived from the weights (lenet.cu): layers whose
    weights are orders of magnitude away from Xavier's must come out within the same tolerance.  conv1 is scaled by s1
    and conv2 by s2 (biases with their layer's output), so the activations between them move over eight decades while the logits stay put."""
    w = {k: v.copy() for k, v in ob.xavier_weights(13, bias_scale=0.2).items()}
    w["conv1_w"] *= np.float32(s1); w["conv1_b"] *= np.float32(s1)
    w["conv2_w"] *= np.float32(s2); w["conv2_b"] *= np.float32(s1 * s2)
    p = make_params(capi.MODE_PREDICTION)
    g = capi.Context(p); o = ob.context(p)
    g.set_lenet_weights(w); o.set_lenet_weights(w)
    rng = np.random.default_rng(29)
    imgs = rng.integers(0, 256, size=(33, 60, 60, 3), dtype=np.uint8)
    imgs[rng.random(imgs.shape) < 0.6] = 0
    imgs[0] = 255
    gl, ol = g.classify_images(imgs), o.classify_images(imgs)
    assert np.all(np.isfinite(gl))
    tol = np.maximum(1e-3, 1e-5 * np.abs(ol))  # s1 * s2 = 1: the logits keep the magnitude of the unscaled network
    assert np.all(np.abs(gl - ol) <= tol), np.abs(gl - ol).max()


@pytest.mark.parametrize("n", [1, 129, 4500])
def test_classify_images_batch_sizes(n):
    """Image-tile / split-K / chunk boundaries of the tensor-core LeNet: one image, one image past a
    128-image tile, and more than one 4096-image chunk.  The batch repeats 43 base images, so the
    oracle classifies 43 and every repeat must reproduce its logits bit for bit."""
    w = ob.xavier_weights(11, bias_scale=0.2)
    p = make_params(capi.MODE_PREDICTION)
    g = capi.Context(p); o = ob.context(p)
    g.set_lenet_weights(w); o.set_lenet_weights(w)
    rng = np.random.default_rng(17)
    base = rng.integers(0, 256, size=(43, 60, 60, 3), dtype=np.uint8)
    base[rng.random(base.shape) < 0.7] = 0  # sparse, like the occupancy planes of real grasp images
    idx = np.arange(n) % 43
    gl = g.classify_images(np.ascontiguousarray(base[idx]))
    ol = o.classify_images(base[: min(n, 43)])
    tol = np.maximum(1e-3, 1e-5 * np.abs(ol))
    assert np.all(np.abs(gl[: len(ol)] - ol) <= tol), np.abs(gl[: len(ol)] - ol).max()
    for r in range(43, n, 43):
        m = min(43, n - r)
        assert np.array_equal(gl[r:r + m], gl[:m]), r
    # ... and whatever the batch they arrive in (a rank's share of the samples holds other images than the whole list)
    alone = g.classify_images(base[:min(n, 7)])
    assert np.array_equal(alone.view(np.uint32), gl[:len(alone)].view(np.uint32))


def test_sample_points_and_precomputed_normals(tiny_scene):
    sc = tiny_scene
    p = make_params(capi.MODE_GEOMETRIC)
    o0 = ob.context(p)
    o0.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); o0.compute_normals()
    nrm = o0.normals()
    rng = np.random.default_rng(9)
    pts = sc["xyz"][sc["samples"][:24]].astype(np.float64) + rng.normal(0, 0.004, (24, 3))
    g, o = both(p, sc, normals=nrm, sample_points=pts)
    g.detect(); o.detect()
    gs, os_ = g.sweep_records(), o.sweep_records()
    for f in ("fingers", "hand", "slot", "depth_step", "num_in_box", "antipodal", "emitted"):
        assert np.array_equal(gs[f], os_[f]), f
    gh, oh = g.hypotheses(), o.hypotheses()
    assert hyp_int_fields_equal(gh, oh) is None
    for f in POSE_FIELDS:
        assert np.allclose(gh[f], oh[f], rtol=0, atol=1e-9), f


def test_plates_known_answer():
    xyz, nrm = scenes.plates_scene(gap=0.05)
    p = make_params(capi.MODE_GEOMETRIC, prune=0, nn_radius_taubin=0.03)
    sc = dict(xyz=xyz, cam_mask=np.ones(len(xyz), np.uint8), num_cams=1, samples=np.array([0], np.int32))
    pts = np.array([[0.0, 0.0, 0.0]])
    g, o = both(p, sc, normals=nrm, sample_points=pts)
    g.detect(); o.detect()
    gh, oh = g.hypotheses(), o.hypotheses()
    assert hyp_int_fields_equal(gh, oh) is None
    assert np.array_equal(g.sweep_records()["fingers"], o.sweep_records()["fingers"])
    assert np.any(gh["full_antipodal"] == 1) and np.allclose(gh[gh["full_antipodal"] == 1]["width"], 0.05, atol=1e-6)


def test_errors_and_empty():
    p = make_params()
    g = capi.Context(p)
    with pytest.raises(capi.Ag2Error) as e:
        g.detect()
    assert e.value.code == capi.AG2_E_EMPTY_CLOUD
    g.set_cloud(np.zeros((10, 3), np.float32) + np.arange(10, dtype=np.float32)[:, None] * 0.001)
    with pytest.raises(capi.Ag2Error) as e:
        g.detect()
    assert e.value.code == capi.AG2_E_NO_SAMPLES
    g.set_sample_indices(np.array([0, 5], np.int32))
    g.detect()
    assert g.num(capi.LIST_HYPOTHESES) >= 0


def _compare_run(g, o, check_normals=True):
    if check_normals:
        gn, on = g.normals(), o.normals()
        assert np.array_equal(np.isnan(gn), np.isnan(on))
        ok = ~np.isnan(on).any(1)
        assert angle(gn[ok], on[ok]).max() <= 1e-4
        assert np.mean(np.all(gn[ok].view(np.uint32) == on[ok].view(np.uint32), axis=1)) > 0.999
    gs, os_ = g.sweep_records(), o.sweep_records()
    for f in ("fingers", "hand", "slot", "depth_step", "num_in_box", "antipodal", "emitted"):
        assert np.array_equal(gs[f], os_[f]), f
    gh, oh = g.hypotheses(capi.LIST_SELECTED), o.hypotheses(capi.LIST_SELECTED)
    assert len(gh) == len(oh) and hyp_int_fields_equal(gh, oh) is None
    for f in POSE_FIELDS:
        assert np.allclose(gh[f], oh[f], rtol=0, atol=1e-9), f
    for k in ("sum_neighbors_hands", "sum_cropped", "sum_in_box", "sum_neighbors_normals"):
        assert g.stats()[k] == o.stats()[k], k


def test_chunked_hand_staging_and_levelled_normals(tiny_scene, monkeypatch):
    """Forces the multi-chunk path of the hand kernel (small staging buffer) and the
    multi-level path of the normals kernel (small first capacity) on the tiny scene:
    results must not change (all accumulators are associative / the ranking is exact)."""
    monkeypatch.setenv("AG2_HANDS_STAGE_CAP", "1664")
    monkeypatch.setenv("AG2_NORMALS_FIRST_TIER", "0")
    p = make_params(capi.MODE_GEOMETRIC)
    g, o = both(p, tiny_scene)
    g.detect(); o.detect()
    _compare_run(g, o)
    n = g.num(capi.LIST_HYPOTHESES)
    for h in range(0, n, max(1, n // 6)):
        assert np.array_equal(g.in_box(capi.LIST_HYPOTHESES, h), o.in_box(capi.LIST_HYPOTHESES, h))


@pytest.mark.parametrize("mode", [capi.MODE_GEOMETRIC, capi.MODE_PREDICTION])
def test_single_launch_and_multi_kernel_lists_agree(tiny_scene, monkeypatch, mode):
    """Small runs build the hypothesis + pruned lists in one single-CTA launch and filter + rank + take the
    selection in another; large ones use flag / scan / compact kernels and tile sorts. Same lists either way."""
    sc = scenes.make_scene("tiny", num_samples=700)  # 5 600 slots: six tiles of the single-launch scan
    p = make_params(mode, min_score_diff=0.0, num_selected=7)
    out = []
    for forced in ("0", "1"):
        monkeypatch.setenv("AG2_LISTS_MULTIKERNEL", forced)
        with capi.Context(p) as g:
            if mode == capi.MODE_PREDICTION:
                g.set_lenet_weights(ob.xavier_weights(3, bias_scale=0.1))
            g.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); g.set_sample_indices(sc["samples"])
            g.detect()
            out.append([g.hypotheses(w).tobytes() for w in
                        (capi.LIST_HYPOTHESES, capi.LIST_PRUNED, capi.LIST_ANTIPODAL, capi.LIST_SELECTED)]
                       + [repr(sorted((k, v) for k, v in g.stats().items() if k.startswith("num_")))])
    sz = capi.HYPOTHESIS_DTYPE.itemsize
    assert len(out[0][3]) == min(7 * sz, len(out[0][2])), (out[0][4], [len(x) // sz for x in out[0][:4]])
    assert out[0] == out[1]


@pytest.mark.parametrize("hook", [("AG2_RENDER_SPILL", "1"), ("AG2_RENDER_SMALL_CAP", "200"),
                                  ("AG2_RENDER_SMALL_CAP", "8")])
def test_render_spill_and_size_class_paths(tiny_scene, monkeypatch, hook):
    """Closing regions too large for shared memory (c4 densities) sort in global memory with a fixed
    set of CTAs walking the images; images above a size boundary go to a second launch with a larger
    capacity. Both forced here on the tiny scene: image bytes equal the oracle's."""
    monkeypatch.setenv(*hook)
    p = make_params(capi.MODE_PREDICTION, min_score_diff=0.0)
    g, o = both(p, tiny_scene, weights=ob.xavier_weights(2))
    for c in (g, o):
        c.generate_hypotheses(); c.prune()
    gi, oi = g.render_images(), o.render_images()
    assert gi.shape == oi.shape and gi.shape[0] > 5 and np.array_equal(gi, oi)


def test_c2_full_size_subset():
    """BASELINE.json config c2 at full cloud size (307 200 points, k_n up to ~500): normals of
    every point and the sweep of 48 samples against the oracle."""
    sc = scenes.make_scene("c2", num_samples=48)
    p = make_params(capi.MODE_GEOMETRIC)
    p.num_threads = 16
    g, o = both(p, sc)
    g.detect(); o.detect()
    _compare_run(g, o)


def test_orientation_and_radius_sweep(tiny_scene):
    """c5: more orientations and a wider hand-search radius."""
    p = make_params(capi.MODE_GEOMETRIC, num_orientations=16, nn_radius_hands=0.12)
    g, o = both(p, tiny_scene)
    g.detect(); o.detect()
    _compare_run(g, o, check_normals=False)


@pytest.mark.parametrize("method", [2, 1])
def test_importance_sampling_rounds(tiny_scene, method):
    """a17: ImportanceSampling::detectGraspPoses — identical drawn samples per round and identical
    accumulated grasp list (rounds reuse the resident grid + normals on the GPU side)."""
    p = make_params(capi.MODE_NONE, num_selected=20)  # mode 0: every pruned hand is kept
    g, o = both(p, tiny_scene)
    ip = g.default_is_params()
    ip.num_iterations, ip.num_samples, ip.seed, ip.method = 3, 40, 5, method
    g.detect_importance(ip)
    o.detect_importance(ip)
    for r in range(3):
        gs, os_ = g.importance_samples(r), o.importance_samples(r)
        assert gs.shape == (40, 3) and np.array_equal(gs, os_), r
    gh, oh = g.hypotheses(capi.LIST_IMPORTANCE), o.hypotheses(capi.LIST_IMPORTANCE)
    assert len(gh) == len(oh) and len(gh) > 20
    assert hyp_int_fields_equal(gh, oh) is None
    for f in POSE_FIELDS:
        assert np.allclose(gh[f], oh[f], rtol=0, atol=1e-9), f


@pytest.mark.parametrize("mode", [capi.MODE_NONE, capi.MODE_PREDICTION])
def test_importance_rounds_driven_by_the_group_host(tiny_scene, mode):
    """The multi-GPU host (ag2_group_detect_importance, csrc/group.cu) runs the rounds itself
    (ag2_draw_importance_samples + one sharded detect per round); with one rank the result is the list
    ag2_detect_importance builds."""
    sc = tiny_scene
    p = make_params(mode, num_selected=12, min_score_diff=-1e9)
    with capi.Context(p) as a, capi.Group(p, rank=0, world=1, device=0) as grp:
        a.init_lenet_xavier(3); grp.init_lenet_xavier(3)
        ip = a.default_is_params()
        ip.num_iterations, ip.num_samples, ip.seed = 3, 40, 9
        a.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); a.set_sample_indices(sc["samples"])
        a.detect_importance(ip)
        want = a.hypotheses(capi.LIST_IMPORTANCE)
        got = grp.detect_importance(sc["xyz"], sc["cam_mask"], sc["num_cams"], sc["samples"], ip)
        assert got.tobytes() == want.tobytes(), (len(got), len(want))
        assert len(want) > 12
        # the drawn batches themselves
        b = grp.local_ctx(0)
        for r in range(3):
            upto = want[want["sample_id"] < len(sc["samples"]) + 40 * r]
            assert np.array_equal(b.draw_importance_samples(ip, r, upto), a.importance_samples(r)), r


@pytest.mark.parametrize("voxelize", [True, False])
def test_preprocess_cloud(tiny_scene, voxelize):
    """f1: CloudCamera::filterWorkspace + voxelizeCloud + subsampleUniformly, bit-exact."""
    p = make_params(capi.MODE_GEOMETRIC)
    g, o = capi.Context(p), ob.context(p)
    xyz = tiny_scene["xyz"].copy()
    xyz[7] = np.nan
    for c in (g, o):
        c.preprocess_cloud(xyz, tiny_scene["cam_mask"], tiny_scene["num_cams"], voxelize=voxelize)
    (gx, gc), (ox, oc) = g.cloud(), o.cloud()
    assert gx.shape == ox.shape and len(gx) > 1000
    assert np.array_equal(gx.view(np.uint32), ox.view(np.uint32))
    assert np.array_equal(gc, oc)
    gi, oi = g.subsample_uniformly(50, seed=9), o.subsample_uniformly(50, seed=9)
    assert np.array_equal(gi, oi)
    g.detect(); o.detect()
    _compare_run(g, o)


@pytest.fixture(scope="module")
def dense_scene():
    return scenes.make_scene("tiny", num_samples=1500)  # enough samples for neighbouring hands to cluster


def test_find_clusters(dense_scene):
    """f2: HandleSearch::findClusters — inlier decisions bit-exact, moved poses / mean scores to 1e-9."""
    p = make_params(capi.MODE_NONE, min_inliers=1)
    g, o = both(p, dense_scene)
    g.detect(); o.detect()
    for c in (g, o):
        c.find_clusters(capi.LIST_HYPOTHESES)
    gc, oc = g.hypotheses(capi.LIST_CLUSTERS), o.hypotheses(capi.LIST_CLUSTERS)
    assert len(gc) == len(oc) and 0 < len(gc) < g.num(capi.LIST_HYPOTHESES)
    assert hyp_int_fields_equal(gc, oc) is None
    for f in POSE_FIELDS + ["score"]:
        assert np.allclose(gc[f], oc[f], rtol=0, atol=1e-9), f


def test_set_min_inliers_changes_the_threshold_of_the_next_clustering(dense_scene):
    """HandleSearch::setMinInliers (handle_search.h:83-86) through ag2_set_min_inliers: the lists of the last call stay,
    the next findClusters uses the new threshold -- the same clusters as a context created with that threshold."""
    a = capi.Context(make_params(capi.MODE_NONE, min_inliers=1))
    b = capi.Context(make_params(capi.MODE_NONE, min_inliers=3))
    for c in (a, b):
        c.set_cloud(dense_scene["xyz"], dense_scene["cam_mask"], dense_scene["num_cams"])
        c.set_sample_indices(dense_scene["samples"])
        c.detect()
    a.find_clusters(capi.LIST_HYPOTHESES)
    loose = a.hypotheses(capi.LIST_CLUSTERS)
    a.set_min_inliers(3)
    assert a.num(capi.LIST_HYPOTHESES) == b.num(capi.LIST_HYPOTHESES) > 0  # nothing was invalidated
    a.find_clusters(capi.LIST_HYPOTHESES)
    b.find_clusters(capi.LIST_HYPOTHESES)
    ga, gb = a.hypotheses(capi.LIST_CLUSTERS), b.hypotheses(capi.LIST_CLUSTERS)
    assert len(ga) == len(gb) and ga.tobytes() == gb.tobytes()
    assert len(ga) < len(loose)
    with pytest.raises(capi.Ag2Error):
        a.set_min_inliers(-1)


def test_detect_with_clustering(dense_scene):
    """grasp_detector.cpp:229-252: clusters of the accepted hands, then top-k over the clusters."""
    w = ob.xavier_weights(3, bias_scale=0.5)
    p = make_params(capi.MODE_PREDICTION, min_score_diff=-1e9, min_inliers=1, num_selected=8)
    g, o = both(p, dense_scene, weights=w)
    g.detect(); o.detect()
    gc, oc = g.hypotheses(capi.LIST_CLUSTERS), o.hypotheses(capi.LIST_CLUSTERS)
    assert len(gc) == len(oc) and len(gc) > 3
    assert hyp_int_fields_equal(gc, oc) is None
    assert np.allclose(gc["bottom"], oc["bottom"], rtol=0, atol=1e-9)
    assert np.all(np.abs(gc["score"] - oc["score"]) <= np.maximum(1e-3, 1e-5 * np.abs(oc["score"])))
    gs, os_ = g.hypotheses(capi.LIST_SELECTED), o.hypotheses(capi.LIST_SELECTED)
    assert len(gs) == len(os_) == min(8, len(gc))
    assert np.all(np.diff(gs["score"]) <= 0)


def test_group_of_one_rank_equals_plain_detect(tiny_scene):
    """ag2_group_detect at world size 1 -- inputs staged in the exchange window, sample shard taken on the
    device, normals written into the window-backed array, selected records read from the result path -- returns
    exactly what ag2_detect + ag2_get_hypotheses return, with host pointers and with device pointers."""
    import torch
    sc = tiny_scene
    p = make_params(capi.MODE_PREDICTION, min_score_diff=-1e9, num_selected=20)
    with capi.Context(p) as a, capi.Group(p, rank=0, world=1, device=0) as grp:
        a.init_lenet_xavier(3); grp.init_lenet_xavier(3)
        a.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); a.set_sample_indices(sc["samples"]); a.detect()
        want = a.hypotheses(capi.LIST_SELECTED)
        got = grp.detect(sc["xyz"], sc["cam_mask"], sc["num_cams"], sc["samples"])
        assert len(want) > 0 and got.tobytes() == want.tobytes()
        dev = torch.device("cuda", 0)
        xd, cd = torch.from_numpy(sc["xyz"]).to(dev), torch.from_numpy(sc["cam_mask"]).to(dev)
        sd = torch.from_numpy(np.asarray(sc["samples"], np.int32)).to(dev)
        torch.cuda.synchronize()
        for _ in range(2):  # the second call reuses windows, grid buffers and tier launches
            got = grp.detect_ptr(xd.data_ptr(), len(sc["xyz"]), cd.data_ptr(), sc["num_cams"], sd.data_ptr(),
                                 len(sc["samples"]), True)
            assert got.tobytes() == want.tobytes()
        assert grp.io_bytes()[0] == 0
        b = grp.local_ctx(0)
        b.n_samples = len(sc["samples"])
        assert np.array_equal(b.sample_indices(), np.asarray(sc["samples"], np.int32))
        # device-side range check of the sample shard
        with pytest.raises(capi.Ag2Error):
            grp.detect(sc["xyz"], sc["cam_mask"], sc["num_cams"], np.array([0, len(sc["xyz"])], np.int32))
        # a smaller and a larger cloud through the same group (window regrowth)
        half = len(sc["xyz"]) // 2
        smp = np.asarray(sc["samples"], np.int32)
        smp = smp[smp < half]
        a.set_cloud(sc["xyz"][:half], sc["cam_mask"][:half], sc["num_cams"]); a.set_sample_indices(smp); a.detect()
        assert grp.detect(sc["xyz"][:half], sc["cam_mask"][:half], sc["num_cams"], smp).tobytes() == \
            a.hypotheses(capi.LIST_SELECTED).tobytes()
    with capi.Context(p) as b:
        b.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"])
        with pytest.raises(capi.Ag2Error):  # device-side range check
            bad = torch.tensor([0, len(sc["xyz"])], dtype=torch.int32, device=dev)
            b.set_sample_indices_device(bad.data_ptr(), 2)


@pytest.mark.skipif("__import__('torch').cuda.device_count() < 2")
@pytest.mark.parametrize("transport", ["peer", "nccl"])
def test_group_of_two_gpus_in_one_process_equals_one_gpu(tiny_scene, monkeypatch, transport):
    """ag2_group_create(n_gpus = 2): two ranks as host threads of this process, exchanges on peer memory (or
    NCCL); the merged selection equals the single-GPU one byte for byte (runs where the box has two GPUs)."""
    sc = tiny_scene
    monkeypatch.setenv("AG2_GROUP_NCCL", "1" if transport == "nccl" else "0")
    p = make_params(capi.MODE_PREDICTION, min_score_diff=-1e9, num_selected=20)
    with capi.Context(p) as a, capi.Group(p, n_gpus=2) as grp:
        a.init_lenet_xavier(3); grp.init_lenet_xavier(3)
        a.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); a.set_sample_indices(sc["samples"]); a.detect()
        want = a.hypotheses(capi.LIST_SELECTED)
        for _ in range(2):
            got = grp.detect(sc["xyz"], sc["cam_mask"], sc["num_cams"], sc["samples"])
            assert got.tobytes() == want.tobytes()
        assert grp.peer_stores() == (transport == "peer")


@pytest.mark.parametrize("world", [1, 3, 8])
def test_normals_shards_assemble_to_the_full_result(tiny_scene, world):
    """Multi-GPU normals on one GPU: every 'rank' computes its interleaved granules into its own
    buffer, the buffers are concatenated (what the all-gather does) and installed; the normals and
    everything downstream are bit-identical to ag2_compute_normals."""
    import torch
    sc = tiny_scene
    p = make_params(capi.MODE_GEOMETRIC)
    a, b = capi.Context(p), capi.Context(p)
    for c in (a, b):
        c.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); c.set_sample_indices(sc["samples"])
    a.compute_normals()
    cap = b.normals_shard_points(world)
    allb = torch.full((world, cap, 4), float("nan"), dtype=torch.float32, device="cuda")
    owned = [b.compute_normals_shard(r, world, allb[r].data_ptr(), cap) for r in range(world)]
    assert sum(owned) == len(a.normals()) and max(owned) == owned[0] <= cap
    b.set_normals_shards(allb.data_ptr(), cap, world)
    assert a.normals().tobytes() == b.normals().tobytes()
    a.detect(); b.detect()
    assert a.hypotheses(capi.LIST_SELECTED).tobytes() == b.hypotheses(capi.LIST_SELECTED).tobytes()


def test_normals_shards_of_a_cloud_smaller_than_the_world(tiny_scene):
    """70 points are three 32-point granules: at world 8 the ranks 3..7 own nothing and must still go through
    the count pass and the per-tier launches; the assembled normals equal the single-GPU ones (most are NaN:
    so few points rarely have three neighbours within the radius)."""
    import torch
    xyz = np.ascontiguousarray(tiny_scene["xyz"][:70])
    p = make_params(capi.MODE_GEOMETRIC)
    a, b = capi.Context(p), capi.Context(p)
    for c in (a, b):
        c.set_cloud(xyz, np.ones(70, np.uint8), 1)
    a.compute_normals()
    cap = b.normals_shard_points(8)
    allb = torch.full((8, cap, 4), float("nan"), dtype=torch.float32, device="cuda")
    owned = [b.compute_normals_shard(r, 8, allb[r].data_ptr(), cap) for r in range(8)]
    assert owned == [32, 32, 6, 0, 0, 0, 0, 0]
    b.set_normals_shards(allb.data_ptr(), cap, 8)
    assert a.normals().tobytes() == b.normals().tobytes()


@pytest.mark.parametrize("aligns", [False, True])
def test_render_point_sets_vs_oracle(aligns):
    """Learning::createGraspImage(points, normals, grasp) (the ClassificationNode image generator): explicit
    unit-box point sets, including an empty set, points outside the image and cells hit many times."""
    rng = np.random.default_rng(23)
    sizes = [0, 1, 37, 900, 5000]
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint32)
    pts = rng.uniform(-0.05, 1.05, (offs[-1], 3))
    pts[offs[3]:offs[3] + 300, :2] = [0.41, 0.52]  # 300 points in one cell: in-run ordering of a long run
    pts[offs[4] + 5] = [1e9, -1e9, 0.0]             # far outside: never matched
    nrm = rng.normal(size=(offs[-1], 3))
    g = capi.Context(make_params(capi.MODE_PREDICTION)); o = ob.context(make_params(capi.MODE_PREDICTION))
    gi, oi = g.render_point_sets(pts, nrm, offs, aligns), o.render_point_sets(pts, nrm, offs, aligns)
    assert gi.shape == oi.shape == (5, 60, 60, 3)
    assert np.array_equal(gi, oi)
    assert not gi[0].any() and gi[3].any()


def _lattice_scene():
    """Two regular 2 mm lattices (a level one and a tilted one): nearly every neighbourhood holds many
    EXACTLY equal float distances, so the (d2, input index) tie order of the radius search is exercised."""
    u, v = np.meshgrid(np.arange(100, dtype=np.float64), np.arange(100, dtype=np.float64), indexing="ij")
    a = np.stack([0.60 + 0.002 * u, -0.10 + 0.002 * v, np.full_like(u, -0.20)], -1).reshape(-1, 3)
    b = np.stack([0.60 + 0.002 * u, 0.15 + 0.002 * v, -0.20 + 0.001 * u], -1).reshape(-1, 3)
    xyz = np.concatenate([a, b]).astype(np.float32)
    rng = np.random.default_rng(5)
    xyz = xyz[rng.permutation(len(xyz))]  # input order != cell order
    return dict(name="lattice", xyz=xyz, cam_mask=np.ones(len(xyz), np.uint8), num_cams=1,
                samples=np.arange(0, len(xyz), 997, dtype=np.int32))


@pytest.mark.parametrize("scene,quad_tiers", [("tiny", None), ("tiny", "8"), ("tiny", "0"), ("lattice", "8"),
                                              ("lattice", None), ("tiny", "cta"), ("lattice", "cta")])
def test_normals_quad_per_point_and_warp_per_point_agree(tiny_scene, monkeypatch, scene, quad_tiers):
    """Points are grouped by neighbourhood size into tiers, one launch per tier, side by side. Sparse tiers
    run four lanes per point (k_normals_quad; up to 208 neighbours when AG2_NORMALS_QUAD_TIERS=8), the
    dense ones (capacity 512 and up; every tier with AG2_NORMALS_CTA_FROM=0) one CTA per point
    (k_normals_cta), the others one warp per point (k_normals); AG2_NORMALS_WARP=1 runs k_normals
    everywhere in chained capacity levels. All reproduce the radius-search order exactly, so the normals are bit-identical --
    and equal to the oracle's."""
    sc = _lattice_scene() if scene == "lattice" else tiny_scene
    if quad_tiers == "cta":  # every tier on the CTA-per-point kernel of the dense tiers
        monkeypatch.setenv("AG2_NORMALS_QUAD_TIERS", "0")
        monkeypatch.setenv("AG2_NORMALS_CTA_FROM", "0")
    elif quad_tiers is not None:
        monkeypatch.setenv("AG2_NORMALS_QUAD_TIERS", quad_tiers)
    monkeypatch.setenv("AG2_NORMALS_FIRST_TIER", "0")  # forced run: first level capacity 64
    p = make_params(capi.MODE_GEOMETRIC)
    out = []
    for forced in ("0", "1"):
        monkeypatch.setenv("AG2_NORMALS_WARP", forced)
        with capi.Context(p) as g:
            load_scene(g, sc)
            g.compute_normals()
            out.append(g.normals().copy())
    assert np.array_equal(out[0].view(np.uint32), out[1].view(np.uint32))
    o = ob.context(p)
    load_scene(o, sc)
    o.compute_normals()
    on = o.normals()
    assert np.array_equal(np.isnan(out[0]), np.isnan(on))
    ok = ~np.isnan(on).any(1)
    assert ok.sum() > 1000
    assert angle(out[0][ok], on[ok]).max() <= 1e-4
    bit_equal = np.mean(np.all(out[0][ok].view(np.uint32) == on[ok].view(np.uint32), axis=1))
    print("%s: bit-equal fraction %.6f" % (scene, bit_equal))
    assert bit_equal > 0.999


# ---- the known-answer scenes of tests/test_oracle_pins.py (hand-derived expectations there), GPU against oracle ----
def _known_answer_scenes():
    from common import cylinder_scene, shell_around, thin_box_scene
    from test_oracle_pins import deepen_scene
    cx, cn, cs = cylinder_scene()
    bx, bn, bs = thin_box_scene()
    sx, sn = shell_around(bs)
    dx, dn, ds = deepen_scene()
    return {
        "cylinder": (cx, cn, cs, dict(nn_radius_taubin=0.01)),
        "thin_box": (bx, bn, bs, dict(nn_radius_taubin=0.03)),
        "below_hand_base": (np.concatenate([bx, sx]), np.concatenate([bn, sn]), bs, dict(nn_radius_taubin=0.03)),
        "deepen_boundary": (dx, dn, ds, dict(nn_radius_taubin=0.03, init_bite=0.01, hand_depth=0.025)),
    }


@pytest.mark.parametrize("name", ["cylinder", "thin_box", "below_hand_base", "deepen_boundary"])
def test_known_answer_scenes_gpu_equals_oracle(name):
    xyz, nrm, sample, kw = _known_answer_scenes()[name]
    p = make_params(capi.MODE_GEOMETRIC, prune=0, **kw)
    g, o = capi.Context(p), ob.context(p)
    for c in (g, o):
        c.set_cloud(xyz, np.ones(len(xyz), np.uint8), 1, normals=nrm)
        c.set_sample_points(sample[None, :])
        c.detect()
    _compare_run(g, o, check_normals=False)
    gf, of = g.frames(), o.frames()
    assert np.array_equal(gf["valid"], of["valid"]) and angle(gf["curvature_axis"], of["curvature_axis"]).max() <= 1e-6
    gh = g.hypotheses()
    if name == "below_hand_base":
        assert len(gh) == 0 and np.all(g.sweep_records()["fingers"] == 0)
    if name == "cylinder":
        assert abs(gh[gh["full_antipodal"] == 1]["width"].max() - 0.06) < 1e-6
    if name == "deepen_boundary":
        assert gh["depth_step"].max() == 3 and gh["hand_top"].max() == 0.025
    g.close()


@pytest.mark.parametrize("aligns", [False, True])
def test_image_cells_on_a_sixtieth_boundary_gpu_equals_oracle(aligns):
    """learning.cpp:152-156 on the device: points on k/60, one ulp below and above, aliasing cells."""
    rng = np.random.default_rng(12)
    pts = []
    for k in range(60):
        pts.append([k / 60.0, ((7 * k) % 60) / 60.0, 0.5])
        pts.append([np.nextafter(k / 60.0, 0.0) if k else 0.0, ((11 * k + 3) % 60) / 60.0, 0.25])
        pts.append([np.nextafter(k / 60.0, 1.0), ((13 * k + 5) % 60) / 60.0, 0.75])
    pts += [[1.0, 0.2, 0.5], [1.0 + 1.0 / 60, 0.3, 0.5], [0.5, 1.0 - 1e-12, 0.5], [0.5, 1.0, 0.5], [-1.0 / 60, 0.5, 0.5]]
    pts = np.array(pts, np.float64)
    nrm = rng.normal(size=pts.shape)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    p = make_params()
    with capi.Context(p) as g:
        got = g.render_point_sets(pts, nrm, [0, len(pts)], aligns=aligns)[0]
    want = ob.context(p).render_point_sets(pts, nrm, [0, len(pts)], aligns=aligns)[0]
    assert np.array_equal(got, want), int(np.sum(got != want))


def test_caffemodel_file_loads_like_set_lenet_weights(tmp_path):
    """Classifier ctor -> net_->CopyTrainedLayersFrom(trained_file) (caffe_classifier.cpp:13-14): a .caffemodel written by
    the hand-made protobuf writer of test_formats_cpu.py (legacy V1 layers, bias of ip2 absent) gives the ctx the same
    blobs, and the same logits bit for bit, as ag2_set_lenet_weights with the arrays."""
    from test_formats_cpu import _caffemodel

    w = {k: v.copy() for k, v in ob.xavier_weights(11, bias_scale=0.2).items()}
    path = tmp_path / "lenet.caffemodel"
    path.write_bytes(_caffemodel(w, True, skip_bias_of=("ip2",)))
    w["ip2_b"] = np.zeros_like(w["ip2_b"])
    imgs = np.random.default_rng(2).integers(0, 256, (9, 60, 60, 3), dtype=np.uint8)
    p = make_params(capi.MODE_PREDICTION)
    with capi.Context(p) as a, capi.Context(p) as b:
        a.load_lenet_caffemodel(str(path))
        b.set_lenet_weights(w)
        got = a.get_lenet_weights()
        for k, _ in capi.LENET_SHAPES:
            assert np.array_equal(np.asarray(got[k]).reshape(w[k].shape), w[k]), k
        assert np.array_equal(a.classify_images(imgs), b.classify_images(imgs))
        with pytest.raises(capi.Ag2Error):
            a.load_lenet_caffemodel(str(tmp_path / "missing.caffemodel"))
