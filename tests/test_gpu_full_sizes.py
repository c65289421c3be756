"""BASELINE.json configurations at FULL size on the GPU.  The oracle is too slow for whole runs at
these sizes, so each case combines (i) the oracle on a bounded slice of the same run — normals of
a few hundred points, the complete sweep of a strided subset of the samples — with (ii) properties
that do not depend on size: determinism, independence of samples (a strided subset run alone
reproduces its records of the full run), multi-GPU normals shards assembling to the full result,
and the invariants every emitted record satisfies."""
import numpy as np
import pytest
import torch

from agile_grasp2_b200 import capi, scenes
from common import angle, both, hyp_int_fields_equal, make_params
from oracle import binding as ob

pytestmark = pytest.mark.gpu
POSE = ["surface", "bottom", "top", "axis", "approach", "binormal", "width"]


def _check_records(h, p):
    """Invariants of emitted / selected records (hand_search.cpp:380-421, grasp_detector.cpp:363-395)."""
    assert np.all(np.isfinite(h["bottom"])) and np.all(np.isfinite(h["surface"])) and np.all(np.isfinite(h["width"]))
    for f in ("axis", "approach", "binormal"):
        assert np.allclose(np.linalg.norm(h[f], axis=1), 1.0, atol=1e-9), f
    assert np.all(np.abs(np.sum(h["axis"] * h["approach"], 1)) < 1e-9)
    assert np.all((h["orientation_id"] >= 0) & (h["orientation_id"] < p.num_orientations))
    assert np.all((h["finger_slot"] >= 0) & (h["finger_slot"] < 10)) and np.all(h["num_in_box"] > 0)
    assert np.all(h["full_antipodal"] <= h["half_antipodal"])


def test_c1_whole_run_against_oracle():
    """c1 (50 k points, 100 samples, geometric mode): small enough for the oracle end to end."""
    sc = scenes.make_scene("c1")
    p = make_params(capi.MODE_GEOMETRIC, num_threads=16)
    g, o = both(p, sc)
    g.detect(); o.detect()
    gs, os_ = g.sweep_records(), o.sweep_records()
    for f in ("fingers", "hand", "slot", "depth_step", "num_in_box", "antipodal", "emitted"):
        assert np.array_equal(gs[f], os_[f]), f
    for which in (capi.LIST_HYPOTHESES, capi.LIST_PRUNED, capi.LIST_ANTIPODAL, capi.LIST_SELECTED):
        gh, oh = g.hypotheses(which), o.hypotheses(which)
        assert len(gh) == len(oh) and hyp_int_fields_equal(gh, oh) is None
        for f in POSE:
            assert np.allclose(gh[f], oh[f], rtol=0, atol=1e-9), f
    gn, on = g.normals(), o.normals()
    assert np.mean(np.all(gn.view(np.uint32) == on.view(np.uint32), axis=1)) > 0.999


def test_c3_full_size():
    """c3: 614 400 points of two views, 10 000 samples x 8 orientations, LeNet scoring."""
    sc = scenes.make_scene("c3")
    S = len(sc["samples"])
    p = make_params(capi.MODE_PREDICTION, num_threads=16, num_selected=50, min_score_diff=-1e9)
    w = ob.xavier_weights(7, bias_scale=0.1)
    g = capi.Context(p); g.set_lenet_weights(w)
    g.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); g.set_sample_indices(sc["samples"])
    g.detect()
    full = g.hypotheses(capi.LIST_HYPOTHESES).copy()
    sel = g.hypotheses(capi.LIST_SELECTED).copy()
    st = g.stats()
    assert st["num_samples"] == S and len(full) == st["num_hypotheses"] > 1000 and len(sel) == 50
    _check_records(full, p)
    assert np.all(np.diff(sel["score"]) <= 0)  # partial_sort order
    # determinism: the same call again gives the same bytes
    g.set_sample_indices(sc["samples"]); g.detect()
    assert g.hypotheses(capi.LIST_HYPOTHESES).tobytes() == full.tobytes()
    assert g.hypotheses(capi.LIST_SELECTED).tobytes() == sel.tobytes()
    # samples are independent: a strided subset alone (with its id map) reproduces its records
    first, stride = 7, 157
    sub = np.ascontiguousarray(sc["samples"][first::stride])
    g.set_sample_indices(sub); g.set_sample_id_map(first, stride); g.generate_hypotheses()
    part = g.hypotheses(capi.LIST_HYPOTHESES).copy()
    part["sample_id"] = first + stride * part["sample_id"]
    want = full[(full["sample_id"] - first) % stride == 0]
    assert len(part) == len(want) > 0
    for f in part.dtype.names:
        if f not in ("score", "full_antipodal"):  # written by the classification stage (grasp_detector.cpp:199-207)
            assert np.array_equal(part[f], want[f]), f
    gsw = g.sweep_records().copy()
    gn = g.normals()
    # the oracle on the same slice: normals of 400 points, the whole sweep of the subset
    o = ob.context(p)
    o.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"])
    pick = np.random.default_rng(1).choice(len(sc["xyz"]), 400, replace=False).astype(np.int32)
    on = ob.point_normals(o, pick)
    ok = ~np.isnan(on).any(1)
    assert np.array_equal(np.isnan(gn[pick]).any(1), ~ok)
    assert angle(gn[pick][ok], on[ok]).max() <= 1e-4
    assert np.mean(np.all(gn[pick][ok].view(np.uint32) == on[ok].view(np.uint32), axis=1)) > 0.99
    o2 = ob.context(p)
    o2.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"], normals=gn)  # GPU normals in: isolates the sweep
    o2.set_sample_indices(sub); o2.set_sample_id_map(first, stride); o2.generate_hypotheses()
    osw = o2.sweep_records()
    for f in ("fingers", "hand", "slot", "depth_step", "num_in_box", "antipodal", "emitted"):
        assert np.array_equal(gsw[f], osw[f]), f
    # multi-GPU normals: 4 shards assemble to the same normals
    b = capi.Context(p)
    b.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"])
    cap = b.normals_shard_points(4)
    allb = torch.empty((4, cap, 4), dtype=torch.float32, device="cuda")
    for r in range(4):
        b.compute_normals_shard(r, 4, allb[r].data_ptr(), cap)
    b.set_normals_shards(allb.data_ptr(), cap, 4)
    assert b.normals().tobytes() == gn.tobytes()


def test_c4_full_size_importance_sampling():
    """c4: 4.9 M fused points (k_n ~ 590, up to ~1 400; K ~ 47 k), 100 000 samples in importance-sampling
    rounds (20 000 initial + 4 rounds x 20 000)."""
    sc = scenes.make_scene("c4")
    p = make_params(capi.MODE_PREDICTION, num_threads=16, num_selected=2000, min_score_diff=-1e9)
    w = ob.xavier_weights(7, bias_scale=0.1)
    g = capi.Context(p); g.set_lenet_weights(w)
    g.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"])
    init = np.ascontiguousarray(sc["samples"][:: len(sc["samples"]) // 20000][:20000])
    ip = g.default_is_params()
    ip.num_init_samples, ip.num_samples, ip.num_iterations, ip.seed = len(init), 20000, 4, 11
    g.set_sample_indices(init)
    g.detect_importance(ip)
    res = g.hypotheses(capi.LIST_IMPORTANCE).copy()
    assert len(res) == 5 * 2000  # every round appends its top num_selected (importance_sampling.cpp:60-99)
    assert g.stats()["num_hypotheses"] > 5000  # of the last round alone
    _check_records(res, p)
    rounds = [g.importance_samples(r) for r in range(4)]
    assert all(len(r) == 20000 for r in rounds)
    lo, hi = sc["xyz"].min(0) - 0.2, sc["xyz"].max(0) + 0.2
    assert all(np.all((r >= lo) & (r <= hi)) for r in rounds)  # drawn around grasps inside the scene
    # determinism of the whole multi-round run
    g.set_sample_indices(init); g.detect_importance(ip)
    assert g.hypotheses(capi.LIST_IMPORTANCE).tobytes() == res.tobytes()
    # the oracle on a slice at this density: normals of 200 points ...
    gn = g.normals()
    o = ob.context(p)
    o.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"])
    pick = np.random.default_rng(2).choice(len(sc["xyz"]), 200, replace=False).astype(np.int32)
    on = ob.point_normals(o, pick)
    ok = ~np.isnan(on).any(1)
    assert angle(gn[pick][ok], on[ok]).max() <= 1e-4
    assert np.mean(np.all(gn[pick][ok].view(np.uint32) == on[ok].view(np.uint32), axis=1)) > 0.99
    # ... and the complete sweep of 24 samples drawn in round 2 (neighbourhoods of ~50 k points: chunked staging)
    pts = rounds[1][:: len(rounds[1]) // 24][:24]
    g2 = capi.Context(make_params(capi.MODE_GEOMETRIC, num_threads=16))
    o2 = ob.context(make_params(capi.MODE_GEOMETRIC, num_threads=16))
    for c in (g2, o2):
        c.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"], normals=gn)
        c.set_sample_points(pts)
        c.generate_hypotheses()
    gsw, osw = g2.sweep_records(), o2.sweep_records()
    for f in ("fingers", "hand", "slot", "depth_step", "num_in_box", "antipodal", "emitted"):
        assert np.array_equal(gsw[f], osw[f]), f
    gh, oh = g2.hypotheses(), o2.hypotheses()
    assert len(gh) == len(oh) and hyp_int_fields_equal(gh, oh) is None
    for f in POSE:
        assert np.allclose(gh[f], oh[f], rtol=0, atol=1e-9), f


@pytest.fixture(scope="module")
def c3_with_normals():
    """The c3 cloud and its normals computed once on the GPU (checked against the oracle by test_c3_full_size)."""
    sc = scenes.make_scene("c3")
    with capi.Context(make_params(capi.MODE_GEOMETRIC)) as g:
        g.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"])
        g.compute_normals()
        gn = g.normals().copy()
    return sc, gn


@pytest.mark.parametrize("R", [8, 16, 32])
@pytest.mark.parametrize("radius", [0.06, 0.08, 0.10, 0.12])
def test_c5_orientation_radius_sweep_on_the_clutter_cloud(c3_with_normals, R, radius):
    """c5 (BASELINE.json configs[4]): num_orientations 8 -> 32 x nn_radius_hands 0.06 -> 0.12 on the c3 cloud.  The
    complete sweep of a 40-sample strided subset against the oracle (the GPU's normals go in on both sides, which
    isolates hand_search.cpp:173-235,319-426 / finger_hand.cpp / antipodal.cpp), every record field bit-exact,
    poses to 1e-9 m.  R = 32 changes the shared-memory carve-up of k_hands (hands_smem_base), r = 0.06 the number of
    cell rows per query and r = 0.12 the staging chunks."""
    sc, gn = c3_with_normals
    sub = np.ascontiguousarray(sc["samples"][11::250])
    assert len(sub) == 40
    p = make_params(capi.MODE_GEOMETRIC, num_threads=16, num_orientations=R, nn_radius_hands=radius)
    g, o = capi.Context(p), ob.context(p)
    for c in (g, o):
        c.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"], normals=gn)
        c.set_sample_indices(sub)
        c.detect()
    gsw, osw = g.sweep_records(), o.sweep_records()
    assert gsw.shape == (40, R)
    for f in ("fingers", "hand", "slot", "depth_step", "num_in_box", "antipodal", "emitted"):
        assert np.array_equal(gsw[f], osw[f]), (f, R, radius)
    for which in (capi.LIST_HYPOTHESES, capi.LIST_PRUNED, capi.LIST_ANTIPODAL, capi.LIST_SELECTED):
        gh, oh = g.hypotheses(which), o.hypotheses(which)
        assert len(gh) == len(oh) and hyp_int_fields_equal(gh, oh) is None, which
        for f in POSE:
            assert np.allclose(gh[f], oh[f], rtol=0, atol=1e-9), f
    assert gsw["emitted"].sum() > 0
    # K grows with the radius: the statistics the bench's roofline uses are the oracle's
    gs, os_ = g.stats(), o.stats()
    for k in ("sum_neighbors_hands", "sum_cropped", "sum_in_box", "num_hypotheses"):
        assert gs[k] == os_[k], k
    g.close()


def test_c5_limits_of_the_orientation_and_depth_tables(tiny_scene=None):
    """The largest tables k_hands accepts: AG2_MAX_ORIENTATIONS = 64 orientations and a deepen list of 61 steps
    (hand_depth 0.31 from init_bite 0.01: shared memory base of ~40 KB per CTA).  Either it runs and matches the
    oracle bit for bit, or it refuses with AG2_E_CAPACITY -- never a silent truncation; one step more than
    AG2_MAX_DEEPEN_STEPS is rejected when the parameters are set."""
    sc = scenes.make_scene("tiny", num_samples=12)
    p = make_params(capi.MODE_GEOMETRIC, num_orientations=64, hand_depth=0.31, nn_radius_hands=0.12)
    g, o = both(p, sc)
    try:
        g.detect()
    except capi.Ag2Error as e:
        assert e.code == capi.AG2_E_CAPACITY
        return
    o.detect()
    gsw, osw = g.sweep_records(), o.sweep_records()
    for f in ("fingers", "hand", "slot", "depth_step", "num_in_box", "antipodal", "emitted"):
        assert np.array_equal(gsw[f], osw[f]), f
    with pytest.raises(capi.Ag2Error):
        capi.Context(make_params(capi.MODE_GEOMETRIC, hand_depth=0.33))
