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
