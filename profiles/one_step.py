#!/usr/bin/env python
"""Profiling driver: `steps` passes of the hot path of one named workload through the C ABI on cuda:0, with no timing
and no baseline work around them, so that an ncu launch list holds exactly steps x (launches of one step).
usage: one_step.py [workload = c3] [steps = 2]      (c1 | c2 | c3 | c5; c4 runs the importance-sampling rounds)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench
from agile_grasp2_b200 import capi, scenes

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
wl = bench.WORKLOADS[name]


class A:
    orientations, radius = 0, 0.0


sc = bench.load_workload(scenes, name)
p = bench.set_params(capi.default_params(), wl, A)
g = capi.Context(p)
if wl["mode"] == capi.MODE_PREDICTION:
    g.init_lenet_xavier(7)
for _ in range(steps):
    g.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"])
    g.set_sample_indices(np.ascontiguousarray(sc["samples"]))
    if wl.get("importance"):
        g.detect_importance(bench.make_is_params(g.default_is_params()))
    else:
        g.detect()
    g.synchronize()
st = g.stats()
print(name, "steps", steps, "hypotheses", st["num_hypotheses"], "pruned", st["num_pruned"], "selected", st["num_selected"])
