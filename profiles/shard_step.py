#!/usr/bin/env python
"""Profiling driver: the work of ONE rank of an N-rank group on one GPU, without the other ranks -- its 1/world of the
normals (ag2_compute_normals_shard), then the whole detect call on its strided 1/world of the samples (the normals of the
other ranks are computed by this GPU as well, outside the part an ncu filter looks at).
usage: shard_step.py [workload = c3] [world = 8] [steps = 2]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench
from agile_grasp2_b200 import capi, scenes

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
world = int(sys.argv[2]) if len(sys.argv) > 2 else 8
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
wl = bench.WORKLOADS[name]


class A:
    orientations, radius = 0, 0.0


sc = bench.load_workload(scenes, name)
g = capi.Context(bench.set_params(capi.default_params(), wl, A))
g.init_lenet_xavier(7)
for _ in range(steps):
    g.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"])
    cap = g.normals_shard_points(world)
    buf = torch.empty((cap, 4), dtype=torch.float32, device="cuda:0")
    torch.cuda.nvtx.range_push("normals_shard")
    g.compute_normals_shard(0, world, buf.data_ptr(), cap)
    torch.cuda.nvtx.range_pop()
    g.compute_normals()  # the rest, so that detect finds complete normals
    g.set_sample_indices(np.ascontiguousarray(sc["samples"][::world]))
    torch.cuda.nvtx.range_push("detect_shard")
    g.detect()
    torch.cuda.nvtx.range_pop()
    g.synchronize()
st = g.stats()
print(name, "world", world, "hypotheses", st["num_hypotheses"], "pruned", st["num_pruned"], st["ms_phase"])
