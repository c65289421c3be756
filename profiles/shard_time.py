#!/usr/bin/env python
"""Wall time of ag2_compute_normals_shard(rank 0 of `world`) on one GPU: the per-rank cost of the sharded normals at
8 GPUs without needing 8 GPUs.  usage: shard_time.py [config=c2|c3] [world=8] [quad_min ...]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from agile_grasp2_b200 import capi, scenes
from common import make_params, load_scene

cfg = sys.argv[1] if len(sys.argv) > 1 else "c2"
worlds = [int(sys.argv[2])] if len(sys.argv) > 2 else [1, 8]
qmins = [int(v) for v in sys.argv[3:]] or [8192]
sc = scenes.make_scene(cfg)
p = make_params(capi.MODE_GEOMETRIC)
for qmin in qmins:
    os.environ["AG2_NORMALS_QUAD_MIN"] = str(qmin)
    with capi.Context(p) as g:
        load_scene(g, sc)
        for world in worlds:
            cap = g.normals_shard_points(world)
            out = torch.empty(cap * 4, dtype=torch.float32, device="cuda")
            ts = []
            for it in range(6):
                torch.cuda.synchronize(); t0 = time.perf_counter()
                g.compute_normals_shard(0, world, out.data_ptr(), cap)
                torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
            print(cfg, "quad_min", qmin, "world", world, "ms %.3f" % min(ts[1:]))
