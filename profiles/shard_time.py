import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from agile_grasp2_b200 import capi, scenes
from common import make_params, load_scene
for cfg in ("c2", "c3"):
    sc = scenes.make_scene(cfg)
    p = make_params(capi.MODE_GEOMETRIC)
    for qmin in (0, 4096, 8192, 16384, 1 << 30):
        os.environ["AG2_NORMALS_QUAD_MIN"] = str(qmin)
        with capi.Context(p) as g:
            load_scene(g, sc)
            for world in (1, 8):
                cap = g.normals_shard_points(world)
                out = torch.empty(cap * 4, dtype=torch.float32, device="cuda")
                ts = []
                for it in range(6):
                    torch.cuda.synchronize(); t0 = time.perf_counter()
                    g.compute_normals_shard(0, world, out.data_ptr(), cap)
                    torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
                print(cfg, "quad_min", qmin, "world", world, "ms %.3f" % min(ts[1:]))
