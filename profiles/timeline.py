#!/usr/bin/env python
"""Device timeline of one step through the C ABI (nsys is not in this image: torch.profiler's CUPTI trace sees the
kernels, copies and synchronisations of libag2.so as well).  Prints every GPU activity of the LAST step with its stream,
start, duration and the idle gap before it, then the host-side synchronisation calls, then a summary.
usage: timeline.py [workload = c3] [world = 1] [steps = 3]     (world > 1: the work of rank 0 of `world`, as shard_step.py)"""
import json, os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench
from agile_grasp2_b200 import capi, scenes

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
world = int(sys.argv[2]) if len(sys.argv) > 2 else 1
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
wl = bench.WORKLOADS[name]


class A:
    orientations, radius = 0, 0.0


sc = bench.load_workload(scenes, name)
g = capi.Context(bench.set_params(capi.default_params(), wl, A))
if wl["mode"] == capi.MODE_PREDICTION:
    g.init_lenet_xavier(7)
samples = np.ascontiguousarray(sc["samples"][::world])
buf = None


def step():
    global buf
    g.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"])
    if world > 1:
        cap = g.normals_shard_points(world)
        if buf is None:
            buf = torch.empty((cap, 4), dtype=torch.float32, device="cuda:0")
        g.compute_normals_shard(0, world, buf.data_ptr(), cap)
        torch.cuda.synchronize()
        g.compute_normals()  # the other ranks' shares (not part of a rank's step)
        torch.cuda.synchronize()
    g.set_sample_indices(samples)
    g.detect()
    g.synchronize()


for _ in range(steps - 1):
    step()
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA, torch.profiler.ProfilerActivity.CPU]) as prof:
    step()
path = os.path.join(tempfile.mkdtemp(), "trace.json")
prof.export_chrome_trace(path)
ev = json.load(open(path))["traceEvents"]
gpu = [e for e in ev if e.get("ph") == "X" and e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
gpu.sort(key=lambda e: e["ts"])
t0 = gpu[0]["ts"]
end = t0
busy = 0.0
print("%-10s %-6s %10s %9s %8s  %s" % ("kind", "stream", "start_us", "dur_us", "gap_us", "name"))
for e in gpu:
    gap = e["ts"] - end
    nm = e["name"].replace("ag2::(anonymous namespace)::", "").split("(")[0][:60]
    print("%-10s %-6s %10.1f %9.1f %8.1f  %s" % (e["cat"], e["args"].get("stream", "?"), e["ts"] - t0, e["dur"], max(gap, 0.0), nm))
    if e["ts"] + e["dur"] > end:
        busy += e["ts"] + e["dur"] - max(end, e["ts"])
        end = e["ts"] + e["dur"]
sync = [e for e in ev if e.get("ph") == "X" and e.get("cat") == "cuda_runtime" and
        ("Synchronize" in e["name"] or e["name"] in ("cudaMemcpy",))]
print("\nhost-side synchronisations (start relative to the first GPU activity):")
for e in sorted(sync, key=lambda e: e["ts"]):
    print("  %10.1f us  %8.1f us  %s" % (e["ts"] - t0, e["dur"], e["name"]))
print("\nspan %.1f us, GPU busy (union over streams) %.1f us, idle %.1f us, %d activities, %d synchronisations" %
      (end - t0, busy, end - t0 - busy, len(gpu), len(sync)))
