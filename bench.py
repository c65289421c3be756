#!/usr/bin/env python
"""bench.py — hand hypotheses/s of the grasp-hypothesis hot path on N B200s.

Contract (driver): `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line.
A "step" is one pass of the hot path (neighbour index -> normals -> frames -> hand search
-> prune -> grasp images -> LeNet -> select) over one synthetic cloud.

  workload  c2 of BASELINE.json: one Kinect-resolution view, 307 200 points, 1 000 samples x 8
            orientations, antipodal_mode 1 (LeNet; seeded Xavier weights — the trained blob is
            absent from the reference checkout).
  value     evaluated (sample, orientation) pairs per second, cloud already resident in HBM.
  e2e       same through the public C ABI with HOST buffers: H2D of cloud + samples and D2H
            of the selected grasp records inside the timed region.
  N > 1     one process per GPU (torchrun); the cloud is broadcast once per step over NCCL,
            every rank evaluates its own block of 1 000 samples (weak scaling: per-GPU work
            fixed), scored grasps are all-gathered.
  --impl reference   times the CPU restatement (oracle/, the only runnable form of the
            reference here) with all host threads on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "hand_hypotheses_per_sec"
UNIT = "hypotheses/s"
R = 8
S_PER_GPU = 1000


def make_params(capi):
    p = capi.default_params()   # the reference's compiled-in defaults (ag2_default_params)
    p.init_bite = 0.01          # launch/file_detect_grasps.launch
    p.antipodal_mode = capi.MODE_PREDICTION
    p.min_score_diff = 0.0
    p.num_selected = 50
    p.num_orientations = R
    p.num_threads = os.cpu_count() or 1
    return p


def make_workload(scenes, n_gpus, name="c2"):
    if name == "c3":  # BASELINE.json configs[2]: fixed 10 000 samples, sharded over the GPUs (strong scaling)
        return scenes.make_scene("c3")
    sc = scenes.make_scene("c2", num_samples=S_PER_GPU * n_gpus)
    return sc


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index
        self.t_begin = self.t_end = None  # wall-clock window of the timed steps; samples outside are dropped

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [r for t, r in self.rows if self.t_begin is None or (self.t_begin <= t <= (self.t_end or t))]
        for r in (inside or [r for _, r in self.rows][-1:]):
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for k, nme in enumerate(names):
                    if r[2 + k].lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return {"hbm_gbs": d.get("hbm_gbs", 6650.0), "bf16_tflops": d.get("bf16_tflops_sustained", 1400.0),
                "src": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "src": "fallback"}


def cpu_reference(ob, capi, scenes, p, steps, warmup, full_S):
    """Times the oracle (kind 'port') on a bounded sample: the whole cloud, S_sub samples.
    tree + normals are per-cloud costs, frames..select per-sample; the full-workload time is
    t_cloud + t_sample * S / S_sub (stated in `sample`)."""
    sc = scenes.make_scene("c2", num_samples=full_S)
    S_sub = min(full_S, 200)
    sub = sc["samples"][:: max(1, full_S // S_sub)][:S_sub]
    w = ob.xavier_weights(1)
    o = ob.context(p)
    o.set_lenet_weights(w)
    times = []
    for it in range(warmup + steps):
        o.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"])
        o.set_sample_indices(sub)
        t0 = time.perf_counter()
        o.detect()
        dt = time.perf_counter() - t0
        ph = o.stats()["ms_phase"]
        t_cloud = (ph["grid"] + ph["normals"]) * 1e-3
        t_sample = max(dt - t_cloud, 0.0)
        if it >= warmup:
            times.append(t_cloud + t_sample * full_S / len(sub))
    t = float(np.mean(times)) if times else float("nan")
    return {"value": full_S * R / t, "unit": UNIT, "cores": int(p.num_threads), "kind": "port",
            "sample": "full 307200-point cloud (kd-tree + normals timed in full), %d of %d samples timed, "
                      "per-sample phases scaled x%.1f" % (len(sub), full_S, full_S / len(sub)),
            "ms_per_step": t * 1e3}


def roofline_for(phase_ms, st, n_points, peaks):
    """Roofline entry of the dominant kernel (largest mean phase time), from ALGORITHMIC work
     / live CUDA-event time of that phase."""
    dom = max(("normals", "hands", "render", "lenet", "grid", "frames"), key=lambda k: phase_ms.get(k, 0.0))
    t = phase_ms[dom] * 1e-3
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(dom)
        except Exception:
            traffic = None
    if dom == "lenet":
        work = 45.41e6 * st["num_pruned"]
        return {"kernel": "lenet (conv1+conv2+ip1+ip2)", "bound": "tensor", "achieved": work / t / 1e12,
                "peak": peaks["bf16_tflops"], "unit": "TFLOP/s", "frac": work / t / 1e12 / peaks["bf16_tflops"],
                "traffic": traffic, "peak_src": peaks["src"], "ms": phase_ms[dom]}
    sm_hz = 148 * 1.965e9
    issue = None
    if dom == "hands":
        work = 25.0 * st["sum_neighbors_hands"] + 192.0 * st["num_hypotheses"]
        # 13 ops per neighbour (distance + crop), 58 per cropped point and orientation,
        # 27 per in-box point of an emitted hand; FP64, nominal 64 FMA/clk/SM
        ops = 13.0 * st["sum_neighbors_hands"] + 58.0 * 8 * st["sum_cropped"] + 27.0 * st["sum_in_box"]
        issue = {"ops": ops, "achieved": ops / t / 1e12, "peak": 2 * 64 * sm_hz / 1e12, "unit": "TFLOP/s (fp64, nominal peak)"}
    elif dom == "normals":
        work = 24.0 * n_points
        # 26 flop per neighbour + ~120 per point (eigen33 + flip); FP32, 128 FMA/clk/SM
        ops = 26.0 * st["sum_neighbors_normals"] + 120.0 * n_points
        issue = {"ops": ops, "achieved": ops / t / 1e12, "peak": 2 * 128 * sm_hz / 1e12, "unit": "TFLOP/s (fp32, nominal peak)"}
        try:  # warp instructions of the phase (ncu, profiles/traffic.json) against 4 issue slots per SM and clock
            winst = json.load(open(tpath)).get("normals_inst")
        except Exception:
            winst = None
        if winst:
            issue["warp_inst"] = winst
            issue["issue_slot_frac"] = winst / t / (4 * sm_hz)
    elif dom == "grid":
        work = 50.0 * n_points
    elif dom == "render":
        work = 24.0 * st["sum_in_box"] + 10800.0 * st["num_pruned"]
    else:
        work = 25.0 * st["num_samples"] * 50
    out = {"kernel": dom, "bound": "hbm", "achieved": work / t / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
           "frac": work / t / 1e9 / peaks["hbm_gbs"], "traffic": traffic, "peak_src": peaks["src"],
           "ms": phase_ms[dom]}
    if issue:
        # the kernel is issue-bound, not HBM-bound (its gathers are L1/L2 hits): the compulsory-byte
        # fraction above is small by construction; this is the same kernel against the SM's FP pipe
        issue["frac"] = issue["achieved"] / issue["peak"]
        out["issue"] = issue
    if phase_ms.get("lenet"):
        tl = phase_ms["lenet"] * 1e-3
        wl = 45.41e6 * st["num_pruned"]
        out["lenet_tensor"] = {"bound": "tensor", "achieved": wl / tl / 1e12, "peak": peaks["bf16_tflops"],
                               "unit": "TFLOP/s", "frac": wl / tl / 1e12 / peaks["bf16_tflops"], "ms": phase_ms["lenet"],
                               "note": "algorithmic flop only; the three tf32 passes of the split are not counted"}
    return out


def run_reference(args):
    from agile_grasp2_b200 import capi, scenes
    from oracle import binding as ob
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    p = make_params(capi)
    full_S = S_PER_GPU * args.gpus
    cb = cpu_reference(ob, capi, scenes, p, max(1, args.steps), max(0, min(args.warmup, 1)), full_S)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args.gpus),
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(n_gpus, name="c2"):
    if name == "c3":
        return {"workload": "c3: synthetic two-view clutter cloud, 614400 points, 10000 samples x %d orientations in "
                            "total (strong scaling), antipodal_mode=1 (LeNet, seeded Xavier weights), num_selected=50" % R,
                "points": 614400, "samples_total": 10000, "orientations": R, "gpus": n_gpus,
                "l2": "flushed between timed steps (256 MiB write)", "parallelism": "sample-sharded x%d" % n_gpus}
    return {"workload": "c2: synthetic Kinect-resolution single view, 307200 points, %d samples x %d orientations "
                        "per GPU, antipodal_mode=1 (LeNet, seeded Xavier weights), num_selected=50" % (S_PER_GPU, R),
            "points": 307200, "samples_per_gpu": S_PER_GPU, "orientations": R, "gpus": n_gpus,
            "l2": "flushed between timed steps (256 MiB write)", "parallelism": "sample-sharded x%d" % n_gpus}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="c2", choices=["c2", "c3"],
                    help="c2 (default, the metric's configuration, weak scaling) or c3 (10 000 samples in total, strong scaling)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return
    import torch
    import torch.distributed as dist

    from agile_grasp2_b200 import capi, scenes, sharded

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libag2 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    W = max(3, args.warmup)
    K = max(1, args.steps)
    p = make_params(capi)
    sc = make_workload(scenes, world, args.workload) if rank == 0 or world == 1 else None
    ctx = capi.Context(p, device=local_rank)
    ctx.init_lenet_xavier(1)
    ext = torch.cuda.ExternalStream(ctx.stream(), device=local_rank)
    runner = sharded.ShardedDetector(ctx, rank=rank, world=world, device=torch.device("cuda", local_rank))
    if rank == 0 or world == 1:
        xyz_h = torch.from_numpy(sc["xyz"]).pin_memory()
        cam_h = torch.from_numpy(sc["cam_mask"]).pin_memory()
        smp_h = torch.from_numpy(sc["samples"]).pin_memory()
        num_cams, n_pts, n_smp = sc["num_cams"], len(sc["xyz"]), len(sc["samples"])
    else:
        xyz_h = cam_h = smp_h = None
        num_cams = n_pts = n_smp = 0
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    gaps = {}

    def timed(fn, iters, warm, clock_window=None):
        phase = {}
        gaps.clear()
        total_ms, launches, last_stats = 0.0, 0, None
        for it in range(warm + iters):
            if it == warm and clock_window is not None:
                clock_window.t_begin = time.time()
            flush.zero_()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(ext)
            t0 = time.perf_counter()
            out = fn()
            e1.record(ext)
            torch.cuda.synchronize()
            wall = (time.perf_counter() - t0) * 1e3
            ms = max(e0.elapsed_time(e1), wall if world > 1 else 0.0)
            if it >= warm:
                if clock_window is not None:
                    clock_window.t_end = time.time()
                st = ctx.stats()
                total_ms += ms
                launches += st["gpu_launches"]
                for k, v in st["ms_phase"].items():
                    phase[k] = phase.get(k, 0.0) + v / iters
                for k, v in st.get("ms_gap", {}).items():
                    gaps[k] = gaps.get(k, 0.0) + v / iters
                last_stats = st
        return total_ms, phase, launches, last_stats, out

    clocks = ClockSampler(local_rank)  # 20 ms period; started early so that nvidia-smi is up when the timed steps run
    clocks.start()
    # resident arm: the cloud sits in HBM on rank 0 before the timed region
    # sizes of the synthetic workload are the same on every rank (one tiny broadcast at set-up)
    meta = torch.tensor([n_pts, n_smp, num_cams], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.broadcast(meta, src=0)
    hint = tuple(int(v) for v in meta.tolist())
    # resident layout: one allocation (xyz | samples | camera mask), so that the per-call exchange is one broadcast
    if rank == 0 or world == 1:
        packed_d = sharded.pack_inputs(xyz_h, cam_h, smp_h, torch.device("cuda", local_rank))[0]
    else:
        packed_d = torch.empty(12 * hint[0] + 4 * hint[1] + hint[0], dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    tot_ms, phase, launches, st, _ = timed(lambda: runner.detect_packed(packed_d, hint[0], hint[1], hint[2]), K, W, clocks)
    clk = clocks.stop()
    gaps_resident = dict(gaps)
    e2e_ms, _, _, _, sel = timed(lambda: runner.detect_host(xyz_h, cam_h, num_cams, smp_h, hint), K, 1)
    per_rank = None
    if world > 1:  # per-rank view of the step: sample blocks differ in cost, the slowest rank sets the pace
        mine = {"rank": rank, "ms_per_step": round(tot_ms / K, 4), "phases_ms": {k: round(v, 4) for k, v in phase.items()},
                "last_call_host_ms": {k: round(v, 4) for k, v in getattr(runner, "last_ms", {}).items()}}
        per_rank = [None] * world
        dist.all_gather_object(per_rank, mine)
    t = torch.tensor([tot_ms, e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tot_ms, e2e_ms = float(t[0]), float(t[1])
    if rank == 0:
        pairs = float(hint[1] * R)  # all samples of the call x orientations
        peaks = measured_peaks()
        line = {"metric": METRIC, "value": pairs * K / (tot_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
                "warmup": W, "ms_per_step": tot_ms / K, "higher_is_better": True,
                "scaling": "strong" if args.workload == "c3" else "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(world, args.workload),
                "clocks": clk, "gpu_launches": int(launches),
                "e2e": {"value": pairs * K / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms / K,
                        "h2d_bytes_per_step": int(n_pts * 13 + n_smp * 4),
                        "d2h_bytes_per_step": int(len(sel) * 192 + 64)},
                "roofline": roofline_for(phase, st, n_pts, peaks),
                "phases_ms": {k: round(v, 4) for k, v in phase.items()},
                "gaps_before_phase_ms": {k: round(v, 4) for k, v in gaps_resident.items()},
                **({"per_rank": per_rank} if per_rank else {}),
                "counts": {k: st[k] for k in ("num_hypotheses", "num_pruned", "num_antipodal", "num_selected",
                                              "sum_neighbors_hands", "sum_cropped", "sum_in_box",
                                              "sum_neighbors_normals")}}
        if world == 1 and not args.no_cpu_baseline and args.workload == "c2":
            from oracle import binding as ob  # cpu_baseline leg: the oracle is the thing timed here, nothing else
            cb = cpu_reference(ob, capi, scenes, p, 1, 1, S_PER_GPU)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
