#!/usr/bin/env python
"""bench.py — hand hypotheses/s of the grasp-hypothesis hot path on N B200s.

Contract (driver): `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line.
A "step" is one pass of the hot path (neighbour index -> normals -> frames -> hand search
-> prune -> grasp images -> LeNet -> select) over one synthetic cloud, through the C ABI's
multi-GPU entry point (ag2_group_detect / ag2_group_detect_importance; one rank at N = 1).

  workload  default c3 of BASELINE.json (the configuration it labels "sharded 1/2/4/8 B200"): two-view
            clutter cloud, 614 400 points, 10 000 samples x 8 orientations IN TOTAL at every N (strong
            scaling), antipodal_mode 1 (LeNet; seeded Xavier weights -- the trained blob is absent from
            the reference checkout).  --workload c1 | c2 | c4 | c5 select the other named configurations.
  value     evaluated (sample, orientation) pairs per second, cloud already resident in HBM on rank 0.
  e2e       same with HOST buffers (pinned): H2D of cloud + samples and D2H of the result window
            inside the timed region.
  N > 1     one process per GPU (torchrun).  torch.distributed only hands the group id to the ranks and
            gathers the report; the data path is libag2's (csrc/group.cu: NVLink peer loads / stores).
            Outside the timed region rank 0 runs the same call on one GPU alone and the line carries
            "sharded_equals_single" (bytes of the merged selection).
  --impl reference   times the CPU restatement (oracle/, the only runnable form of the reference
            here) with all host threads on the same workload; does not load libag2.so.
"""
import argparse
import hashlib
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
MODE_PREDICTION, MODE_GEOMETRIC = 1, 2  # include/ag2.h AG2_MODE_* (kept here so that the reference arm needs no capi call)

# BASELINE.json configs[0..4]
WORKLOADS = {
    "c1": dict(scene="c1", mode=MODE_GEOMETRIC, R=8, radius=0.10,
               text="c1: synthetic tabletop cloud (box + cylinder), 50000 points, 100 samples x 8 orientations, "
                    "antipodal_mode=2 (geometric), num_selected=50"),
    "c2": dict(scene="c2", mode=MODE_PREDICTION, R=8, radius=0.10,
               text="c2: synthetic Kinect-resolution single view, 307200 points, 1000 samples x 8 orientations in total, "
                    "antipodal_mode=1 (LeNet, seeded Xavier weights), num_selected=50"),
    "c3": dict(scene="c3", mode=MODE_PREDICTION, R=8, radius=0.10,
               text="c3: synthetic two-view clutter cloud, 614400 points, 10000 samples x 8 orientations in total, "
                    "antipodal_mode=1 (LeNet, seeded Xavier weights), num_selected=50"),
    "c4": dict(scene="c4", mode=MODE_PREDICTION, R=8, radius=0.10, importance=True,
               text="c4: synthetic 16-view fused scene, ~4.9M points, importance sampling: 50000 initial samples + 5 rounds "
                    "x 10000 (70% Gaussian, 30% random) = 100000 samples x 8 orientations, antipodal_mode=1, num_selected=50 per round"),
    "c5": dict(scene="c3", mode=MODE_PREDICTION, R=32, radius=0.12,
               text="c5: c3 cloud, 10000 samples, hand-search scaling study (num_orientations and nn_radius_hands as given), "
                    "antipodal_mode=1, num_selected=50"),
}
C4_INIT, C4_ROUNDS, C4_PER_ROUND = 50000, 5, 10000


def set_params(p, wl, args):
    p.init_bite = 0.01          # launch/file_detect_grasps.launch
    p.antipodal_mode = wl["mode"]
    p.min_score_diff = 0.0
    p.num_selected = 50
    p.num_orientations = args.orientations or wl["R"]
    p.nn_radius_hands = args.radius or wl["radius"]
    p.num_threads = os.cpu_count() or 1
    return p


def make_is_params(ip):
    ip.num_iterations, ip.num_samples = C4_ROUNDS, C4_PER_ROUND
    ip.method, ip.prob_rand_samples, ip.radius, ip.seed = 2, 0.3, 0.02, 4
    return ip


def load_workload(scenes, name):
    sc = scenes.make_scene(WORKLOADS[name]["scene"])
    if name == "c4":
        sc = dict(sc)
        sc["samples"] = np.ascontiguousarray(sc["samples"][:: len(sc["samples"]) // C4_INIT][:C4_INIT])
    return sc


def total_samples(name, n_samples):
    return C4_INIT + C4_ROUNDS * C4_PER_ROUND if name == "c4" else n_samples


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
                "src": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "src": "fallback (B200_PROFILING.md)"}


def cpu_reference(name, args, steps, warmup):
    """Times the oracle (kind 'port': the reference itself cannot be built here) with all host
    threads on the workload.  c1 / c2 / c3 / c5 are timed in full; c4 (minutes of CPU work) on a bounded sample:
    the whole cloud for kd-tree + normals, 1/20 of the samples of every round, per-sample phases scaled."""
    from agile_grasp2_b200 import scenes
    from oracle import binding as ob
    wl = WORKLOADS[name]
    p = set_params(ob.reference_default_params(), wl, args)
    sc = load_workload(scenes, name)
    R = int(p.num_orientations)
    o = ob.context(p)
    if wl["mode"] == MODE_PREDICTION:
        o.set_lenet_weights(ob.xavier_weights(1))
    scale, sample_txt = 1.0, "the whole workload, every sample timed"
    samples = sc["samples"]
    ip = None
    if name == "c4":
        scale = 20.0
        samples = np.ascontiguousarray(samples[::20])
        ip = make_is_params(o.default_is_params())
        ip.num_samples = C4_PER_ROUND // 20
        sample_txt = ("whole %d-point cloud (kd-tree + normals timed in full); %d of %d initial samples and %d of %d per "
                      "round timed, per-sample phases scaled x20" % (len(sc["xyz"]), len(samples), C4_INIT, ip.num_samples,
                                                                     C4_PER_ROUND))
    times = []
    for it in range(warmup + steps):
        o.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"])
        o.set_sample_indices(samples)
        t0 = time.perf_counter()
        if ip is not None:
            o.detect_importance(ip)
        else:
            o.detect()
        dt = time.perf_counter() - t0
        if scale != 1.0:
            ph = o.stats()["ms_phase"]
            t_cloud = (ph["grid"] + ph["normals"]) * 1e-3
            dt = t_cloud + max(dt - t_cloud, 0.0) * scale
        if it >= warmup:
            times.append(dt)
    t = float(np.mean(times))
    S = total_samples(name, len(sc["samples"]))
    return {"value": S * R / t, "unit": UNIT, "cores": int(p.num_threads), "kind": "port", "sample": sample_txt,
            "ms_per_step": t * 1e3, "points": len(sc["xyz"]), "samples_total": S, "orientations": R}


def _sha(paths):
    h = hashlib.sha256()
    for p in paths:
        with open(os.path.join(ROOT, "agile_grasp2_b200", "csrc", p), "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


PHASE_SOURCES = {"normals": ["normals.cu", "dev_common.cuh"], "hands": ["hands.cu", "hands_kernel.cuh", "dev_common.cuh"],
                 "lenet": ["lenet.cu", "tc_ptx.cuh"], "render": ["render.cu"], "grid": ["grid.cu"], "frames": ["frames.cu"]}


def ncu_counts(workload, phase):
    """DRAM bytes / warp instructions of one step of `phase` from the committed ncu capture
    (profiles/traffic.json), or None when the capture is older than the kernel's source files."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        ent = json.load(open(path))[workload][phase]
    except Exception:
        return None
    ent = dict(ent)
    ent["stale"] = ent.get("sources_sha16") != _sha(PHASE_SOURCES[phase])
    return ent


def roofline_for(workload, phase_ms, st, n_points, R, peaks, fp):
    """Roofline entry of the dominant kernel (largest mean phase time), from ALGORITHMIC work
     / live CUDA-event time of that phase on the ctx stream."""
    dom = max(("normals", "hands", "render", "lenet", "grid", "frames"), key=lambda k: phase_ms.get(k, 0.0))
    t = phase_ms[dom] * 1e-3
    cnt = ncu_counts(workload, dom)
    traffic = None if (cnt is None or cnt["stale"]) else cnt.get("dram_bytes")
    extra = {"peak_src": peaks["src"], "ms": phase_ms[dom],
             "traffic_capture": None if cnt is None else {k: cnt.get(k) for k in ("commit", "sources_sha16", "stale")}}
    if dom == "lenet":
        work = 45.41e6 * st["num_pruned"]
        return {"kernel": "lenet (k_conv1_tc + k_conv2_h + k_ip1_tc + k_ip2)", "bound": "tensor", "achieved": work / t / 1e12,
                "peak": peaks["bf16_tflops"], "unit": "TFLOP/s", "frac": work / t / 1e12 / peaks["bf16_tflops"],
                "traffic": traffic, **extra}
    if dom == "hands":
        hbm_bytes = 25.0 * st["sum_neighbors_hands"] + 192.0 * st["num_hypotheses"]
        # 13 ops per neighbour (distance + crop), 58 per cropped point and orientation,
        # 27 per in-box point of an emitted hand; FP64
        ops, peak, what = (13.0 * st["sum_neighbors_hands"] + 58.0 * R * st["sum_cropped"] + 27.0 * st["sum_in_box"],
                           fp["fp64_tflops"], "fp64")
    elif dom == "normals":
        hbm_bytes = 24.0 * n_points
        # 26 flop per neighbour + ~120 per point (eigen33 + flip); FP32
        ops, peak, what = 26.0 * st["sum_neighbors_normals"] + 120.0 * n_points, fp["fp32_tflops"], "fp32"
    else:
        hbm_bytes = {"grid": 50.0 * n_points, "render": 24.0 * st["sum_in_box"] + 10800.0 * st["num_pruned"],
                     "frames": 25.0 * st["num_samples"] * 50}[dom]
        return {"kernel": dom, "bound": "hbm", "achieved": hbm_bytes / t / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": hbm_bytes / t / 1e9 / peaks["hbm_gbs"], "traffic": traffic, **extra}
    # normals and hands are bound by the SM's issue rate (their gathers are L1/L2 hits: a few MB of DRAM traffic):
    # algorithmic flop / time against the FMA peak measured in this process (ag2_measure_fp_peaks)
    out = {"kernel": dom, "bound": "issue", "achieved": ops / t / 1e12, "peak": peak,
           "unit": "TFLOP/s (%s, peak = same-run FMA microbenchmark)" % what, "frac": ops / t / 1e12 / peak,
           "traffic": traffic, "ops": ops,
           "hbm": {"achieved": hbm_bytes / t / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                   "frac": hbm_bytes / t / 1e9 / peaks["hbm_gbs"], "note": "compulsory bytes only; not the bound"},
           **extra}
    if cnt is not None and not cnt["stale"] and cnt.get("warp_inst"):
        out["warp_inst"] = cnt["warp_inst"]
        out["issue_slot_frac"] = cnt["warp_inst"] / t / (4 * 148 * fp["sm_hz"])
    return out


def workload_config(name, n_gpus, n_points, n_samples, R, radius):
    return {"workload": WORKLOADS[name]["text"], "name": name, "points": n_points, "samples_total": n_samples,
            "orientations": R, "nn_radius_hands": radius, "gpus": n_gpus,
            "l2": "flushed between timed steps (256 MiB write)", "parallelism": "sample-sharded x%d" % n_gpus}


def run_reference(args):
    """Reference arm: the CPU path on the box's host cores (oracle port; libag2.so is not loaded)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = args.workload
    cb = cpu_reference(name, args, max(1, args.steps), 1 if args.warmup > 0 else 0)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(name, args.gpus, cb["points"], cb["samples_total"], cb["orientations"],
                                      args.radius or WORKLOADS[name]["radius"]),
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS),
                    help="c3 (default: the configuration BASELINE.json labels 1/2/4/8 GPUs) or another named configuration")
    ap.add_argument("--orientations", type=int, default=0, help="c5: num_orientations (8..32)")
    ap.add_argument("--radius", type=float, default=0.0, help="c5: nn_radius_hands (0.06..0.12)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return
    import torch
    import torch.distributed as dist

    from agile_grasp2_b200 import capi, scenes

    name = args.workload
    wl = WORKLOADS[name]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libag2 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W = max(3, args.warmup)
    K = max(1, args.steps)
    p = set_params(capi.default_params(), wl, args)  # the reference's compiled-in defaults (ag2_default_params)
    R = int(p.num_orientations)
    # the group id travels through torch.distributed; everything after that is libag2's
    uid_t = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0 and world > 1:
        uid_t.copy_(torch.frombuffer(bytearray(capi.group_unique_id()), dtype=torch.uint8))
    if world > 1:
        dist.broadcast(uid_t, src=0)
    grp = capi.Group(p, uid=bytes(uid_t.cpu().numpy().tobytes()), rank=rank, world=world, device=local_rank)
    if wl["mode"] == MODE_PREDICTION:
        grp.init_lenet_xavier(1)
    ctx = grp.local_ctx(0)
    ext = torch.cuda.ExternalStream(ctx.stream(), device=local_rank)
    ip = make_is_params(ctx.default_is_params()) if wl.get("importance") else None
    meta = torch.zeros(3, dtype=torch.int64, device=dev)
    xyz_h = cam_h = smp_h = xyz_d = cam_d = smp_d = None
    if rank == 0:
        sc = load_workload(scenes, name)
        xyz_h = torch.from_numpy(sc["xyz"]).pin_memory()
        cam_h = torch.from_numpy(sc["cam_mask"]).pin_memory()
        smp_h = torch.from_numpy(np.ascontiguousarray(sc["samples"], dtype=np.int32)).pin_memory()
        xyz_d, cam_d, smp_d = xyz_h.to(dev), cam_h.to(dev), smp_h.to(dev)  # the resident arm's inputs
        meta[0], meta[1], meta[2] = len(sc["xyz"]), len(sc["samples"]), sc["num_cams"]
    if world > 1:
        dist.broadcast(meta, src=0)
    n_pts, n_smp, num_cams = (int(v) for v in meta.tolist())
    S_total = total_samples(name, n_smp)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def ptrs(t3):
        return tuple(0 if t is None else t.data_ptr() for t in t3)

    def call(t3, on_device):
        a, b, c_ = ptrs(t3)
        if ip is not None:
            return grp.detect_importance_ptr(a, n_pts, b, num_cams, c_, n_smp, on_device, ip)
        return grp.detect_ptr(a, n_pts, b, num_cams, c_, n_smp, on_device)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    gaps = {}
    segs = {}

    def timed(fn, iters, warm, clock_window=None):
        phase = {}
        gaps.clear()
        segs.clear()
        total_ms, launches, last_stats, out = 0.0, 0, None, None
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
            # the call returns with the results on the host: the device interval and the host interval agree;
            # with several ranks the host interval also covers the waits for the other ranks
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
                if ip is None:
                    for k, v in grp.last_timing().items():
                        segs[k] = segs.get(k, 0.0) + v / iters
                last_stats = st
        return total_ms, phase, launches, last_stats, out

    clocks = ClockSampler(local_rank)  # 20 ms period; started early so that nvidia-smi is up when the timed steps run
    clocks.start()
    fp32, fp64 = ctx.measure_fp_peaks()
    tot_ms, phase, launches, st, sel = timed(lambda: call((xyz_d, cam_d, smp_d), True), K, W, clocks)
    clk = clocks.stop()
    gaps_resident = dict(gaps)
    segs_resident = {k: round(v, 4) for k, v in segs.items()}
    e2e_ms, _, _, _, _ = timed(lambda: call((xyz_h, cam_h, smp_h), False), K, 1)
    io_h2d, io_d2h = grp.io_bytes()
    # ---- result check (outside the timed region): the sharded call against one GPU alone, byte for byte
    equal = None
    if world > 1:
        if rank == 0:
            with capi.Context(p, device=local_rank) as one:
                if wl["mode"] == MODE_PREDICTION:
                    one.init_lenet_xavier(1)
                one.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"])
                one.set_sample_indices(sc["samples"])
                if ip is not None:
                    one.detect_importance(ip)
                    ref_sel = one.hypotheses(capi.LIST_IMPORTANCE)
                else:
                    one.detect()
                    ref_sel = one.hypotheses(capi.LIST_SELECTED)
            equal = bool(len(ref_sel) == len(sel) and np.array_equal(ref_sel.view(np.uint8), sel.view(np.uint8)))
        barrier()
    per_rank = None
    if world > 1:  # per-rank view of the step: sample shards differ in cost, the slowest rank sets the pace
        mine = {"rank": rank, "ms_per_step": round(tot_ms / K, 4), "phases_ms": {k: round(v, 4) for k, v in phase.items()},
                "segments_ms": segs_resident}
        per_rank = [None] * world
        dist.all_gather_object(per_rank, mine)
    t = torch.tensor([tot_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tot_ms, e2e_ms = float(t[0]), float(t[1])
    if rank == 0:
        pairs = float(S_total * R)  # all samples of the call x orientations
        peaks = measured_peaks()
        fp = {"fp32_tflops": fp32, "fp64_tflops": fp64, "sm_hz": (clk.get("sm_mhz") or 1965.0) * 1e6}
        line = {"metric": METRIC, "value": pairs * K / (tot_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
                "warmup": W, "ms_per_step": tot_ms / K, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": workload_config(name, world, n_pts, S_total, R, float(p.nn_radius_hands)),
                "clocks": clk, "gpu_launches": int(launches),
                "e2e": {"value": pairs * K / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms / K,
                        "h2d_bytes_per_step": int(io_h2d), "d2h_bytes_per_step": int(io_d2h)},
                "roofline": roofline_for(name, phase, st, n_pts, R, peaks, fp),
                "fp_peaks_measured": {"fp32_tflops": round(fp32, 2), "fp64_tflops": round(fp64, 2),
                                      "how": "ag2_measure_fp_peaks: 8 FMA chains x 256 threads x 8 CTAs per SM, best of 3, this process"},
                "phases_ms": {k: round(v, 4) for k, v in phase.items()},
                "gaps_before_phase_ms": {k: round(v, 4) for k, v in gaps_resident.items()},
                "segments_ms": segs_resident,
                "transport": ("peer memory (NVLink loads/stores, device-side flags)" if grp.peer_stores() else "nccl") if world > 1 else "single GPU",
                **({"sharded_equals_single": equal} if equal is not None else {}),
                **({"per_rank": per_rank} if per_rank else {}),
                "counts": {k: st[k] for k in ("num_hypotheses", "num_pruned", "num_antipodal", "num_selected",
                                              "sum_neighbors_hands", "sum_cropped", "sum_in_box",
                                              "sum_neighbors_normals")}}
        if phase.get("lenet"):
            tl, wlf = phase["lenet"] * 1e-3, 45.41e6 * st["num_pruned"]
            line["lenet_tensor"] = {"bound": "tensor", "achieved": wlf / tl / 1e12, "peak": peaks["bf16_tflops"],
                                    "unit": "TFLOP/s", "frac": wlf / tl / 1e12 / peaks["bf16_tflops"], "ms": phase["lenet"],
                                    "note": "algorithmic flop only; the extra MMAs of the hi / lo operand split are not counted"}
        if world == 1 and not args.no_cpu_baseline:
            # cpu_baseline leg: the oracle is the thing timed here, nothing else
            cb = cpu_reference(name, args, 1, 1 if name in ("c1", "c2") else 0)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line))
    grp.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
