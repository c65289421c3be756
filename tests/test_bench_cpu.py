"""bench.py's reference arm on CPU: the JSON line of the contract (keys, types, `impl`, zero-copy `e2e`), and that the
arm does not load libag2.so (the driver lists the shared objects the reference process mapped)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_the_contract_line_without_loading_libag2():
    code = ("import runpy, sys; sys.argv = ['bench.py', '--impl', 'reference', '--workload', 'c1', '--steps', '1', '--warmup', '1'];\n"
            "runpy.run_path(%r, run_name='__main__')\n"
            "maps = open('/proc/self/maps').read()\n"
            "print('MAPPED_LIBAG2', 'libag2.so' in maps)\n") % os.path.join(ROOT, "bench.py")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, r.stdout[-2000:]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "hand_hypotheses_per_sec" and d["unit"] == "hypotheses/s"
    assert d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 1 and d["higher_is_better"] is True
    assert d["vs_baseline"] is None and d["data"] == "synthetic" and d["config"]["name"] == "c1"
    assert d["value"] > 0 and abs(d["value"] - 800 / (d["ms_per_step"] * 1e-3)) < 1e-6 * d["value"]  # 100 samples x 8
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    e = d["e2e"]
    assert e["value"] == d["value"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
    assert "MAPPED_LIBAG2 False" in r.stdout


def test_roofline_entry_and_traffic_staleness():
    """roofline_for: the dominant phase decides the bound; a traffic.json entry is used only while the kernel sources it
    was captured from are unchanged (hash of the files), and is reported as stale otherwise."""
    sys.path.insert(0, ROOT)
    import bench
    st = {"sum_neighbors_hands": 1.0e8, "num_hypotheses": 8000, "sum_cropped": 3.0e7, "sum_in_box": 2.0e7,
          "sum_neighbors_normals": 1.1e7, "num_pruned": 4500, "num_samples": 10000}
    peaks = {"hbm_gbs": 6500.0, "bf16_tflops": 1400.0, "src": "test"}
    fp = {"fp32_tflops": 70.0, "fp64_tflops": 34.0, "sm_hz": 1.9e9}
    ph = {"grid": 0.1, "normals": 2.2, "frames": 0.3, "hands": 3.5, "render": 0.8, "lenet": 1.2}
    r = bench.roofline_for("c3", ph, st, 614400, 8, peaks, fp)
    assert r["kernel"] == "hands" and r["bound"] == "issue" and r["peak"] == 34.0
    ops = 13.0 * 1.0e8 + 58.0 * 8 * 3.0e7 + 27.0 * 2.0e7
    assert abs(r["achieved"] - ops / 3.5e-3 / 1e12) < 1e-9 and abs(r["frac"] - r["achieved"] / 34.0) < 1e-12
    assert r["hbm"]["peak"] == 6500.0 and r["traffic_capture"] is not None
    committed = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["c3"]["hands"]
    fresh = committed["sources_sha16"] == bench._sha(bench.PHASE_SOURCES["hands"])
    assert r["traffic_capture"]["stale"] == (not fresh)
    assert (r["traffic"] == committed["dram_bytes"]) if fresh else (r["traffic"] is None)
    r = bench.roofline_for("c3", dict(ph, lenet=9.0), st, 614400, 8, peaks, fp)
    assert r["bound"] == "tensor" and abs(r["achieved"] - 45.41e6 * 4500 / 9.0e-3 / 1e12) < 1e-9
    r = bench.roofline_for("no such workload", dict(ph, render=9.0), st, 614400, 8, peaks, fp)
    assert r["bound"] == "hbm" and r["traffic"] is None and r["traffic_capture"] is None
