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
