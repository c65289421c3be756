#!/usr/bin/env python
"""Builds profiles/traffic.json[workload] from an ncu launch list of profiles/one_step.py
   ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum
       --clock-control none --csv --log-file L.csv python profiles/one_step.py <workload> <steps>
Kernels are attributed to a phase by the source file that defines them; per phase the entry holds, PER STEP,
dram_bytes (read + write), warp_inst, the summed launch durations, the commit and a hash of the phase's sources
(bench.py drops an entry whose hash no longer matches the sources it runs).
usage: make_traffic.py L.csv workload steps [commit]"""
import collections, csv, json, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

path, workload, steps = sys.argv[1], sys.argv[2], int(sys.argv[3])
commit = sys.argv[4] if len(sys.argv) > 4 else subprocess.run(
    ["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
CS = os.path.join(ROOT, "agile_grasp2_b200", "csrc")
owner = {}
for phase, files in bench.PHASE_SOURCES.items():
    for f in files:
        if f == "dev_common.cuh" or f == "tc_ptx.cuh":
            continue
        src = open(os.path.join(CS, f)).read()
        for m in re.finditer(r"__global__", src):
            k = re.search(r"\b(k_\w+)\s*\(", src[m.end():m.end() + 300])
            if k:
                owner[k.group(1)] = phase
rows = list(csv.reader(open(path)))
i = [k for k, r in enumerate(rows) if r and r[0] == "ID"][0]
h = rows[i]
acc = collections.defaultdict(lambda: collections.Counter())
kernels = collections.defaultdict(set)
for r in rows[i + 1:]:
    d = dict(zip(h, r))
    name = d["Kernel Name"].replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
    name = re.split(r"[<(]", name)[0].split("::")[-1].split()[-1]
    ph = owner.get(name)
    if ph is None:
        continue
    v = float(d["Metric Value"].replace(",", ""))
    unit = d.get("Metric Unit", "")
    m = d["Metric Name"]
    if m.startswith("dram__bytes"):
        v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
        acc[ph]["dram_bytes"] += v
    elif m == "smsp__inst_executed.sum":
        acc[ph]["warp_inst"] += v
    elif m == "gpu__time_duration.sum":
        v *= {"ns": 1e-3, "us": 1, "ms": 1e3, "usecond": 1, "nsecond": 1e-3, "msecond": 1e3}.get(unit, 1e-3)
        acc[ph]["us_under_ncu"] += v
        acc[ph]["launches"] += 1
    kernels[ph].add(name)
out_path = os.path.join(ROOT, "profiles", "traffic.json")
try:
    cur = json.load(open(out_path))
    if not all(isinstance(v, dict) for v in cur.values()):
        cur = {}
except Exception:
    cur = {}
ent = {}
for ph, c in acc.items():
    ent[ph] = {"dram_bytes": int(c["dram_bytes"] / steps), "warp_inst": int(c["warp_inst"] / steps),
               "us_under_ncu": round(c["us_under_ncu"] / steps, 1), "launches": c["launches"] / steps,
               "kernels": sorted(kernels[ph]), "commit": commit, "sources_sha16": bench._sha(bench.PHASE_SOURCES[ph]),
               "capture": os.path.basename(path)}
cur[workload] = ent
json.dump(cur, open(out_path, "w"), indent=1, sort_keys=True)
print(json.dumps(ent, indent=1))
