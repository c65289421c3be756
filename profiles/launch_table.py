#!/usr/bin/env python
"""Per-step kernel table from an ncu launch list (--metrics gpu__time_duration.sum,smsp__inst_executed.sum --csv).
usage: launch_table.py launches.csv [kernel_launched_once_per_step]   (default k_hands: it fixes the step count)"""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
once = sys.argv[2] if len(sys.argv) > 2 else "k_hands"
i = [k for k, r in enumerate(rows) if r and r[0] == "ID"][0]
h = rows[i]
t, n, inst = collections.OrderedDict(), collections.Counter(), collections.Counter()
for r in rows[i + 1:]:
    d = dict(zip(h, r))
    name = d["Kernel Name"].split("(")[0].split("::")[-1]
    v = float(d["Metric Value"].replace(",", ""))
    if d["Metric Name"] == "gpu__time_duration.sum":
        t[name] = t.get(name, 0) + v
        n[name] += 1
    else:
        inst[name] += v
steps = n[once]
tot = sum(t.values())
print("steps captured: %d" % steps)
print("| kernel | launches/step | us/step | share | M warp inst/step |\n|---|---|---|---|---|")
for k, v in sorted(t.items(), key=lambda kv: -kv[1]):
    print("| %s | %.1f | %.1f | %.1f%% | %.2f |" % (k, n[k] / steps, v / steps / 1000, 100 * v / tot, inst[k] / steps / 1e6))
print("| **sum** | %.0f | %.1f | | %.1f |" % (sum(n.values()) / steps, tot / steps / 1000, sum(inst.values()) / steps / 1e6))
