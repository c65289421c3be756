#!/usr/bin/env python
"""Per-source-line instruction / stall-sample shares of one kernel from an .ncu-rep.
usage: ncu_lines.py report.ncu-rep kernel_regex [top]"""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
fname, hdr, agg = "", None, {}
launch = 0
for r in rows:
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; continue
    if hdr and r[0].isdigit() and len(r) > 8:
        key = (fname, int(r[0]), r[1].strip()[:100])
        a = agg.setdefault(key, [0, 0])
        try:
            a[0] += int(r[7]); a[1] += int(r[6])
        except ValueError:
            pass
ti = sum(a[0] for a in agg.values()) or 1
ts = sum(a[1] for a in agg.values()) or 1
print("kernel %s: warp-instructions %d, samples %d (all captured launches summed)" % (kern, ti, ts))
for (f, ln, src), (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%5.1f%% inst %5.1f%% smp  %s:%d  %s" % (100.0 * n / ti, 100.0 * s / ts, f, ln, src))
