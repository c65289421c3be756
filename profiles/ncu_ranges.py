#!/usr/bin/env python
"""Instruction / stall-sample shares of source-line ranges of one kernel (sections of k_hands by default).
usage: ncu_ranges.py report.ncu-rep kernel_regex file.cuh marker1 marker2 ...   (a range starts at the first line containing the marker)"""
import csv, subprocess, sys
rep, kern, path = sys.argv[1], sys.argv[2], sys.argv[3]
markers = sys.argv[4:]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k", "regex:" + kern],
                     capture_output=True, text=True).stdout
fname, hdr, agg = "", None, {}
for r in csv.reader(txt.splitlines()):
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr and r[0].isdigit() and len(r) > 8:
        try:
            a = agg.setdefault((fname, int(r[0])), [0, 0]); a[0] += int(r[7]); a[1] += int(r[6])
        except ValueError:
            pass
ti = sum(a[0] for a in agg.values()) or 1
ts = sum(a[1] for a in agg.values()) or 1
base = path.split("/")[-1]
src = open(path).read().splitlines()
starts = []
for m in markers:
    for n, l in enumerate(src, 1):
        if m in l:
            starts.append((n, m)); break
starts.sort()
print("kernel %s: %d warp instructions, %d samples" % (kern, ti, ts))
for k, (n, m) in enumerate(starts):
    hi = starts[k + 1][0] - 1 if k + 1 < len(starts) else 10 ** 9
    i = sum(a[0] for (f, l), a in agg.items() if f == base and n <= l <= hi)
    s = sum(a[1] for (f, l), a in agg.items() if f == base and n <= l <= hi)
    print("  %-40s lines %4d-%-6s inst %5.1f%%  samples %5.1f%%" % (m[:40], n, hi if hi < 10 ** 9 else "end", 100 * i / ti, 100 * s / ts))
for f in sorted(set(f for f, _ in agg)):
    if f != base:
        i = sum(a[0] for (ff, l), a in agg.items() if ff == f); s = sum(a[1] for (ff, l), a in agg.items() if ff == f)
        print("  %-40s %-17s inst %5.1f%%  samples %5.1f%%" % (f, "(inlined helpers)", 100 * i / ti, 100 * s / ts))
