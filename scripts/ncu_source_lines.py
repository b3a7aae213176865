#!/usr/bin/env python
"""Executed warp instructions, stall samples and active lanes per CUDA source line, from an .ncu-rep captured with
--import-source on:  python scripts/ncu_source_lines.py report.ncu-rep [top] > profiles/xyz_source_lines.txt"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
cur, hdr, out = None, None, []
for r in rows:
    if r and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) >= len(hdr) and r[0].isdigit():
        d = dict(zip(hdr, r))
        if not (d["Instructions Executed"] or "0").isdigit():
            continue  # a source line with commas in it
        out.append((int(d["Instructions Executed"] or 0), int(d["# Samples"] or 0),
                    int(d["Thread Instructions Executed"] or 0), cur, r[0], r[1].strip()[:90]))
ti, ts, tt = (sum(o[i] for o in out) for i in range(3))
print("# %s" % rep.split("/")[-1])
print("# warp instructions %d, stall samples %d, active lanes per instruction %.1f" % (ti, ts, tt / max(ti, 1)))
byfile = collections.Counter()
for o in out:
    byfile[o[3]] += o[0]
print("# by file: " + ", ".join("%s %.1f %%" % (k, 100.0 * v / ti) for k, v in byfile.most_common()))
print("#  inst %  samples %  lanes  file:line  source")
for o in sorted(out, reverse=True)[:top]:
    print("%7.1f %9.1f %6.1f  %s:%s  %s" % (100.0 * o[0] / ti, 100.0 * o[1] / ts, o[2] / max(o[0], 1), o[3], o[4], o[5]))
