#!/usr/bin/env python
"""Summarise an `ncu --page source --csv --print-source cuda,sass` dump per source line:
instructions executed and stall samples.  usage: ncu_lines.py dump.csv [topN]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = None
lines = []
tot_inst = tot_samp = 0
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if len(r) < 8 or r[0] in ("Line No", "Function Name"):
        continue
    if r[0] != "":  # a source line summary row
        try:
            inst = int(r[7]); samp = int(r[4])
        except ValueError:
            continue
        lines.append((inst, samp, cur_file, r[0], r[1].strip()[:90]))
        tot_inst += inst; tot_samp += samp
print("total inst", tot_inst, "samples", tot_samp)
print("--- by instructions")
for inst, samp, f, ln, src in sorted(lines, reverse=True)[:top]:
    print(f"{inst:12d} {100*inst/tot_inst:5.1f}%  samp {100*samp/max(tot_samp,1):5.1f}%  {f}:{ln}  {src}")
print("--- by stall samples")
for inst, samp, f, ln, src in sorted(lines, key=lambda x: -x[1])[:top]:
    print(f"{samp:8d} {100*samp/max(tot_samp,1):5.1f}%  inst {100*inst/tot_inst:5.1f}%  {f}:{ln}  {src}")
