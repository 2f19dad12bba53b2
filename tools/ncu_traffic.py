#!/usr/bin/env python
"""Write profiles/traffic.json from an `ncu --set full` raw page
(`ncu -i x.ncu-rep --page raw --csv > raw.csv`): per kernel, DRAM bytes of one launch
(dram__bytes_read.sum + dram__bytes_write.sum, mean over the captured launches) and the number
of blocks the launch processed (given on the command line: the capture is made with
tools/prof_run.py at bench.py's size).
usage: ncu_traffic.py raw.csv <blocks per launch> <out.json> [kernel-name-substring ...]"""
import csv
import json
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
blocks = int(sys.argv[2])
want = sys.argv[4:]
SCALE = {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}


def num(r, name):
    return float(r[idx[name]].replace(",", "")) * SCALE[units[idx[name]].lower()]


acc = {}
for r in rows[2:]:
    name = r[idx["Kernel Name"]].split("(")[0].split("<")[0].replace("void ", "").strip()
    if want and not any(w in name for w in want):
        continue
    acc.setdefault(name, []).append(num(r, "dram__bytes_read.sum") + num(r, "dram__bytes_write.sum"))
out = {k: {"dram_bytes_per_launch": int(sum(v) / len(v)), "launches": len(v), "blocks": blocks,
           "source": "ncu --set full --clock-control none, tools/prof_run.py %d 0" % blocks}
       for k, v in acc.items()}
json.dump(out, open(sys.argv[3], "w"), indent=1)
print(json.dumps(out, indent=1))
