#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv` per CUDA source line:
stall samples, warp instructions and average active threads.  usage: ncu_lines.py file.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
cur, hdr, out = None, None, []
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        continue
    if hdr and len(r) == len(hdr) and r[0] not in ("", "-"):
        try:
            out.append((int(r[6]), cur.split("/")[-1], r[0], r[1][:100], int(r[7]), int(r[8])))
        except ValueError:
            pass
tot = sum(o[0] for o in out) or 1
insts = sum(o[4] for o in out) or 1
print(f"total samples {tot}, warp insts {insts}, thread insts {sum(o[5] for o in out)}")
byfile = {}
for o in out:
    f = byfile.setdefault(o[1], [0, 0, 0])
    f[0] += o[0]; f[1] += o[4]; f[2] += o[5]
for k, v in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print(f"  {k:18s} samples {100*v[0]/tot:5.1f}%  warp-inst {100*v[1]/insts:5.1f}%  avg thr {v[2]/max(1,v[1]):5.1f}")
for o in sorted(out, reverse=True)[:top]:
    print(f"{100*o[0]/tot:5.2f}% {o[1]:16s}:{o[2]:>4s} inst {100*o[4]/insts:5.2f}% thr {o[5]/max(1,o[4]):5.1f} | {o[3]}")
