#!/usr/bin/env python
"""per-pixel completion times from an ER_WF_TIMELINE dump: when did each pixel finish, which finish last
   pixel_done.py dump.bin"""
import sys

import numpy as np

raw = np.fromfile(sys.argv[1], dtype=np.uint32)
grid, n = int(raw[0]), int(raw[1])
x_res = int(raw[2]) >> 8
body = raw[4:]
done = body[grid * n * 8:].astype(np.int64)
h = done.size // x_res
done = done.reshape(h, x_res)
m = done > 0
t0 = done[m].min()
ms = (done - t0) * 1.024e-3
ms[~m] = np.nan
rows = np.nonzero(m.any(axis=1))[0]
print(f"rows {rows.min()}..{rows.max()}, pixels {int(m.sum())}, last pixel done {np.nanmax(ms):.1f} ms after the first")
qs = [10, 25, 50, 75, 90, 95, 99, 99.9, 99.99, 100]
print("completion time percentiles (ms):", {q: round(float(np.nanpercentile(ms, q)), 1) for q in qs})
late = np.argwhere(ms > np.nanpercentile(ms, 99.9))
print("pixels in the last 0.1 %:", len(late), "rows", np.bincount(late[:, 0])[rows.min():rows.max() + 1].tolist())
print("  x histogram (64-px bins):", np.histogram(late[:, 1], bins=np.arange(0, x_res + 64, 64))[0].tolist())
# spatial clustering: fraction of late pixels with a late 8-neighbour
ls = set(map(tuple, late.tolist()))
nb = sum(any((y + dy, x + dx) in ls for dy in (-1, 0, 1) for dx in (-1, 0, 1) if dy or dx) for y, x in ls)
print(f"  with a late neighbour: {nb / max(1, len(ls)):.2f}")
for r in (rows.min(), (rows.min() + rows.max()) // 2, rows.max()):
    print(f"row {r}: mean {np.nanmean(ms[r]):.0f} ms, max {np.nanmax(ms[r]):.0f}")
