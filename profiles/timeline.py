#!/usr/bin/env python
"""Reads an ER_WF_TIMELINE dump: per time bucket, over all CTAs, the live paths, queue sizes, and the tickets handed
out (throughput over the launch).   timeline.py dump.bin [samples_per_ticket]"""
import sys

import numpy as np

raw = np.fromfile(sys.argv[1], dtype=np.uint32)
grid, n, shift, work_total = (int(v) for v in raw[:4])
tl = raw[4:].reshape(grid, n, 8).astype(np.int64)
valid = tl[..., 0] > 0
last = int(np.nonzero(valid.any(axis=0))[0].max())
ms = (1 << shift) / 1.965e6
print(f"{grid} CTAs, bucket {ms:.2f} ms, launch ~{(last + 1) * ms:.0f} ms, {work_total} pixel tickets")
print("   t ms  CTAs  live/CTA  min   QT    QH    QE   tickets   tickets/ms")
prev_t, prev_k = 0, 0
step = max(1, (last + 1) // 40)
for k in range(0, last + 1, step):
    v = valid[:, k]
    if not v.any():
        continue
    live = tl[v, k, 0] - 1
    tick = int(tl[v, k, 4].max())
    rate = (tick - prev_t) / max(1e-9, (k - prev_k) * ms) if k else 0.0
    print(f"{k * ms:7.1f} {int(v.sum()):5d} {live.mean():8.1f} {int(live.min()):5d} {tl[v, k, 1].mean():5.0f} {tl[v, k, 2].mean():5.0f} "
          f"{tl[v, k, 3].mean():5.0f} {tick:9d} {rate:10.1f}")
    prev_t, prev_k = tick, k
