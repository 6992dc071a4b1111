#!/usr/bin/env python
"""static SASS instruction count per source line of one kernel.
usage: nvdisasm -g -c X.cubin > dis.txt; sass_lines.py dis.txt <kernel-substring> [top]"""
import re
import sys

fn, cur, cnt = None, None, {}
key = sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
for line in open(sys.argv[1]):
    m = re.match(r'\s*\.section\s+\.text\.(\S+?),', line)
    if m:
        fn = m.group(1)
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (m.group(1), int(m.group(2)))
        continue
    if fn and key in fn and re.match(r'\s+/\*[0-9a-f]{4,6}\*/', line):
        cnt[cur] = cnt.get(cur, 0) + 1
byf = {}
for (f, l), c in cnt.items():
    byf[f.split('/')[-1]] = byf.get(f.split('/')[-1], 0) + c
print("total", sum(cnt.values()), sorted(byf.items(), key=lambda kv: -kv[1]))
src = {}
for (f, l), c in sorted(cnt.items(), key=lambda kv: -kv[1])[:top]:
    try:
        if f not in src:
            src[f] = open(f).read().split('\n')
        t = src[f][l - 1].strip()[:90]
    except Exception:
        t = ''
    print(c, f.split('/')[-1], l, t)
