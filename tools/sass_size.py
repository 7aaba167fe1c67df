#!/usr/bin/env python
"""Static SASS size per source line: `nvdisasm -g -c <cubin>` -> instructions attributed to each file:line
of one kernel (code size matters: the instruction cache is a first-order limit for these kernels)."""
import collections, re, subprocess, sys
cubin, fn = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(['nvdisasm', '-g', '-c', cubin], capture_output=True, text=True).stdout
cnt = collections.Counter(); cur = None; inside = False; total = 0
for line in txt.splitlines():
    if line.startswith('.text.'):
        inside = fn in line
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    if re.match(r'\s*/\*[0-9a-f]{4,}\*/', line):
        cnt[cur] += 1; total += 1
print(f'{total} instructions = {total * 16 / 1024:.0f} KB')
byfile = collections.Counter()
for (f, l), c in cnt.items():
    byfile[f] += c
print(dict(byfile))
for (f, l), c in cnt.most_common(top):
    print(f'{c:6d} {100.0 * c / total:5.1f}%  {f}:{l}')
