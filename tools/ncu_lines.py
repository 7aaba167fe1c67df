"""Join an ncu SASS source page (per-instruction counters) with nvdisasm line info so that instruction
counts and stall samples can be read per CUDA source line.
usage: ncu_lines.py <report.ncu-rep> <kernel regex> <cubin.sass from `nvdisasm -g -c`> <function substring> [N]
NCU_COL=<column> (e.g. stall_long_sb, stall_no_inst) ranks by that sampling column instead of '# Samples'."""
import collections
import csv
import os
import re
import subprocess
import sys

rep, kre, sass, fn = sys.argv[1:5]
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', f'regex:{kre}'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = [i for i, r in enumerate(rows) if 'Instructions Executed' in r][0]
hdr = rows[hi]
ci, si, ai = hdr.index('Instructions Executed'), hdr.index(os.environ.get('NCU_COL', '# Samples')), hdr.index('Address')
xi = hdr.index(os.environ['NCU_SUB']) if os.environ.get('NCU_SUB') else None   # NCU_SUB=stall_barrier: samples minus that column
insts = []
for r in rows[hi + 1:]:
    try:
        insts.append((int(r[ci]), int(r[si]) - (int(r[xi]) if xi is not None else 0)))
    except (ValueError, IndexError):
        pass
# nvdisasm: instructions of the function in order, each preceded by "//## File ..., line N" markers
lines = open(sass).read().splitlines()
start = [i for i, l in enumerate(lines) if l.startswith('.text.') and fn in l][0]
cur = None
seq = []
for l in lines[start + 1:]:
    if l.startswith('.text.') or l.startswith('.section'):
        if seq:
            break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/', l):
        seq.append(cur)
print(f'{len(insts)} profiled instructions, {len(seq)} disassembled', file=sys.stderr)
n = min(len(insts), len(seq))
agg = collections.defaultdict(lambda: [0, 0])
for (c, s), loc in zip(insts[:n], seq[:n]):
    agg[loc][0] += c
    agg[loc][1] += s
tot_c = sum(v[0] for v in agg.values()) or 1
tot_s = sum(v[1] for v in agg.values()) or 1
src = {}
for loc, (c, s) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[5]) if len(sys.argv) > 5 else 40]:
    if loc is None:
        print(f'{100*c/tot_c:5.1f}% inst {100*s/tot_s:5.1f}% samples  <no line>')
        continue
    f, ln = loc
    if f not in src:
        try:
            src[f] = open(f'/root/repo/magical-1_b200/csrc/{f}').read().splitlines()
        except OSError:
            try:
                src[f] = open(f'/root/repo/include/{f}').read().splitlines()
            except OSError:
                src[f] = []
    text = src[f][ln - 1].strip()[:100] if ln - 1 < len(src[f]) else ''
    print(f'{100*c/tot_c:5.1f}% inst {100*s/tot_s:5.1f}% samples  {f}:{ln}  {text}')
