#!/usr/bin/env python
"""Join an `ncu --page source --csv` SASS dump with `nvdisasm -g` line info: per source line share of
executed warp instructions, lane efficiency and stall samples.  usage: sass_lines.py ncu_src.csv wf.sass <mangled kernel>"""
import csv, re, sys, collections
src_csv, sass, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 50
# 1. line info per instruction (in order) from nvdisasm
lines = open(sass).read().split('\n')
start = next(i for i, l in enumerate(lines) if l.startswith('.text.' + kern + ':'))
cur = None; seq = []
for l in lines[start + 1:]:
    if l.startswith('.text.') or l.startswith('.section'):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        inl = re.findall(r'inlined at "([^"]+)", line (\d+)', l)
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        seq.append((cur, m.group(2)))
# 2. ncu rows of the first launch
rows = list(csv.reader(open(src_csv)))
blocks = []; b = None
for r in rows:
    if r and r[0] == 'Kernel Name': b = {'h': None, 'r': []}; blocks.append(b); continue
    if b is None: continue
    if b['h'] is None: b['h'] = r; continue
    b['r'].append(r)
b = blocks[0]; h = b['h']
iI, iT, iS = h.index('Instructions Executed'), h.index('Thread Instructions Executed'), h.index('# Samples')
if len(b['r']) != len(seq):
    print(f"warning: ncu has {len(b['r'])} instructions, nvdisasm {len(seq)}", file=sys.stderr)
agg = collections.defaultdict(lambda: [0, 0, 0])
for (loc, txt), r in zip(seq, b['r']):
    a = agg[loc]; a[0] += int(r[iI]); a[1] += int(r[iT]); a[2] += int(r[iS])
ti = sum(a[0] for a in agg.values()); ts = sum(a[2] for a in agg.values()); tt = sum(a[1] for a in agg.values())
print(f"total warp inst {ti}, thread inst {tt}, lanes/inst {tt / ti:.2f}, samples {ts}")
for loc, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{100 * a[0] / ti:5.1f}% inst  lanes={a[1] / max(1, a[0]):5.1f}  {100 * a[2] / max(1, ts):5.1f}% samples  {loc}")
print("---- by thread instructions")
for loc, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{100 * a[1] / tt:5.1f}% thr-inst  {100 * a[0] / ti:5.1f}% warp-inst lanes={a[1] / max(1, a[0]):5.1f}  {loc}")
files = collections.defaultdict(lambda: [0, 0, 0])
for loc, a in agg.items():
    f = files[loc[0] if loc else None]
    for k in range(3): f[k] += a[k]
print("---- by file")
for f, a in sorted(files.items(), key=lambda kv: -kv[1][0]):
    print(f"{100 * a[0] / ti:5.1f}% warp-inst {100 * a[1] / tt:5.1f}% thr-inst {100 * a[2] / max(1, ts):5.1f}% samples lanes={a[1] / max(1, a[0]):5.1f} {f}")
