"""Per-source-line instruction counts from `ncu --page source --csv --print-source cuda,sass`,
one table per kernel:  python tools/ncu_source_lines.py src.csv [top_n] [kernel-substring]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
only = sys.argv[3] if len(sys.argv) > 3 else None
sections, cur = [], None
for r in rows:
    if r and r[0] == 'Function Name':
        cur = {'name': r[1], 'hdr': None, 'lines': []}
        sections.append(cur)
    elif r and r[0] == 'Line No' and cur is not None:
        cur['hdr'] = r
    elif cur is not None and cur['hdr'] and r and r[0].isdigit() and r[2] == '-':
        h = cur['hdr']
        ie, sa, te = h.index('Instructions Executed'), h.index('# Samples'), h.index('Thread Instructions Executed')
        cur['lines'].append((int(r[ie] or 0), int(r[sa] or 0), int(r[te] or 0), int(r[0]), r[1].strip()))
for s in sections:
    if only and only not in s['name']:
        continue
    lines = s['lines']
    tot = sum(l[0] for l in lines) or 1
    tots = sum(l[1] for l in lines) or 1
    tt = sum(l[2] for l in lines)
    print(f"## {s['name']}: warp instructions {tot}, thread instructions {tt} ({tt / tot:.1f} threads/inst), samples {tots}")
    for n, sm, t, ln, src in sorted(lines, reverse=True)[:top]:
        print(f"{n / tot * 100:5.1f}% inst {sm / tots * 100:5.1f}% smp {t / max(n, 1):5.1f} thr  L{ln:4d}  {src[:100]}")
