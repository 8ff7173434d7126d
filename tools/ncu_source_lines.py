"""Per-source-line instruction counts from `ncu --page source --csv --print-source cuda,sass`:
python tools/ncu_source_lines.py src.csv [top_n]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = next(r for r in rows if r and r[0] == 'Line No')
ie, sa = hdr.index('Instructions Executed'), hdr.index('# Samples')
lines = []
for r in rows:
    if len(r) > ie and r[0].isdigit() and r[2] == '-':
        lines.append((int(r[ie] or 0), int(r[sa] or 0), int(r[0]), r[1].strip()))
tot = sum(l[0] for l in lines) or 1
tots = sum(l[1] for l in lines) or 1
print(f"total warp instructions {tot}, samples {tots}")
for n, s, ln, src in sorted(lines, reverse=True)[:top]:
    print(f"{n / tot * 100:5.1f}% inst {s / tots * 100:5.1f}% smp  L{ln:4d}  {src[:110]}")
