"""Instruction shares of primary_kernel by code section (line ranges of kernels.cu) from `ncu --page source --csv`:
    python tools/ncu_source_groups.py src.csv [kernel-substring]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
only = sys.argv[2] if len(sys.argv) > 2 else "primary_kernel"
GROUPS = [  # (first line, last line, label) in kernels.cu as of rc_build_id 71fcfeb4 (round 2); anything else: "other"
    (37, 107, "fp helpers, det_sin / det_exp"),
    (112, 220, "cursor, top table, wrap_coord, lookup"),
    (221, 317, "fetch_region / region_word"),
    (318, 327, "occlusion_state"),
    (329, 386, "build_ray, init_delta_pos, count_cell"),
    (387, 430, "dda_step (loop B)"),
    (431, 445, "dda_step_bl"),
    (446, 470, "dda_step_exact_bl (replay)"),
    (471, 485, "left_region / enter_region"),
    (486, 511, "enter_region_a (column regions)"),
    (512, 563, "after_lookup, colours, sky"),
    (564, 627, "sky_gate_distance, sky_drop_distance, counter_slot"),
    (648, 733, "shade_pixel: set-up, camera ray, first lookup"),
    (734, 794, "loop A: gate tests, empty-cell set-up"),
    (795, 807, "loop A: empty-cell step loop"),
    (808, 832, "loop A: sibling walk, position recovery"),
    (833, 850, "loop A: brick walk, word"),
    (851, 878, "free flight hand-over"),
    (879, 905, "ratios + replay loop"),
    (906, 990, "loop B"),
    (991, 1058, "shading"),
    (1059, 1106, "shadow ray"),
    (1107, 1132, "store_pixel"),
    (1672, 1777, "trace_tile: stores, seeds, counters"),
    (1778, 1797, "kernel prologue"),
]
sections, cur, path = [], None, None
for r in rows:
    if r and r[0] == 'File Path':
        path = r[1]
    elif r and r[0] == 'Function Name':
        cur = {'name': r[1], 'path': path, 'hdr': None, 'lines': []}
        sections.append(cur)
    elif r and r[0] == 'Line No' and cur is not None:
        cur['hdr'] = r
    elif cur is not None and cur['hdr'] and r and r[0].isdigit() and r[2] == '-':
        h = cur['hdr']
        ie, sa, te = h.index('Instructions Executed'), h.index('# Samples'), h.index('Thread Instructions Executed')
        cur['lines'].append((int(r[0]), int(r[ie] or 0), int(r[sa] or 0), int(r[te] or 0)))
tot = sum(l[1] for s in sections if only in s['name'] for l in s['lines']) or 1
smp = sum(l[2] for s in sections if only in s['name'] for l in s['lines']) or 1
acc = {}
for s in sections:
    if only not in s['name']:
        continue
    base = s['path'].split('/')[-1]
    for ln, n, sm, t in s['lines']:
        label = base
        if base == 'kernels.cu':
            label = 'kernels.cu: other'
            for a, b, g in GROUPS:
                if a <= ln <= b:
                    label = g
                    break
        e = acc.setdefault(label, [0, 0, 0])
        e[0] += n
        e[1] += sm
        e[2] += t
print(f"{only}: {tot} warp instructions, {smp} samples")
for label, (n, sm, t) in sorted(acc.items(), key=lambda kv: -kv[1][0]):
    print(f"{n / tot * 100:5.1f}% inst {sm / smp * 100:5.1f}% smp {t / max(n, 1):5.1f} thr  {n / 1e6:7.2f} M  {label}")
