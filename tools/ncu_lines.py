#!/usr/bin/env python
"""Per-source-line totals of an `ncu --page source --csv --print-source cuda,sass` export (instructions executed, active lanes,
stall samples), for one kernel whose name contains <substr>:  python tools/ncu_lines.py <export.csv> <substr> [min_pct] [warps]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2]; minpct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.8; warps = float(sys.argv[4]) if len(sys.argv) > 4 else 0
func = hdr = fpath = None; out = {}
for r in rows:
    if not r: continue
    if r[0] == 'Function Name': func = r[1]; continue
    if r[0] == 'File Path': fpath = r[1]; continue
    if r[0] == 'Line No': hdr = r; continue
    if r[0] and r[0].isdigit() and hdr and func and want in func:
        d = dict(zip(hdr, r))
        try: ie = int(d['Instructions Executed']); te = int(d['Thread Instructions Executed']); sm = int(d['# Samples'])
        except (KeyError, ValueError): continue
        k = (fpath.split('/')[-1], int(r[0])); o = out.get(k, (0, 0, 0, ''))
        out[k] = (o[0] + ie, o[1] + te, o[2] + sm, r[1][:100])
tot = sum(v[0] for v in out.values()); tsm = sum(v[2] for v in out.values()); tthr = sum(v[1] for v in out.values())
print(f"instructions {tot}  samples {tsm}  avg lanes {tthr / max(tot, 1):.1f}" + (f"  per warp {tot / warps:.0f}" if warps else ""))
for k, v in sorted(out.items()):
    if v[0] > tot * minpct / 100 or v[2] > tsm * minpct / 100:
        print(f"{k[0]}:{k[1]:5d} instr {100 * v[0] / tot:5.1f}% samp {100 * v[2] / max(tsm, 1):5.1f}% lanes {v[1] / max(v[0], 1):5.1f} | {v[3]}")
