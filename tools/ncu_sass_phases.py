"""Executed-instruction shares between barriers from `ncu --page source --csv --print-source sass`:
python tools/ncu_sass_phases.py sass.csv   (phases = stretches of SASS between BAR / SYNCS instructions)"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1], errors='replace')))
fn = None; hdr = None; out = {}
for r in rows:
    if not r:
        continue
    if r[0] == "Kernel Name":
        fn = r[1].split("(")[0][-30:]; out[fn] = []; continue
    if r[0] == "Address":
        hdr = r; continue
    if hdr and fn and len(r) > 6 and r[0].startswith('0x'):
        try:
            out[fn].append((r[1].strip(), int(r[hdr.index('Instructions Executed')]), int(r[hdr.index('# Samples')])))
        except ValueError:
            pass
for fn, l in out.items():
    tot = sum(x[1] for x in l); ts = sum(x[2] for x in l)
    print("===", fn, len(l), "sass instrs; executed", tot)
    seg = segs = n = start = 0
    for i, (t, c, s) in enumerate(l):
        seg += c; segs += s; n += 1
        if t.startswith('BAR') or i == len(l) - 1:
            print(f"  sass[{start:5d}-{i:5d}] n={n:5d} exec={seg:9d} ({100*seg/tot:5.1f}%) samples={100*segs/max(ts,1):5.1f}%  ends: {t[:40]}")
            seg = segs = n = 0; start = i + 1
