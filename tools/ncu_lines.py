"""Per-source-line instruction / stall-sample totals from `ncu --page source --csv --print-source cuda,sass`.
usage: python tools/ncu_lines.py file.csv [top_n]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1], errors="replace")))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
fn = None; path = None; hdr = None
agg = collections.defaultdict(lambda: collections.defaultdict(lambda: [0, 0, ""]))
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        path = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        fn = r[1].split("(")[0][-40:]; continue
    if r[0] == "Line No":
        hdr = r; continue
    if hdr is None or len(r) < 8 or r[2] != "-":
        continue
    try:
        ins = int(r[hdr.index("Instructions Executed")]); smp = int(r[hdr.index("# Samples")])
    except ValueError:
        continue
    a = agg[fn][(path, int(r[0]))]
    a[0] += ins; a[1] += smp; a[2] = r[1].strip()[:90]
for fn, d in agg.items():
    ti = sum(v[0] for v in d.values()); ts = sum(v[1] for v in d.values())
    print(f"=== {fn}: instr={ti} samples={ts}")
    for (pth, ln), v in sorted(d.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"  {pth}:{ln:4d} instr={100*v[0]/max(ti,1):5.1f}% samples={100*v[1]/max(ts,1):5.1f}%  {v[2]}")
