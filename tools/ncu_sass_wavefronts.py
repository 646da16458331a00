"""Shared-memory wavefronts, executed instructions and stall samples per phase (stretch of SASS between barriers) from
`ncu -i rep --page source --csv --print-source sass`:   python tools/ncu_sass_wavefronts.py sass.csv [kernel substring] [top N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1], errors='replace')))
pick = sys.argv[2] if len(sys.argv) > 2 else ""
top = int(sys.argv[3]) if len(sys.argv) > 3 else 12
hdr = None; fn = None; data = {}
for r in rows:
    if not r: continue
    if r[0] == "Kernel Name": fn = r[1].split("(")[0][-40:]; data[fn] = []; continue
    if r[0] == "Address": hdr = r; continue
    if hdr and fn and r[0].startswith('0x'): data[fn].append(r)
ix = {h: i for i, h in enumerate(hdr)}
def I(r, k):
    try: return int(r[ix[k]])
    except ValueError: return 0
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
for fn, l in data.items():
    if pick not in fn: continue
    tot = sum(I(r, 'Instructions Executed') for r in l); wf = sum(I(r, 'L1 Wavefronts Shared') for r in l); ts = sum(I(r, '# Samples') for r in l)
    print(f"=== {fn}: {len(l)} SASS, executed {tot}, shared wavefronts {wf} (ideal {sum(I(r,'L1 Wavefronts Shared Ideal') for r in l)}), samples {ts}")
    start = 0; seg = []
    for i, r in enumerate(l):
        seg.append(r)
        if r[1].strip().startswith('BAR') or i == len(l) - 1:
            ex = sum(I(x, 'Instructions Executed') for x in seg); w = sum(I(x, 'L1 Wavefronts Shared') for x in seg); s = sum(I(x, '# Samples') for x in seg)
            st = sorted(((sum(I(x, k) for x in seg), k) for k in stalls), reverse=True)[:4]
            print(f"  sass[{start:5d}-{i:5d}] exec {ex:9d} ({100*ex/tot:5.1f}%)  smem wf {w:8d} ({100*w/max(wf,1):5.1f}%)  samples {100*s/max(ts,1):5.1f}%  " +
                  " ".join(f"{k[6:]}={100*v/max(s,1):.0f}%" for v, k in st))
            start = i + 1; seg = []
    print("  -- top shared-memory instructions by wavefronts")
    for r in sorted(l, key=lambda r: -I(r, 'L1 Wavefronts Shared'))[:top]:
        print(f"     {l.index(r):5d} {r[1].strip()[:70]:70s} exec {I(r,'Instructions Executed'):8d} wf {I(r,'L1 Wavefronts Shared'):8d} ideal {I(r,'L1 Wavefronts Shared Ideal'):8d}")
