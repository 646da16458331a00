"""Counts the SASS mnemonics that show which hardware paths every kernel of libhmm_b200.so uses (no GPU needed):
python tools/sass_summary.py > profiles/r2_sass_summary.txt
UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st (TMEM), UTMALDG = TMA tensor load, UBLKCP = cp.async.bulk (1-D TMA),
HMMA = mma.sync (TF32 here), FFMA2 / FMUL2 / FADD2 = packed fp32 pairs, LDGSTS = cp.async, SYNCS = mbarrier."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "hmm_layer_b200", "libhmm_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
WANT = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UBLKPF", "HMMA", "FFMA2", "FMUL2", "FADD2", "FFMA", "LDGSTS", "SYNCS",
        "SHFL", "BAR", "MUFU", "LDS", "STS", "LDG", "STG", "ATOM", "RED", "NANOSLEEP"]
cur, counts, total = None, collections.OrderedDict(), {}
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[cur], total[cur] = collections.Counter(), 0
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        total[cur] += 1
        op = m.group(1)
        for w in WANT:
            if op == w or op.startswith(w + "."):
                counts[cur][w] += 1
print(f"# SASS mnemonic counts per kernel of {os.path.relpath(so, ROOT)} (cuobjdump -sass; static instruction counts)")
print(f"# {__doc__.splitlines()[2]}\n# {__doc__.splitlines()[3]}")
for k, c in counts.items():
    print(f"{k}: {total[k]} instructions; " + ", ".join(f"{w} {c[w]}" for w in WANT if c[w]))
