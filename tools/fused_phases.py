"""Per-phase cycle counts of the single-launch gene-HMM kernel on config 2 (clock64 of thread 0 of every CTA)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, synthetic as syn, _lib
from tests.helpers import gene_params

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
T = int(sys.argv[2]) if len(sys.argv) > 2 else 9999
if len(sys.argv) > 3:
    ops.set_tuning(int(sys.argv[3]), 0)
x = syn.gene_inputs(1, B, T, seed=1234).cuda()
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
out = torch.empty((1, B, T, 15), device="cuda")
for _ in range(3):
    ops.genepred_run(x, A, pi, Bfull, tab, out={"post": out})
nct = 4096
dbg = torch.zeros((2, nct, 16), dtype=torch.int64, device="cuda")
_lib.load().hmm_b200_debug_phases(dbg.data_ptr())
ops.genepred_run(x, A, pi, Bfull, tab, out={"post": out})
torch.cuda.synchronize()
_lib.load().hmm_b200_debug_phases(None)
fine = dbg.cpu().numpy().reshape(-1, 16)[4096:]
d = dbg.cpu().numpy().reshape(-1, 16)
nc = int((d[:4096, 0] > 0).sum())
d = d[:nc]
names = ["params", "emission+coef", "rows", "blocks", "tree", "publish+wait+scan", "down-sweep", "recursion", "output"]
t = d[:, :10]
dt = np.diff(t, axis=1)
print(f"fused kernel: {nc} CTAs")
for j, nm in enumerate(names):
    print(f"   {nm:22s} median {np.median(dt[:, j]):9.0f}  max {dt[:, j].max():9.0f}")
print(f"   per-CTA total          median {np.median(t[:, -1] - t[:, 0]):9.0f}")
w0, w1, sm = d[:, 14], d[:, 15], d[:, 13]
base = w0.min()
st, en = np.sort(w0 - base), np.sort(w1 - base)
print(f"   wall clock ns: starts p50 {st[len(st)//2]} max {st[-1]}; ends p0 {en[0]} p50 {en[len(en)//2]} max {en[-1]}; CTA wall median {np.median(w1 - w0):.0f} ns")
pub, got = d[:, 10] - base, d[:, 11] - base
P = nc // B
pub_s = pub.reshape(B, P)
print(f"   publish time (ns after first CTA start): p0 {pub.min()} p50 {int(np.median(pub))} max {pub.max()};  per-sequence spread (max-min) median {int(np.median(pub_s.max(1) - pub_s.min(1)))} max {int((pub_s.max(1) - pub_s.min(1)).max())}")
print(f"   wait for the sequence's tiles (ns): median {int(np.median(got - pub))} max {int((got - pub).max())}")
cnt = np.bincount(sm.astype(int)); alone = np.isin(sm, np.where(cnt == 1)[0])
if alone.any() and (~alone).any():
    print(f"   publish time of CTAs alone on an SM: median {int(np.median(pub[alone]))}; sharing: median {int(np.median(pub[~alone]))} p90 {int(np.percentile(pub[~alone], 90))}")
print(f"   SMs used {len(np.unique(sm))}, max CTAs on one SM {np.bincount(sm.astype(int)).max()}")

if fine[:, 9].any():
    f = fine[:nc]
    seq = np.stack([f[:, 8], f[:, 9], f[:, 0], f[:, 1], f[:, 2], f[:, 3], f[:, 4], d[:, 7]], 1)
    print("   fine marks (cycles): prologue, level 1, 2, 4, 8, 16, final barrier:", [int(np.median(x)) for x in np.diff(seq, axis=1).T])
    # thread 0's mark 6 is read in front of the barrier that ends the top-level scan (warp 1): the gap to the first fine mark
    # is the scan itself, so "publish+wait+scan" above is publish + wait + staging and "down-sweep" includes the scan
    print("   top-level scan (mark 6 -> first fine mark):", int(np.median(f[:, 8] - d[:, 6])), " two adjacent fine marks:", int(np.median(f[:, 12] - f[:, 8])))
