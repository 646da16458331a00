"""Per-phase cycle counts of the gene-HMM kernels on config 2 (debug instrumentation, clock64 of thread 0)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, synthetic as syn, _lib
from tests.helpers import gene_params

L = int(sys.argv[1]) if len(sys.argv) > 1 else 0
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
x = syn.gene_inputs(1, B, 9999, seed=1234).cuda()
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
out = torch.empty((1, B, 9999, 15), device="cuda")
ops.set_tuning(L, 0)
for _ in range(3):
    ops.genepred_run(x, A, pi, Bfull, tab, out={"post": out})
nct = 4096
dbg = torch.zeros((2, nct, 16), dtype=torch.int64, device="cuda")
_lib.load().hmm_b200_debug_phases(dbg.data_ptr())
ops.genepred_run(x, A, pi, Bfull, tab, out={"post": out})
torch.cuda.synchronize()
_lib.load().hmm_b200_debug_phases(None)
d = dbg.cpu().numpy().reshape(-1, 16)
nc = int((d[:, 0] > 0).sum()) // 2
d = d[:2 * nc].reshape(2, nc, 16)
# mark index order per kernel (see HMM_MARK in gene15.cu)
order = [[0, 9, 10, 11, 7, 8, 1, 2, 3, 4, 5, 6], [0, 1, 2, 5, 6, 7]]
names = [["bulk issue", "Aw loads", "tables", "barrier", "input wait", "codes", "emission+coef", "rows", "blocks", "tree", "store"],
         ["prologue", "tree wait+down-sweep", "E wait+recursion", "z", "out"]]
for k in range(2):
    t = d[k][:, order[k]]
    print(f"kernel {k}: {nc} CTAs; wall span = {(t[:, -1].max() - t[:, 0].min())} cycles")
    dt = np.diff(t, axis=1)
    for j, nm in enumerate(names[k]):
        print(f"   {nm:22s} median {np.median(dt[:, j]):9.0f}  max {dt[:, j].max():9.0f}")
    print(f"   per-CTA total          median {np.median(t[:, -1] - t[:, 0]):9.0f}")
    w0, w1, sm = d[k][:, 14], d[k][:, 15], d[k][:, 13]
    base = d[0][:, 14].min()
    st, en = np.sort(w0 - base), np.sort(w1 - base)
    print(f"   wall clock (ns after the first CTA of kernel 0 started): starts p0 {st[0]} p25 {st[len(st)//4]} p50 {st[len(st)//2]} p75 {st[3*len(st)//4]} max {st[-1]};"
          f" ends p0 {en[0]} p50 {en[len(en)//2]} max {en[-1]}; CTA wall median {np.median(w1 - w0):.0f} ns")
    print(f"   SMs used {len(np.unique(sm))}, max CTAs on one SM {np.bincount(sm.astype(int)).max()}")
