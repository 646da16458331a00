import sys, os
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from hmm_layer_b200 import ops, synthetic as syn, _lib
from tests.helpers import gene_params
B, T = 32, 9999
x = syn.gene_inputs(1, B, T, seed=1234).cuda()
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
out = torch.empty((1, B, T, 15), device="cuda")
for _ in range(3):
    ops.genepred_run(x, A, pi, Bfull, tab, out={"post": out})
for rep in range(2):
    dbg = torch.zeros((2, 4096, 16), dtype=torch.int64, device="cuda")
    _lib.load().hmm_b200_debug_phases(dbg.data_ptr())
    ops.genepred_run(x, A, pi, Bfull, tab, out={"post": out})
    torch.cuda.synchronize()
    _lib.load().hmm_b200_debug_phases(None)
    d = dbg.cpu().numpy().reshape(-1, 16)[:288]
    sm = d[:, 13]
    by = {}
    for cta, s in enumerate(sm):
        by.setdefault(int(s), []).append(cta)
    pairs = [v for v in by.values() if len(v) == 2]
    diffs = sorted(set(v[1] - v[0] for v in pairs))
    par = sum(((v[0] // 148) & 1) != ((v[1] // 148) & 1) for v in pairs)
    par2 = sum((v[0] & 1) != (v[1] & 1) for v in pairs)
    print(f"rep {rep}: {len(pairs)} SMs with two CTAs; differences between the two CTA ids: {diffs[:12]}{'...' if len(diffs) > 12 else ''}; pairs split by (id // 148) parity: {par}, by id parity: {par2}")
    print("   first pairs:", pairs[:8])
