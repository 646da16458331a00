"""Minimal config-2 driver for ncu: a few fused forward+posterior passes (no timing, no oracle)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, synthetic as syn
from tests.helpers import gene_params

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4
L = int(sys.argv[2]) if len(sys.argv) > 2 else 0
x = syn.gene_inputs(1, 32, 9999, seed=1234).cuda()
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
out = torch.empty((1, 32, 9999, 15), device="cuda")
ops.set_tuning(L, 0)
for _ in range(n):
    ops.genepred_run(x, A, pi, Bfull, tab, out={"post": out})
torch.cuda.synchronize()
print("ok", float(out[0, 0, 0, 0]))
