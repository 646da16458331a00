"""Config-2 drivers for ncu: a few gradient and Viterbi passes (no timing, no oracle)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, synthetic as syn
from tests.helpers import gene_params

x = syn.gene_inputs(1, 32, 9999, seed=1234).cuda()
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
tab = ops.prepare_codon(tab)
w = torch.ones((1, 32), device="cuda")
for _ in range(2):
    r = ops.genepred_run(x, A, pi, Bfull, tab, want_post=True, want_alpha=True, alpha_normalized=True)
    ops.genepred_grad(x, A, pi, Bfull, tab, r["post"], r["alpha"], w)
    ops.genepred_viterbi(x, A, pi, Bfull, tab)
torch.cuda.synchronize()
print("ok")
