"""Config-2 Viterbi decoding time (emission kernel + viterbi kernel), CUDA events."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, synthetic as syn
from tests.helpers import gene_params

B, T = int(sys.argv[1]) if len(sys.argv) > 1 else 32, 9999
x = syn.gene_inputs(1, B, T, seed=1234).cuda()
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
tab = ops.prepare_codon(tab)
for _ in range(3):
    ops.genepred_viterbi(x, A, pi, Bfull, tab)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
n = 20
for _ in range(n):
    ops.genepred_viterbi(x, A, pi, Bfull, tab)
e1.record(); torch.cuda.synchronize()
print("viterbi B=%d T=%d: %.1f us per call" % (B, T, e0.elapsed_time(e1) / n * 1e3))
