"""Config-2 timing of the single-launch path against the two-kernel path (CUDA events, rotating buffers > L2)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, synthetic as syn
from tests.helpers import gene_params

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
T = int(sys.argv[2]) if len(sys.argv) > 2 else 9999
NS = 8
xs = [syn.gene_inputs(1, B, T, seed=1234 + i).cuda() for i in range(NS)]
outs = [torch.empty((1, B, T, 15), device="cuda") for _ in range(NS)]
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
tab = ops.prepare_codon(tab)
for name, on in (("fused", True), ("two-kernel", False)):
    ops.set_fused(on)
    for i in range(20):
        ops.genepred_run(xs[i % NS], A, pi, Bfull, tab, out={"post": outs[i % NS]})
    torch.cuda.synchronize()
    ts = []
    for rep in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(100):
            ops.genepred_run(xs[i % NS], A, pi, Bfull, tab, out={"post": outs[i % NS]})
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 10)
    print(f"{name}: B={B} T={T} median {np.median(ts):.1f} us  best {min(ts):.1f} us per pass")
ops.set_fused(True)
