"""Chunk-length / CTA-width sweep on a gene-HMM shape: per-kernel CUDA-event times (ms) from the library's profiler.
usage: python tools/sweep_L.py [B] [T] [L list comma] [k2 threads list comma]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, synthetic as syn
from tests.helpers import gene_params

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
T = int(sys.argv[2]) if len(sys.argv) > 2 else 9999
Ls = [int(v) for v in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0, 8, 12, 14, 16, 18, 20, 22, 24, 28, 32]
NTs = [int(v) for v in sys.argv[4].split(",")] if len(sys.argv) > 4 else [0]
NBUF = 4
xs = [syn.gene_inputs(1, B, T, seed=1234 + i).cuda() for i in range(NBUF)]
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
tab = ops.prepare_codon(tab)
outs = [torch.empty((1, B, T, 15), device="cuda") for _ in range(NBUF)]
for L in Ls:
    for NT in NTs:
        ops.set_tuning(L, NT)
        for i in range(10):
            ops.genepred_run(xs[i % NBUF], A, pi, Bfull, tab, out={"post": outs[i % NBUF]})
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        n = 50
        for i in range(n):
            ops.genepred_run(xs[i % NBUF], A, pi, Bfull, tab, out={"post": outs[i % NBUF]})
        e1.record()
        torch.cuda.synchronize()
        tot = e0.elapsed_time(e1) / n
        ops.profile_enable(True)
        ker = None
        for i in range(20):
            ops.genepred_run(xs[i % NBUF], A, pi, Bfull, tab, out={"post": outs[i % NBUF]})
            k = ops.profile_read()
            ker = k if ker is None else [a + b for a, b in zip(ker, k)]
        ops.profile_enable(False)
        print(f"L={L:3d} NT={NT:3d}  step {tot*1e3:7.1f} us   kernels " + " ".join(f"{v/20*1e3:6.1f}" for v in ker), flush=True)
ops.set_tuning(0, 0)
