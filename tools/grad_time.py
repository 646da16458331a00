"""Config-2 training step of the gene HMM: differentiable forward (fused evaluation keeping log posterior + normalised
forward vectors) and the gradient kernels, CUDA-event times; plus the achieved gradient errors on a small case."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, synthetic as syn
from tests.helpers import gene_params

B, T = 32, 9999
x = syn.gene_inputs(1, B, T, seed=1234).cuda()
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
tab = ops.prepare_codon(tab)
w = torch.ones((1, B), device="cuda")

def ev():
    return torch.cuda.Event(enable_timing=True)

def timeit(fn, n=50):
    for _ in range(5): fn()
    e0, e1 = ev(), ev()
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3

print("forward loglik only            %7.1f us" % timeit(lambda: ops.genepred_run(x, A, pi, Bfull, tab, want_post=False)))
print("forward posterior              %7.1f us" % timeit(lambda: ops.genepred_run(x, A, pi, Bfull, tab)))
r = ops.genepred_run(x, A, pi, Bfull, tab, want_post=True, want_alpha=True, alpha_normalized=True)
print("forward post + alpha-hat       %7.1f us" % timeit(lambda: ops.genepred_run(x, A, pi, Bfull, tab, want_post=True, want_alpha=True, alpha_normalized=True)))
print("gradient kernels (all grads)   %7.1f us" % timeit(lambda: ops.genepred_grad(x, A, pi, Bfull, tab, r["post"], r["alpha"], w)))
print("gradient kernels (params only) %7.1f us" % timeit(lambda: ops.genepred_grad(x, A, pi, Bfull, tab, r["post"], r["alpha"], w, want_inputs=False)))
xg = x.clone().requires_grad_(True)
Ag, pig, Bg = A.clone().requires_grad_(True), pi.clone().requires_grad_(True), Bfull.clone().requires_grad_(True)
def step():
    for t in (xg, Ag, pig, Bg): t.grad = None
    ops.GenePredLoglik.apply(xg, Ag, pig, Bg, tab, None, False).sum().backward()
print("autograd forward + backward    %7.1f us" % timeit(step))

# achieved errors against fp64 autograd over the reference port (small case; needs oracle/, test infrastructure)
from tests.test_gpu_grad import oracle_grads, fold_b13, EDGE_SRC, EDGE_DST
Bs, Ts = 4, 1000
xs = syn.gene_inputs(1, Bs, Ts, seed=9)
An, pin, Bn, tabn = gene_params(pi_random=True)
ref = oracle_grads(xs.numpy(), An, pin, Bn, tabn, np.ones((1, Bs)))
d = lambda v: torch.as_tensor(np.asarray(v, np.float32)).cuda()
xg = xs.cuda().requires_grad_(True); Ag, pig, Bg = d(An).requires_grad_(True), d(pin).requires_grad_(True), d(Bn).requires_grad_(True)
ops.GenePredLoglik.apply(xg, Ag, pig, Bg, d(tabn), None, False).sum().backward()
e = (0, EDGE_SRC, EDGE_DST)
rel = lambda g, r: float(np.abs(g - r).max() / np.abs(r).max())
print("max-norm relative errors (B=4, T=1000): dx %.2e  dA %.2e  dpi %.2e  dB %.2e" % (
    rel(xg.grad.cpu().numpy()[..., :15], ref["x"][..., :15]), rel(Ag.grad.cpu().numpy()[e], ref["A"][e]),
    rel(pig.grad.cpu().numpy(), ref["pi"]), rel(fold_b13(Bg.grad.cpu().numpy()), ref["B13"])))
