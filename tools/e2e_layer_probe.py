"""Host-side profile of MsaHmmLayer.state_posterior_log_probs(CPU tensor) on config 2: where the time outside the C call goes."""
import sys, os, time, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hmm_layer_b200 as H
from hmm_layer_b200 import synthetic as syn

dev = torch.device("cuda:0")
em = H.GenePredHMMEmitter(**syn.GENE_CODONS)
em.build((1, 32, 9999, 15))
with torch.no_grad():
    em.emission_kernel.copy_(syn.gene_emission_kernel())
tr = H.GenePredMultiHMMTransitioner(initial_exon_len=200, initial_intron_len=4500, initial_ir_len=10000)
cell = H.HmmCell([15], 15, em, tr).to(dev)
layer = H.MsaHmmLayer(cell, use_prior=False)
x = syn.gene_inputs(1, 32, 9999, seed=1234).pin_memory()
for _ in range(5):
    layer.state_posterior_log_probs(x)
torch.cuda.synchronize()
n = 50
t0 = time.perf_counter()
for _ in range(n):
    out = layer.state_posterior_log_probs(x)
torch.cuda.synchronize()
print("layer call %.3f ms" % ((time.perf_counter() - t0) / n * 1e3))
pr = cProfile.Profile()
pr.enable()
for _ in range(n):
    out = layer.state_posterior_log_probs(x)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)

from hmm_layer_b200 import ops
gp = cell._gene_params
def tm(fn, keep, n=50):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    held = None
    for _ in range(n):
        r = fn()
        if keep: held = r
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
print("ops discard  %.3f" % tm(lambda: ops.genepred_run_host(x, *gp), False))
print("ops keep     %.3f" % tm(lambda: ops.genepred_run_host(x, *gp), True))
print("layer discard %.3f" % tm(lambda: layer.state_posterior_log_probs(x), False))
print("layer keep    %.3f" % tm(lambda: layer.state_posterior_log_probs(x), True))
print("ops discard  %.3f" % tm(lambda: ops.genepred_run_host(x, *gp), False))
