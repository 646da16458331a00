"""Per-call wall time of the first host-buffer layer calls in a fresh process (config 2): how long the transient lasts."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hmm_layer_b200 as H
from hmm_layer_b200 import synthetic as syn

dev = torch.device("cuda:0")
em = H.GenePredHMMEmitter(**syn.GENE_CODONS)
em.build((1, 32, 9999, 15))
tr = H.GenePredMultiHMMTransitioner(initial_exon_len=200, initial_intron_len=4500, initial_ir_len=10000)
cell = H.HmmCell([15], 15, em, tr).to(dev)
layer = H.MsaHmmLayer(cell, use_prior=False)
x = syn.gene_inputs(1, 32, 9999, seed=1234).pin_memory()
ts = []
for i in range(120):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out = layer.state_posterior_log_probs(x)
    torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
print(" ".join("%.2f" % t for t in ts))
