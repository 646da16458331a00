"""Config 2 through the compact wire format: device-resident pass and host-buffer end to end, next to the fp32 format."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, synthetic as syn
from tests.helpers import gene_params

x = syn.gene_inputs(1, 32, 9999, seed=1234)
cls16, codes = ops.pack_compact(x)
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
tab = ops.prepare_codon(tab)
xd, cd, nd = x.cuda(), cls16.cuda(), codes.cuda()
xp, cp, npn = x.pin_memory(), cls16.pin_memory(), codes.pin_memory()

def dev_time(fn, n=100):
    for _ in range(10): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3

def wall(fn, n=60):
    for _ in range(30): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    keep = None
    for _ in range(n): keep = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3

print("device, fp32 (80 B/pos in)       %.1f us" % dev_time(lambda: ops.genepred_run(xd, A, pi, Bfull, tab)))
print("device, compact (31 B/pos in)    %.1f us" % dev_time(lambda: ops.genepred_run_compact(cd, nd, A, pi, Bfull, tab)))
for rep in range(2):
    print("host e2e, fp32    (25.6 MB up)   %.3f ms" % wall(lambda: ops.genepred_run_host(xp, A, pi, Bfull, tab)))
    print("host e2e, compact ( 9.9 MB up)   %.3f ms" % wall(lambda: ops.genepred_run_compact(cp, npn, A, pi, Bfull, tab)))
