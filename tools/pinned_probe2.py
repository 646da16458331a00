"""hmm_b200_genepred_run_host time as a function of when its pinned output was allocated relative to device allocations."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, synthetic as syn, _lib
from tests.helpers import gene_params

x = syn.gene_inputs(1, 32, 9999, seed=1234).pin_memory()
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
codon = ops.prepare_codon(tab)
lib = _lib.load()
scratch = torch.empty(lib.hmm_b200_genepred_host_scratch_size(1, 32, 9999), dtype=torch.uint8, device="cuda")
ll_h = torch.empty((1, 32)).pin_memory()
st = torch.cuda.current_stream().cuda_stream
def run(out_h, n=40):
    def raw():
        rc = lib.hmm_b200_genepred_run_host(x.data_ptr(), 1, 32, 9999, A.data_ptr(), pi.data_ptr(), Bfull.data_ptr(), codon.data_ptr(), None, 0,
                                            out_h.data_ptr(), None, None, ll_h.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
        assert rc == 0
        torch.cuda.current_stream().synchronize()
    for _ in range(10): raw()
    t0 = time.perf_counter()
    for _ in range(n): raw()
    return (time.perf_counter() - t0) / n * 1e3
o1 = torch.empty((1, 32, 9999, 15), pin_memory=True)
print("out1 (pinned after all device allocs)     %.3f ms" % run(o1))
dummy = torch.empty(64_000_001, device="cuda")
print("out1 after a cudaMalloc                   %.3f ms" % run(o1))
o2 = torch.empty((1, 32, 9999, 15), pin_memory=True)
print("out2 (new pinned)                         %.3f ms" % run(o2))
print("out1 again                                %.3f ms" % run(o1))
del dummy; torch.cuda.empty_cache()
print("out2 after cudaFree                       %.3f ms" % run(o2))
o3 = torch.empty((1, 32, 9999, 15), pin_memory=True)
print("out3 (new pinned)                         %.3f ms" % run(o3))
print("out3, second loop                         %.3f ms" % run(o3))
x2 = x.clone().pin_memory()
def run_x(xx, out_h, n=40):
    global x
    keep = x; x = xx
    try: return run(out_h, n)
    finally: x = keep
print("new pinned INPUT, out2                    %.3f ms" % run_x(x2, o2))
