"""Where the end-to-end time of the host-buffer path goes (config 2): raw PCIe copies, pinned allocation, run_host."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, synthetic as syn, _lib
from tests.helpers import gene_params

x = syn.gene_inputs(1, 32, 9999, seed=1234).pin_memory()
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
codon = ops.prepare_codon(tab)
out_h = torch.empty((1, 32, 9999, 15)).pin_memory()
out_d = torch.empty((1, 32, 9999, 15), device="cuda")
xd = torch.empty_like(x, device="cuda")

def timeit(fn, n=20):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3

print("H2D 25.6 MB           %.3f ms" % timeit(lambda: xd.copy_(x, non_blocking=True)))
print("D2H 19.2 MB           %.3f ms" % timeit(lambda: out_h.copy_(out_d, non_blocking=True)))
s2 = torch.cuda.Stream()
def both():
    xd.copy_(x, non_blocking=True)
    with torch.cuda.stream(s2):
        out_h.copy_(out_d, non_blocking=True)
print("H2D || D2H            %.3f ms" % timeit(both))
print("pinned alloc 19.2 MB  %.3f ms" % timeit(lambda: torch.empty((1, 32, 9999, 15), pin_memory=True)))
print("device run            %.3f ms" % timeit(lambda: ops.genepred_run(xd, A, pi, Bfull, codon, out={"post": out_d})))
print("run_host (ops)        %.3f ms" % timeit(lambda: ops.genepred_run_host(x, A, pi, Bfull, codon)))
lib = _lib.load()
scratch = torch.empty(lib.hmm_b200_genepred_host_scratch_size(1, 32, 9999), dtype=torch.uint8, device="cuda")
ll_h = torch.empty((1, 32)).pin_memory()
st = torch.cuda.current_stream().cuda_stream
def raw():
    rc = lib.hmm_b200_genepred_run_host(x.data_ptr(), 1, 32, 9999, A.data_ptr(), pi.data_ptr(), Bfull.data_ptr(), codon.data_ptr(), None, 0,
                                        out_h.data_ptr(), None, None, ll_h.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
    assert rc == 0
print("run_host (C ABI, reused pinned output) %.3f ms" % timeit(raw))
def raw_ll():
    rc = lib.hmm_b200_genepred_run_host(x.data_ptr(), 1, 32, 9999, A.data_ptr(), pi.data_ptr(), Bfull.data_ptr(), codon.data_ptr(), None, 0,
                                        None, None, None, ll_h.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
    assert rc == 0
print("run_host loglik only (H2D + kernels)   %.3f ms" % timeit(raw_ll))
def raw_sync():
    raw(); torch.cuda.synchronize()
print("run_host + sync each call              %.3f ms" % timeit(raw_sync))
