"""Why bench.py's e2e leg (1.08 ms) is slower than tools/e2e_layer_probe.py (0.81 ms): same layer call, after the
device-resident leg's allocations; per-call wall times in several situations."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from hmm_layer_b200 import ops, synthetic as syn

if len(sys.argv) > 1 and sys.argv[1] == "affinity":
    print("affinity:", bench.bind_to_gpu_numa_node(0), sorted(os.sched_getaffinity(0)))
dev = torch.device("cuda", 0)
layer, cell = bench.build_layer(dev)
x = syn.gene_inputs(1, 32, 9999, seed=1234).pin_memory()

def loop(n=60, tag=""):
    for _ in range(30):
        layer.state_posterior_log_probs(x)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n):
        out = layer.state_posterior_log_probs(x)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / n * 1e3
    import pynvml as nv
    nv.nvmlInit(); h = nv.nvmlDeviceGetHandleByIndex(0)
    print("%-40s %.3f ms   [SM %d MHz, mem %d MHz, pstate %s]" % (tag, dt, nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM),
          nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_MEM), nv.nvmlDeviceGetPerformanceState(h)), flush=True)

loop(tag="fresh process")
xs = [x.to(dev, non_blocking=True).clone() for _ in range(8)]
outs = [torch.empty((1, 32, 9999, 15), device=dev) for _ in range(8)]
loop(tag="after allocating 8 in/out sets")
em, tr = cell.emitter[0], cell.transitioner
A, pi = tr.A.detach().contiguous(), cell.init_dist.detach().reshape(1, 15).contiguous()
Bfull, codon = em.make_B_full().detach().contiguous(), ops.prepare_codon(em.codon_probs)
for i in range(300):
    ops.genepred_run(xs[i % 8], A, pi, Bfull, codon, want_post=True, out={"post": outs[i % 8]})
torch.cuda.synchronize()
loop(tag="after 300 device-resident steps")
host_out = torch.empty((1, 32, 9999, 15)).pin_memory()
loop(tag="after one more pinned allocation")
del xs, outs
torch.cuda.empty_cache()
loop(tag="after freeing the device sets")

# ---- split the slow state: allocation vs call ----
host_out2 = torch.empty((1, 32, 9999, 15)).pin_memory()     # provoke the slow state again
loop(tag="slow state?")
def t(fn, n=60):
    fn(); torch.cuda.synchronize(); t0 = time.perf_counter()
    keep = None
    for _ in range(n):
        keep = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
print("  pinned alloc per call (keep pattern)   %.3f ms" % t(lambda: torch.empty((1, 32, 9999, 15), pin_memory=True)))
gp = cell._gene_params
print("  ops.genepred_run_host                  %.3f ms" % t(lambda: ops.genepred_run_host(x, *gp)))
from hmm_layer_b200 import _lib
lib = _lib.load()
scratch = torch.empty(lib.hmm_b200_genepred_host_scratch_size(1, 32, 9999), dtype=torch.uint8, device="cuda")
ll_h = torch.empty((1, 32), pin_memory=True)
fixed = torch.empty((1, 32, 9999, 15), pin_memory=True)
st = torch.cuda.current_stream().cuda_stream
def raw():
    rc = lib.hmm_b200_genepred_run_host(x.data_ptr(), 1, 32, 9999, gp[0].data_ptr(), gp[1].data_ptr(), gp[2].data_ptr(), gp[3].data_ptr(), None, 0,
                                        fixed.data_ptr(), None, None, ll_h.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
    torch.cuda.current_stream().synchronize()
print("  raw C call, fixed pinned output        %.3f ms" % t(raw))
print("  raw C call, output = host_out2         %.3f ms" % t(lambda: (lib.hmm_b200_genepred_run_host(x.data_ptr(), 1, 32, 9999, gp[0].data_ptr(), gp[1].data_ptr(), gp[2].data_ptr(), gp[3].data_ptr(), None, 0,
                                        host_out2.data_ptr(), None, None, ll_h.data_ptr(), scratch.data_ptr(), scratch.numel(), st), torch.cuda.current_stream().synchronize())))
loop(tag="layer again")
print("  ops, keep only post                    %.3f ms" % t(lambda: ops.genepred_run_host(x, *gp)["post"]))
print("  ops, keep dict                         %.3f ms" % t(lambda: ops.genepred_run_host(x, *gp)))
def with_ri():
    cell.recurrent_init()
    return ops.genepred_run_host(x, *gp)
print("  recurrent_init + ops, keep dict        %.3f ms" % t(with_ri))
from hmm_layer_b200 import MsaHMMLayer as ML
print("  _gene_run                              %.3f ms" % t(lambda: ML._gene_run(x, cell, None, False, want_post=True, no_loglik=False)))
print("  _gene_run post only                    %.3f ms" % t(lambda: ML._gene_run(x, cell, None, False, want_post=True, no_loglik=False)["post"]))
print("  impl                                   %.3f ms" % t(lambda: ML._state_posterior_log_probs_impl(x, cell)))
print("  layer                                  %.3f ms" % t(lambda: layer.state_posterior_log_probs(x)))
