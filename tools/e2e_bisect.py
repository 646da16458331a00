"""Which part of bench.py's flow puts the host-buffer layer call into its slow mode (1.09 ms instead of 0.80 ms)?
usage: python tools/e2e_bisect.py [skip flags: sampler spin device sets pinned events dist]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from hmm_layer_b200 import ops, synthetic as syn

skip = set(sys.argv[1:])
if "dist" not in skip:
    import torch.distributed as dist  # noqa: F401
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
layer, cell = bench.build_layer(dev)
em, tr = cell.emitter[0], cell.transitioner
cell.recurrent_init()
A, pi = tr.A.detach().contiguous(), cell.init_dist.detach().reshape(1, 15).contiguous()
Bfull, codon = em.make_B_full().detach().contiguous(), ops.prepare_codon(em.codon_probs)
host_in = syn.gene_inputs(1, 32, 9999, seed=1234).pin_memory()
NBUF = 1 if "sets" in skip else 8
xs = [host_in.to(dev, non_blocking=True).clone() for _ in range(NBUF)]
outs = [torch.empty((1, 32, 9999, 15), dtype=torch.float32, device=dev) for _ in range(NBUF)]
if "pinned" not in skip:
    host_out = torch.empty((1, 32, 9999, 15), dtype=torch.float32).pin_memory()
    host_ll = torch.empty((1, 32), dtype=torch.float32).pin_memory()
def step(i):
    j = i % NBUF
    return ops.genepred_run(xs[j], A, pi, Bfull, codon, want_post=True, out={"post": outs[j]})
if "sampler" not in skip:
    sampler = bench.ClockSampler(0); sampler.start()
if "spin" not in skip:
    t_end = time.perf_counter() + 0.3
    i = 0
    while time.perf_counter() < t_end:
        step(i); i += 1
    torch.cuda.synchronize()
if "device" not in skip:
    for i in range(220):
        step(i)
    torch.cuda.synchronize()
for _ in range(30):
    layer.state_posterior_log_probs(host_in)
n = 50
if "events" in skip:
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n):
        host_post = layer.state_posterior_log_probs(host_in)
    torch.cuda.synchronize(); ms = (time.perf_counter() - t0) / n * 1e3
else:
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); f0.record()
    for _ in range(n):
        host_post = layer.state_posterior_log_probs(host_in)
    f1.record(); torch.cuda.synchronize(); ms = f0.elapsed_time(f1) / n
print("skip=%-40s e2e %.3f ms" % (",".join(sorted(skip)) or "-", ms), flush=True)
os._exit(0)
