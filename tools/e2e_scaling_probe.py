"""Which resource bounds the end-to-end (host-buffer) path when N ranks run at once?  Launch with torch.distributed.run.
Every rank, simultaneously: (i) H2D only (25.6 MB pinned -> device), (ii) D2H only (19.2 MB), (iii) both at once on two
streams, (iv) the layer's host path (fp32 layout), (v) the compact wire format host path (9.9 MB up).  Prints per-rank
time (max over ranks) and the aggregate GB/s; rank 0 first runs the same alone (N = 1 reference on the same box)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import hmm_layer_b200 as H
from hmm_layer_b200 import ops, synthetic as syn

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
k, B, T = 1, 32, 9999
x = syn.gene_inputs(k, B, T, seed=1234 + rank).pin_memory()
xd = torch.empty_like(x, device=dev)
od = torch.empty((k, B, T, 15), device=dev)
oh = torch.empty((k, B, T, 15)).pin_memory()
em = H.GenePredHMMEmitter(**syn.GENE_CODONS); em.build((1, B, T, 15))
with torch.no_grad():
    em.emission_kernel.copy_(syn.gene_emission_kernel())
tr = H.GenePredMultiHMMTransitioner(initial_exon_len=200, initial_intron_len=4500, initial_ir_len=10000)
cell = H.HmmCell([15], 15, em, tr).to(dev)
layer = H.MsaHmmLayer(cell, use_prior=False)
cell.recurrent_init()
A, pi, Bf, cod = tr.A.detach().contiguous(), cell.init_dist.detach().reshape(1, 15).contiguous(), em.make_B_full().detach().contiguous(), em.codon5(dev)
cls16, codes = ops.pack_compact(x)
cls16, codes = cls16.pin_memory(), codes.pin_memory()
s2 = torch.cuda.Stream()

def timed(fn, n=30, group=True):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    if group and world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n], dtype=torch.float64, device=dev)
    if group and world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())

def both():
    xd.copy_(x, non_blocking=True)
    with torch.cuda.stream(s2):
        oh.copy_(od, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s2)

cases = [("H2D 25.6 MB", lambda: xd.copy_(x, non_blocking=True), 25.6), ("D2H 19.2 MB", lambda: oh.copy_(od, non_blocking=True), 19.2),
         ("H2D + D2H at once", both, 44.8), ("layer host path (fp32 layout)", lambda: layer.state_posterior_log_probs(x), 44.8),
         ("compact host path (9.9 MB up)", lambda: ops.genepred_run_compact(cls16, codes, A, pi, Bf, cod, want_post=True), 29.1)]
if world > 1 and rank == 0:
    print("--- rank 0 alone (others idle)")
    for name, fn, mb in cases:
        ms = timed(fn, group=False)
        print(f"   {name:32s} {ms:7.3f} ms  {mb / ms:6.1f} GB/s")
if world > 1:
    dist.barrier()
if rank == 0:
    print(f"--- {world} ranks at once (max over ranks; aggregate GB/s)")
for name, fn, mb in cases:
    ms = timed(fn)
    if rank == 0:
        print(f"   {name:32s} {ms:7.3f} ms  {world * mb / ms:6.1f} GB/s")
if rank == 0:
    try:
        print("numa nodes of the GPUs:", sorted(set(open(f).read().strip() for f in __import__("glob").glob("/sys/bus/pci/devices/*/numa_node"))), "cpus:", os.cpu_count())
    except Exception as e:
        print(e)
if world > 1:
    dist.destroy_process_group()
