"""One rank's time of the sequence-sharded path at config 5 (1 Mbp over `world` ranks; the all-gather is replaced by a
stack of this rank's own block) for several chunk lengths: python tools/c5_shard_time.py [world] [L list]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, sharded, synthetic as syn
from tests.helpers import gene_params

world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
Ls = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0, 16, 18, 20, 24, 27, 30, 35]
x = syn.gene_inputs(1, 1, 1_000_000, seed=1234)
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
codon = ops.prepare_codon(tab)
rank = world // 2
shard, halo, off, tot = sharded.slice_with_halo(x, rank, world)
shard, halo = shard.cuda(), halo.cuda()
def once():
    block, ws = ops.genepred_shard_reduce(shard, halo, off, tot, A, pi, Bfull, codon)
    allb = torch.stack([block] * world)
    return ops.genepred_shard_finalize(shard, halo, off, tot, A, pi, Bfull, codon, allb, rank, world, ws, want_post=True)
for L in Ls:
    ops.set_tuning(L, 0)
    for _ in range(5): once()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    n = 30
    for _ in range(n): once()
    e1.record(); torch.cuda.synchronize()
    print("world %d  L=%2d  %.1f us per rank pass" % (world, L, e0.elapsed_time(e1) / n * 1e3), flush=True)
ops.set_tuning(0, 0)
