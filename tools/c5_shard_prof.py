"""One rank's kernels of the sequence-sharded path at config 5 (1 Mbp over `world` ranks), for an ncu launch list:
python tools/c5_shard_prof.py [world] [reps]   (the all-gather is replaced by a stack of this rank's own block)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, sharded, synthetic as syn
from tests.helpers import gene_params

world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
x = syn.gene_inputs(1, 1, 1_000_000, seed=1234)
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
codon = ops.prepare_codon(tab)
rank = world // 2
shard, halo, off, tot = sharded.slice_with_halo(x, rank, world)
shard, halo = shard.cuda(), halo.cuda()
for _ in range(reps):
    block, ws = ops.genepred_shard_reduce(shard, halo, off, tot, A, pi, Bfull, codon)
    allb = torch.stack([block] * world)
    r = ops.genepred_shard_finalize(shard, halo, off, tot, A, pi, Bfull, codon, allb, rank, world, ws, want_post=True)
torch.cuda.synchronize()
print("ok", tuple(r["post"].shape))
