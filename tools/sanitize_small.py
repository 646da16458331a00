"""Small invocations of every kernel family for compute-sanitizer (memcheck / racecheck / synccheck), e.g.
    compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import _lib, ops, sharded, synthetic as syn
from tests.helpers import gene_params

A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params(pi_random=True)]
tab5 = ops.prepare_codon(tab)
which = sys.argv[1:] or ["fused", "two", "groups", "viterbi", "shard", "dense"]
if "fused" in which:      # single-launch kernel, ragged tile ends
    x = syn.gene_inputs(1, 3, 2777, seed=1, p_n=0.02).cuda()
    r = ops.genepred_run(x, A, pi, Bfull, tab5, want_post=True)
    print("fused", float(r["loglik"].sum()))
if "two" in which:        # two-kernel path with alpha / beta
    x = syn.gene_inputs(1, 2, 1501, seed=2, p_n=0.02).cuda()
    r = ops.genepred_run(x, A, pi, Bfull, tab5, want_post=True, want_alpha=True, want_beta=True)
    print("two-kernel", float(r["loglik"].sum()))
if "groups" in which:     # long tile list: boundary_groups_kernel (chunk length 4 -> 128 positions per tile)
    ops.set_tuning(4, 0)
    x = syn.gene_inputs(1, 1, 20001, seed=3).cuda()
    r = ops.genepred_run(x, A, pi, Bfull, tab5, want_post=True, want_alpha=True)
    ops.set_tuning(0, 0)
    print("groups", float(r["loglik"].sum()))
if "viterbi" in which:
    x = syn.gene_inputs(1, 2, 2049, seed=4).cuda()
    p, s = ops.genepred_viterbi(x, A, pi, Bfull, tab5)
    x = syn.gene_inputs(1, 2, 300, seed=5).cuda()
    p2, s2 = ops.genepred_viterbi(x, A, pi, Bfull, tab5)
    print("viterbi", float(s.sum()), float(s2.sum()))
if "shard" in which:      # one-launch shard step on two virtual ranks of this device
    B, T, world = 1, 9001, 2
    x = syn.gene_inputs(1, B, T, seed=6).cuda()
    shards = [sharded.slice_with_halo(x, r, world) for r in range(world)]
    n_blk, n_flg = 2 * world * B * 256, 2 * world * B
    tables = [torch.zeros(n_blk + n_flg, dtype=torch.float32, device="cuda") for _ in range(world)]
    arr = ctypes.c_ulonglong * world
    pb, pf = arr(*[t.data_ptr() for t in tables]), arr(*[t.data_ptr() + 4 * n_blk for t in tables])
    lib = _lib.load()
    wss = []
    for r, (s, h, off, tot) in enumerate(shards):
        ws = torch.zeros(lib.hmm_b200_genepred_workspace_size(1, B, s.shape[2]), dtype=torch.uint8, device="cuda")
        ops.genepred_shard_reduce_x(s, h, off, tot, A, pi, Bfull, tab5, ws, 1, peer_blocks=pb, peer_flags=pf, rank=r, world=world)
        wss.append(ws)
    for r, (s, h, off, tot) in enumerate(shards):
        res = ops.genepred_shard_run_x(s, h, off, tot, A, pi, Bfull, tab5, wss[r], 1, pb, pf, r, world)
    print("shard", float(res["loglik"].sum()))
if "dense" in which:
    rng = np.random.default_rng(0)
    k, Bn, T, q = 1, 4, 40, 67
    E = torch.as_tensor(rng.uniform(0.01, 1.0, size=(k, Bn, T, q)).astype(np.float32)).cuda()
    Ad = torch.as_tensor(rng.dirichlet(np.ones(q), size=(k, q)).astype(np.float32)).cuda()
    pid = torch.as_tensor(rng.dirichlet(np.ones(q), size=(k,)).astype(np.float32)).cuda()
    r = ops.dense_run(E, Ad, pid, want_post=True)
    print("dense", float(r["loglik"].sum()))
torch.cuda.synchronize()
print("done")
