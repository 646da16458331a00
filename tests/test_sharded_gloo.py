"""Host logic of the sequence-sharded path on CPU with two gloo ranks: shard bounds, halo slicing, the single
all-gather, and the fold of the gathered per-shard transfer blocks into the whole-sequence log-likelihood and the
boundary vectors of every rank — with the per-shard matrices produced by the numpy model of the device algorithm
(tests/chunk_algo_model.py) and checked against the fp64 oracle.  No CUDA call is made here."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import hmm_oracle as O
from hmm_layer_b200 import sharded, synthetic as syn
from tests import chunk_algo_model as M
from tests.helpers import gene_params


def test_shard_bounds_and_halo():
    assert sharded.shard_bounds(10, 3) == [(0, 4), (4, 3), (7, 3)]
    assert sharded.shard_bounds(8, 8) == [(i, 1) for i in range(8)]
    with pytest.raises(ValueError):
        sharded.shard_bounds(3, 4)
    x = syn.gene_inputs(2, 3, 50, seed=1)
    pieces = [sharded.slice_with_halo(x, r, 4) for r in range(4)]
    assert torch.equal(torch.cat([p[0] for p in pieces], dim=2), x)
    nuc = x[..., 15:].reshape(6, 50, 5)
    for shard, halo, off, tot in pieces:
        assert tot == 50 and halo.shape == (6, 4, 5)
        n = shard.shape[2]
        for slot, t in enumerate((off - 2, off - 1, off + n, off + n + 1)):
            want = nuc[:, t] if 0 <= t < 50 else torch.zeros(6, 5)
            assert torch.equal(halo[:, slot], want)


def _shard_block(E, A, off, n):
    """Transfer matrix of positions [off, off+n) as one model block (chunk length = the shard)."""
    k1, k2 = M.position_coefficients(E, A)
    return M.block(*M.chunk_matrix(E, A, k1, k2, off, off + n - 1, off == 0))


def _worker(rank, world, port, T, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        x = syn.gene_inputs(1, 1, T, seed=99)
        A, pi, Bfull, tab = gene_params(pi_random=True)
        E = np.maximum(O.genepred_emission(x.numpy().astype(np.float64), Bfull, tab)[0, 0].astype(np.float32), M.EPS)
        shard, halo, off, tot = sharded.slice_with_halo(x, rank, world)
        blk = torch.as_tensor(_shard_block(E, A[0].astype(np.float32), off, shard.shape[2])).reshape(1, 256)
        allb = sharded.all_gather_blocks(blk)                       # the one collective
        assert allb.shape == (world, 1, 256)
        # fold (TotalProbabilityCell semantics across ranks): forward boundary entering this rank, total log-likelihood
        v, e, lg = pi[0].astype(np.float32), 0, 0.0
        entering = None
        for r in range(world):
            if r == rank:
                entering = (v.copy(), e, lg)
            v, e, m = M.fwd_matvec(v, e, allb[r, 0].numpy().reshape(16, 16))
            lg += float(m)
        ll = lg + e * np.log(2.0) + np.log(float(v.sum()))
        ret[rank] = (ll, entering, off)
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_fold_matches_oracle():
    T, world = 600, 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, T, ret), nprocs=world, join=True)
    x = syn.gene_inputs(1, 1, T, seed=99)
    A, pi, Bfull, tab = gene_params(pi_random=True)
    ref = O.posterior(O.genepred_emission(x.numpy().astype(np.float64), Bfull, tab), A, pi)
    for rank in range(world):
        ll, (v, e, lg), off = ret[rank]
        assert abs(ll - ref["loglik"][0, 0]) / abs(ref["loglik"][0, 0]) < 1e-6
        if off > 0:
            # entering vector = alpha at the last position of the previous shard, up to scale
            la = ref["log_alpha"][0, 0, off - 1]
            p_ref = np.exp(la - la.max()); p_ref /= p_ref.sum()
            p = v / v.sum()
            np.testing.assert_allclose(p, p_ref, atol=1e-5)
    assert ret[0][0] == ret[1][0]
