"""Tests that need at least two GPUs (skipped otherwise; `gpurun --gpus 2`):
  * the sequence-sharded path on REAL ranks over NCCL / NVLink: every shard's posterior against the fp64 oracle, for
    the exchange routes (peer stores into symmetric memory inside the kernels, as ONE launch per step where the shard
    fits and as four launches; the NCCL all-gather fallback) and the plain three-call route (shard_reduce / all_gather /
    shard_finalize);
  * one process driving two devices in turn (the library's per-device state)."""
import os
import socket

import numpy as np
import pytest
import torch

from oracle import hmm_oracle as O
from hmm_layer_b200 import ops, sharded, synthetic as syn
from tests.helpers import gene_params, oracle_gene

pytestmark = pytest.mark.gpu
needs2 = pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two CUDA devices")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, B, T, seed, outdir, chunk_len):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    if chunk_len:
        ops.set_tuning(chunk_len, 0)
    x = syn.gene_inputs(1, B, T, seed=seed, p_n=0.01)
    A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).to(dev) for v in gene_params(pi_random=True)]
    shard, halo, off, tot = sharded.slice_with_halo(x, rank, world)
    shard, halo = shard.to(dev), halo.to(dev)
    res = {}
    for name, peer, one in (("peer", True, True), ("peer4", True, False), ("nccl", False, False)):
        ctx = sharded.ShardContext(shard, halo, off, tot, A, pi, Bfull, tab, use_peer_memory=peer, one_launch=one)
        for _ in range(3):                      # repeated steps: tags, table halves, the self-resetting counter
            r = ctx.run(shard, want_post=True)
        res[name] = (r["post"].cpu().numpy(), r["loglik"].cpu().numpy(), ctx.exchange, ctx.launches_per_step)
    r = sharded.sequence_sharded_run(shard, halo, off, tot, A, pi, Bfull, tab, want_post=True)
    res["plain"] = (r["post"].cpu().numpy(), r["loglik"].cpu().numpy(), "three calls", 8)
    torch.cuda.synchronize()
    np.savez(os.path.join(outdir, f"rank{rank}.npz"), **{f"{k}_{i}": v[i] for k, v in res.items() for i in (0, 1)},
             info=np.array([f"{k}: {v[2]} ({v[3]} launches)" for k, v in res.items()]))
    dist.barrier()
    dist.destroy_process_group()


@needs2
@pytest.mark.parametrize("B,T,chunk_len,max_world", [(2, 40_000, 0, 8), (1, 300_001, 0, 8), (1, 300_001, 4, 2)])
def test_sequence_sharded_on_real_ranks_matches_oracle(tmp_path, B, T, chunk_len, max_world):
    """chunk_len = 4 on two ranks: 1172 tiles per rank, beyond the 32 x 32 tiles the in-kernel scans cover -- the
    general boundary kernels do the scans, the exchange stays in-kernel."""
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), max_world)
    port = _free_port()
    mp.spawn(_rank_main, args=(world, port, B, T, 900 + B, str(tmp_path), chunk_len), nprocs=world, join=True)
    x = syn.gene_inputs(1, B, T, seed=900 + B, p_n=0.01)
    A, pi, Bfull, tab = gene_params(pi_random=True)
    ref = oracle_gene(x.numpy(), A, pi, Bfull, tab)
    parts = [dict(np.load(tmp_path / f"rank{r}.npz")) for r in range(world)]
    print(parts[0]["info"])
    for route in ("peer", "peer4", "nccl", "plain"):
        post = np.concatenate([p[f"{route}_0"] for p in parts], axis=2)
        for p in parts:
            np.testing.assert_array_equal(p[f"{route}_1"], parts[0][f"{route}_1"])      # the same log-likelihood on every rank
        rep = O.parity_report(post, parts[0][f"{route}_1"], ref["log_post"], ref["loglik"])
        assert rep["argmax_mismatch"] == 0 and rep["max_abs_dp"] < 1e-4 and rep["loglik_rel"] < 2e-6, (route, rep)


@needs2
def test_one_process_two_devices():
    """genepred_run / dense_run on cuda:0, then cuda:1, then cuda:0 again from one process: the raised shared-memory
    limits, SM counts and profiling events are kept per device."""
    A, pi, Bfull, tab = gene_params()
    x = syn.gene_inputs(1, 4, 3000, seed=3)
    ref = oracle_gene(x.numpy(), A, pi, Bfull, tab)
    Ad, pid, Bm = syn.profile_model(1, 40)
    xd = syn.profile_inputs(1, 4, 50)
    E = torch.einsum("kbts,kqs->kbtq", xd, Bm)
    refd = O.posterior(E.numpy().astype(np.float64), Ad.numpy().astype(np.float64), pid.numpy().astype(np.float64))
    for d in (0, 1, 0):
        dev = torch.device("cuda", d)
        t = lambda v: torch.as_tensor(np.asarray(v, np.float32)).to(dev)
        for kw in (dict(want_post=True), dict(want_post=True, want_alpha=True, want_beta=True)):      # both kernel families
            r = ops.genepred_run(x.to(dev), t(A), t(pi), t(Bfull), t(tab), **kw)
            rep = O.parity_report(r["post"].cpu().numpy(), r["loglik"].cpu().numpy(), ref["log_post"], ref["loglik"])
            assert rep["argmax_mismatch"] == 0 and rep["max_abs_dp"] < 1e-4 and rep["loglik_rel"] < 2e-6, (d, rep)
        rd = ops.dense_run(E.to(dev), Ad.to(dev), pid.to(dev), want_post=True)
        rep = O.parity_report(rd["post"].cpu().numpy(), rd["loglik"].cpu().numpy(), refd["log_post"], refd["loglik"])
        assert rep["max_abs_dp"] < 1e-4 and rep["loglik_rel"] < 2e-6, (d, rep)
