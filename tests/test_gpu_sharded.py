"""GPU parity of the sequence-sharded path (BASELINE config 5 shape class): one sequence cut into contiguous shards,
one stored block per shard exchanged, every shard finalised on its own.  Virtual ranks on one device run the same
kernels in the same order as the multi-rank path (hmm_layer_b200/sharded.py); the all-gather is a stack."""
import numpy as np
import pytest
import torch

from oracle import hmm_oracle as O
from hmm_layer_b200 import ops, sharded, synthetic as syn
from tests.helpers import gene_params, oracle_gene
from tests.test_gpu_genepred import check_ab, check_post, dev, run_gpu

pytestmark = pytest.mark.gpu


def run_sharded(x, world, A, pi, Bfull, tab, **kw):
    xd = x.cuda()
    shards = [sharded.slice_with_halo(xd, r, world) for r in range(world)]
    hints = kw.pop("end_hints", None)
    if hints is not None:
        kw["end_hints"] = dev(hints)
    res = sharded.sequence_sharded_run_local(shards, dev(A), dev(pi), dev(Bfull), ops.prepare_codon(dev(tab)), **kw)
    torch.cuda.synchronize()
    out = {k: np.concatenate([r[k].cpu().numpy() for r in res], axis=2) for k in res[0] if k != "loglik"}
    lls = [r["loglik"].cpu().numpy() for r in res]
    for ll in lls[1:]:
        assert np.array_equal(ll, lls[0])          # the log-likelihood of the whole sequence, identical on every rank
    out["loglik"] = lls[0]
    return out


@pytest.mark.parametrize("B,T,world", [(2, 4000, 2), (1, 4001, 3), (3, 2500, 8), (2, 64, 4), (1, 17, 8), (2, 1000, 1)])
def test_sharded_matches_oracle_and_unsharded(B, T, world):
    x = syn.gene_inputs(1, B, T, seed=500 + T, p_n=0.01)
    A, pi, Bfull, tab = gene_params(pi_random=True)
    ref = oracle_gene(x.numpy(), A, pi, Bfull, tab)
    got = run_sharded(x, world, A, pi, Bfull, tab, want_post=True, want_alpha=True, want_beta=True)
    check_post(got, ref)
    check_ab(got, ref)
    one = run_gpu(x, A, pi, Bfull, tab, want_post=True)
    np.testing.assert_allclose(got["loglik"], one["loglik"], rtol=1e-6)
    np.testing.assert_allclose(np.exp(got["post"]), np.exp(one["post"]), atol=2e-5)


def test_sharded_hints_flags_and_n_at_shard_edges():
    """end_hints belong to the first / last position of the WHOLE sequence; N and soft nucleotides sit exactly on the
    shard edges so that the 3-mer windows of both neighbours read them through the halo."""
    B, T, world = 2, 1203, 4
    x = syn.gene_inputs(1, B, T, seed=8, p_n=0.0)
    for off, n in sharded.shard_bounds(T, world):
        x[0, 0, max(off - 1, 0), 15:] = torch.tensor([0., 0, 0, 0, 1])
        x[0, 1, min(off + n, T - 1), 15:] = torch.tensor([.1, .2, .3, .4, 0.])
    A, pi, Bfull, tab = gene_params()
    hints = np.random.RandomState(1).rand(1, B, 2, 15) + 0.01
    for kw in (dict(end_hints=hints), dict(training=True), dict(end_hints=hints, training=True)):
        ref = oracle_gene(x.numpy(), A, pi, Bfull, tab, **kw)
        got = run_sharded(x, world, A, pi, Bfull, tab, want_post=True, **kw)
        check_post(got, ref)


def test_config5_shape_properties():
    """One ~1 Mbp sequence over 8 shards: the fp64 oracle on the WHOLE sequence (posterior at every position, argmax,
    log-likelihood; log alpha on a prefix), normalisation everywhere, loglik equal to the unsharded evaluation."""
    T, world = 999_900, 8
    x = syn.gene_inputs(1, 1, T, seed=1234)
    A, pi, Bfull, tab = gene_params()
    got = run_sharded(x, world, A, pi, Bfull, tab, want_post=True, want_alpha=True)
    assert np.isfinite(got["post"]).all()
    np.testing.assert_allclose(np.exp(got["post"].astype(np.float64)).sum(-1), 1.0, atol=1e-4)
    one = run_gpu(x, A, pi, Bfull, tab, want_post=True)
    np.testing.assert_allclose(got["loglik"], one["loglik"], rtol=1e-6)
    assert (np.argmax(got["post"], -1) != np.argmax(one["post"], -1)).mean() < 1e-5
    ref = oracle_gene(x.numpy(), A, pi, Bfull, tab)
    for res in (got, one):
        rep = O.parity_report(res["post"], res["loglik"], ref["log_post"], ref["loglik"])
        assert rep["argmax_mismatch"] == 0 and rep["max_abs_dp"] < 1e-4 and rep["loglik_rel"] < 1e-4, rep
    n = 3000
    r, g = ref["log_alpha"][:, :, :n], got["alpha"][:, :, :n]
    big = r - r.max(-1, keepdims=True) > -14
    assert (np.abs(g - r) / np.maximum(1.0, np.abs(r)))[big].max() < 1e-4


def test_shard_argument_validation():
    A, pi, Bfull, tab = gene_params()
    x = syn.gene_inputs(1, 1, 100).cuda()
    with pytest.raises(ValueError):
        ops.genepred_shard_reduce(x[:, :, :50].contiguous(), None, 0, 100, dev(A), dev(pi), dev(Bfull), dev(tab))   # no halo
    with pytest.raises(ValueError):
        ops.genepred_shard_reduce(x, None, 10, 100, dev(A), dev(pi), dev(Bfull), dev(tab))                          # outside


def test_two_level_boundary_scan_long_tile_lists():
    """More than 96 tiles per sequence (here forced with chunk length 4: 128 positions per tile) take the two-level
    boundary scan (group_reduce_kernel / boundary_kernel on the group matrices / group_down_kernel), unsharded and
    sharded (config 5's route at 1 Mbp)."""
    A, pi, Bfull, tab = gene_params(pi_random=True)
    ops.set_tuning(4, 0)
    try:
        x = syn.gene_inputs(1, 2, 20011, seed=77, p_n=0.01)            # 157 tiles per sequence
        ref = oracle_gene(x.numpy(), A, pi, Bfull, tab)
        got = run_gpu(x, A, pi, Bfull, tab, want_post=True, want_alpha=True, want_beta=True)
        check_post(got, ref)
        check_ab(got, ref)
        only = run_gpu(x, A, pi, Bfull, tab, want_post=False)
        np.testing.assert_allclose(only["loglik"], got["loglik"], rtol=1e-6)
        xs = syn.gene_inputs(1, 1, 30000, seed=78, p_n=0.01)           # 2 shards x 118 tiles
        refs = oracle_gene(xs.numpy(), A, pi, Bfull, tab)
        gs = run_sharded(xs, 2, A, pi, Bfull, tab, want_post=True)
        check_post(gs, refs)
    finally:
        ops.set_tuning(0, 0)
    same = run_gpu(x, A, pi, Bfull, tab, want_post=True)                # default chunk length: single-level scan
    np.testing.assert_allclose(same["loglik"], got["loglik"], rtol=1e-6)
    np.testing.assert_allclose(np.exp(same["post"]), np.exp(got["post"]), atol=2e-5)


def _virtual_tables(world, NB):
    """Block / flag tables of `world` virtual ranks on this device (the layout ShardContext keeps in symmetric memory:
    [2 halves][world][NB] blocks of 256 floats, then the same of flags), as ctypes pointer arrays in rank order."""
    import ctypes
    n_blk, n_flg = 2 * world * NB * 256, 2 * world * NB
    tables = [torch.zeros(n_blk + n_flg, dtype=torch.float32, device="cuda") for _ in range(world)]
    arr = ctypes.c_ulonglong * world
    return tables, arr(*[t.data_ptr() for t in tables]), arr(*[t.data_ptr() + 4 * n_blk for t in tables])


@pytest.mark.parametrize("B,T,world,hints", [(2, 4000, 2, False), (1, 40_001, 3, True), (3, 2500, 8, False), (1, 150_000, 2, False),
                                             (2, 64, 4, True), (2, 1000, 1, False)])
def test_one_launch_shard_step_matches_oracle(B, T, world, hints):
    """hmm_b200_genepred_shard_run_x on virtual ranks of one device: the four-launch route first publishes every rank's
    block for the step tag (so no launch waits for a kernel that is not running), then each rank's one-launch step of the
    SAME tag re-publishes its block and folds ranks -> groups -> tiles inside the kernel.  150 000 positions on two ranks =
    several tile groups per rank (the cross-group fold); T = 64 on 4 ranks = shards shorter than a chunk."""
    x = syn.gene_inputs(1, B, T, seed=700 + T, p_n=0.01)
    A, pi, Bfull, tab = gene_params(pi_random=True)
    eh = None
    if hints:
        rng = np.random.default_rng(5)
        eh = rng.uniform(0.2, 1.0, size=(1, B, 2, 15)).astype(np.float32)
    ref = oracle_gene(x.numpy(), A, pi, Bfull, tab, end_hints=eh)
    Ad, pid, Bd, tabd = dev(A), dev(pi), dev(Bfull), ops.prepare_codon(dev(tab))
    ehd = dev(eh) if eh is not None else None
    xd = x.cuda()
    shards = [sharded.slice_with_halo(xd, r, world) for r in range(world)]
    tables, pb, pf = _virtual_tables(world, B)
    import hmm_layer_b200._lib as L
    lib = L.load()
    posts, lls = [None] * world, [None] * world
    for tag in (1, 2, 3):                       # both table halves, stale flags of tag - 2 in place
        wss = []
        for r, (s, h, off, tot) in enumerate(shards):
            ws = torch.zeros(max(lib.hmm_b200_genepred_workspace_size(1, B, s.shape[2]), 256), dtype=torch.uint8, device="cuda")
            ops.genepred_shard_reduce_x(s, h, off, tot, Ad, pid, Bd, tabd, ws, tag, peer_blocks=pb, peer_flags=pf, rank=r,
                                        world=world, end_hints=ehd)
            wss.append(ws)
        for r, (s, h, off, tot) in enumerate(shards):
            if ops.genepred_shard_fused_chunk(1, B, s.shape[2], world) == 0:
                pytest.skip("shard does not fit one launch on this device")
            res = ops.genepred_shard_run_x(s, h, off, tot, Ad, pid, Bd, tabd, wss[r], tag, pb, pf, r, world, end_hints=ehd)
            posts[r], lls[r] = res["post"], res["loglik"]
        torch.cuda.synchronize()
    got = {"post": np.concatenate([p.cpu().numpy() for p in posts], axis=2), "loglik": lls[0].cpu().numpy()}
    for ll in lls[1:]:
        assert np.array_equal(ll.cpu().numpy(), got["loglik"])
    check_post(got, ref)


def test_one_launch_shard_step_rejects_what_does_not_fit():
    from hmm_layer_b200._lib import HmmB200Error
    assert ops.genepred_shard_fused_chunk(1, 1, 2_000_000, 1) == 0
    B, T = 1, 2_000_000
    x = torch.zeros((1, B, T, 20), device="cuda")
    A, pi, Bfull, tab = gene_params()
    tables, pb, pf = _virtual_tables(1, B)
    import hmm_layer_b200._lib as L
    ws = torch.zeros(L.load().hmm_b200_genepred_workspace_size(1, B, T), dtype=torch.uint8, device="cuda")
    with pytest.raises(HmmB200Error, match="do not fit one launch"):
        ops.genepred_shard_run_x(x, None, 0, T, dev(A), dev(pi), dev(Bfull), ops.prepare_codon(dev(tab)), ws, 1, pb, pf, 0, 1)
