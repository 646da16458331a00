"""The single-launch forward+posterior kernel (csrc/gene15_fused.cu) against the fp64 oracle and against the two-kernel
path (products_kernel + finalize_kernel) on the same inputs: posterior-only calls take it whenever every tile of the
call fits the device at once."""
import numpy as np
import pytest
import torch

from oracle import hmm_oracle as O
from hmm_layer_b200 import ops, synthetic as syn
from tests.helpers import gene_params, oracle_gene
from tests.test_gpu_genepred import check_post, dev, run_gpu

pytestmark = pytest.mark.gpu


def both_paths(x, A, pi, Bfull, tab, **kw):
    fused = run_gpu(x, A, pi, Bfull, tab, want_post=True, **kw)
    ops.set_fused(False)
    try:
        two = run_gpu(x, A, pi, Bfull, tab, want_post=True, **kw)
    finally:
        ops.set_fused(True)
    return fused, two


@pytest.mark.parametrize("B,T", [(2, 1000), (3, 999), (1, 64), (2, 65), (4, 7), (1, 1), (2, 2), (5, 333), (1, 16), (1, 17), (3, 513),
                                 (7, 4099), (32, 2000)])
def test_fused_posterior_matches_oracle_and_two_kernel_path(B, T):
    x = syn.gene_inputs(1, B, T, seed=170 + T, p_n=0.01)
    A, pi, Bfull, tab = gene_params(pi_random=True)
    ref = oracle_gene(x.numpy(), A, pi, Bfull, tab)
    fused, two = both_paths(x, A, pi, Bfull, tab)
    check_post(fused, ref)
    check_post(two, ref)
    np.testing.assert_allclose(fused["loglik"], two["loglik"], rtol=1e-6)
    np.testing.assert_allclose(np.exp(fused["post"]), np.exp(two["post"]), atol=2e-5)


def test_fused_flags_hints_soft_nucleotides_and_models():
    B, T = 3, 1203
    x = syn.gene_inputs(2, B, T, seed=91, p_n=0.02)
    x[0, 0, 0, 15:] = torch.tensor([0., 0, 0, 0, 1])                 # N at the very first position
    x[0, 1, T - 1, 15:] = torch.tensor([0., 0, 0, 0, 1])             # and at the last
    x[1, 2, 600, 15:] = torch.tensor([.1, .2, .3, .4, 0.])           # soft nucleotide rows (neither one-hot nor N)
    x[1, 0, 4, 15:] = torch.tensor([.25, .25, .5, 0., 0.])
    x[0, 2, 5 * 37, 15:] = torch.tensor([.5, .5, 0., 0., 0.])        # on a run boundary of the emission phase
    A, pi, Bfull, tab = gene_params(num_models=2, pi_random=True)
    hints = np.random.RandomState(3).rand(2, B, 2, 15) + 0.01
    for kw in (dict(), dict(end_hints=hints), dict(training=True), dict(end_hints=hints, training=True, no_loglik=True)):
        ref = oracle_gene(x.numpy(), A, pi, Bfull, tab, **kw)
        fused, two = both_paths(x, A, pi, Bfull, tab, **kw)
        # no_loglik: the output is log gamma + loglik ~ -1.3e4, one fp32 ulp there is 1e-3 (the reference's fp32 output has
        # the same resolution), so probabilities recovered from it are only good to ~5e-4
        nl = bool(kw.get("no_loglik"))
        check_post(fused, ref, tol_dp=1e-3 if nl else 1e-4)
        check_post(two, ref, tol_dp=1e-3 if nl else 1e-4)
        if not nl:
            np.testing.assert_allclose(np.exp(fused["post"]), np.exp(two["post"]), atol=2e-5)


@pytest.mark.parametrize("L", [16, 23, 32, 35, 38])
def test_fused_chunk_lengths(L):
    """Every emission-run length (1 .. 5 positions per thread) and partial last tiles / chunks."""
    x = syn.gene_inputs(1, 3, 2 * 32 * L + 11, seed=7 + L, p_n=0.01)
    A, pi, Bfull, tab = gene_params()
    ref = oracle_gene(x.numpy(), A, pi, Bfull, tab)
    ops.set_tuning(L, 0)
    try:
        got = run_gpu(x, A, pi, Bfull, tab, want_post=True)
    finally:
        ops.set_tuning(0, 0)
    check_post(got, ref)


def test_fused_repeated_calls_are_bit_identical_and_config2_full_size():
    x = syn.gene_inputs(1, 32, 9999, seed=1234)
    A, pi, Bfull, tab = gene_params()
    got = run_gpu(x, A, pi, Bfull, tab, want_post=True)
    for _ in range(3):
        again = run_gpu(x, A, pi, Bfull, tab, want_post=True)
        assert np.array_equal(again["post"], got["post"]) and np.array_equal(again["loglik"], got["loglik"])
    ref = oracle_gene(x.numpy(), A, pi, Bfull, tab)
    rep = check_post(got, ref)
    assert rep["argmax_mismatch"] == 0


def test_run_under_cuda_graph_capture_replays_correctly():
    """hmm_b200_genepred_run captured into a CUDA graph and replayed on new inputs (the single-launch kernel's flags carry a
    per-launch tag that a graph would replay unchanged: a captured launch clears them first) -- posterior and
    log-likelihood of every replay equal the eager call's, for the single-launch route and the two-kernel route."""
    A, pi, Bfull, tab = [dev(v) for v in gene_params(pi_random=True)]
    tab5 = ops.prepare_codon(tab)
    B, T = 4, 6000
    xs = [syn.gene_inputs(1, B, T, seed=40 + i, p_n=0.01).cuda() for i in range(3)]
    for fused in (True, False):
        ops.set_fused(fused)
        try:
            eager = []
            for x in xs:
                r = ops.genepred_run(x, A, pi, Bfull, tab5, want_post=True)
                eager.append((r["post"].clone(), r["loglik"].clone()))
            static_x = xs[0].clone()
            out = {"post": torch.empty((1, B, T, 15), device="cuda")}
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                r = ops.genepred_run(static_x, A, pi, Bfull, tab5, want_post=True, out=out)
            for rep in range(2):
                for x, (post, ll) in zip(xs, eager):
                    static_x.copy_(x)
                    g.replay()
                    torch.cuda.synchronize()
                    assert torch.equal(r["post"], post), (fused, rep)
                    assert torch.equal(r["loglik"], ll), (fused, rep)
        finally:
            ops.set_fused(True)


def test_two_streams_run_the_single_launch_kernel_concurrently():
    """Two streams issue config-2-sized calls at the same time.  Every CTA of the single-launch kernel waits inside the
    kernel for the other tiles of its sequence, so a launch must be resident as a whole: the launches are cooperative
    (the driver places a grid completely or not yet), and concurrent callers neither deadlock nor disturb each other."""
    A, pi, Bfull, tab = [dev(v) for v in gene_params(pi_random=True)]
    tab5 = ops.prepare_codon(tab)
    B, T = 32, 9999
    xs = [syn.gene_inputs(1, B, T, seed=70 + i, p_n=0.01).cuda() for i in range(2)]
    ref = [ops.genepred_run(x, A, pi, Bfull, tab5, want_post=True) for x in xs]
    ref = [(r["post"].clone(), r["loglik"].clone()) for r in ref]
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    outs = [[None] * 6, [None] * 6]
    for it in range(6):
        for k, st in enumerate(streams):
            with torch.cuda.stream(st):     # (a workspace per stream: ops keys its cache by stream)
                outs[k][it] = ops.genepred_run(xs[k], A, pi, Bfull, tab5, want_post=True)
    torch.cuda.synchronize()
    for k in range(2):
        for it in range(6):
            assert torch.equal(outs[k][it]["post"], ref[k][0]), (k, it)
            assert torch.equal(outs[k][it]["loglik"], ref[k][1]), (k, it)


def test_refused_cooperative_launch_falls_back_to_the_two_kernel_path(monkeypatch):
    """When the device cannot hold the whole grid at once (reduced SM share under MPS, a smaller shared-memory carve-out) the
    cooperative launch is refused for size: the call then runs the two-kernel path, which needs no co-residency.  The refusal
    is simulated through the library's test hook."""
    A, pi, Bfull, tab = [dev(v) for v in gene_params(pi_random=True)]
    tab5 = ops.prepare_codon(tab)
    x = syn.gene_inputs(1, 6, 5000, seed=91, p_n=0.01).cuda()
    ops.set_fused(False)
    try:
        two = ops.genepred_run(x, A, pi, Bfull, tab5, want_post=True)
        two = (two["post"].clone(), two["loglik"].clone())
    finally:
        ops.set_fused(True)
    monkeypatch.setenv("HMM_B200_TEST_REFUSE_COOP", "1")
    got = ops.genepred_run(x, A, pi, Bfull, tab5, want_post=True)
    torch.cuda.synchronize()
    assert torch.equal(got["post"], two[0]) and torch.equal(got["loglik"], two[1])
    monkeypatch.delenv("HMM_B200_TEST_REFUSE_COOP")
    fused = ops.genepred_run(x, A, pi, Bfull, tab5, want_post=True)
    torch.cuda.synchronize()
    assert (fused["post"].exp() - two[0].exp()).abs().max() < 2e-5
