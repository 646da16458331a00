"""GPU parity of the gene-HMM log-likelihood gradient (hmm_b200_genepred_grad through ops.GenePredLoglik and
MsaHmmLayer.forward) against torch autograd over the fp64 CPU port of the reference path (oracle/ref_port.py: the
reference itself has no backward code, autograd traces its Python loop — MsaHmmCell.py:73-106).

Tolerance: every gradient tensor within 1e-4 of the oracle's, relative to that tensor's largest magnitude (the kernels
read fp32 log posteriors; the log-likelihood itself is held to 1e-4 relative elsewhere).  At BASELINE config-2 size the
check is through size-independent identities of expected counts."""
import numpy as np
import pytest
import torch

from oracle import ref_port as RP
from hmm_layer_b200 import ops, synthetic as syn
import hmm_layer_b200 as H
from hmm_layer_b200.gene_pred_hmm_transitioner import make_transition_matrix_from_indices
from tests.helpers import gene_params

pytestmark = pytest.mark.gpu

EDGE_SRC = [0, 0, 14, 7, 5, 4, 4, 8, 1, 1, 11, 5, 5, 9, 2, 2, 12, 6, 6, 10, 3, 3, 13]
EDGE_DST = [0, 7, 0, 5, 14, 5, 8, 1, 1, 11, 4, 6, 9, 2, 2, 12, 5, 4, 10, 3, 3, 13, 6]


def dev(x):
    return torch.as_tensor(np.asarray(x, np.float32)).cuda()


def b13_of(Bfull):
    """(k,15,15) rows per state -> the 13 kernel rows (introns share one row, gene_pred_hmm_emitter.py:115-116)."""
    return np.concatenate([Bfull[:, :2], Bfull[:, 4:]], axis=1)


def fold_b13(gB):
    return np.concatenate([gB[:, :1], gB[:, 1:4].sum(1, keepdims=True), gB[:, 4:]], axis=1)


def oracle_grads(x, A, pi, Bfull, tab, w, end_hints=None, training=False):
    """fp64 autograd over the reference port: d sum(w * loglik) / d (class probabilities, A, pi, B13)."""
    xt = torch.tensor(np.asarray(x, np.float64), requires_grad=True)
    At = torch.tensor(A, dtype=torch.float64, requires_grad=True)
    pit = torch.tensor(pi, dtype=torch.float64, requires_grad=True)
    Bt = torch.tensor(b13_of(Bfull), dtype=torch.float64, requires_grad=True)
    hints = None if end_hints is None else torch.tensor(end_hints, dtype=torch.float64)
    E = RP.genepred_emission(xt, Bt, torch.tensor(tab, dtype=torch.float64), hints, training)
    k, b, T, q = E.shape
    f0 = pit.reshape(k, 1, q).expand(k, b, q).reshape(-1, q)
    _, ll = RP.scan(E.reshape(k * b, T, q), f0, At, k, q, False)
    ll = ll.reshape(k, b)
    (ll * torch.tensor(w, dtype=torch.float64)).sum().backward()
    return dict(loglik=ll.detach().numpy(), x=xt.grad.numpy(), A=At.grad.numpy(), pi=pit.grad.numpy(), B13=Bt.grad.numpy())


def close(got, ref, tol=1e-4, what=""):
    scale = np.abs(ref).max()
    err = np.abs(got - ref).max() / scale
    assert np.isfinite(got).all(), what
    assert err < tol, (what, err, scale)
    return err


@pytest.mark.parametrize("B,T,training", [(3, 400, False), (2, 257, True), (4, 3, False), (2, 2, False), (1, 1300, False)])
def test_grad_matches_fp64_autograd(B, T, training):
    x = syn.gene_inputs(1, B, T, seed=500 + T)
    A, pi, Bfull, tab = gene_params(pi_random=True)
    w = np.random.RandomState(T).uniform(0.5, 1.5, size=(1, B))
    ref = oracle_grads(x.numpy(), A, pi, Bfull, tab, w, training=training)
    xg = x.cuda().requires_grad_(True)
    Ag, pig, Bg = dev(A).requires_grad_(True), dev(pi).requires_grad_(True), dev(Bfull).requires_grad_(True)
    ll = ops.GenePredLoglik.apply(xg, Ag, pig, Bg, dev(tab), None, training)
    (ll * dev(w)).sum().backward()
    torch.cuda.synchronize()
    assert np.abs(ll.detach().cpu().numpy() - ref["loglik"]).max() / np.abs(ref["loglik"]).max() < 1e-5
    gx = xg.grad.cpu().numpy()
    assert (gx[..., 15:] == 0).all()            # nucleotide flags are data
    close(gx[..., :15], ref["x"][..., :15], what="d/d class probabilities")
    e = (0, EDGE_SRC, EDGE_DST)
    gA = Ag.grad.cpu().numpy()
    if T > 1:
        close(gA[e], ref["A"][e], what="d/dA on the 23 edges")
    mask = np.ones_like(gA, bool); mask[e] = False
    assert (gA[mask] == 0).all()
    close(pig.grad.cpu().numpy(), ref["pi"], what="d/dpi")
    close(fold_b13(Bg.grad.cpu().numpy()), ref["B13"], what="d/dB")


def test_grad_with_end_hints():
    B, T = 3, 300
    x = syn.gene_inputs(1, B, T, seed=901)
    A, pi, Bfull, tab = gene_params()
    hints = np.random.RandomState(5).uniform(0.2, 1.0, size=(1, B, 2, 15))
    w = np.ones((1, B))
    ref = oracle_grads(x.numpy(), A, pi, Bfull, tab, w, end_hints=hints)
    xg = x.cuda().requires_grad_(True)
    Ag = dev(A).requires_grad_(True)
    ll = ops.GenePredLoglik.apply(xg, Ag, dev(pi), dev(Bfull), dev(tab), dev(hints), False)
    ll.sum().backward()
    close(xg.grad.cpu().numpy()[..., :15], ref["x"][..., :15], what="d/dx with hints")
    close(Ag.grad.cpu().numpy()[(0, EDGE_SRC, EDGE_DST)], ref["A"][(0, EDGE_SRC, EDGE_DST)], what="d/dA with hints")


def make_layer():
    em = H.GenePredHMMEmitter(**syn.GENE_CODONS)
    em.build((1, 1, 1, 15))
    with torch.no_grad():
        em.emission_kernel.copy_(syn.gene_emission_kernel())
    np.random.seed(0)
    tr = H.GenePredMultiHMMTransitioner(initial_exon_len=200, initial_intron_len=4500, initial_ir_len=10000)
    with torch.no_grad():
        tr.transition_kernel.add_(1e-3)
        tr.starting_distribution_kernel.copy_(torch.randn(1, 1, 15, generator=torch.Generator().manual_seed(3)))
    cell = H.HmmCell([15], 15, em, tr).cuda()
    return H.MsaHmmLayer(cell, use_prior=False), em, tr


def test_layer_parameter_gradients_match_oracle_chain():
    """loglik_mean.backward() through MsaHmmLayer.forward: parameter gradients equal fp64 autograd through the same
    parameter -> (A, pi, B) construction followed by the reference port."""
    B, T = 4, 350
    layer, em, tr = make_layer()
    x = syn.gene_inputs(1, B, T, seed=77)
    for _ in range(2):   # second pass: a fresh graph per call, gradients accumulate like any nn.Module
        loglik, mean = layer(x.cuda())
        mean.backward()
    torch.cuda.synchronize()
    assert loglik.requires_grad
    # oracle chain in fp64
    ek = em.emission_kernel.detach().cpu().double().requires_grad_(True)
    tk = tr.transition_kernel.detach().cpu().double().requires_grad_(True)
    sk = tr.starting_distribution_kernel.detach().cpu().double().requires_grad_(True)
    A = make_transition_matrix_from_indices(tr.indices[:, 1:], tk.reshape(-1), 15).unsqueeze(0)
    pi = torch.softmax(sk, -1).reshape(1, 15)
    B13 = torch.softmax(ek, -1)
    E = RP.genepred_emission(x.double(), B13, em.codon_probs.cpu().double())
    _, ll = RP.scan(E.reshape(B, T, 15), pi.expand(B, 15), A, 1, 15, False)
    ll.mean().backward()
    close(em.emission_kernel.grad.cpu().numpy() / 2, ek.grad.numpy(), what="emission kernel")
    close(tr.transition_kernel.grad.cpu().numpy() / 2, tk.grad.numpy(), what="transition kernel")
    close(tr.starting_distribution_kernel.grad.cpu().numpy() / 2, sk.grad.numpy(), what="starting distribution kernel")
    # inference calls keep using the cached, detached parameters
    with torch.no_grad():
        l2, _ = layer(x.cuda())
    assert not l2.requires_grad
    assert torch.allclose(l2, loglik.detach(), rtol=1e-6)


def test_grad_identities_config2_size():
    """BASELINE config 2 (batch 32 x 9999): expected-count identities that hold for any size —
    sum_c x_t[c] dx_t[c] = w_n at every position, sum_edges A dA = (T-1) sum w, sum pi dpi = sum w, sum B dB = T sum w."""
    B, T = 32, 9999
    x = syn.gene_inputs(1, B, T, seed=1234)
    A, pi, Bfull, tab = gene_params()
    w = np.random.RandomState(1).uniform(0.5, 1.5, size=(1, B))
    xg = x.cuda().requires_grad_(True)
    Ag, pig, Bg = dev(A).requires_grad_(True), dev(pi).requires_grad_(True), dev(Bfull).requires_grad_(True)
    ll = ops.GenePredLoglik.apply(xg, Ag, pig, Bg, dev(tab), None, False)
    (ll * dev(w)).sum().backward()
    torch.cuda.synchronize()
    per_pos = (xg.grad[..., :15] * xg.detach()[..., :15]).sum(-1).cpu().numpy()
    assert np.abs(per_pos / w[..., None] - 1).max() < 1e-4
    sw = w.sum()
    assert abs((Ag.grad.double() * Ag.detach().double()).sum().item() / ((T - 1) * sw) - 1) < 1e-4
    assert abs((pig.grad.double() * pig.detach().double()).sum().item() / sw - 1) < 1e-4
    assert abs((Bg.grad.double() * Bg.detach().double()).sum().item() / (T * sw) - 1) < 1e-4
    # deterministic: a second evaluation gives bit-identical gradients
    g1 = [t.grad.clone() for t in (xg, Ag, pig, Bg)]
    for t in (xg, Ag, pig, Bg):
        t.grad = None
    ll = ops.GenePredLoglik.apply(xg, Ag, pig, Bg, dev(tab), None, False)
    (ll * dev(w)).sum().backward()
    for a, b in zip(g1, (xg, Ag, pig, Bg)):
        assert torch.equal(a, b.grad)


def test_fused_parameter_construction_matches_torch_construction():
    """ops.GeneParams (one launch forward, one backward) against the PyTorch construction the classes expose
    (make_A / make_initial_distribution / make_B_full): values and gradients."""
    layer, em, tr = make_layer()
    g = torch.Generator().manual_seed(11)
    with torch.no_grad():
        tr.transition_kernel.add_(0.3 * torch.randn(tr.transition_kernel.shape, generator=g).cuda())
    wA, wpi, wB = [torch.randn(s, generator=g).cuda() for s in ((1, 15, 15), (1, 15), (1, 15, 15))]
    A, pi, Bf = ops.GeneParams.apply(tr.transition_kernel, tr.starting_distribution_kernel, em.emission_kernel)
    ((A * wA).sum() + (pi * wpi).sum() + (Bf * wB).sum()).backward()
    got = [p.grad.clone() for p in (tr.transition_kernel, tr.starting_distribution_kernel, em.emission_kernel)]
    for p in (tr.transition_kernel, tr.starting_distribution_kernel, em.emission_kernel):
        p.grad = None
    em.recurrent_init()
    A2, pi2, B2 = tr.make_A(), tr.make_initial_distribution().reshape(1, 15), em.make_B_full()
    ((A2 * wA).sum() + (pi2 * wpi).sum() + (B2 * wB).sum()).backward()
    ref = [p.grad for p in (tr.transition_kernel, tr.starting_distribution_kernel, em.emission_kernel)]
    assert torch.allclose(A, A2, rtol=1e-6, atol=1e-9) and torch.allclose(pi, pi2, rtol=1e-6) and torch.allclose(Bf, B2, rtol=1e-6)
    assert ((A == 0) == (A2 == 0)).all()
    for a, b in zip(got, ref):
        assert tuple(a.shape) == tuple(b.shape)
        assert (a - b).abs().max() <= 1e-5 * b.abs().max() + 1e-9
