"""Emitters / transitioners outside the 15-state single-copy case: SimpleGenePredHMMEmitter (7 states) and
GenePredHMMEmitter / GenePredMultiHMMTransitioner with 2 and 3 gene-model copies (reference
gene_pred_hmm_emitter.py:93-121, 231-263; gene_pred_hmm_transitioner.py:132-148, 263-303).
CPU part: the oracle against the reference-object fixture (tests/golden/ref_multicopy.npz, make_golden.py multicopy).
GPU part: the general emission kernel against fixture + oracle, and the models through MsaHmmLayer (dense recurrence)."""
import numpy as np
import pytest
import torch

import hmm_layer_b200 as H
from hmm_layer_b200 import ops, synthetic as syn
from oracle import hmm_oracle as O


def oracle_E(g, nc, hints=None, training=False):
    Bfull = O.expand_class_rows(O.softmax(g[f"kernel{nc}"].astype(np.float64), -1), nc)
    tab = O.make_codon_tables(**syn.GENE_CODONS)
    return O.genepred_emission(g["inputs"].astype(np.float64), Bfull, tab, num_copies=nc, end_hints=hints, training=training)


@pytest.mark.parametrize("nc", [2, 3])
def test_oracle_multicopy_emission_matches_reference_fixture(golden, nc):
    g = golden("ref_multicopy.npz")
    np.testing.assert_allclose(oracle_E(g, nc), g[f"E{nc}"], rtol=2e-5, atol=1e-12)
    np.testing.assert_allclose(oracle_E(g, nc, g[f"hints{nc}"].astype(np.float64), True), g[f"E{nc}_hints_training"], rtol=2e-5, atol=1e-12)


def test_oracle_simple_model_matches_reference_fixture(golden):
    g = golden("ref_multicopy.npz")
    Bfull = O.expand_class_rows(O.softmax(g["kernel7"].astype(np.float64), -1), 1)
    np.testing.assert_allclose(O.class_emission(g["inputs7"].astype(np.float64), Bfull), g["E7"], rtol=2e-6)
    np.testing.assert_allclose(O.class_emission(g["inputs7"].astype(np.float64), Bfull, g["hints7"].astype(np.float64)), g["E7_hints"], rtol=2e-6)
    # the 7-state transition matrix of the mirror class equals the reference's
    tr = H.SimpleGenePredHMMTransitioner()
    with torch.no_grad():
        tr.transition_kernel.copy_(torch.as_tensor(g["transition_kernel7"]))
    np.testing.assert_allclose(tr.make_A().detach().numpy(), g["A7"], rtol=1e-6, atol=1e-12)
    np.testing.assert_allclose(tr.make_initial_distribution().detach().numpy()[0], g["pi7"], rtol=1e-6)


def test_multicopy_transitioner_matches_oracle_edges():
    for k in (2, 3):
        np.random.seed(5)
        tr = H.GenePredMultiHMMTransitioner(k=k, initial_exon_len=200, initial_intron_len=4500, initial_ir_len=10000)
        A = tr.make_A().detach().numpy()[0]
        assert A.shape == (1 + 14 * k, 1 + 14 * k)
        np.testing.assert_allclose(A.sum(-1), 1.0, rtol=1e-6)
        edges = O.genepred_transition_indices(k)
        want = O.transition_matrix_from_edges(edges, tr.transition_kernel.detach().numpy().reshape(-1).astype(np.float64), 1 + 14 * k)
        np.testing.assert_allclose(A, want, rtol=1e-5, atol=1e-12)
        assert set(map(tuple, np.argwhere(A > 0))) == set(map(tuple, edges))


# ---------------------------------------------------------------------------------------------------- GPU

@pytest.mark.gpu
@pytest.mark.parametrize("nc", [2, 3])
def test_general_emission_kernel_multicopy(golden, nc):
    g = golden("ref_multicopy.npz")
    em = H.GenePredHMMEmitter(num_copies=nc, **syn.GENE_CODONS)
    em.build((1, 2, 40, 15))
    with torch.no_grad():
        em.emission_kernel.copy_(torch.as_tensor(g[f"kernel{nc}"]))
    em = em.cuda()
    x = torch.as_tensor(g["inputs"]).cuda()
    hints = torch.as_tensor(g[f"hints{nc}"]).cuda()
    with torch.no_grad():                       # the CUDA kernel
        E = em(x).cpu().numpy()
        Eh = em(x, end_hints=hints, training=True).cpu().numpy()
    np.testing.assert_allclose(E, g[f"E{nc}"], rtol=2e-5, atol=1e-12)
    np.testing.assert_allclose(Eh, g[f"E{nc}_hints_training"], rtol=2e-5, atol=1e-12)
    Ed = em(x, end_hints=hints, training=True)   # grad enabled: the differentiable form (same values, carries a graph)
    assert Ed.requires_grad
    np.testing.assert_allclose(Ed.detach().cpu().numpy(), Eh, rtol=2e-6, atol=1e-12)
    # N and soft nucleotides (not in the reference fixture: its second k-mer call is wrong at N, D5): the oracle decides
    x2 = syn.gene_inputs(1, 3, 257, seed=5, p_n=0.05)
    x2[0, 1, 100, 15:] = torch.tensor([.1, .2, .3, .4, 0.])
    x2[0, 2, 0, 15:] = torch.tensor([.25, .25, .5, 0., 0.])
    Bfull = O.expand_class_rows(O.softmax(g[f"kernel{nc}"].astype(np.float64), -1), nc)
    want = O.genepred_emission(x2.numpy().astype(np.float64), Bfull, O.make_codon_tables(**syn.GENE_CODONS), num_copies=nc)
    with torch.no_grad():
        np.testing.assert_allclose(em(x2.cuda()).cpu().numpy(), want, rtol=2e-5, atol=1e-12)


@pytest.mark.gpu
def test_simple_emitter_kernel(golden):
    g = golden("ref_multicopy.npz")
    em = H.SimpleGenePredHMMEmitter()
    em.build((1, 2, 40, 7))
    with torch.no_grad():
        em.emission_kernel.copy_(torch.as_tensor(g["kernel7"]))
    em = em.cuda()
    x = torch.as_tensor(g["inputs7"]).cuda()
    with torch.no_grad():
        np.testing.assert_allclose(em(x).cpu().numpy(), g["E7"], rtol=2e-6)
        np.testing.assert_allclose(em(x, end_hints=torch.as_tensor(g["hints7"]).cuda()).cpu().numpy(), g["E7_hints"], rtol=2e-6)
    np.testing.assert_allclose(em(x).detach().cpu().numpy(), g["E7"], rtol=2e-6)


def _layer_vs_oracle(cell, x, E_ref):
    layer = H.MsaHmmLayer(cell, use_prior=False)
    post = layer.state_posterior_log_probs(x.cuda())
    la, ll = layer.forward_recursion(x.cuda())
    lb = layer.backward_recursion(x.cuda())
    loglik = layer(x.cuda())[0].detach()
    torch.cuda.synchronize()
    tr = getattr(cell.transitioner, "base", cell.transitioner)
    A = tr.make_A().detach().cpu().numpy().astype(np.float64)
    pi = tr.make_initial_distribution().detach().cpu().numpy().reshape(1, -1).astype(np.float64)
    ref = O.posterior(E_ref, A, pi)
    rep = O.parity_report(post.cpu().numpy(), loglik.cpu().numpy(), ref["log_post"], ref["loglik"])
    assert rep["loglik_rel"] < 1e-5 and rep["max_abs_dp"] < 1e-4 and rep["argmax_mismatch"] == 0, rep
    np.testing.assert_allclose(ll.cpu().numpy(), ref["loglik"], rtol=1e-5)
    for got, key in ((la, "log_alpha"), (lb, "log_beta")):
        r = ref[key]
        big = r - r.max(-1, keepdims=True) > -14
        assert (np.abs(got.cpu().numpy() - r) / np.maximum(1.0, np.abs(r)))[big].max() < 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("nc", [2, 3])
def test_multicopy_model_through_the_layer(nc):
    """GenePredHMMEmitter(num_copies) + GenePredMultiHMMTransitioner(k) through MsaHmmLayer: emission kernel + the
    persistent dense recurrence (q = 29 / 43) against the fp64 oracle."""
    B, T = 4, 700
    np.random.seed(3)
    em = H.GenePredHMMEmitter(num_copies=nc, **syn.GENE_CODONS)
    em.build((1, B, T, 15))
    with torch.no_grad():
        em.emission_kernel.copy_(syn.gene_emission_kernel(1, seed=50 + nc, rows=em.emission_kernel.shape[1]))
    tr = H.GenePredMultiHMMTransitioner(k=nc, initial_exon_len=200, initial_intron_len=4500, initial_ir_len=10000)
    cell = H.HmmCell([1 + 14 * nc], 15, em, tr).cuda()
    x = syn.gene_inputs(1, B, T, seed=600 + nc, p_n=0.01)
    Bfull = em.make_B_full().detach().cpu().numpy().astype(np.float64)
    E = O.genepred_emission(x.numpy().astype(np.float64), Bfull, em.codon_probs.cpu().numpy().astype(np.float64), num_copies=nc)
    _layer_vs_oracle(cell, x, E)


@pytest.mark.gpu
def test_simple_7_state_model_through_the_layer_and_trains():
    B, T = 5, 600
    em = H.SimpleGenePredHMMEmitter()
    em.build((1, B, T, 7))
    with torch.no_grad():
        em.emission_kernel.copy_(torch.randn(em.emission_kernel.shape, generator=torch.Generator().manual_seed(8)))
    tr = H.SimpleGenePredHMMTransitioner()
    cell = H.HmmCell([7], 7, em, tr).cuda()
    x = torch.softmax(2 * torch.randn(1, B, T, 7, generator=torch.Generator().manual_seed(9)), -1)
    Bfull = em.make_B_full().detach().cpu().numpy().astype(np.float64)
    _layer_vs_oracle(cell, x, O.class_emission(x.numpy().astype(np.float64), Bfull))
    # training through the layer: gradients reach the emission and transition kernels
    layer = H.MsaHmmLayer(cell, use_prior=False)
    ll, mean = layer(x.cuda())
    mean.backward()
    assert em.emission_kernel.grad is not None and float(em.emission_kernel.grad.abs().sum()) > 0
    assert tr.transition_kernel.grad is not None and float(tr.transition_kernel.grad.abs().sum()) > 0
    # against fp64 autograd over the oracle-shaped torch recursion
    ek = em.emission_kernel.detach().cpu().double().requires_grad_(True)
    Bk = torch.softmax(ek, -1)
    Bf = torch.cat([Bk[:, :2], Bk[:, 1:2], Bk[:, 1:2], Bk[:, 2:]], 1)
    E = torch.einsum("kbts,kqs->kbtq", x.double(), Bf)
    A = tr.make_A().detach().cpu().double()
    pi = tr.make_initial_distribution().detach().cpu().double().reshape(1, -1)
    a = pi[:, None, :] * torch.clamp(E[:, :, 0], min=1e-16)
    lls = torch.log(a.sum(-1)); a = a / a.sum(-1, keepdim=True)
    for t in range(1, T):
        a = torch.clamp(E[:, :, t], min=1e-16) * torch.clamp(torch.einsum("kbi,kij->kbj", a, A), min=1e-16)
        s = a.sum(-1); lls = lls + torch.log(s); a = a / s[..., None]
    lls.mean().backward()
    np.testing.assert_allclose(em.emission_kernel.grad.cpu().numpy(), ek.grad.numpy(), rtol=2e-3, atol=1e-5)
