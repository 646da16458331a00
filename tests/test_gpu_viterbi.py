"""GPU parity of the gene-HMM Viterbi kernels (hmm_b200_genepred_viterbi: the chunked (max,+) scan of gene15_viterbi.cu for
1024 <= T <= 131072, the one-warp-per-sequence kernel otherwise) against the fp64 oracle (oracle/hmm_oracle.py
viterbi: max-product form of the reference recursion; the reference itself ships no Viterbi).  State paths must match
EXACTLY (BASELINE north_star) on the seeded cases; at BASELINE config-2 size the check is through the path's own log
joint probability recomputed in fp64 (equal to the kernel's score, never above the posterior log-likelihood, every step
an edge of the model)."""
import numpy as np
import pytest
import torch

from oracle import hmm_oracle as O
from hmm_layer_b200 import ops, synthetic as syn
import hmm_layer_b200 as H
from tests.helpers import gene_params, oracle_gene

pytestmark = pytest.mark.gpu


def dev(x):
    return torch.as_tensor(np.asarray(x, np.float32)).cuda()


@pytest.mark.parametrize("B,T", [(4, 1000), (3, 2049), (2, 4097), (5, 1), (3, 2), (2, 17), (1, 6000), (2, 1024), (2, 1025),
                                 (3, 9999), (1, 32768), (1, 32769), (2, 1023), (1, 100_003), (1, 131_073)])
def test_viterbi_path_exact(B, T):
    x = syn.gene_inputs(1, B, T, seed=300 + T)
    A, pi, Bfull, tab = gene_params(pi_random=True)
    ref = oracle_gene(x.numpy(), A, pi, Bfull, tab)
    path_ref, score_ref = O.viterbi(ref["E"], A, pi)
    path, score = ops.genepred_viterbi(x.cuda(), dev(A), dev(pi), dev(Bfull), dev(tab))
    torch.cuda.synchronize()
    path, score = path.cpu().numpy().astype(np.int64), score.cpu().numpy()
    assert np.abs(score - score_ref).max() / np.abs(score_ref).max() < 2e-6
    mism = int((path != path_ref).sum())
    if mism:   # a different path is acceptable only as an exact-score tie in fp64
        assert np.abs(O.path_score(path, ref["E"], A, pi) - score_ref).max() < 1e-9, mism
    assert mism == 0


def test_viterbi_with_hints_and_layer_api():
    B, T = 3, 700
    x = syn.gene_inputs(1, B, T, seed=11)
    A, pi, Bfull, tab = gene_params()
    hints = np.random.RandomState(5).uniform(0.01, 1.0, size=(1, B, 2, 15))
    ref = oracle_gene(x.numpy(), A, pi, Bfull, tab, end_hints=hints)
    path_ref, _ = O.viterbi(ref["E"], A, pi)
    path, _ = ops.genepred_viterbi(x.cuda(), dev(A), dev(pi), dev(Bfull), dev(tab), end_hints=dev(hints))
    assert np.array_equal(path.cpu().numpy(), path_ref)
    em = H.GenePredHMMEmitter(**syn.GENE_CODONS)
    em.build((1, B, T, 15))
    tr = H.GenePredMultiHMMTransitioner(initial_exon_len=200, initial_intron_len=4500, initial_ir_len=10000)
    layer = H.MsaHmmLayer(H.HmmCell([15], 15, em, tr).cuda(), use_prior=False)
    p, sc = layer.viterbi(x.cuda(), return_score=True)
    assert p.dtype == torch.int64 and tuple(p.shape) == (1, B, T) and tuple(sc.shape) == (1, B)
    loglik = layer(x.cuda())[0].detach()
    assert (sc <= loglik + 1e-3).all()


def test_viterbi_config2_size():
    B, T = 32, 9999
    x = syn.gene_inputs(1, B, T, seed=1234)
    A, pi, Bfull, tab = gene_params()
    path, score = ops.genepred_viterbi(x.cuda(), dev(A), dev(pi), dev(Bfull), dev(tab))
    loglik = ops.genepred_run(x.cuda(), dev(A), dev(pi), dev(Bfull), dev(tab), want_post=False)["loglik"]
    torch.cuda.synchronize()
    path, score = path.cpu().numpy().astype(np.int64), score.cpu().numpy()
    E = O.genepred_emission(x.numpy().astype(np.float64), Bfull, tab)
    mine = O.path_score(path, E, A, pi)
    assert np.isfinite(mine).all()                       # every step is an edge of the model
    assert np.abs(mine - score).max() / np.abs(mine).max() < 2e-6
    assert (score <= loglik.cpu().numpy() + 1e-2).all()
    # optimality on a subset: the fp64 oracle's best score equals the kernel path's score
    _, best = O.viterbi(E[:, :4], A, pi)
    assert np.abs(best - mine[:, :4]).max() < 1e-6 * np.abs(best).max()
