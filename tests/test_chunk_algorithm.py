"""The device algorithm (modelled in numpy float32, tests/chunk_algo_model.py) against the fp64 oracle."""
import numpy as np
import pytest

from oracle import hmm_oracle as O
from hmm_layer_b200 import synthetic as syn
from tests import chunk_algo_model as M


def gene_problem(B, T, seed=1234, p_n=0.001):
    x = syn.gene_inputs(1, B, T, seed=seed, p_n=p_n).numpy().astype(np.float64)
    Bfull = O.expand_class_rows(O.softmax(syn.gene_emission_kernel().numpy(), -1))
    tab = O.make_codon_tables(**syn.GENE_CODONS)
    edges = O.genepred_transition_indices(1)
    A = O.transition_matrix_from_edges(edges, O.genepred_transition_init(edges, 1, 200, 4500, 10000) + 1e-3, 15)[None]
    pi = np.full((1, 15), 1 / 15)
    return O.genepred_emission(x, Bfull, tab), A, pi


@pytest.mark.parametrize("T,L,G,OCT", [(1000, 32, 8, 4), (999, 35, 4, 2), (130, 128, 4, 2), (64, 64, 4, 2), (65, 64, 2, 1),
                                       (7, 64, 4, 2), (301, 17, 8, 4), (1, 8, 4, 2), (2, 1, 4, 2), (40, 1, 8, 4)])
def test_model_matches_oracle(T, L, G, OCT):
    E, A, pi = gene_problem(2, T)
    ref = O.posterior(E, A, pi)
    for b in range(2):
        got = M.run(E[0, b], A[0], pi[0], L, G, OCT)
        rep = O.parity_report(got["log_post"], got["loglik"], ref["log_post"][0, b], ref["loglik"][0, b])
        assert rep["loglik_rel"] < 1e-6, rep
        assert rep["max_abs_dp"] < 1e-4 and rep["argmax_mismatch"] == 0 and rep["max_norm_dev"] < 1e-5, rep
        big = ref["log_alpha"][0, b] - ref["log_alpha"][0, b].max(-1, keepdims=True) > -14
        assert np.abs(got["log_alpha"] - ref["log_alpha"][0, b])[big].max() / max(T, 50) < 1e-5
        big = ref["log_beta"][0, b] - ref["log_beta"][0, b].max(-1, keepdims=True) > -14
        assert np.abs(got["log_beta"] - ref["log_beta"][0, b])[big].max() / max(T, 50) < 1e-5
