"""Shared problem builders for the tests (oracle-side parameter construction)."""
import numpy as np

from oracle import hmm_oracle as O
from hmm_layer_b200 import synthetic as syn


def gene_params(num_models=1, seed=1236, kernel_scale=1.0, pi_random=False):
    """A (k,15,15), pi (k,15), Bfull (k,15,15), codon (2,9,64) as float64 numpy, built by the oracle."""
    ker = syn.gene_emission_kernel(num_models, seed).numpy().astype(np.float64) * kernel_scale
    Bfull = O.expand_class_rows(O.softmax(ker, -1))
    tab = O.make_codon_tables(**syn.GENE_CODONS)
    edges = O.genepred_transition_indices(1)
    rng = np.random.RandomState(seed)
    As = []
    for m in range(num_models):
        kern = O.genepred_transition_init(edges, 1, 200, 4500, 10000) + 1e-3 + (0.3 * rng.randn(len(edges)) if m else 0)
        As.append(O.transition_matrix_from_edges(edges, kern, 15))
    A = np.stack(As)
    pi = O.softmax(rng.randn(num_models, 15), -1) if pi_random else np.full((num_models, 15), 1.0 / 15)
    return A, pi, Bfull, tab


def oracle_gene(x, A, pi, Bfull, tab, end_hints=None, training=False, no_loglik=False):
    E = O.genepred_emission(np.asarray(x, np.float64), Bfull, tab, end_hints=end_hints, training=training)
    r = O.posterior(E, A, pi, no_loglik=no_loglik)
    r["E"] = E
    return r
