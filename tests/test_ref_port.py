"""The float32 torch port (the timed CPU baseline) against the reference-object fixtures and the fp64 oracle."""
import numpy as np
import torch

from oracle import hmm_oracle as O, ref_port as P


def test_port_matches_reference_fixture(golden):
    g = golden("ref_genepred.npz")
    t = lambda x: torch.as_tensor(x)
    E = P.genepred_emission(t(g["inputs"]), t(g["B13"]), t(g["codon_probs"]))
    np.testing.assert_allclose(E.numpy(), g["E"], rtol=1e-6, atol=0)
    Eh = P.genepred_emission(t(g["inputs"]), t(g["B13"]), t(g["codon_probs"]), end_hints=t(g["end_hints"]))
    np.testing.assert_allclose(Eh.numpy(), g["E_hints"], rtol=1e-6, atol=0)
    r = P.posterior(t(g["E"]), t(g["A"]), t(g["pi"]))
    np.testing.assert_allclose(r["loglik"].numpy(), g["loglik"], rtol=1e-6)
    np.testing.assert_allclose(r["log_alpha"].numpy()[0], g["fwd_out"][..., :15] + g["fwd_out"][..., 15:], rtol=1e-5, atol=1e-4)
    np.testing.assert_allclose(r["log_beta"].numpy()[0], g["bwd_out"][..., :15] + g["bwd_out"][..., 15:], rtol=1e-5, atol=1e-4)


def test_port_matches_oracle(golden):
    g = golden("ref_genepred.npz")
    r = P.posterior(torch.as_tensor(g["E"]), torch.as_tensor(g["A"]), torch.as_tensor(g["pi"]))
    o = O.posterior(g["E"].astype(np.float64), g["A"], g["pi"])
    rep = O.parity_report(r["log_post"].numpy(), r["loglik"].numpy(), o["log_post"], o["loglik"])
    assert rep["loglik_rel"] < 1e-6 and rep["max_abs_dp"] < 1e-4 and rep["argmax_mismatch"] == 0   # fp32 log alpha ~ -550 has 6e-5 resolution


def test_reference_objects_driver_matches_oracle():
    """oracle/ref_driver.py (the timed CPU arm when oracle/_ref exists: the reference's OWN classes) against the fp64 oracle."""
    import os
    import numpy as np
    import pytest
    import torch
    from oracle import hmm_oracle as O, ref_driver
    from hmm_layer_b200 import synthetic as syn
    ref_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")
    if not os.path.isdir(os.path.join(ref_dir, "hmm_layer")):
        pytest.skip("oracle/_ref not populated (oracle/make_ref.sh runs where /root/reference exists)")
    run = ref_driver.make_genepred_posterior(ref_dir, syn.GENE_CODONS, syn.gene_emission_kernel(), 2, 400)
    x = syn.gene_inputs(1, 2, 400, seed=5, p_n=0.0)      # no N: the shipped second k-mer call is wrong there (D5)
    r = run(x)
    Bfull = O.expand_class_rows(O.softmax(syn.gene_emission_kernel().numpy().astype(np.float64), -1))
    edges = O.genepred_transition_indices(1)
    A = O.transition_matrix_from_edges(edges, O.genepred_transition_init(edges, 1, 200, 4500, 10000) + 1e-3, 15)[None]
    E = O.genepred_emission(x.numpy().astype(np.float64), Bfull, O.make_codon_tables(**syn.GENE_CODONS))
    ref = O.posterior(E, A, np.full((1, 15), 1 / 15))
    rep = O.parity_report(r["log_post"].numpy(), r["loglik"].numpy(), ref["log_post"], ref["loglik"])
    assert rep["argmax_mismatch"] == 0 and rep["max_abs_dp"] < 2e-4 and rep["loglik_rel"] < 1e-5, rep
