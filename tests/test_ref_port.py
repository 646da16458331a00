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
