"""The oracle against every known-answer vector there is for this path: the TF k-mer outputs saved in the
reference's tests/test_tf.ipynb, and outputs of the reference's own objects (tests/golden/make_golden.py)."""
import numpy as np

from oracle import hmm_oracle as O

CODONS = dict(start_codons=[("ATG", 1.)], stop_codons=[("TAG", .34), ("TAA", .33), ("TGA", .33)],
              intron_begin_pattern=[("NGT", .99), ("NGC", .005), ("NAT", .005)],
              intron_end_pattern=[("AGN", .99), ("ACN", .01)])


def test_kmers_match_tf_golden(golden):
    g = golden("kmer_tf_golden.npz")
    seq = g["sequences"].astype(np.float64)
    before = seq.copy()
    np.testing.assert_array_equal(O.make_k_mers(seq, 3, True), g["k_mers_left"])
    np.testing.assert_array_equal(O.make_k_mers(seq, 3, False), g["k_mers_right"])
    np.testing.assert_array_equal(seq, before)           # input untouched (the torch port mutates it, D5)
    np.testing.assert_array_equal(O.encode_kmer_string("ACGN", True), g["encoded_kmer_left"])
    np.testing.assert_array_equal(O.encode_kmer_string("ACGN", False), g["encoded_kmer_right"])
    # facts quoted in SURVEY §4
    assert g["k_mers_left"][0, 0, 6, 0] == 1 and g["k_mers_left"][0, 0].sum() == 1
    np.testing.assert_array_equal(g["k_mers_right"][0, 4, 14], [.25] * 4)


def test_codon_tables_match_reference(golden):
    g = golden("ref_genepred.npz")
    tab = O.make_codon_tables(**CODONS)
    np.testing.assert_allclose(tab, g["codon_probs"], rtol=1e-6, atol=0)
    assert ((tab > 0).sum(-1) == np.array([[64, 1, 12, 12, 12, 64, 64, 64, 64], [61, 64, 64, 61, 64, 8, 8, 8, 3]])).all()


def test_transition_matrix_matches_reference(golden):
    g = golden("ref_genepred.npz")
    edges = O.genepred_transition_indices(1)
    np.testing.assert_array_equal(edges, g["transition_indices"])
    A = O.transition_matrix_from_edges(edges, g["transition_kernel"][0], 15)
    np.testing.assert_allclose(A, g["A"][0], rtol=2e-6, atol=1e-12)
    assert (A > 0).sum() == 23
    init = O.genepred_transition_init(edges, 1, 200, 4500, 10000)
    np.testing.assert_allclose(init + 1e-3, g["transition_kernel"][0], rtol=1e-6, atol=1e-7)
    # as-shipped quirk D6: exact-zero kernel entries vanish -> 9 edges
    A9 = O.transition_matrix_from_edges(edges, init, 15, quirk_zero_is_absent=True)
    assert (A9 > 0).sum() == 9


def _Bfull(g):
    return O.expand_class_rows(O.softmax(g["emission_kernel"], -1))


def test_emission_matches_reference(golden):
    g = golden("ref_genepred.npz")
    tab = O.make_codon_tables(**CODONS)
    np.testing.assert_allclose(O.softmax(g["emission_kernel"], -1), g["B13"], rtol=1e-6)
    E = O.genepred_emission(g["inputs"], _Bfull(g), tab)
    np.testing.assert_allclose(E, g["E"], rtol=2e-6, atol=1e-12)
    Eh = O.genepred_emission(g["inputs"], _Bfull(g), tab, end_hints=g["end_hints"])
    np.testing.assert_allclose(Eh, g["E_hints"], rtol=2e-6, atol=1e-12)
    Et = O.genepred_emission(g["inputs"], _Bfull(g), tab, training=True)
    np.testing.assert_allclose(Et, g["E_training"], rtol=2e-6, atol=1e-12)
    assert (g["E"] == 0).any()                           # codon-constrained states emit exact zeros


def test_sequential_forward_backward_match_reference(golden):
    g = golden("ref_genepred.npz")
    E = g["E"].astype(np.float64)
    f = O.forward_scaled(E, g["A"], g["pi"])
    np.testing.assert_allclose(f["scaled"][0], g["fwd_out"][..., :15], rtol=0, atol=2e-5)
    np.testing.assert_allclose(f["cum"][0], g["fwd_out"][..., 15], rtol=2e-6)
    np.testing.assert_allclose(f["loglik"], g["loglik"], rtol=2e-6)
    b = O.backward_scaled(E, g["A"])
    np.testing.assert_allclose(b["scaled"][0], g["bwd_out"][..., :15], rtol=0, atol=2e-5)
    np.testing.assert_allclose(b["cum"][0], g["bwd_out"][..., 15], rtol=2e-6, atol=1e-6)
    # float32 mode of the oracle reproduces the reference's own float32 almost to the bit
    f32 = O.forward_scaled(g["E"], g["A"], g["pi"], dtype=np.float32)
    np.testing.assert_allclose(f32["cum"][0], g["fwd_out"][..., 15], rtol=1e-6)


def test_chunked_forward_matches_reference(golden):
    g = golden("ref_genepred.npz")
    E = g["E"].astype(np.float64)
    r = O.chunked_posterior(E, g["A"], g["pi"], pf=4)
    k, B, T, q = E.shape
    ref = g["chunk_fwd_out"].reshape(B, 4, T // 4, 240)
    ref_cond = ref[..., :225].reshape(B, 4, T // 4, q, q) + ref[..., 225:, None]
    big = ref_cond > -30                                     # floor-level rows differ only in float noise
    np.testing.assert_allclose(r["cond_fwd"][big], ref_cond[big], rtol=0, atol=5e-5)
    np.testing.assert_allclose(r["log_alpha"], g["chunk_total_fwd"], rtol=1e-6, atol=1e-4)
    np.testing.assert_allclose(r["loglik"], g["chunk_loglik"], rtol=2e-6)
    # chunked and sequential agree (in probability space; floors differ by construction)
    s = O.posterior(E, g["A"], g["pi"])
    np.testing.assert_allclose(r["loglik"], s["loglik"], rtol=1e-12)
    np.testing.assert_allclose(np.exp(r["log_post"]), np.exp(s["log_post"]), rtol=0, atol=1e-9)


def test_posterior_is_normalised(golden):
    g = golden("ref_genepred.npz")
    s = O.posterior(g["E"].astype(np.float64), g["A"], g["pi"])
    np.testing.assert_allclose(np.exp(s["log_post"]).sum(-1), 1.0, rtol=0, atol=1e-9)


def test_dense67_matches_reference(golden):
    g = golden("ref_dense67.npz")
    np.testing.assert_array_equal(np.unique(O.profile_hmm_edges(32), axis=0), np.unique(g["edges"], axis=0))
    E = np.einsum("kbts,kqs->kbtq", g["inputs"].astype(np.float64), g["B"])
    np.testing.assert_allclose(E, g["E"], rtol=1e-6)
    f = O.forward_scaled(E, g["A"], g["pi"])
    np.testing.assert_allclose(f["loglik"], g["loglik"], rtol=2e-6)
    np.testing.assert_allclose(f["scaled"][0], g["fwd_out"][..., :67], rtol=0, atol=2e-5)
