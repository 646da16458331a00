"""The reference's one runnable entry point through the drop-in: the body of the reference's
tests/parallel_rnn_forward.py:12-82 with only the import lines changed (hmm_layer -> hmm_layer_b200) and valid
probabilities as inputs (the script feeds torch.randn, which is not a probability vector).  The wiring is the script's:
a CPU-resident cell, a reverse cell from make_reverse_direction_offspring, BaseRNN(reverse_cell, reverse=True) inside
Bidirectional, TotalProbabilityCell scans, parallel_factor=99, training=True, return_prior=True."""
import warnings

import numpy as np
import pytest
import torch

from hmm_layer_b200.MsaHmmCell import HmmCell
from hmm_layer_b200.gene_pred_hmm_emitter import SimpleGenePredHMMEmitter, GenePredHMMEmitter
from hmm_layer_b200.gene_pred_hmm_transitioner import SimpleGenePredHMMTransitioner, GenePredMultiHMMTransitioner
from hmm_layer_b200.BaseRNN import BaseRNN
from hmm_layer_b200.Bidirectional import Bidirectional
from hmm_layer_b200.TotalProbabilityCell import TotalProbabilityCell
from hmm_layer_b200.MsaHMMLayer import _state_posterior_log_probs_impl
from hmm_layer_b200 import synthetic as syn
from oracle import hmm_oracle as O

pytestmark = pytest.mark.gpu


def parallel_rnn_forward(dim, batch, seq_len, parallel_factor):
    num_states = [dim]
    if dim == 15:
        stacked_inputs = syn.gene_inputs(1, batch, seq_len, seed=4242)          # valid class probabilities ‖ one-hot bases
        embedding_inputs = stacked_inputs[..., :15]
    else:
        g = torch.Generator().manual_seed(4243)
        embedding_inputs = torch.softmax(2.0 * torch.randn(1, batch, seq_len, dim, generator=g), -1)
        stacked_inputs = embedding_inputs

    emitter = GenePredHMMEmitter(
        start_codons=[("ATG", 1.)],
        stop_codons=[("TAG", .34), ("TAA", 0.33), ("TGA", 0.33)],
        intron_begin_pattern=[("NGT", 0.99), ("NGC", 0.005), ("NAT", 0.005)],
        intron_end_pattern=[("AGN", 0.99), ("ACN", 0.01)],
        initial_variance=0.05,
        temperature=100.,
    )
    transitioner = GenePredMultiHMMTransitioner(
        initial_exon_len=200,
        initial_intron_len=4500,
        initial_ir_len=10000,
        starting_distribution_init="zeros"
    )
    if dim == 7:
        emitter = SimpleGenePredHMMEmitter()
        transitioner = SimpleGenePredHMMTransitioner()

    emitter.build(embedding_inputs.shape)

    cell = HmmCell(
        num_states=num_states,
        dim=dim,
        emitter=emitter,
        transitioner=transitioner,
    )

    reverse_cell = cell.make_reverse_direction_offspring()
    rnn = BaseRNN(cell, batch_first=True, return_sequences=True, return_state=True)
    reverse_rnn = BaseRNN(reverse_cell, batch_first=True, return_sequences=True, return_state=True, reverse=True)

    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        bidirectional_rnn = Bidirectional(rnn, merge_mode='concat', backward_layer=reverse_rnn)
    bidirectional_rnn.forward_layer = rnn
    bidirectional_rnn.backward_layer = reverse_rnn

    total_prob_cell = TotalProbabilityCell(cell=cell, )
    total_prob_rnn = BaseRNN(total_prob_cell, batch_first=True, return_sequences=True, return_state=True)
    total_prob_cell_rev = TotalProbabilityCell(cell=reverse_cell)
    total_prob_rnn_rev = BaseRNN(total_prob_cell_rev, batch_first=True, return_sequences=True, return_state=True, reverse=True)

    log_post, prior, _ = _state_posterior_log_probs_impl(
        inputs=stacked_inputs,
        cell=cell,
        reverse_cell=reverse_cell,
        bidirectional_rnn=bidirectional_rnn,
        total_prob_rnn=total_prob_rnn,
        total_prob_rnn_rev=total_prob_rnn_rev,
        parallel_factor=parallel_factor,
        training=True,
        return_prior=True
    )
    return log_post, prior, stacked_inputs, cell, emitter, transitioner


def test_reference_script_gene_model_full_size():
    """The script as shipped: 15-state model, 32 x 9999, parallel_factor=99 -- oracle on every sequence."""
    log_post, prior, x, cell, em, tr = parallel_rnn_forward(15, 32, 9999, 99)
    assert tuple(log_post.shape) == (1, 32, 9999, 15)
    assert not log_post.is_cuda                      # host tensors in, host tensors out, like the reference
    A = tr.make_A().detach().numpy().astype(np.float64)
    pi = tr.make_initial_distribution().detach().numpy().reshape(1, 15).astype(np.float64)
    Bfull = em.make_B_full().detach().numpy().astype(np.float64)
    E = O.genepred_emission(x.numpy().astype(np.float64), Bfull, em.codon_probs.numpy(), training=True)
    ref = O.posterior(E, A, pi)
    ll = ref["loglik"]
    rep = O.parity_report(log_post.numpy(), ll, ref["log_post"], ll)
    assert rep["argmax_mismatch"] == 0 and rep["max_abs_dp"] < 1e-4, rep
    assert rep["max_norm_dev"] < 1e-4, rep           # "- loglik" applied: rows are normalised log probabilities
    assert float(torch.as_tensor(prior).abs().sum()) == 0.0


def test_reference_script_simple_7_state_model():
    """The script's dim == 7 branch: SimpleGenePredHMMEmitter / SimpleGenePredHMMTransitioner."""
    log_post, prior, x, cell, em, tr = parallel_rnn_forward(7, 3, 495, 99)
    assert tuple(log_post.shape) == (1, 3, 495, 7)
    A = tr.make_A().detach().cpu().numpy().astype(np.float64)
    pi = tr.make_initial_distribution().detach().cpu().numpy().reshape(1, 7).astype(np.float64)
    Bfull = em.make_B_full().detach().cpu().numpy().astype(np.float64)
    E = np.einsum("kbts,kqs->kbtq", x.numpy().astype(np.float64), Bfull)
    ref = O.posterior(E, A, pi)
    rep = O.parity_report(log_post.cpu().numpy(), ref["loglik"], ref["log_post"], ref["loglik"])
    assert rep["argmax_mismatch"] == 0 and rep["max_abs_dp"] < 1e-4 and rep["max_norm_dev"] < 1e-4, rep
