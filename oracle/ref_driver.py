"""Drives the reference's OWN objects (copied to oracle/_ref by oracle/make_ref.sh) for the timed CPU baseline.
TEST / BASELINE INFRASTRUCTURE ONLY (never imported by hmm_layer_b200/).

The wiring is the one tests/golden/make_golden.py uses and tests/test_ref_port.py pins: separate transitioner instances
for the forward and the reverse cell (the shipped shared `reverse` flag is defect D1), inputs cloned before every
emitter call (the emitter writes into its input, D5), sequential parallel_factor = 1 route:
  emission  GenePredHMMEmitter.forward            gene_pred_hmm_emitter.py:231-263
  forward   HmmCell.forward inside BaseRNN        MsaHmmCell.py:73-106, BaseRNN.py:174-248, MsaHMMLayer.py:247-272
  backward  reverse HmmCell inside BaseRNN        MsaHMMLayer.py:343-371
  posterior alpha + beta - loglik                 MsaHMMLayer.py:501-514"""
import sys

import numpy as np
import torch


def _import(ref_dir):
    import os
    for d in (os.path.join(ref_dir, "hmm_layer"), ref_dir):      # the package's modules import each other by bare name
        if d not in sys.path:
            sys.path.insert(0, d)
    from hmm_layer.MsaHmmCell import HmmCell
    from hmm_layer.gene_pred_hmm_emitter import GenePredHMMEmitter
    from hmm_layer.gene_pred_hmm_transitioner import GenePredMultiHMMTransitioner
    from hmm_layer.BaseRNN import BaseRNN
    return HmmCell, GenePredHMMEmitter, GenePredMultiHMMTransitioner, BaseRNN


def make_genepred_posterior(ref_dir, codons, emission_kernel, nseq, T):
    """Returns run(inputs (1,nseq,T,20)) -> dict(log_post, loglik) computed by the reference's objects."""
    HmmCell, Emitter, Transitioner, BaseRNN = _import(ref_dir)
    q = 15

    def make_cell(reverse):
        np.random.seed(0)
        em = Emitter(**codons)
        em.build((1, nseq, T, 15))
        with torch.no_grad():
            em.emission_kernel.copy_(emission_kernel)
        tr = Transitioner(initial_exon_len=200, initial_intron_len=4500, initial_ir_len=10000)
        with torch.no_grad():
            tr.transition_kernel += 1e-3           # keep all 23 edges (the shipped dense == 0 test drops exact zeros, D6)
        cell = HmmCell([q], 15, em, tr)
        if reverse:
            cell.reverse_direction()
        cell.recurrent_init()
        return cell

    cell, rcell = make_cell(False), make_cell(True)
    rnn = BaseRNN(cell, batch_first=True, return_sequences=True, return_state=True)
    rrnn = BaseRNN(rcell, batch_first=True, return_sequences=True, return_state=True, reverse=True)

    def run(inputs):
        with torch.no_grad():
            B = inputs.shape[1]
            E = cell.emission_probs(inputs.clone())
            Ef = E.reshape(B, T, q)
            st0 = cell.get_initial_state(batch_size=B, parallel_factor=1)
            f1, st1 = cell(Ef[:, 0], st0, init=True)
            if T > 1:
                fr, stT = rnn(Ef[:, 1:], st1)
                fwd = torch.cat([f1.unsqueeze(1), fr], 1)
            else:
                fwd, stT = f1.unsqueeze(1), st1
            rs0 = rcell.get_initial_state(batch_size=B, parallel_factor=1)
            b1, rs1 = rcell(Ef[:, -1], rs0, init=True)
            if T > 1:
                br, _ = rrnn(Ef[:, :-1], rs1)
                bwd = torch.flip(torch.cat([b1.unsqueeze(1), br], 1), [1])
            else:
                bwd = b1.unsqueeze(1)
            loglik = stT[1].reshape(1, B)
            log_alpha = (fwd[..., :-1] + fwd[..., -1:]).reshape(1, B, T, q)
            log_beta = (bwd[..., :-1] + bwd[..., -1:]).reshape(1, B, T, q)
            return dict(log_post=log_alpha + log_beta - loglik[..., None, None], loglik=loglik)

    return run
