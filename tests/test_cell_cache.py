"""HmmCell.recurrent_init: rebuild-per-call by default (reference MsaHmmCell.py:41-50) and the opt-in
freeze_derived() cache; get_initial_state(parallel_factor > 1) against the reference fixture (MsaHmmCell.py:121-142).
CPU only: parameter -> matrix construction is host-side torch."""
import numpy as np
import pytest
import torch

import hmm_layer_b200 as H
from hmm_layer_b200 import synthetic as syn


def make_cell():
    em = H.GenePredHMMEmitter(**syn.GENE_CODONS)
    em.build((1, 2, 8, 15))
    with torch.no_grad():
        em.emission_kernel.copy_(syn.gene_emission_kernel())
    tr = H.GenePredMultiHMMTransitioner(initial_exon_len=200, initial_intron_len=4500, initial_ir_len=10000)
    return H.HmmCell([15], 15, em, tr), em, tr


def fresh(cell, tr, em):
    assert torch.allclose(tr.A, tr.make_A(), atol=0, rtol=0)
    assert torch.allclose(em.B, em.make_B(), atol=0, rtol=0)
    assert torch.allclose(cell.init_dist, tr.make_initial_distribution(), atol=0, rtol=0)


def test_default_rebuilds_every_call_even_after_data_writes():
    cell, em, tr = make_cell()
    tr.transition_kernel.data.add_(0.37)            # no version bump
    em.emission_kernel.data.mul_(0.5)
    cell.recurrent_init()
    fresh(cell, tr, em)
    assert not cell.derived_fresh()


def test_frozen_cache_sees_versioned_updates_and_plugin_swaps():
    cell, em, tr = make_cell()
    cell.freeze_derived(True)
    cell.recurrent_init()
    A0 = tr.A
    cell.recurrent_init()
    assert tr.A is A0 and cell.derived_fresh()      # unchanged parameters: nothing rebuilt
    # optimizer step (in-place, versioned)
    opt = torch.optim.SGD([tr.transition_kernel, em.emission_kernel], lr=0.1)
    tr.transition_kernel.grad = torch.ones_like(tr.transition_kernel) * torch.arange(23.)
    em.emission_kernel.grad = torch.randn_like(em.emission_kernel)
    opt.step()
    assert not cell.derived_fresh()
    cell.recurrent_init()
    assert tr.A is not A0
    fresh(cell, tr, em)
    # load_state_dict
    sd = {k: v.clone() for k, v in cell.state_dict().items()}
    sd["transitioner.transition_kernel"] += 0.25
    cell.load_state_dict(sd)
    cell.recurrent_init()
    fresh(cell, tr, em)
    # .data edit: NOT seen (documented) until invalidate()
    A1 = tr.A
    tr.transition_kernel.data.add_(0.5)
    cell.recurrent_init()
    assert tr.A is A1
    cell.invalidate()
    cell.recurrent_init()
    fresh(cell, tr, em)
    # emitter swapped inside the ModuleList
    em2 = H.GenePredHMMEmitter(**syn.GENE_CODONS)
    em2.build((1, 2, 8, 15))
    with torch.no_grad():
        em2.emission_kernel.copy_(syn.gene_emission_kernel(seed=99))
    cell.emitter[0] = em2
    cell.recurrent_init()
    assert em2.B is not None and torch.equal(em2.B, em2.make_B())
    # unfreeze: back to rebuild-per-call
    cell.freeze_derived(False)
    tr.transition_kernel.data.add_(0.1)
    cell.recurrent_init()
    fresh(cell, tr, em2)


@pytest.mark.parametrize("frozen", [False, True])
def test_reverse_cell_follows_transition_updates(frozen):
    """The reverse cell holds a reversed *view* of the transitioner whose base is not a registered sub-module: an update
    of transition_kernel alone must still reach rev.recurrent_init()."""
    cell, em, tr = make_cell()
    rev = cell.make_reverse_direction_offspring()
    rev.freeze_derived(frozen)
    rev.recurrent_init()
    with torch.no_grad():
        tr.transition_kernel.add_(torch.linspace(-1, 1, 23).reshape(1, -1))     # versioned in-place update
    rev.recurrent_init()
    assert torch.equal(rev.transitioner.A_transposed, tr.make_A().transpose(1, 2))
    assert torch.equal(rev.log_A_dense, tr.make_log_A())


def test_chunked_initial_state_matches_reference_fixture(golden):
    """get_initial_state(parallel_factor=4): forward = identity for the first chunk, rows of A for the others
    (reference MsaHmmCell.py:121-142).  The fixture's first chunked step (init=True: R = the initial state) gives
    chunk_fwd_out[:, 0] = log(E_0 * max(init, eps) / S) || log S, so the initial state is pinned through one line of
    numpy; the reverse orientation is diag(e_next) A^T with the identity at the last chunk (the D2-free version)."""
    g = golden("ref_genepred.npz")
    B, T, q, pf = 3, 48, 15, 4
    cell, em, tr = make_cell()
    with torch.no_grad():
        em.emission_kernel.copy_(torch.as_tensor(g["emission_kernel"]))
        tr.transition_kernel.copy_(torch.as_tensor(g["transition_kernel"]))
    cell.recurrent_init()
    np.testing.assert_allclose(tr.A.detach().numpy(), g["A"], rtol=1e-6, atol=1e-12)
    init, ll0 = cell.get_initial_state(batch_size=B * pf, parallel_factor=pf)
    assert init.shape == (B * pf, q * q) and ll0.shape == (B * pf, q) and float(ll0.abs().max()) == 0.0
    st = init.reshape(B, pf, q, q).numpy()
    A = g["A"][0]
    for c in range(pf):
        np.testing.assert_allclose(st[:, c], np.broadcast_to(np.eye(q) if c == 0 else A, (B, q, q)), rtol=1e-6, atol=1e-12)
    E0 = g["E"].reshape(B * pf, T // pf, q)[:, 0].astype(np.float64)
    R = np.maximum(init.reshape(B * pf, q, q).numpy().astype(np.float64), 1e-16)
    sf = np.maximum(E0, 1e-16)[:, None, :] * R
    S = sf.sum(-1, keepdims=True)
    want = np.concatenate([np.log(sf / S).reshape(B * pf, q * q), np.log(S).reshape(B * pf, q)], -1)
    np.testing.assert_allclose(want, g["chunk_fwd_out"][:, 0], rtol=2e-5, atol=2e-5)
    # reverse orientation
    rev = cell.make_reverse_direction_offspring()
    Ein = torch.as_tensor(g["E"]).reshape(B * pf, T // pf, q)
    rinit, rll = rev.get_initial_state(inputs=Ein, batch_size=B * pf, parallel_factor=pf)
    rs = rinit.reshape(B, pf, q, q).numpy()
    e_first = g["E"].reshape(B, pf, T // pf, q)[:, :, 0]
    for c in range(pf - 1):
        np.testing.assert_allclose(rs[:, c], e_first[:, c + 1][:, :, None] * A.T[None], rtol=1e-6, atol=1e-12)
    np.testing.assert_allclose(rs[:, pf - 1], np.broadcast_to(np.eye(q), (B, q, q)), atol=0)
    assert rll.shape == (B * pf, q)
