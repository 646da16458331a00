"""CPU checks of the reference-interface wrappers that are plain tensor plumbing (no CUDA kernel involved):
TotalProbabilityCell / BaseRNN / Bidirectional against the reference-object fixture and the numpy oracle."""
import numpy as np
import pytest
import torch

import hmm_layer_b200 as H
from oracle import hmm_oracle as O


class _CellStub(torch.nn.Module):
    """What TotalProbabilityCell needs from an HmmCell: num_models, max_num_states, device, transitioner."""
    def __init__(self, pi):
        super().__init__()
        self.num_models, self.max_num_states = pi.shape[0], pi.shape[1]
        self.device = torch.device("cpu")
        cell = self

        class T:
            def make_initial_distribution(self):
                return torch.as_tensor(pi, dtype=torch.float32).unsqueeze(0)
        self.transitioner = T()


def test_total_probability_cell_reproduces_reference_chunk_combine(golden):
    """chunk_fwd_out (reference HmmCell, parallel_factor=4) -> total forward, as MsaHMMLayer.py:285-319 does it."""
    g = golden("ref_genepred.npz")
    B, T, q, pf = 3, 48, 15, 4
    cf = torch.as_tensor(g["chunk_fwd_out"])                       # (B*pf, T/pf, q*q + q)
    cond = (cf[..., :-q].reshape(B * pf, T // pf, q, q) + cf[..., -q:].unsqueeze(-1)).reshape(B, pf, T // pf, q, q)
    tp = H.TotalProbabilityCell(_CellStub(g["pi"]))
    rnn = H.BaseRNN(tp, batch_first=True, return_sequences=True, return_state=True)
    last = cond[:, :, -1].reshape(B, pf, q * q)
    totals, (final, loglik) = rnn(last, tp.get_initial_state(batch_size=B, dtype=torch.float32))
    np.testing.assert_allclose(loglik.numpy(), g["chunk_loglik"].reshape(-1), rtol=2e-6)
    init = torch.log(torch.as_tensor(g["pi"]).expand(B, q) + 1e-16)
    Tc = torch.cat([init.unsqueeze(1), totals[:, :-1]], 1)          # total entering every chunk
    alpha = torch.logsumexp(cond + Tc[:, :, None, :, None], dim=-2).reshape(1, B, T, q)
    ref = g["chunk_total_fwd"]
    big = ref - ref.max(-1, keepdims=True) > -14
    assert np.abs(alpha.numpy() - ref)[big].max() < 2e-4
    # and the oracle's boundary scan agrees with the class
    tot_o, ll_o = O.total_prob_scan(last.reshape(B, pf, q, q).numpy().astype(np.float64), np.log(np.broadcast_to(g["pi"], (B, q))))
    np.testing.assert_allclose(totals.numpy(), tot_o, rtol=0, atol=2e-4)
    np.testing.assert_allclose(loglik.numpy(), ll_o, rtol=2e-6)


def test_total_probability_cell_reverse_initial_state():
    tp = H.TotalProbabilityCell(_CellStub(np.full((1, 5), 0.2, np.float32)), reverse=True)
    v, ll = tp.get_initial_state(batch_size=4, dtype=torch.float32)
    assert v.shape == (4, 5) and float(v.abs().max()) == 0.0 and ll.shape == (4,)


class _SumCell(torch.nn.Module):
    """Toy tuple-state cell: state = (running sum, step count); output = running sum."""
    def forward(self, x, st):
        s, n = st
        s = s + x
        return s, [s, n + 1]


def test_base_rnn_and_bidirectional_wiring():
    x = torch.arange(2 * 5 * 3, dtype=torch.float32).reshape(2, 5, 3)
    st = lambda: [torch.zeros(2, 3), torch.zeros(2, 1)]
    f = H.BaseRNN(_SumCell(), batch_first=True, return_sequences=True, return_state=True)
    out, (s, n) = f(x, st())
    assert torch.equal(out, torch.cumsum(x, 1)) and float(n[0, 0]) == 5
    r = H.BaseRNN(_SumCell(), batch_first=True, return_sequences=True, return_state=True, reverse=True)
    out_r, _ = r(x, st())
    assert torch.equal(out_r, torch.cumsum(torch.flip(x, [1]), 1))
    last = H.BaseRNN(_SumCell(), batch_first=True, return_sequences=False, return_state=False)(x, st())
    assert torch.equal(last[:, 0], x.sum(1))
    b = H.BaseRNN(_SumCell(), batch_first=True, return_sequences=True, return_state=True)
    bi = H.Bidirectional(f, b, merge_mode="concat")
    o, fs, fn, bs, bn = bi(x, initial_state=(*st(), *st()))
    assert o.shape == (2, 5, 6)
    assert torch.equal(o[..., :3], torch.cumsum(x, 1))
    assert torch.equal(o[..., 3:], torch.flip(torch.cumsum(torch.flip(x, [1]), 1), [1]))     # suffix sums, aligned
    assert torch.equal(H.Bidirectional(f, b, merge_mode="sum")(x, initial_state=(*st(), *st()))[0], o[..., :3] + o[..., 3:])
    with pytest.raises(ValueError):
        H.Bidirectional(f, b, merge_mode="mul")
    # the reference script's wiring (tests/parallel_rnn_forward.py:57-63): a backward layer built with reverse=True is
    # accepted with a warning and scans right-to-left once (the reference flips twice, defect D3)
    with pytest.warns(UserWarning):
        bi_r = H.Bidirectional(f, backward_layer=r, merge_mode="concat")
    o_r = bi_r(x, initial_state=(*st(), *st()))[0]
    assert torch.equal(o_r, o)
    with pytest.raises(ValueError):
        H.Bidirectional(f, H.BaseRNN(_SumCell(), batch_first=False, return_sequences=True, return_state=True))
    with pytest.raises(ValueError):
        H.Bidirectional("not a module", b)
