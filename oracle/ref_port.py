"""CPU port of the reference's PyTorch implementation of the path, float32, op for op.  TEST / BASELINE
INFRASTRUCTURE ONLY (never imported by hmm_layer_b200/).

The reference is pure Python + torch and lives at /root/reference, which does not exist on the GPU box, so the
timed CPU baseline (`bench.py --impl reference`, `cpu_baseline.kind == "port"`) is this restatement.  It keeps
the reference's cost structure on purpose: one Python-level time loop with the same ~14 small torch ops per step
(MsaHmmCell.py:73-106 inside BaseRNN.py:216-227), the same einsum / k-mer intermediates in the emitter
(gene_pred_hmm_emitter.py:93-121,231-263; kmer.py:3-47), run in both directions like Bidirectional.py:137-150,
and posterior = alpha + beta - loglik (MsaHMMLayer.py:501-514).  Wiring is the intended one (separate A / A^T per
direction, no in-place aliasing), i.e. what the reference's own components compute when driven correctly.
It is validated against the reference-object fixtures in tests/test_ref_port.py."""
import torch

EPS = 1e-16


def make_k_mers(seq, k, pivot_left=True):
    p = seq[..., :-1] + 0.25 * (seq[..., -1:] == 1).to(seq.dtype)           # kmer.py:21-25 without the aliasing
    L = p.shape[-2]
    pad = torch.ones_like(p[:, :k - 1, :]) / 4
    if pivot_left:
        padded = torch.cat([p, pad], dim=-2)
        k_mers = padded[:, :L, None, :]
        rng = range(1, k)
    else:
        padded = torch.cat([pad, p], dim=-2)
        k_mers = padded[:, k - 1:L + k - 1, None, :]
        rng = range(k - 2, -1, -1)
    for i in rng:
        shift = padded[:, i:L + i, None, :, None]
        k_mers = (k_mers[..., None, :] * shift)
        k_mers = k_mers.reshape(list(k_mers.shape[:-3]) + [-1, 4])
    return k_mers


def genepred_emission(inputs, B13, codon_probs, end_hints=None, training=False):
    """inputs (k,b,T,20), B13 (k,13,15) = softmax(kernel), codon_probs (2,9,64) -> (k,b,T,15)."""
    nuc, cls = inputs[..., -5:], inputs[..., :-5]
    emit = torch.einsum("k...s,kqs->k...q", cls, B13)
    emit = torch.cat([emit[..., :2], emit[..., 1:2], emit[..., 1:2], emit[..., 2:]], dim=-1)
    if end_hints is not None:
        emit = torch.cat([end_hints[..., :1, :] * emit[..., :1, :], emit[..., 1:-1, :],
                          end_hints[..., 1:, :] * emit[..., -1:, :]], dim=-2)
    k, b, T = nuc.shape[:3]
    flat = nuc.reshape(-1, T, 5)
    left = make_k_mers(flat, 3, True).reshape(k, b, T, 64)
    right = make_k_mers(flat, 3, False).reshape(k, b, T, 64)
    both = torch.stack([left, right], dim=-2)
    cod = torch.einsum("k...rs,rqs->k...rq", both, codon_probs).prod(dim=-2)
    cod = torch.cat([torch.ones_like(cod[..., :6]) / 4096., cod], dim=-1)
    if training:
        cod = cod + 1e-7
    return emit * cod


def cell_step(E_t, scaled, loglik, A_or_At, k, q, reverse, init):
    """HmmCell.forward (MsaHmmCell.py:73-106)."""
    old = scaled.view(k, -1, q)
    R = old if init else torch.matmul(old, A_or_At)
    E = E_t.view(k, -1, q)
    qc = R.shape[1] // E.shape[1]
    R = R.view(k, -1, qc, q)
    E = E.view(k, -1, 1, q)
    old_ll = loglik.view(k, -1, qc, 1)
    E = torch.maximum(E, torch.tensor(EPS))
    R = torch.maximum(R, torch.tensor(EPS))
    sf = E * R
    S = torch.sum(sf, dim=-1, keepdim=True)
    ll = old_ll + torch.log(S)
    sf = sf / S
    sf = sf.view(-1, qc * q)
    ll = ll.view(-1, qc)
    if reverse:
        out = torch.cat([torch.log(R).view(-1, qc * q), old_ll.view(-1, qc)], dim=-1)
    else:
        out = torch.cat([torch.log(sf), ll], dim=-1)
    return out, sf, ll


def scan(E, init_scaled, A_or_At, k, q, reverse):
    """BaseRNN.forward over (N,T,q) (BaseRNN.py:194-233) with the first step done with init=True
    (MsaHMMLayer.py:259-262).  Returns outputs in processing order and the final loglik."""
    x = E.transpose(0, 1)
    if reverse:
        x = torch.flip(x, [0])
    scaled, ll = init_scaled, torch.zeros((init_scaled.shape[0], 1))
    outs = []
    for t in range(x.shape[0]):
        o, scaled, ll = cell_step(x[t], scaled, ll, A_or_At, k, q, reverse, init=(t == 0))
        outs.append(o)
    return torch.stack(outs, dim=0).transpose(0, 1), ll


def posterior(E, A, pi, no_loglik=False):
    """E (k,b,T,q) -> dict(log_post, loglik, log_alpha, log_beta), sequential (parallel_factor = 1) route."""
    k, b, T, q = E.shape
    Ef = E.reshape(k * b, T, q)
    f0 = pi.reshape(k, 1, q).expand(k, b, q).reshape(-1, q)
    fwd, ll = scan(Ef, f0, A, k, q, False)
    bwd, _ = scan(Ef, torch.ones(k * b, q), A.transpose(1, 2), k, q, True)
    bwd = torch.flip(bwd, [1])
    log_alpha = (fwd[..., :-1] + fwd[..., -1:]).reshape(k, b, T, q)
    log_beta = (bwd[..., :-1] + bwd[..., -1:]).reshape(k, b, T, q)
    loglik = ll.reshape(k, b)
    post = log_alpha + log_beta
    if not no_loglik:
        post = post - loglik[..., None, None]
    return dict(log_post=post, loglik=loglik, log_alpha=log_alpha, log_beta=log_beta)


def genepred_posterior(inputs, A, pi, B13, codon_probs, end_hints=None, training=False, no_loglik=False):
    with torch.no_grad():
        E = genepred_emission(inputs, B13, codon_probs, end_hints, training)
        return posterior(E, A, pi, no_loglik)
