"""GPU parity of the dense-path log-likelihood gradient (ops.DenseLoglik, MsaHmmLayer.forward with profile-HMM cells)
against fp64 torch autograd over the reference port's recursion (oracle/ref_port.py scan = HmmCell.forward in a loop,
MsaHmmCell.py:73-106).  Tolerance: every gradient tensor within 1e-4 of the oracle's, relative to its largest magnitude."""
import numpy as np
import pytest
import torch

from oracle import ref_port as RP
from hmm_layer_b200 import ops
import hmm_layer_b200 as H

pytestmark = pytest.mark.gpu


def close(got, ref, tol=1e-4, what=""):
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    err = np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30)
    assert np.isfinite(got).all() and err < tol, (what, err)


def oracle(E, A, pi, w):
    E, A, pi = [torch.tensor(np.asarray(v, np.float64), requires_grad=True) for v in (E, A, pi)]
    k, b, T, q = E.shape
    f0 = pi.reshape(k, 1, q).expand(k, b, q).reshape(-1, q)
    _, ll = RP.scan(E.reshape(k * b, T, q), f0, A, k, q, False)
    ll = ll.reshape(k, b)
    (ll * torch.tensor(w)).sum().backward()
    gA = torch.zeros_like(A) if A.grad is None else A.grad       # T = 1: no transition
    return ll.detach().numpy(), E.grad.numpy(), gA.numpy(), pi.grad.numpy()


@pytest.mark.parametrize("k,B,T,q", [(2, 4, 50, 23), (1, 16, 100, 67), (1, 3, 1, 9), (1, 8, 40, 260)])
def test_dense_grad_matches_fp64_autograd(k, B, T, q):
    g = torch.Generator().manual_seed(q + T)
    E = torch.rand((k, B, T, q), generator=g) + 0.01
    E[..., 0] = 0.0                                     # an exactly-zero emission column: floored, no gradient
    mask = (torch.rand((k, q, q), generator=g) < 0.3) | torch.eye(q, dtype=torch.bool)
    A = torch.rand((k, q, q), generator=g) * mask
    A = A / A.sum(-1, keepdim=True)
    pi = torch.softmax(torch.randn((k, q), generator=g), -1)
    w = np.random.RandomState(1).uniform(0.5, 1.5, size=(k, B))
    ll_ref, gE_ref, gA_ref, gpi_ref = oracle(E.numpy(), A.numpy(), pi.numpy(), w)
    Eg, Ag, pig = [t.cuda().requires_grad_(True) for t in (E, A, pi)]
    ll = ops.DenseLoglik.apply(Eg, Ag, pig)
    (ll * torch.tensor(w, dtype=torch.float32).cuda()).sum().backward()
    torch.cuda.synchronize()
    close(ll.detach().cpu().numpy(), ll_ref, 1e-5, "loglik")
    close(Eg.grad.cpu().numpy(), gE_ref, what="dE")
    assert (Eg.grad[..., 0] == 0).all()
    if T > 1:
        sup = A.numpy() > 0
        close(Ag.grad.cpu().numpy()[sup], gA_ref[sup], what="dA on the support of A")
    close(pig.grad.cpu().numpy(), gpi_ref, what="dpi")


def test_profile_hmm_layer_trains_through():
    """MsaHmmCell with its default plugins: loglik_mean.backward() reaches every trainable kernel and matches the fp64
    chain (same torch construction of A, pi, B run in float64 on the CPU, then the reference port's recursion)."""
    torch.manual_seed(7)
    cell = H.MsaHmmCell([6, 9])
    with torch.no_grad():
        for p in cell.emitter[0].emission_kernel:
            p.normal_(0, 1)
    twin = H.MsaHmmCell([6, 9])
    twin.load_state_dict(cell.state_dict())
    twin = twin.double()
    cell = cell.cuda()
    layer = H.MsaHmmLayer(cell, use_prior=False)
    x = torch.nn.functional.one_hot(torch.randint(0, 23, (2, 5, 30), generator=torch.Generator().manual_seed(2)), 24).float()
    loglik, mean = layer(x.cuda())
    mean.backward()
    torch.cuda.synchronize()
    twin.recurrent_init(fresh=True)
    E = twin.emission_probs(x.double())
    k, b, T, q = E.shape
    pi = twin.init_dist.reshape(k, q)
    f0 = pi.reshape(k, 1, q).expand(k, b, q).reshape(-1, q)
    _, ll = RP.scan(E.reshape(k * b, T, q), f0, twin.transitioner.A, k, q, False)
    ll.reshape(k, b).mean().backward()
    close(loglik.detach().cpu().numpy(), ll.reshape(k, b).detach().numpy(), 1e-5, "loglik")
    named, named_twin = dict(cell.named_parameters()), dict(twin.named_parameters())
    checked = 0
    for name, p in named.items():
        if not p.requires_grad:
            continue
        assert p.grad is not None, name
        ref = named_twin[name].grad
        if ref is None or float(ref.abs().max()) == 0.0:
            assert float(p.grad.abs().max()) < 1e-6, name
            continue
        close(p.grad.cpu().numpy(), ref.numpy(), 2e-4, name)
        checked += 1
    assert checked >= 20
