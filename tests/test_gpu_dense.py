"""GPU parity of the generic dense path (any q) and of the step-by-step cell API against the fp64 oracle and the
reference-object fixture.  Tolerances as in test_gpu_genepred.py (loglik 1e-4 relative, we assert 2e-6;
posterior 1e-4 absolute in probability space; log alpha / log beta 1e-4 relative on the entries that matter)."""
import numpy as np
import pytest
import torch

import hmm_layer_b200 as H
from oracle import hmm_oracle as O
from hmm_layer_b200 import ops, synthetic as syn

pytestmark = pytest.mark.gpu


def dev(x):
    return torch.as_tensor(np.asarray(x, np.float32)).cuda()


def check_ab(got, ref):
    for name, key in (("alpha", "log_alpha"), ("beta", "log_beta")):
        r = ref[key]
        big = r - r.max(-1, keepdims=True) > -14
        d = np.abs(got[name] - r) / np.maximum(1.0, np.abs(r))
        assert d[big].max() < 1e-4, (name, d[big].max())
        assert np.isfinite(got[name]).all()


def test_reference_fixture_dense67(golden):
    g = golden("ref_dense67.npz")
    r = ops.dense_run(dev(g["E"]), dev(g["A"]), dev(g["pi"]), want_alpha=True)
    torch.cuda.synchronize()
    np.testing.assert_allclose(r["loglik"].cpu().numpy(), g["loglik"], rtol=2e-6)
    ref_alpha = g["fwd_out"][..., :67] + g["fwd_out"][..., 67:]
    got = r["alpha"].cpu().numpy()[0]
    big = ref_alpha - ref_alpha.max(-1, keepdims=True) > -14
    assert (np.abs(got - ref_alpha) / np.maximum(1, np.abs(ref_alpha)))[big].max() < 1e-5


@pytest.mark.parametrize("k,L,B,T", [(1, 32, 16, 100), (2, 5, 3, 17), (1, 1, 1, 1), (1, 64, 70, 9), (1, 500, 3, 6)])
def test_dense_posterior_matches_oracle(k, L, B, T):
    """(1,32,16,100) is BASELINE config 1's shape; (1,500,3,6) is config 4's state count (q = 1003)."""
    A, pi, Bm = syn.profile_model(k, L)
    x = syn.profile_inputs(k, B, T)
    E = torch.einsum("kbts,kqs->kbtq", x, Bm)
    ref = O.posterior(E.numpy().astype(np.float64), A.numpy().astype(np.float64), pi.numpy().astype(np.float64))
    r = ops.dense_run(E.cuda(), A.cuda(), pi.cuda(), want_post=True, want_alpha=True, want_beta=True)
    torch.cuda.synchronize()
    got = {n: v.cpu().numpy() for n, v in r.items()}
    rep = O.parity_report(got["post"], got["loglik"], ref["log_post"], ref["loglik"])
    assert rep["loglik_rel"] < 2e-6 and rep["max_abs_dp"] < 1e-4 and rep["argmax_mismatch"] == 0, rep
    check_ab(got, ref)
    only = ops.dense_run(E.cuda(), A.cuda(), pi.cuda())
    assert np.array_equal(only["loglik"].cpu().numpy(), got["loglik"])
    nl = ops.dense_run(E.cuda(), A.cuda(), pi.cuda(), want_post=True, no_loglik=True)["post"].cpu().numpy()
    np.testing.assert_allclose(nl - got["loglik"][..., None, None], got["post"], rtol=0, atol=2e-3)


@pytest.mark.parametrize("reverse", [False, True])
@pytest.mark.parametrize("qc_full", [False, True])
@pytest.mark.parametrize("init", [False, True])
def test_cell_step_matches_oracle(reverse, qc_full, init):
    k, n, q = 2, 5, 15
    rng = np.random.RandomState(3)
    A = rng.rand(k, q, q); A /= A.sum(-1, keepdims=True)
    Aor = A.transpose(0, 2, 1) if reverse else A
    qc = q if qc_full else 1
    E = rng.rand(k * n, q) * (rng.rand(k * n, q) > 0.2)              # exact zeros exercise the floor
    sc = rng.rand(k * n, qc * q); ll = rng.randn(k * n, qc)
    out, ns, nl = O.cell_step(E, sc, ll, A, k, reverse=reverse, init=init)
    go, gs, gl = ops.cell_step(dev(E), dev(sc), dev(ll), dev(Aor), k, reverse=reverse, init=init)
    torch.cuda.synchronize()
    np.testing.assert_allclose(go.cpu().numpy(), out, rtol=2e-5, atol=2e-5)
    np.testing.assert_allclose(gs.cpu().numpy(), ns, rtol=2e-5, atol=1e-12)
    np.testing.assert_allclose(gl.cpu().numpy(), nl, rtol=2e-5, atol=2e-5)


def make_profile_layer(k, L, pf=1):
    A, pi, Bm = syn.profile_model(k, L)
    cell = H.MsaHmmCell([L] * k, dim=24, emitter=H.DenseEmitter(Bm), transitioner=H.DenseTransitioner(A, pi)).cuda()
    return H.MsaHmmLayer(cell, use_prior=False, parallel_factor=pf), (A, pi, Bm)


def test_msa_hmm_layer_config1_forward_loglik():
    """BASELINE config 1: MsaHmmLayer forward log-likelihood, 1 head, model length 32, batch 16 x len 100."""
    layer, (A, pi, Bm) = make_profile_layer(1, 32)
    x = syn.profile_inputs(1, 16, 100)
    loglik, mean = layer(x.cuda())
    E = torch.einsum("kbts,kqs->kbtq", x, Bm).numpy().astype(np.float64)
    ref = O.forward_scaled(E, A.numpy().astype(np.float64), pi.numpy().astype(np.float64))
    np.testing.assert_allclose(loglik.cpu().numpy(), ref["loglik"], rtol=2e-6)
    assert abs(float(mean) - ref["loglik"].mean()) < 1e-3
    la, ll2 = layer.forward_recursion(x.cuda())
    lb = layer.backward_recursion(x.cuda())
    post = layer.state_posterior_log_probs(x.cuda())
    full = O.posterior(E, A.numpy().astype(np.float64), pi.numpy().astype(np.float64))
    rep = O.parity_report(post.cpu().numpy(), ll2.cpu().numpy(), full["log_post"], full["loglik"])
    assert rep["loglik_rel"] < 2e-6 and rep["max_abs_dp"] < 1e-4 and rep["argmax_mismatch"] == 0, rep
    check_ab(dict(alpha=la.cpu().numpy(), beta=lb.cpu().numpy()), full)


def test_step_by_step_cell_with_base_rnn_matches_fused_path():
    """HmmCell driven through BaseRNN (one kernel launch per step), the way the reference drives it
    (MsaHMMLayer.py:247-272), equals the fused scan."""
    layer, (A, pi, Bm) = make_profile_layer(2, 6)
    cell = layer.cell
    x = syn.profile_inputs(2, 3, 12).cuda()
    cell.recurrent_init()
    E = cell.emission_probs(x)
    Ef = E.reshape(2 * 3, 12, cell.max_num_states)
    rnn = H.BaseRNN(cell, batch_first=True, return_sequences=True, return_state=True)
    st0 = cell.get_initial_state(batch_size=3, parallel_factor=1)
    f1, st1 = cell(Ef[:, 0], st0, init=True)
    fr, stT = rnn(Ef[:, 1:], st1)
    fwd = torch.cat([f1.unsqueeze(1), fr], 1)
    log_alpha = (fwd[..., :-1] + fwd[..., -1:]).reshape(2, 3, 12, -1)
    la, ll = layer.forward_recursion(x)
    np.testing.assert_allclose(stT[1].reshape(2, 3).cpu().numpy(), ll.cpu().numpy(), rtol=2e-6)
    big = (la - la.max(-1, keepdim=True).values > -14).cpu().numpy()
    assert np.abs((log_alpha - la).cpu().numpy())[big].max() < 1e-4


@pytest.fixture
def tensor_core_path():
    ops.set_tuning(0, 0, dense_path=2)
    yield
    ops.set_tuning(0, 0, dense_path=0)


@pytest.mark.parametrize("k,L,B,T", [(1, 32, 16, 20), (2, 100, 130, 9), (1, 500, 3, 6), (2, 500, 140, 12), (1, 160, 1, 1),
                                     (1, 700, 5, 3)])
def test_tensor_core_step_matches_oracle(tensor_core_path, k, L, B, T):
    """tcgen05 step kernel (dense_tc.cu) forced on: padded rows (B not a multiple of 128), padded columns (q not a
    multiple of 64 / 256), one and several column tiles, several heads; (2,500,140,12) is config 4's state count."""
    A, pi, Bm = syn.profile_model(k, L)
    x = syn.profile_inputs(k, B, T)
    E = torch.einsum("kbts,kqs->kbtq", x, Bm)
    ref = O.posterior(E.numpy().astype(np.float64), A.numpy().astype(np.float64), pi.numpy().astype(np.float64))
    r = ops.dense_run(E.cuda(), A.cuda(), pi.cuda(), want_post=True, want_alpha=True, want_beta=True)
    torch.cuda.synchronize()
    got = {n: v.cpu().numpy() for n, v in r.items()}
    rep = O.parity_report(got["post"], got["loglik"], ref["log_post"], ref["loglik"])
    assert rep["loglik_rel"] < 2e-6 and rep["max_abs_dp"] < 1e-4 and rep["argmax_mismatch"] == 0, rep
    check_ab(got, ref)
    only = ops.dense_run(E.cuda(), A.cuda(), pi.cuda())["loglik"].cpu().numpy()       # forward direction alone
    assert np.array_equal(only, got["loglik"])
    bonly = ops.dense_run(E.cuda(), A.cuda(), pi.cuda(), want_beta=True)              # loglik + backward
    assert np.array_equal(bonly["beta"].cpu().numpy(), got["beta"])
    # the CUDA-core step on the same inputs: same recurrence, fp32 FMA instead of the bf16-pair tensor-core product
    ops.set_tuning(0, 0, dense_path=1)
    cc = ops.dense_run(E.cuda(), A.cuda(), pi.cuda(), want_post=True)
    np.testing.assert_allclose(cc["loglik"].cpu().numpy(), got["loglik"], rtol=2e-6)
    dp = np.abs(np.exp(cc["post"].cpu().numpy()) - np.exp(got["post"])).max()
    assert dp < 1e-4, dp


def test_config4_full_size_properties():
    """BASELINE config 4 at full size (8 heads x 1003 states, batch 512 x len 600) on the tensor-core path: finite,
    posteriors sum to one, the fp64 oracle on one full 128-sequence row tile of one head (every position), and the
    CUDA-core step on a slice of the batch."""
    k, L, B, T = 8, 500, 512, 600
    q = 2 * L + 3
    A, pi, Bm = syn.profile_model(k, L)
    g = torch.Generator(device="cuda").manual_seed(2100)
    idx = torch.randint(0, 24, (k, B, T), device="cuda", generator=g)
    E = Bm.cuda().transpose(1, 2)[torch.arange(k, device="cuda")[:, None, None], idx].contiguous()     # (k,B,T,q)
    r = ops.dense_run(E, A.cuda(), pi.cuda(), want_post=True)
    torch.cuda.synchronize()
    post, ll = r["post"], r["loglik"]
    assert bool(torch.isfinite(post).all()) and bool(torch.isfinite(ll).all())
    s = torch.exp(post[:, ::37].double()).sum(-1)
    assert float((s - 1).abs().max()) < 2e-3, float((s - 1).abs().max())
    h, sl = 3, slice(128, 256)
    ref = O.posterior(E[h:h + 1, sl].cpu().numpy().astype(np.float64), A[h:h + 1].numpy().astype(np.float64),
                      pi[h:h + 1].numpy().astype(np.float64))
    rep = O.parity_report(post[h:h + 1, sl].cpu().numpy(), ll[h:h + 1, sl].cpu().numpy(), ref["log_post"], ref["loglik"])
    assert rep["loglik_rel"] < 1e-5 and rep["max_abs_dp"] < 1e-4 and rep["argmax_mismatch"] == 0, rep
    try:
        ops.set_tuning(0, 0, dense_path=1)
        cc = ops.dense_run(E[:, :16].contiguous(), A.cuda(), pi.cuda())["loglik"]
    finally:
        ops.set_tuning(0, 0, dense_path=0)
    np.testing.assert_allclose(cc.cpu().numpy(), ll[:, :16].cpu().numpy(), rtol=1e-5)


@pytest.mark.parametrize("k,L,B,T,tc", [(1, 32, 16, 100, False), (2, 20, 5, 33, False), (1, 90, 3, 300, False), (1, 500, 130, 9, True),
                                        (2, 500, 140, 12, True), (1, 100, 130, 7, True)])
def test_onehot_route_equals_materialised_emissions(k, L, B, T, tc):
    """hmm_b200_dense_run_onehot (column gather inside seq_kernel's ring fill / chain_kernel's epilogue) against the same
    engines fed with the materialised tensor: the values entering the recurrence are identical, so are the results."""
    A, pi, Bm = syn.profile_model(k, L)
    x = syn.profile_inputs(k, B, T)
    E = torch.einsum("kbts,kqs->kbtq", x, Bm)
    sym = x.argmax(-1).to(torch.uint8).cuda()
    Bt = Bm.transpose(1, 2).contiguous().cuda()
    ops.set_tuning(0, 0, dense_path=2 if tc else 0)
    try:
        ref = ops.dense_run(E.cuda(), A.cuda(), pi.cuda(), want_post=True, want_alpha=True, want_beta=True)
        got = ops.dense_run_onehot(sym, Bt, A.cuda(), pi.cuda(), want_post=True, want_alpha=True, want_beta=True)
        only = ops.dense_run_onehot(sym, Bt, A.cuda(), pi.cuda())
    finally:
        ops.set_tuning(0, 0, dense_path=0)
    torch.cuda.synchronize()
    for key in ("loglik", "post", "alpha", "beta"):
        assert torch.equal(ref[key], got[key]), key
    assert torch.equal(only["loglik"], ref["loglik"])
    full = O.posterior(E.numpy().astype(np.float64), A.numpy().astype(np.float64), pi.numpy().astype(np.float64))
    rep = O.parity_report(got["post"].cpu().numpy(), got["loglik"].cpu().numpy(), full["log_post"], full["loglik"])
    assert rep["loglik_rel"] < 2e-6 and rep["max_abs_dp"] < 1e-4 and rep["argmax_mismatch"] == 0, rep


def test_onehot_route_through_the_layer_and_fallback_engine():
    """cell.onehot_inputs = True (float one-hot inputs) and integer symbols through MsaHmmLayer; the per-step engine has
    no gather and expands the table inside the workspace."""
    layer, (A, pi, Bm) = make_profile_layer(1, 32)
    x = syn.profile_inputs(1, 16, 100)
    ll_gemm = layer(x.cuda())[0]
    post_gemm = layer.state_posterior_log_probs(x.cuda())
    layer.cell.onehot_inputs = True
    ll_1h = layer(x.cuda())[0]
    post_1h = layer.state_posterior_log_probs(x.cuda())
    sym = x.argmax(-1).cuda()
    ll_sym = layer(sym)[0]
    la_sym, _ = layer.forward_recursion(sym)
    assert torch.equal(ll_gemm, ll_1h) and torch.equal(ll_1h, ll_sym) and torch.equal(post_gemm, post_1h)
    assert la_sym.shape == (1, 16, 100, 67)
    ops.set_tuning(0, 0, dense_path=1)
    try:
        ll_step = layer(sym)[0]
    finally:
        ops.set_tuning(0, 0, dense_path=0)
    np.testing.assert_allclose(ll_step.cpu().numpy(), ll_gemm.cpu().numpy(), rtol=2e-6)
