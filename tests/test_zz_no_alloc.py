"""hmm_b200_genepred_run_noalloc (include/hmm_b200.h): the forward + posterior entry point whose scratch is a static
device array inside the library -- no counterpart in the reference, added for callers that cannot allocate.  Called
through the C ABI (ctypes, raw pointers) and compared with hmm_b200_genepred_run on the same inputs, two calls in a row
(the second with other inputs and a smaller shape, so that the array still holds the first call's contents), for the
single-launch route (posterior only) and the two-kernel route (alpha / beta requested: the arrival counters live in the
scratch).  Collected last on purpose (the file name)."""
import ctypes

import numpy as np
import pytest
import torch

from hmm_layer_b200 import _lib, ops, synthetic as syn
from tests.helpers import gene_params

pytestmark = pytest.mark.gpu

NOALLOC_BYTES = 64 << 20          # HMM_B200_NOALLOC_SCRATCH_BYTES


def _params():
    A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda().contiguous() for v in gene_params(pi_random=True)]
    return A, pi.reshape(1, 15).contiguous(), Bfull, ops.prepare_codon(tab)


def _call(lib, fn_name, x, params, want_ab, ws=None):
    """One call through the C ABI; outputs are allocated HERE, before the call (the caller still passes inputs and outputs)."""
    A, pi, Bfull, tab5 = params
    k, B, T, _ = x.shape
    post = torch.full((k, B, T, 15), float("nan"), device="cuda")
    alpha = torch.full((k, B, T, 15), float("nan"), device="cuda") if want_ab else None
    beta = torch.full((k, B, T, 15), float("nan"), device="cuda") if want_ab else None
    ll = torch.full((k, B), float("nan"), device="cuda")
    p = lambda t: None if t is None else ctypes.c_void_p(t.data_ptr())   # noqa: E731
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    head = (p(x), k, B, T, p(A), p(pi), p(Bfull), p(tab5), None, 0, p(post), p(alpha), p(beta), p(ll))
    if fn_name == "hmm_b200_genepred_run":
        rc = lib.hmm_b200_genepred_run(*head, p(ws), ws.numel(), stream)
    else:
        rc = lib.hmm_b200_genepred_run_noalloc(*head, stream)
    _lib.check(rc)
    torch.cuda.synchronize()
    out = {"post": post, "loglik": ll}
    if want_ab:
        out.update(alpha=alpha, beta=beta)
    return {n: v.cpu().numpy() for n, v in out.items()}


def _same(got, ref):
    """The tolerance of the parity tests (tests/test_gpu_genepred.py: 1e-4 on posteriors, 2e-6 relative on the
    log-likelihood); the two entry points run the same kernels, so in practice the outputs are identical."""
    assert np.isfinite(got["loglik"]).all()
    np.testing.assert_allclose(got["loglik"], ref["loglik"], rtol=2e-6, atol=0)
    assert not np.isnan(got["post"]).any()
    assert np.abs(np.exp(got["post"]) - np.exp(ref["post"])).max() < 1e-4
    for name in ("alpha", "beta"):
        if name in ref:
            r, g = ref[name], got[name]
            assert not np.isnan(g).any()
            big = r - r.max(-1, keepdims=True) > -14
            assert (np.abs(g - r) / np.maximum(1.0, np.abs(r)))[big].max() < 1e-4, name


@pytest.mark.parametrize("want_ab", [False, True])
def test_noalloc_entry_point_matches_run_on_two_calls_in_a_row(want_ab):
    lib = _lib.load()
    params = _params()
    shapes = [(3, 9000, 11), (2, 4001, 12)]      # second call: other inputs, smaller shape -> stale scratch behind it
    for B, T, seed in shapes:
        x = syn.gene_inputs(1, B, T, seed=seed, p_n=0.01).cuda().contiguous()
        nbytes = lib.hmm_b200_genepred_workspace_size(1, B, T)
        assert nbytes <= NOALLOC_BYTES
        ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        ref = _call(lib, "hmm_b200_genepred_run", x, params, want_ab, ws=ws)
        got = _call(lib, "hmm_b200_genepred_run_noalloc", x, params, want_ab)
        _same(got, ref)
    # a third call, same shape as the second: the device's free memory does not move across it (nothing allocated, and
    # every lazily loaded kernel is loaded by now)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    got = _call(lib, "hmm_b200_genepred_run_noalloc", x, params, want_ab)
    free1, _ = torch.cuda.mem_get_info()
    _same(got, ref)
    assert free1 >= free0 - (1 << 20), (free0, free1)   # (the outputs of _call come from torch's cached blocks)


def test_noalloc_entry_point_rejects_sizes_beyond_the_static_array():
    lib = _lib.load()
    params = _params()
    B, T = 32, 9999                                      # BASELINE config 2: 186 MB of scratch
    assert lib.hmm_b200_genepred_workspace_size(1, B, T) > NOALLOC_BYTES
    x = torch.zeros((1, B, T, 20), device="cuda")
    with pytest.raises(_lib.HmmB200Error, match="static array holds"):
        _call(lib, "hmm_b200_genepred_run_noalloc", x, params, False)
