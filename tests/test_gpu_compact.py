"""Compact wire format (hmm_b200_genepred_run_compact / _run_host_compact): bf16 class probabilities +
uint8 nucleotide codes give BIT-IDENTICAL results to the fp32 entry point fed with the fp32 expansion of the same
values, on the device path and on the pipelined host path; and stay within the usual tolerances of the fp64 oracle."""
import numpy as np
import pytest
import torch

from hmm_layer_b200 import ops, synthetic as syn
from tests.helpers import gene_params, oracle_gene
from oracle import hmm_oracle as O

pytestmark = pytest.mark.gpu


def dev(x):
    return torch.as_tensor(np.asarray(x, np.float32)).cuda()


def expand(cls16, codes):
    x = torch.zeros(cls16.shape[:-1] + (20,), dtype=torch.float32)
    x[..., :15] = cls16.float()
    x[..., 15:] = torch.nn.functional.one_hot(codes.long(), 5).float()
    return x


@pytest.mark.parametrize("B,T", [(3, 1000), (2, 1), (4, 37), (32, 9999)])
def test_compact_equals_fp32_path(B, T):
    x = syn.gene_inputs(1, B, T, seed=40 + T)
    cls16, codes = ops.pack_compact(x)
    assert cls16.dtype == torch.bfloat16 and codes.dtype == torch.uint8 and int(codes.max()) <= 4
    xr = expand(cls16, codes)
    A, pi, Bfull, tab = [dev(v) for v in gene_params()]
    # bit-identity holds against the kernels the compact format shares (products + finalize); posterior-only fp32 calls
    # normally take the single-launch kernel, whose arithmetic order differs in the last bits
    ops.set_fused(False)
    try:
        ref = ops.genepred_run(xr.cuda(), A, pi, Bfull, tab, want_post=True, want_alpha=(T < 5000), want_beta=(T < 5000))
        host_ref = ops.genepred_run_host(xr.pin_memory(), A, pi, Bfull, tab, want_post=True)
    finally:
        ops.set_fused(True)
    fused = ops.genepred_run(xr.cuda(), A, pi, Bfull, tab, want_post=True)
    got = ops.genepred_run_compact(cls16.cuda(), codes.cuda(), A, pi, Bfull, tab, want_post=True, want_alpha=(T < 5000),
                                   want_beta=(T < 5000))
    host = ops.genepred_run_compact(cls16.pin_memory(), codes.pin_memory(), A, pi, Bfull, tab, want_post=True)
    torch.cuda.synchronize()
    for key in ref:
        assert torch.equal(ref[key], got[key]), key
    # the host path cuts the batch into slices (its own chunk length), so its twin is the fp32 HOST path
    assert not host["post"].is_cuda
    assert torch.equal(host["post"], host_ref["post"]) and torch.equal(host["loglik"], host_ref["loglik"])
    assert torch.allclose(host["post"].exp(), ref["post"].cpu().exp(), atol=2e-5)
    assert torch.allclose(fused["post"].exp(), ref["post"].exp(), atol=2e-5)
    assert torch.allclose(fused["loglik"], ref["loglik"], rtol=1e-6)


def test_compact_against_oracle_and_rejects_soft_rows():
    B, T = 3, 800
    x = syn.gene_inputs(1, B, T, seed=5)
    cls16, codes = ops.pack_compact(x)
    A, pi, Bfull, tab = gene_params()
    ref = oracle_gene(expand(cls16, codes).numpy(), A, pi, Bfull, tab)
    got = ops.genepred_run_compact(cls16.cuda(), codes.cuda(), dev(A), dev(pi), dev(Bfull), dev(tab))
    rep = O.parity_report(got["post"].cpu().numpy(), got["loglik"].cpu().numpy(), ref["log_post"], ref["loglik"])
    assert rep["loglik_rel"] < 2e-6 and rep["max_abs_dp"] < 1e-4 and rep["argmax_mismatch"] == 0, rep
    soft = x.clone()
    soft[0, 0, 3, 15:] = torch.tensor([0.5, 0.5, 0, 0, 0])
    with pytest.raises(ValueError):
        ops.pack_compact(soft)
    with pytest.raises(ValueError):
        ops.genepred_run_compact(cls16.float().cuda(), codes.cuda(), dev(A), dev(pi), dev(Bfull), dev(tab))
