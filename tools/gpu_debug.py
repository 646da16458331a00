"""Quick GPU-side diagnosis: parity of the fused gene-HMM path against the oracle for a list of shapes/tunings,
plus CUDA-event timings.  Prints one line per case; never raises (so one gpurun call shows everything)."""
import sys, os, time, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from oracle import hmm_oracle as O
from hmm_layer_b200 import ops, synthetic as syn
from tests.helpers import gene_params, oracle_gene


def dev(x):
    return torch.as_tensor(np.asarray(x, np.float32)).cuda()


def case(B, T, L=0, NT=0, full=False, seed=7, **kw):
    tag = f"B={B} T={T} L={L} NT={NT} full={full} {kw}"
    try:
        x = syn.gene_inputs(1, B, T, seed=seed + T)
        A, pi, Bfull, tab = gene_params()
        ref = oracle_gene(x.numpy(), A, pi, Bfull, tab, **kw)
        ops.set_tuning(L, NT)
        r = ops.genepred_run(x.cuda(), dev(A), dev(pi), dev(Bfull), dev(tab), want_post=True, want_alpha=full, want_beta=full, **kw)
        torch.cuda.synchronize()
        got = {k: v.cpu().numpy() for k, v in r.items()}
        rep = O.parity_report(got["post"], got["loglik"], ref["log_post"], ref["loglik"])
        msg = f"ll_rel={rep['loglik_rel']:.2e} dp={rep['max_abs_dp']:.2e} norm={rep['max_norm_dev']:.2e} argmis={rep['argmax_mismatch']} dlog={rep['max_abs_dlog_big']:.2e} finite={np.isfinite(got['post']).all()}"
        if full:
            for name, key in (("alpha", "log_alpha"), ("beta", "log_beta")):
                rr = ref[key]; big = rr - rr.max(-1, keepdims=True) > -14
                d = np.abs(got[name] - rr) / np.maximum(1.0, np.abs(rr))
                msg += f" {name}={d[big].max():.2e}"
        print("CASE", tag, msg, flush=True)
    except Exception as e:  # noqa: BLE001
        print("CASE", tag, "EXC", repr(e), flush=True)
        traceback.print_exc()
    finally:
        ops.set_tuning(0, 0)


def timing():
    x = syn.gene_inputs(1, 32, 9999, seed=1234).cuda()
    A, pi, Bfull, tab = [dev(v) for v in gene_params()]
    out = torch.empty((1, 32, 9999, 15), device="cuda")
    for L in (0, 24, 32, 35, 40, 46):
        try:
            ops.set_tuning(L, 0)
            for _ in range(5):
                ops.genepred_run(x, A, pi, Bfull, tab, out={"post": out})
            torch.cuda.synchronize()
            ops.profile_enable(True)
            acc = None
            for _ in range(20):
                ops.genepred_run(x, A, pi, Bfull, tab, out={"post": out})
                t = ops.profile_read()
                acc = t if acc is None else [a + b for a, b in zip(acc, t)]
            ops.profile_enable(False)
            print("TIME L=%d kernels(us)=%s total=%.1f us" % (L, ["%.1f" % (1e3 * v / 20) for v in acc], 1e3 * sum(acc) / 20), flush=True)
        except Exception as e:  # noqa: BLE001
            print("TIME L=%d EXC %r" % (L, e), flush=True)
        finally:
            ops.set_tuning(0, 0)


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0))
    case(1, 1); case(2, 2); case(1, 7); case(2, 64); case(2, 65)
    case(2, 1000); case(3, 999, full=True); case(2, 1500, L=1); case(2, 1500, L=2, full=True); case(2, 1500, L=5, NT=256)
    case(2, 1500, L=33, NT=160, full=True); case(2, 1500, L=48); case(3, 500, no_loglik=True); case(3, 500, training=True, full=True)
    case(4, 9999); case(2, 9999, full=True); case(1, 100000); case(1, 100000, L=8)
    timing()
