"""Quick check of the tcgen05 dense step against the CUDA-core step: python tools/tc_debug.py [k L B T]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, synthetic as syn

k, L, B, T = [int(v) for v in sys.argv[1:5]] if len(sys.argv) >= 5 else (1, 100, 130, 4)
A, pi, Bm = syn.profile_model(k, L)
x = syn.profile_inputs(k, B, T)
E = torch.einsum("kbts,kqs->kbtq", x, Bm).cuda()
A = A.cuda(); pi = pi.cuda()
ops.set_tuning(0, 0, dense_path=1)
ref = ops.dense_run(E, A, pi, want_alpha=True, want_beta=True, want_post=True)
torch.cuda.synchronize()
ops.set_tuning(0, 0, dense_path=2)
got = ops.dense_run(E, A, pi, want_alpha=True, want_beta=True, want_post=True)
torch.cuda.synchronize()
print("loglik ref", ref["loglik"].flatten()[:4].tolist())
print("loglik tc ", got["loglik"].flatten()[:4].tolist())
for n in ("alpha", "beta", "post"):
    r, g = ref[n].cpu().numpy(), got[n].cpu().numpy()
    big = r - r.max(-1, keepdims=True) > -14
    d = np.abs(r - g)
    print(n, "finite", bool(np.isfinite(g).all()), "max abs diff (big)", float(d[big].max()), "per-t", [float(d[:, :, t][big[:, :, t]].max()) for t in range(min(T, 6))])
