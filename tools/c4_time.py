"""Times the generic dense path at BASELINE config 4's shape (8 heads x 1003 states, batch 512): python tools/c4_time.py [T] [path] [reps]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hmm_layer_b200 import ops, synthetic as syn

T = int(sys.argv[1]) if len(sys.argv) > 1 else 600
path = int(sys.argv[2]) if len(sys.argv) > 2 else 0
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
k, L, B = 8, 500, 512
q = 2 * L + 3
A, pi, Bm = syn.profile_model(k, L)
A = A.cuda(); pi = pi.cuda(); Bm = Bm.cuda()
g = torch.Generator(device="cuda").manual_seed(2100)
idx = torch.randint(0, 24, (k, B, T), device="cuda", generator=g)
E = Bm.transpose(1, 2)[torch.arange(k, device="cuda")[:, None, None], idx]      # (k,B,T,q) = B[:, :, symbol]
E = E.contiguous()
ops.set_tuning(0, 0, dense_path=path)
modes = sys.argv[4].split(",") if len(sys.argv) > 4 else ("loglik", "post")
for mode in modes:
    for r in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = ops.dense_run(E, A, pi, want_post=(mode == "post"), want_alpha=(mode == "alpha"), want_beta=(mode == "beta"))
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        cells = float(k) * B * T * q * q
        flops = 2 * cells * (2 if mode in ("post", "beta") else 1)
        print(f"path={path} T={T} {mode}: {ms:.2f} ms  {cells / ms * 1e3:.3e} cells/s  {flops / ms * 1e-9:.1f} TFLOP/s algorithmic  loglik[0,0]={float(out['loglik'][0, 0]):.4f}")
        del out
