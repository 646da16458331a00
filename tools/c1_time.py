"""Config-1 shape (1 head, q = 67, batch 16 x 100) and neighbours on the dense path: persistent seq_kernel (default for
q < 192) against the per-step engine (set_tuning dense_path = 1)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hmm_layer_b200 import ops

def timeit(fn, n=50):
    for _ in range(5): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3

for (k, B, T, q) in [(1, 16, 100, 67), (1, 512, 100, 67), (4, 64, 600, 131), (2, 64, 300, 183), (1, 32, 9999, 29)]:
    g = torch.Generator().manual_seed(1)
    E = (torch.rand((k, B, T, q), generator=g) + 0.01).cuda()
    A = torch.rand((k, q, q), generator=g); A = (A / A.sum(-1, keepdim=True)).cuda()
    pi = torch.softmax(torch.randn((k, q), generator=g), -1).cuda()
    res = {}
    for path in (0, 1):
        ops.set_tuning(0, 0, path)
        r = ops.dense_run(E, A, pi, want_post=True)
        res[path] = r
        t_ll = timeit(lambda: ops.dense_run(E, A, pi))
        t_post = timeit(lambda: ops.dense_run(E, A, pi, want_post=True))
        print("k=%d B=%d T=%d q=%d  path %d: loglik %.1f us  posterior %.1f us" % (k, B, T, q, path, t_ll, t_post), flush=True)
    d = (res[0]["post"].exp() - res[1]["post"].exp()).abs().max().item()
    dl = ((res[0]["loglik"] - res[1]["loglik"]).abs() / res[1]["loglik"].abs()).max().item()
    print("   engines agree: max |dp| %.2e, loglik rel %.2e" % (d, dl))
ops.set_tuning(0, 0, 0)
