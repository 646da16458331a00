"""Phase times of the one-launch sharded step on ONE GPU (world = 1: the shard's block goes into its own table, so every
wait is local): kernel time by CUDA events for several chunk lengths, then wall-clock marks (globaltimer, ns after the
first CTA start) of the tile CTAs and the reducer CTAs.   python tools/shard_phases.py [T_local] [chunk ...]"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import ops, synthetic as syn, _lib
from tests.helpers import gene_params

T = int(sys.argv[1]) if len(sys.argv) > 1 else 125_000
chunks = [int(v) for v in sys.argv[2:]] or [0]
lib = _lib.load()
x = [syn.gene_inputs(1, 1, T, seed=1234 + i).cuda() for i in range(6)]   # 6 x 10 MB: rotating, not L2-resident together
A, pi, Bfull, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda() for v in gene_params()]
tab = ops.prepare_codon(tab)
table = torch.zeros(2 * 256 + 2, dtype=torch.float32, device="cuda")
arr = ctypes.c_ulonglong * 1
pb, pf = arr(table.data_ptr()), arr(table.data_ptr() + 4 * 2 * 256)
ws = torch.zeros(lib.hmm_b200_genepred_workspace_size(1, 1, T), dtype=torch.uint8, device="cuda")
out = {"post": torch.empty((1, 1, T, 15), device="cuda"), "loglik": torch.empty((1, 1), device="cuda")}
tag = 0


def step(i):
    global tag
    tag += 1
    return ops.genepred_shard_run_x(x[i % len(x)], None, 0, T, A, pi, Bfull, tab, ws, tag, pb, pf, 0, 1, out=out)


for L in chunks:
    ops.set_tuning(L, 0)
    got = ops.genepred_shard_fused_chunk(1, 1, T, 1)
    if got == 0:
        print(f"chunk {L}: does not fit")
        continue
    for i in range(5):
        step(i)
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for i in range(50):
            step(i)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / 50)
    P = -(-(-(-T // got)) // 32)
    PG = -(-P // 32)
    print(f"chunk {got:2d}: {P} tiles + {PG} + 1 reducers, {np.median(ts) * 1e3:7.1f} us per step")
    dbg = torch.zeros((2, 4096, 16), dtype=torch.int64, device="cuda")
    lib.hmm_b200_debug_phases(dbg.data_ptr())
    step(0)
    torch.cuda.synchronize()
    lib.hmm_b200_debug_phases(None)
    d = dbg.cpu().numpy().reshape(-1, 16)[:P + PG + 1]
    t, r, blk = d[:P], d[P:P + PG], d[P + PG]
    base = t[:, 14].min()
    print(f"   tile CTAs (ns after the first start): start p50 {int(np.median(t[:, 14] - base))} max {int((t[:, 14] - base).max())}; "
          f"matrix published p50 {int(np.median(t[:, 10] - base))} max {int((t[:, 10] - base).max())}; "
          f"boundary vectors seen p50 {int(np.median(t[:, 11] - base))} max {int((t[:, 11] - base).max())}; "
          f"end p50 {int(np.median(t[:, 15] - base))} max {int((t[:, 15] - base).max())}")
    cyc = np.diff(t[:, :10], axis=1)
    names = ["params", "emission+coef", "rows", "blocks", "tree", "publish+wait", "down-sweep", "recursion", "output"]
    print("   tile phases (cycles, median): " + ", ".join(f"{n} {int(np.median(cyc[:, j]))}" for j, n in enumerate(names)))
    rn = ["start", "tiles ready", "group published", "groups ready", "ranks ready", "root vectors", "vectors published"]
    for g in range(PG):
        print(f"   reducer {g}: " + ", ".join(f"{n} {int(r[g, j] - base)}" for j, n in enumerate(rn)))
    print("   block reducer: " + ", ".join(f"{n} {int(blk[j] - base)}" for j, n in enumerate(["start", "groups ready", "product", "stored"])))
ops.set_tuning(0, 0)
