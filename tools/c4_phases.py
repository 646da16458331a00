"""Phase timestamps (globaltimer) of CTA (0,0) of the persistent dense chain kernel at config 4's shape."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hmm_layer_b200 import ops, synthetic as syn, _lib

T = int(sys.argv[1]) if len(sys.argv) > 1 else 300
mode = sys.argv[2] if len(sys.argv) > 2 else "post"
k, L, B = 8, 500, 512
q = 2 * L + 3
A, pi, Bm = syn.profile_model(k, L)
A = A.cuda(); pi = pi.cuda(); Bm = Bm.cuda()
idx = torch.randint(0, 24, (k, B, T), device="cuda")
E = Bm.transpose(1, 2)[torch.arange(k, device="cuda")[:, None, None], idx].contiguous()
ops.set_tuning(0, 0, dense_path=2)
ops.dense_run(E, A, pi, want_post=(mode == "post"))
dbg = torch.zeros(3 * 8 * 8, dtype=torch.int64, device="cuda")
_lib.load().hmm_b200_debug_phases(dbg.data_ptr())
ops.dense_run(E, A, pi, want_post=(mode == "post"))
torch.cuda.synchronize()
_lib.load().hmm_b200_debug_phases(None)
d = dbg.cpu().reshape(3, 8, 8)
t0 = int(d[1, 0, 0])
names = {0: ["flag wait start", "flag ok", "last kb issued"], 1: ["item start", "tmem free", "first stage full", "all MMAs issued"],
         2: ["item start", "E staged+flag", "accumulator full", "chunks done", "published"]}
for it in range(8):
    print(f"item {it}:")
    for role, rn in ((0, "producer"), (1, "mma"), (2, "epilogue")):
        vals = [(int(d[role, it, s]) - t0) / 1000.0 if int(d[role, it, s]) else None for s in range(len(names[role]))]
        print(f"   {rn:9s}", "  ".join(f"{n}={v:.2f}us" if v is not None else f"{n}=-" for n, v in zip(names[role], vals)))
