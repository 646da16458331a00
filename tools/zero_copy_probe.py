"""Does the single-launch kernel run well straight on pinned HOST buffers (no copy engines: the SMs read the inputs and write
the log-posterior over PCIe, UVA pointers of cudaHostAlloc'ed memory)?  Config 2 on one GPU, each variant complete
(synchronised) per call like the layer's host path:
  whole batch in one launch (host in / host out, host in / device out, device in / host out), and the batch cut into n
  slices issued round-robin on s streams (a slice's upload overlaps its neighbours' compute / download, no events between)."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from hmm_layer_b200 import _lib, ops, synthetic as syn
from tests.helpers import gene_params

lib = _lib.load()
k, B, T = 1, 32, 9999
x = syn.gene_inputs(k, B, T, seed=1234).pin_memory()
xd = x.cuda()
oh = torch.empty((k, B, T, 15)).pin_memory()
od = torch.empty((k, B, T, 15), device="cuda")
ll = torch.empty((k, B), device="cuda")
A, pi, Bf, tab = [torch.as_tensor(np.asarray(v, np.float32)).cuda().contiguous() for v in gene_params()]
pi = pi.reshape(1, 15).contiguous()
tab5 = ops.prepare_codon(tab)
NS = 8
streams = [torch.cuda.Stream() for _ in range(NS)]
wss = [torch.empty(lib.hmm_b200_genepred_workspace_size(1, B, T), dtype=torch.uint8, device="cuda") for _ in range(NS)]
P = ctypes.c_void_p


def run(inp, out, b0, nb, ws, stream):
    rc = lib.hmm_b200_genepred_run(P(inp.data_ptr() + b0 * T * 80), 1, nb, T, P(A.data_ptr()), P(pi.data_ptr()), P(Bf.data_ptr()),
                                   P(tab5.data_ptr()), None, 0, P(out.data_ptr() + b0 * T * 60), None, None, P(ll.data_ptr() + 4 * b0),
                                   P(ws.data_ptr()), ws.numel(), P(stream.cuda_stream))
    _lib.check(rc)


def sliced(inp, out, n, s):
    main = torch.cuda.current_stream()
    ev = torch.cuda.Event(); ev.record(main)
    nb = B // n
    for st in streams[:s]:
        st.wait_event(ev)
    for i in range(n):
        run(inp, out, i * nb, nb, wss[i % s], streams[i % s])
    for st in streams[:s]:
        main.wait_stream(st)
    main.synchronize()


def timed(fn, n=20):
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


main = torch.cuda.current_stream()
ref = None
for name, inp, out in (("device in, device out", xd, od), ("host in, host out", x, oh), ("host in, device out", x, od), ("device in, host out", xd, oh)):
    def one():
        run(inp, out, 0, B, wss[0], main)
        main.synchronize()
    ms = timed(one)
    got = out.cpu() if out.is_cuda else out.clone()
    if ref is None:
        ref = got
    print(f"   one launch, {name:24s} {ms:7.3f} ms   identical to device run: {bool(torch.equal(got, ref))}")
for n, s in ((2, 2), (4, 2), (4, 3), (4, 4), (8, 2), (8, 3), (8, 4), (16, 3), (16, 4), (8, 8), (16, 8), (32, 8)):
    oh.zero_()
    ms = timed(lambda: sliced(x, oh, n, s))
    print(f"   host in / host out, {n:2d} slices round-robin on {s} streams   {ms:7.3f} ms   identical: {bool(torch.equal(oh, ref))}")
