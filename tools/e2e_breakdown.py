"""Where the end-to-end (host buffers) time of config 2 goes on ONE GPU.  Times, with CUDA events around the whole call and
host clocks around its pieces:
  raw copies (H2D, D2H, both at once), the library's host pipeline through ops.genepred_run_host (time until the C call
  returns = enqueue cost; time until complete), the layer call, and a copies-only replica of the pipeline (same slices and
  dependencies, no kernels) -- the floor the platform gives this schedule.
HMM_HOST_SLICES / HMM_HOST_RAMP select the schedule inside the library (read once per process): run this script once per
setting, e.g.  for s in 2 4 8; do HMM_HOST_SLICES=$s python tools/e2e_breakdown.py; done"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hmm_layer_b200 as H
from hmm_layer_b200 import ops, synthetic as syn

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
k, B, T = 1, 32, 9999
x = syn.gene_inputs(k, B, T, seed=1234).pin_memory()
xd = torch.empty_like(x, device=dev)
od = torch.empty((k, B, T, 15), device=dev)
oh = torch.empty((k, B, T, 15)).pin_memory()
em = H.GenePredHMMEmitter(**syn.GENE_CODONS); em.build((1, B, T, 15))
with torch.no_grad():
    em.emission_kernel.copy_(syn.gene_emission_kernel())
tr = H.GenePredMultiHMMTransitioner(initial_exon_len=200, initial_intron_len=4500, initial_ir_len=10000)
cell = H.HmmCell([15], 15, em, tr).to(dev)
layer = H.MsaHmmLayer(cell, use_prior=False)
cell.recurrent_init()
A, pi, Bf, cod = (tr.A.detach().contiguous(), cell.init_dist.detach().reshape(1, 15).contiguous(),
                  em.make_B_full().detach().contiguous(), em.codon5(dev))
s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, n=40):
    for _ in range(8):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.perf_counter()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, (time.perf_counter() - w0) * 1e3 / n


def both():
    xd.copy_(x, non_blocking=True)
    with torch.cuda.stream(s_out):
        oh.copy_(od, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s_out)


def replica(sizes, kernels):
    """the library's schedule rebuilt from torch calls: H2D of slice i on s_in, kernel on the main stream, D2H on s_out; two
    device buffers each way"""
    main = torch.cuda.current_stream()
    ev_k, ev_out = {}, {}
    b0 = 0
    for i, nb in enumerate(sizes):
        sl = slice(b0, b0 + nb)
        with torch.cuda.stream(s_in):
            if i >= 2:
                s_in.wait_event(ev_k[i - 2])
            xd[:, sl].copy_(x[:, sl], non_blocking=True)
            ev_in = torch.cuda.Event(); ev_in.record(s_in)
        main.wait_event(ev_in)
        if i >= 2:
            main.wait_event(ev_out[i - 2])
        if kernels:
            ops.genepred_run(xd[:, sl], A, pi, Bf, cod, want_post=True, out={"post": od[:, sl]})
        ev_k[i] = torch.cuda.Event(); ev_k[i].record(main)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev_k[i])
            oh[:, sl].copy_(od[:, sl], non_blocking=True)
            ev_out[i] = torch.cuda.Event(); ev_out[i].record(s_out)
        b0 += nb
    main.wait_stream(s_out)


def host_call():
    r = ops.genepred_run_host(x, A, pi, Bf, cod, want_post=True)
    return r


print(f"HMM_HOST_SLICES={os.environ.get('HMM_HOST_SLICES', '-')} HMM_HOST_RAMP={os.environ.get('HMM_HOST_RAMP', '-')}")
for name, fn in (("H2D 25.6 MB", lambda: xd.copy_(x, non_blocking=True)), ("D2H 19.2 MB", lambda: oh.copy_(od, non_blocking=True)),
                 ("H2D + D2H at once", both),
                 ("device kernel, 32 sequences", lambda: ops.genepred_run(xd, A, pi, Bf, cod, want_post=True, out={"post": od})),
                 ("device kernel, 8 sequences", lambda: ops.genepred_run(xd[:, :8], A, pi, Bf, cod, want_post=True, out={"post": od[:, :8]})),
                 ("device kernel, 2 sequences", lambda: ops.genepred_run(xd[:, :2], A, pi, Bf, cod, want_post=True, out={"post": od[:, :2]})),
                 ("library host pipeline (ops.genepred_run_host)", host_call),
                 ("layer.state_posterior_log_probs(pinned CPU)", lambda: layer.state_posterior_log_probs(x))):
    ev, wall = timed(fn)
    print(f"   {name:48s} events {ev:7.3f} ms   wall {wall:7.3f} ms")
if os.environ.get("E2E_PROFILE"):
    import cProfile, pstats
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(200):
        layer.state_posterior_log_probs(x)
    pr.disable()
    pstats.Stats(pr).sort_stats("tottime").print_stats(22)
    sys.exit(0)
if os.environ.get("E2E_SHORT"):
    sys.exit(0)
for sizes in ([8] * 4, [4] * 8, [2, 6, 8, 8, 6, 2], [2, 4, 8, 8, 6, 4], [1, 3, 6, 8, 8, 4, 2]):
    for kern in (False, True):
        ev, wall = timed(lambda: replica(sizes, kern), n=20)
        print(f"   torch replica {str(sizes):28s} {'with kernels' if kern else 'copies only '}  events {ev:7.3f} ms   wall {wall:7.3f} ms")
