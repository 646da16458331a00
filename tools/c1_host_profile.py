"""Host-side cost of one config-1 call through MsaHmmLayer (the step is launch / Python bound: ~150 us per call against
~55 us of kernels): cProfile over 2000 calls, top functions by cumulative time.  python tools/c1_host_profile.py"""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hmm_layer_b200 as H
from hmm_layer_b200 import synthetic as syn

dev = torch.device("cuda", 0)
k, B, T, L = 1, 16, 100, 32
torch.manual_seed(2000)
cell = H.MsaHmmCell(L)
with torch.no_grad():
    cell.emitter[0].emission_kernel[0].normal_(0, 1)
cell = cell.to(dev)
layer = H.MsaHmmLayer(cell, use_prior=False)
x = syn.profile_inputs(k, B, T, seed=2100).to(dev)
n = 2000
with torch.no_grad():
    for _ in range(50):
        layer(x)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        layer(x)
    torch.cuda.synchronize()
    print(f"wall per call: {(time.perf_counter() - t0) / n * 1e6:.1f} us")
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(n):
        layer(x)
    pr.disable()
    torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(35)
