"""Generates tests/golden/ref_profile.npz by IMPORTING the reference (run in the build container only):
    python tests/golden/make_golden_profile.py
Only functions of /root/reference/hmm_layer that run unmodified are used: the explicit / implicit
edge lists (Transitioner.py:429-503) and make_transition_matrix_from_indices (:337-380) on seeded random kernels.
The delete-state folding (:206-258) and the emission matrix (Emitter.py:97-117: F.one_hot(..., dtype=) raises) cannot
be run as shipped (D9) and are not part of the fixture."""
import sys
import types

import numpy as np
import torch

sys.path[:0] = ["/root/reference", "/root/reference/hmm_layer"]
sys.modules.setdefault("learnMSA", types.ModuleType("learnMSA"))
import warnings
warnings.filterwarnings("ignore")
from hmm_layer import Transitioner as RT   # noqa: E402

out = {}
rng = np.random.RandomState(42)
for L in (3, 5, 32):
    parts = RT._make_explicit_transition_kernel_parts(L)
    idx_e = RT._make_sparse_transition_indices_explicit(L)
    idx_i = RT._make_sparse_transition_indices_implicit(L)
    for name, _ in parts:
        out[f"L{L}_explicit_{name}"] = np.asarray(idx_e[name], dtype=np.int64).reshape(-1, 2)
    for name, _ in RT._make_implicit_transition_parts(L):
        out[f"L{L}_implicit_{name}"] = np.asarray(idx_i[name], dtype=np.int64).reshape(-1, 2)
    kernel = rng.randn(sum(n for _, n in parts)).astype(np.float32)
    indices = np.concatenate([np.asarray(idx_e[name], dtype=np.int64).reshape(-1, 2) for name, _ in parts], axis=0)
    dense = RT.make_transition_matrix_from_indices(indices, torch.tensor(kernel), 3 * L + 5)
    out[f"L{L}_kernel"] = kernel
    out[f"L{L}_explicit_probs_dense"] = dense.numpy()
np.savez_compressed("tests/golden/ref_profile.npz", **out)
print("wrote tests/golden/ref_profile.npz with", len(out), "arrays")
