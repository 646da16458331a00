"""Deterministic synthetic workloads.  Everything is drawn on the CPU with seeded torch generators
so that tests, bench.py and the CPU baseline see bit-identical inputs; callers move the tensors to the device."""
import math

import numpy as np
import torch

GENE_CODONS = dict(
    start_codons=[("ATG", 1.)],
    stop_codons=[("TAG", .34), ("TAA", .33), ("TGA", .33)],
    intron_begin_pattern=[("NGT", .99), ("NGC", .005), ("NAT", .005)],
    intron_end_pattern=[("AGN", .99), ("ACN", .01)],
)


def gene_inputs(num_models, batch, T, seed=1234, p_n=0.001):
    """(k,B,T,20) float32: softmax(2*randn) class probabilities ‖ one-hot nucleotides over A,C,G,T,N."""
    g = torch.Generator().manual_seed(seed)
    cls = torch.softmax(2.0 * torch.randn(num_models, batch, T, 15, generator=g), -1)
    g2 = torch.Generator().manual_seed(seed + 1)
    nuc = torch.randint(0, 4, (num_models, batch, T), generator=g2)
    is_n = torch.rand(num_models, batch, T, generator=g2) < p_n
    nuc = torch.where(is_n, torch.full_like(nuc, 4), nuc)
    return torch.cat([cls, torch.nn.functional.one_hot(nuc, 5).float()], -1).contiguous()


def gene_emission_kernel(num_models=1, seed=1236, rows=13):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(num_models, rows, 15, generator=g)


def profile_hmm_edges(L):
    """(from,to) edges of the implicit profile HMM with 2L+3 states: LEFT_FLANK, M1..ML, I1..I(L-1),
    UNANNOTATED, RIGHT_FLANK, TERMINAL (pattern of the reference's Transitioner.py:429-465)."""
    lf, fi, us, rf, tm = 0, L + 1, 2 * L, 2 * L + 1, 2 * L + 2
    M = list(range(1, L + 1))
    e = [(lf, lf), (lf, rf), (lf, us), (lf, tm)] + [(lf, m) for m in M]
    e += [(m, m + 1) for m in M[:-1]]
    e += [(i, j) for i in range(1, L - 1) for j in range(i + 2, L + 1)]
    for m in M:
        e += [(m, us), (m, rf), (m, tm)]
    for i in range(L - 1):
        e += [(i + 1, fi + i), (fi + i, i + 2), (fi + i, fi + i)]
    e += [(us, m) for m in M] + [(us, us), (us, rf), (us, tm), (rf, rf), (rf, tm), (tm, tm)]
    return np.array(sorted(set(e)), dtype=np.int64)


def profile_model(num_models, L, s=24, seed=2000):
    """Random profile-HMM shaped model: A (k,q,q) row-stochastic on the profile pattern, pi (k,q), B (k,q,s)."""
    q = 2 * L + 3
    edges = profile_hmm_edges(L)
    A = torch.zeros(num_models, q, q)
    for h in range(num_models):
        g = torch.Generator().manual_seed(seed + h)
        A[h, edges[:, 0], edges[:, 1]] = torch.rand(len(edges), generator=g) + 0.05
    A = A / A.sum(-1, keepdim=True)
    g = torch.Generator().manual_seed(seed + 50)
    pi = torch.softmax(torch.randn(num_models, q, generator=g), -1)
    B = torch.softmax(torch.randn(num_models, q, s, generator=g), -1)
    return A.contiguous(), pi.contiguous(), B.contiguous()


def profile_inputs(num_models, batch, T, s=24, seed=2100):
    g = torch.Generator().manual_seed(seed)
    return torch.nn.functional.one_hot(torch.randint(0, s, (num_models, batch, T), generator=g), s).float().contiguous()


def cells(num_models, batch, T, q):
    """BASELINE.json's unit of work: k*B*T*S^2 per complete evaluation."""
    return float(num_models) * batch * T * q * q
