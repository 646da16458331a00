"""CPU oracle for the profile-HMM parameter -> (A, pi, B) construction.  TEST INFRASTRUCTURE ONLY (never imported by
hmm_layer_b200/).

The reference folds the silent states (BEGIN, END, D1..DL) of the explicit profile HMM into an implicit transition
matrix with closed-form log-space sums (Transitioner.py:206-258).  This oracle does NOT restate those sums: it builds the
explicit (3L+5)-state matrix from the kernels (softmax over out-edges, Transitioner.py:173-195,337-380; edge list
:464-503) and eliminates the silent states by linear algebra,
    A_implicit = A_ee + A_es (I - A_ss)^-1 A_se,
which is what "folding" means; float64 numpy.  The initial distribution is: LEFT_FLANK with the flank probability,
otherwise start at BEGIN (Transitioner.py:123-160).  Emission matrix: Emitter.py:97-117.
Parity pinning: the edge lists and the out-edge softmax are checked against the reference's own runnable functions
through tests/golden/ref_profile.npz; the shipped folding code cannot run and masks the wrong triangle
(:221), so for the folding itself parity is defined by this elimination ("parity unpinned" by reference tests)."""
import numpy as np

PARTS = ["begin_to_match", "match_to_end", "match_to_match", "match_to_insert", "insert_to_match", "insert_to_insert",
         "match_to_delete", "delete_to_match", "delete_to_delete", "left_flank_loop", "left_flank_exit",
         "unannotated_segment_loop", "unannotated_segment_exit", "right_flank_loop", "right_flank_exit",
         "end_to_unannotated_segment", "end_to_right_flank", "end_to_terminal"]


def explicit_edges(L):
    """dict part -> (n, 2) [from, to] in the explicit numbering LEFT_FLANK, M1..ML, I1..I(L-1), UNANNOTATED, RIGHT_FLANK,
    TERMINAL, BEGIN, END, D1..DL."""
    M = lambda i: i                      # 1..L
    I = lambda i: L + i                  # 1..L-1
    US, RF, T, BEG, END = 2 * L, 2 * L + 1, 2 * L + 2, 2 * L + 3, 2 * L + 4
    D = lambda i: 2 * L + 4 + i          # 1..L
    e = {p: [] for p in PARTS}
    for j in range(1, L + 1):
        e["begin_to_match"].append((BEG, M(j)))
        e["match_to_end"].append((M(j), END))
    for i in range(1, L):
        e["match_to_match"].append((M(i), M(i + 1)))
        e["match_to_insert"].append((M(i), I(i)))
        e["insert_to_match"].append((I(i), M(i + 1)))
        e["insert_to_insert"].append((I(i), I(i)))
        e["delete_to_delete"].append((D(i), D(i + 1)))
    e["match_to_delete"].append((BEG, D(1)))
    for i in range(1, L):
        e["match_to_delete"].append((M(i), D(i + 1)))
    for i in range(1, L):
        e["delete_to_match"].append((D(i), M(i + 1)))
    e["delete_to_match"].append((D(L), END))
    e["left_flank_loop"] = [(0, 0)]; e["left_flank_exit"] = [(0, BEG)]
    e["unannotated_segment_loop"] = [(US, US)]; e["unannotated_segment_exit"] = [(US, BEG)]
    e["right_flank_loop"] = [(RF, RF)]; e["right_flank_exit"] = [(RF, T)]
    e["end_to_unannotated_segment"] = [(END, US)]; e["end_to_right_flank"] = [(END, RF)]; e["end_to_terminal"] = [(END, T)]
    return {p: np.asarray(v, dtype=np.int64).reshape(-1, 2) for p, v in e.items()}


def explicit_matrix(kernels, L):
    """kernels: dict part -> 1-D array.  (3L+5, 3L+5) row-stochastic explicit matrix (TERMINAL absorbing)."""
    n = 3 * L + 5
    edges = explicit_edges(L)
    W = np.full((n, n), -np.inf)
    for p in PARTS:
        k = np.asarray(kernels[p], np.float64).reshape(-1)
        assert len(k) == len(edges[p]), (p, len(k), len(edges[p]))
        W[edges[p][:, 0], edges[p][:, 1]] = np.maximum(k, -999.0)
    W[2 * L + 2, 2 * L + 2] = 0.0
    W = W - W.max(1, keepdims=True)
    P = np.exp(W)
    return P / P.sum(1, keepdims=True)


def implicit_model(kernels, L, flank_init_kernel=0.0):
    """-> A (2L+3, 2L+3), pi (2L+3,) by eliminating BEGIN, END and the delete states."""
    P = explicit_matrix(kernels, L)
    q = 2 * L + 3
    ee, es, se, ss = P[:q, :q], P[:q, q:], P[q:, :q], P[q:, q:]
    reach = np.linalg.solve(np.eye(ss.shape[0]) - ss, se)      # silent -> first emitting state reached
    A = ee + es @ reach
    flank = 1.0 / (1.0 + np.exp(-float(np.asarray(flank_init_kernel).reshape(-1)[0])))
    pi = (1.0 - flank) * reach[0]                               # row 0 of the silent block is BEGIN
    pi[0] += flank
    return A, pi


def emission_matrix(em_kernel, ins_kernel, L):
    """(2L+3, s+1): LEFT_FLANK, M1..ML, I1..I(L-1), UNANNOTATED, RIGHT_FLANK rows, then the TERMINAL row."""
    em, ins = np.asarray(em_kernel, np.float64), np.asarray(ins_kernel, np.float64)
    s = em.shape[-1]
    rows = np.concatenate([ins[None], em, np.repeat(ins[None], L + 1, 0)], 0)
    rows = np.exp(rows - rows.max(-1, keepdims=True))
    rows = rows / rows.sum(-1, keepdims=True)
    B = np.zeros((2 * L + 3, s + 1))
    B[:-1, :s] = rows
    B[-1, s] = 1.0
    return B
