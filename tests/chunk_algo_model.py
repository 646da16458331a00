"""numpy float32 model of the three-kernel device algorithm (products -> boundary scan -> finalize), used to
check the *derivation* (boundary conventions, scale bookkeeping) against the oracle on CPU before any GPU time
is spent.  It mirrors hmm_layer_b200/csrc/gene15.cu step for step; it is test infrastructure, not product."""
import numpy as np

EPS = np.float32(1e-16)


def chunk_products(E, A, L):
    """K_A: per chunk c the row-normalised transfer matrix and a common log scale.
    Row i of chunk c>0 conditions on state i at position t0-1 (transition first); chunk 0 conditions on the
    state AT position 0 (no transition).  Returns Mn (C,q,q) with max row scale 1, g (C,)."""
    T, q = E.shape
    C = (T + L - 1) // L
    Mn = np.zeros((C, q, q), np.float32); g = np.zeros(C, np.float32)
    for c in range(C):
        a = np.eye(q, dtype=np.float32); ls = np.zeros(q, np.float32)
        for t in range(c * L, min(T, (c + 1) * L)):
            R = a if (c == 0 and t == 0) else a @ A
            a = np.maximum(E[t], EPS)[None] * np.maximum(R, EPS)
            S = a.sum(1, keepdims=True, dtype=np.float32)
            a = a / S; ls = ls + np.log(S[:, 0])
        g[c] = ls.max()
        Mn[c] = a * np.exp(ls - g[c])[:, None]
    return Mn, g


def boundary_scan(Mn, g, pi):
    """K_B: fvec[c] = normalised alpha at t0-1 (pi for c=0), fcum[c] its log scale; rvec[c] = normalised R at the
    last position of chunk c (ones for the last chunk), rcum[c] its log scale; loglik."""
    C, q, _ = Mn.shape
    fvec = np.zeros((C, q), np.float32); fcum = np.zeros(C, np.float64)
    rvec = np.zeros((C, q), np.float32); rcum = np.zeros(C, np.float64)
    f = pi.astype(np.float32); F = 0.0
    for c in range(C):
        fvec[c] = f; fcum[c] = F
        f = f @ Mn[c]; S = f.sum(dtype=np.float32); f = f / S; F += float(g[c]) + np.log(float(S))
    r = np.ones(q, np.float32); Rc = 0.0
    for c in range(C - 1, -1, -1):
        rvec[c] = r; rcum[c] = Rc
        if c > 0:
            r = Mn[c] @ r; S = r.sum(dtype=np.float32); r = r / S; Rc += float(g[c]) + np.log(float(S))
    return fvec, fcum, rvec, rcum, F


def finalize(E, A, L, fvec, fcum, rvec, rcum, loglik, no_loglik=False):
    """K_C: exact sequential recursions inside each chunk from the boundary vectors; locally normalised
    posterior log(alpha-hat * R-hat / z)."""
    T, q = E.shape
    C = fvec.shape[0]
    post = np.zeros((T, q), np.float32); la = np.zeros((T, q), np.float32); lb = np.zeros((T, q), np.float32)
    for c in range(C):
        t0, t1 = c * L, min(T, (c + 1) * L)
        ah = np.zeros((t1 - t0, q), np.float32); rh = np.zeros((t1 - t0, q), np.float32)
        a = fvec[c]; cs = fcum[c]
        for t in range(t0, t1):
            R = a if (c == 0 and t == 0) else a @ A
            a = np.maximum(E[t], EPS) * np.maximum(R, EPS)
            S = a.sum(dtype=np.float32); a = a / S; cs += np.log(float(S))
            ah[t - t0] = a; la[t] = np.log(a) + np.float32(cs)
        r = rvec[c]; cs = rcum[c]
        for t in range(t1 - 1, t0 - 1, -1):
            R = np.maximum(r, EPS)
            rh[t - t0] = R; lb[t] = np.log(R) + np.float32(cs)
            b = np.maximum(E[t], EPS) * R
            S = b.sum(dtype=np.float32); b = b / S; cs += np.log(float(S))
            r = A @ b
        z = (ah * rh).sum(1, keepdims=True, dtype=np.float32)
        post[t0:t1] = np.log(ah) + np.log(rh) - np.log(z)
    if no_loglik:
        post = post + np.float32(loglik)
    return post, la, lb


def run(E, A, pi, L, no_loglik=False):
    E = np.asarray(E, np.float32); A = np.asarray(A, np.float32)
    Mn, g = chunk_products(E, A, L)
    fvec, fcum, rvec, rcum, ll = boundary_scan(Mn, g, np.asarray(pi, np.float32))
    post, la, lb = finalize(E, A, L, fvec, fcum, rvec, rcum, ll, no_loglik)
    return dict(log_post=post, log_alpha=la, log_beta=lb, loglik=ll)
