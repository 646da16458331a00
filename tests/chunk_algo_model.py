"""numpy float32 model of the device algorithm of hmm_layer_b200/csrc/gene15.cu (products -> boundary scan ->
exact in-chunk recursions), used to check the *derivation* (two-step recurrence over the 7 core states,
power-of-two lazy scaling, boundary conventions, scale bookkeeping) against the fp64 oracle on the CPU before any
GPU time is spent.  It mirrors the kernels step for step; it is test infrastructure, not product.

State order (gene_pred_hmm_transitioner.py:187-188): 0 Ir | 1-3 I0-2 | 4-6 E0-2 | 7 START | 8-10 EI0-2 |
11-13 IE0-2 | 14 STOP.  "core" = 0..6, "pass" = 7..14: every pass state has exactly one predecessor (a core
state) and exactly one successor (a core state), and there is no pass->pass edge
(gene_pred_hmm_transitioner.py:200-221)."""
import numpy as np

F = np.float32
EPS = F(1e-16)
Q = 15
ESRC = [0, 0, 14, 7, 5, 4, 4, 8, 1, 1, 11, 5, 5, 9, 2, 2, 12, 6, 6, 10, 3, 3, 13]
EDST = [0, 7, 0, 5, 14, 5, 8, 1, 1, 11, 4, 6, 9, 2, 2, 12, 5, 4, 10, 3, 3, 13, 6]
SRC1 = [0, 1, 2, 3, 6, 4, 5]                       # core predecessor of core state j
PRED = {7: 0, 8: 4, 9: 5, 10: 6, 11: 1, 12: 2, 13: 3, 14: 5}
SUCC = {7: 5, 8: 1, 9: 2, 10: 3, 11: 4, 12: 5, 13: 6, 14: 0}
LN2 = F(np.log(2.0))


def pow2_floor_exp(x):
    """exponent e with 2^e <= x < 2^(e+1) for positive normal float32 (the kernel reads the exponent bits)."""
    bits = np.asarray(x, F).view(np.int32)
    return ((bits >> 23) & 0xff).astype(np.int32) - 127


def ldexp32(x, e):
    return (np.asarray(x, F) * np.exp2(np.asarray(e, np.float64)).astype(F)).astype(F)


# ------------------------------------------------------------------------------------------------
# K1: chunk transfer matrices by the two-step recurrence
# ------------------------------------------------------------------------------------------------

def position_coefficients(E, A):
    """k1 (T,7): E_t[j]*A[src1(j),j];  k2 (T,8) by pass state p: E_t[succ p]*A[p,succ p]*E_{t-1}[p]*A[pred p,p]."""
    T = E.shape[0]
    k1 = np.zeros((T, 7), F); k2 = np.zeros((T, 8), F)
    for j in range(7):
        k1[:, j] = E[:, j] * A[SRC1[j], j]
    for p in range(7, 15):
        k2[1:, p - 7] = (E[1:, SUCC[p]] * A[p, SUCC[p]]) * (E[:-1, p] * A[PRED[p], p])
    return k1, k2


def chunk_matrix(E, A, k1, k2, t0, t1, first_chunk, rescale_every=4):
    """Row-normalised transfer matrix of positions t0..t1 (inclusive) and per-row log scales.
    Row i conditions on state i at t0-1 (transition first); for the first chunk on the state AT position 0."""
    Mh = np.zeros((Q, Q), F); ls = np.zeros(Q, F)
    ts = t0 + (1 if first_chunk else 0)
    for i in range(Q):
        xm1 = np.zeros(7, F); ym1 = np.zeros(8, F)
        init = E[0, i] if first_chunk else F(1)
        if i < 7:
            xm1[i] = init
        else:
            ym1[i - 7] = init
        esum = 0
        if ts > t1:                                   # first chunk of length 1
            vec = np.concatenate([xm1, ym1])
        else:
            # first step: full 23-edge form from (x, y) at ts-1
            x = np.zeros(7, F)
            for j in range(7):
                x[j] = k1[ts, j] * xm1[SRC1[j]]
            for p in range(7, 15):
                x[SUCC[p]] += (E[ts, SUCC[p]] * A[p, SUCC[p]]) * ym1[p - 7]
            xp = xm1                                   # x_{t-1}
            for t in range(ts + 1, t1 + 1):
                xn = np.zeros(7, F)
                for j in range(7):
                    xn[j] = k1[t, j] * x[SRC1[j]]
                for p in range(7, 15):
                    xn[SUCC[p]] += k2[t, p - 7] * xp[PRED[p]]
                xp, x = x, xn
                if (t - ts) % rescale_every == rescale_every - 1:
                    s = x.sum(dtype=F) + xp.sum(dtype=F)
                    if s > 0:
                        e = int(pow2_floor_exp(s))
                        x = ldexp32(x, -e); xp = ldexp32(xp, -e); esum += e
            y = np.zeros(8, F)
            for p in range(7, 15):
                y[p - 7] = (E[t1, p] * A[PRED[p], p]) * xp[PRED[p]]
            vec = np.concatenate([x, y])
        S = max(vec.sum(dtype=F), F(1e-37))
        Mh[i] = vec / S
        ls[i] = np.log(S) + F(esum) * LN2
    return Mh, ls


def block(Mh, ls):
    """16x16 storage block: rows 0-14 = [M-hat row | ls], row 15 = [column sums of diag(s')M-hat | m]."""
    m = ls.max()
    sp = np.exp(ls - m).astype(F)
    blk = np.zeros((16, 16), F)
    blk[:15, :15] = Mh; blk[:15, 15] = ls
    blk[15, :15] = (sp[:, None] * Mh).sum(0, dtype=F); blk[15, 15] = m
    return blk


def matmul_blocks(a, b):
    """Ordered product a.b of two stored blocks, result in the same storage format."""
    Ah, la = a[:15, :15], a[:15, 15]
    Bh, lb, mb = b[:15, :15], b[:15, 15], b[15, 15]
    sb = np.exp(lb - mb).astype(F)
    C = (Ah @ (sb[:, None] * Bh)).astype(F)
    r = np.maximum(C.sum(1, dtype=F), F(1e-37))
    return block((C / r[:, None]).astype(F), (la + mb + np.log(r)).astype(F))


def fwd_matvec(v, ve, blk):
    """(v, ve): vector with power-of-two exponent bookkeeping, true = v * 2^ve * e^(external).  Returns the new
    (v, ve, dlog) with dlog = the block's max log scale m (added to the running natural-log scale)."""
    Mh, ls, m = blk[:15, :15], blk[:15, 15], blk[15, 15]
    u = v * np.exp(ls - m).astype(F)
    out = (u @ Mh).astype(F)
    s = u.sum(dtype=F)
    e = int(pow2_floor_exp(s)) if s > 0 else 0
    return ldexp32(out, -e), ve + e, m


def bwd_matvec(r, re, blk):
    Mh, ls, m, cs = blk[:15, :15], blk[:15, 15], blk[15, 15], blk[15, :15]
    out = (np.exp(ls - m).astype(F) * (Mh @ r)).astype(F)
    s = (cs * r).sum(dtype=F)
    e = int(pow2_floor_exp(s)) if s > 0 else 0
    return ldexp32(out, -e), re + e, m


# ------------------------------------------------------------------------------------------------
# K2: exact in-chunk recursions with lazily (power-of-two) scaled state
# ------------------------------------------------------------------------------------------------

def seq_forward(E, A, t0, t1, fv, first_chunk):
    """alpha-hat_t up to a power of two: returns rows v_t (sum in [1,2)) and integer exponents n_t with
    alpha_t(true, relative to the boundary vector's scale) = v_t * 2^(n_t)."""
    a = fv.astype(F).copy(); nsum = 0
    rows = []; exps = []
    g_prev = None
    for t in range(t0, t1 + 1):
        kap = a.sum(dtype=F)                       # the device uses the lag-1 sum g (same value)
        if first_chunk and t == 0:
            R = a
        else:
            R = np.zeros(Q, F)
            for e_ in range(23):
                R[EDST[e_]] += a[ESRC[e_]] * A[ESRC[e_], EDST[e_]]
        b = E[t] * np.maximum(R, EPS * kap)
        g = b.sum(dtype=F)
        e = int(pow2_floor_exp(g))
        rows.append(ldexp32(b, -e)); exps.append(nsum + e)
        if (t - t0) % 2 == 1:                      # rescale the propagated state every second step
            a = ldexp32(b, -e); nsum += e
        else:
            a = b
    return np.array(rows, F), np.array(exps, np.int64)


def seq_backward(E, A, t0, t1, rv):
    """R_t (beta before the emission at t, floored) up to a power of two: rows (sum in [1,2)) and exponents."""
    bh = None; nsum = 0
    rows = []; exps = []
    cnt = 0
    for t in range(t1, t0 - 1, -1):
        if bh is None:
            R = rv.astype(F).copy()               # boundary vector holds R itself; scale convention: bkap
            Rf = np.maximum(R, EPS * bkap(R))
        else:
            kap = bh.sum(dtype=F)
            R = np.zeros(Q, F)
            for e_ in range(23):
                R[ESRC[e_]] += A[ESRC[e_], EDST[e_]] * bh[EDST[e_]]
            Rf = np.maximum(R, EPS * kap)
        s = Rf.sum(dtype=F)
        e = int(pow2_floor_exp(s))
        rows.append(ldexp32(Rf, -e)); exps.append(nsum + e)
        b = E[t] * Rf
        if cnt % 2 == 1:
            g = b.sum(dtype=F); e2 = int(pow2_floor_exp(g))
            bh = ldexp32(b, -e2); nsum += e2
        else:
            bh = b
        cnt += 1
    return np.array(rows[::-1], F), np.array(exps[::-1], np.int64)


def bkap(R):
    """Scale of a boundary R vector: the boundary vectors handed to the recursions are normalised so that the
    vector an exact recursion would hold (sum_q b-hat = 1 pushed through A^T, or all-ones at the sequence end)
    has this scale.  The model passes vectors with max entry in [0.5,1]; the floor only matters at 1e-16."""
    return F(1)


# ------------------------------------------------------------------------------------------------
# whole pipeline
# ------------------------------------------------------------------------------------------------

def run(E, A, pi, L, G=4, OCT=2, no_loglik=False):
    """E (T,15) unfloored emissions, A (15,15), pi (15).  L = chunk length, G chunks per tile, OCT chunks per
    group inside a tile (the device uses G = 32, OCT = 8)."""
    E = np.maximum(np.asarray(E, F), EPS); A = np.asarray(A, F); pi = np.asarray(pi, F)
    T = E.shape[0]
    C = (T + L - 1) // L
    k1, k2 = position_coefficients(E, A)
    blocks = []
    for c in range(C):
        t0, t1 = c * L, min(T, (c + 1) * L) - 1
        blocks.append(block(*chunk_matrix(E, A, k1, k2, t0, t1, c == 0)))
    # up-sweep: group products and tile products
    ntile = (C + G - 1) // G
    groups = {}; tiles = []
    for p in range(ntile):
        cs = list(range(p * G, min(C, (p + 1) * G)))
        gp = []
        for o in range(0, len(cs), OCT):
            acc = blocks[cs[o]]
            for c in cs[o + 1:o + OCT]:
                acc = matmul_blocks(acc, blocks[c])
            gp.append(acc)
        groups[p] = gp
        acc = gp[0]
        for gmat in gp[1:]:
            acc = matmul_blocks(acc, gmat)
        tiles.append(acc)
    # top-level scans over tiles (natural-log scale kept in float64 like the device's double accumulator)
    fv = pi.copy(); fe = 0; fl = 0.0
    tile_f = []
    for p in range(ntile):
        tile_f.append((fv, fe, fl))
        fv, fe, m = fwd_matvec(fv, fe, tiles[p]); fl += float(m)
    loglik = fl + fe * float(np.log(2.0)) + float(np.log(float(fv.sum(dtype=F))))
    rv = np.ones(Q, F); re = 0; rl = 0.0
    tile_r = [None] * ntile
    for p in range(ntile - 1, -1, -1):
        tile_r[p] = (rv, re, rl)
        if p > 0:
            rv, re, m = bwd_matvec(rv, re, tiles[p]); rl += float(m)
    # per tile: group scan, chunk scan, exact recursions
    post = np.zeros((T, Q), F); la = np.zeros((T, Q), F); lb = np.zeros((T, Q), F)
    for p in range(ntile):
        cs = list(range(p * G, min(C, (p + 1) * G)))
        ng = len(groups[p])
        gf = [None] * ng; gr = [None] * ng
        cur = tile_f[p]
        for o in range(ng):
            gf[o] = cur
            v, e, m = fwd_matvec(cur[0], cur[1], groups[p][o]); cur = (v, e, cur[2] + float(m))
        cur = tile_r[p]
        for o in range(ng - 1, -1, -1):
            gr[o] = cur
            v, e, m = bwd_matvec(cur[0], cur[1], groups[p][o]); cur = (v, e, cur[2] + float(m))
        for o in range(ng):
            ccs = cs[o * OCT:(o + 1) * OCT]
            cf = {}; cr = {}
            cur = gf[o]
            for c in ccs:
                cf[c] = cur
                v, e, m = fwd_matvec(cur[0], cur[1], blocks[c]); cur = (v, e, cur[2] + float(m))
            cur = gr[o]
            for c in reversed(ccs):
                cr[c] = cur
                v, e, m = bwd_matvec(cur[0], cur[1], blocks[c]); cur = (v, e, cur[2] + float(m))
            for c in ccs:
                t0, t1 = c * L, min(T, (c + 1) * L) - 1
                fvec, fexp, flog = cf[c]; rvec, rexp, rlog = cr[c]
                # hand the recursions vectors whose sum is in [1,2)
                ef = int(pow2_floor_exp(fvec.sum(dtype=F))); fvec = ldexp32(fvec, -ef); fexp += ef
                er = int(pow2_floor_exp(rvec.max())); rvec = ldexp32(rvec, -er); rexp += er
                fr, fx = seq_forward(E, A, t0, t1, fvec, c == 0)
                rr, rx = seq_backward(E, A, t0, t1, rvec)
                g = np.maximum(fr * rr, F(1e-37))
                z = g.sum(1, keepdims=True, dtype=F)
                post[t0:t1 + 1] = np.log(g) - np.log(z)
                la[t0:t1 + 1] = np.log(fr) + ((fx + fexp)[:, None] * float(np.log(2.0)) + flog).astype(F)
                lb[t0:t1 + 1] = np.log(rr) + ((rx + rexp)[:, None] * float(np.log(2.0)) + rlog).astype(F)
    if no_loglik:
        post = post + F(loglik)
    return dict(log_post=post, log_alpha=la, log_beta=lb, loglik=loglik)
