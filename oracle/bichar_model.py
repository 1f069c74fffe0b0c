"""TEST INFRASTRUCTURE (oracle side): numpy statement of the factorised bi-char
denominator exactly as the CUDA sweeps compute it (csrc/kernels_bisweep.cu), so that the
recurrences, the power-of-two rescaling, the checkpoint indexing and the combine formula can
be checked on the CPU against the reference algorithm (oracle/engine.path_logsumexp on the
matrices of build_ctc_bigram_decoding_fst, fst_utils.py:759-817) before any GPU time is spent.

State (c, l) = (context, last letter); class of the state cls(c, l) = c*S + l, except that
all (c, 0) share class 0 unless blanks are contextual.  With e_t = exp(x_t - max x_t):

  forward  sweep  A'[c][l] = e[c][l] (P[c] + m(c,l) A[c][l])          l >= 1
                  P'[l]    = sum_c A'[c][l] + e0[l] P[l]              (P'[0] = e0[0] P[0])
  backward sweep  F'[c][l] = e[c][l] (Q[l] + m(c,l) F[c][l])          l >= 1
                  Q'[c]    = sum_{l>=1} F'[c][l] + e0[c] Q[c]
  posterior       gamma_t[cls(c,l)] = e_t (P_t[c] + m A_t) (Q_{t+1}[l] + m F_{t+1}) / Z   l >= 1
                  gamma_t[cls(c,0)] += e0_t[c] P_t[c] Q_{t+1}[c] / Z

m(c,l) = allow_nonblank_selfloops and l != 0 and l != c.  Only tests import this module.
"""
from __future__ import annotations

import numpy as np

TARGET = 60
BAND = 20


def _flags(allow_nonblank_selfloops=True, use_contextual_blanks=False, loop_using_symbol_repetitions=False):
    loops, ctx = bool(allow_nonblank_selfloops), bool(use_contextual_blanks)
    if loop_using_symbol_repetitions:
        loops, ctx = False, True
    return loops, ctx


def _exponent(v):
    return int(np.floor(np.log2(v)))


def sweep(x, S, loops, ctx, K, reverse, dtype=np.float32):
    """One sweep over one utterance ``x`` (len, V).  Returns the spilled vectors R_t (the value
    *before* the step at frame t), the per-frame rescale exponents k_t, the cumulative exponent
    of the spilled vector, the checkpoints of the per-state values and the final vector."""
    n, V = x.shape
    m = x.max(axis=1)
    mask = np.zeros((S, S), dtype)
    if loops:
        for c in range(S):
            for l in range(1, S):
                if l != c:
                    mask[c, l] = 1
    val = np.zeros((S, S), dtype)           # A[c][l] or F[c][l], column l = 0 unused
    R = np.zeros(S, dtype)
    if reverse:
        R[:] = dtype(2.0) ** TARGET
    else:
        R[0] = dtype(2.0) ** TARGET
    spill = np.zeros((n, S), dtype)
    kexp = np.zeros(n, np.int64)
    cum = np.zeros(n, np.int64)             # exponent of spill[t]
    ckpt = {}
    Kcum = TARGET
    k_pending = [0, 0]                      # k of this frame and the next (decided with a lag of two)
    order = range(n - 1, -1, -1) if reverse else range(n)
    for step, t in enumerate(order):
        # lagged decision for the frame after next, from the vector entering this frame
        mx = float(R.max())
        knext2 = 0
        if mx > 0 and np.isfinite(mx):
            ex = _exponent(mx) + k_pending[0] + k_pending[1]
            if ex > TARGET + BAND or ex < TARGET - BAND:
                knext2 = max(-120, min(120, TARGET - ex))
        k = k_pending[0]
        k_pending = [k_pending[1], knext2]
        spill[t] = R
        cum[t] = Kcum
        kexp[t] = k
        if loops and ((t % K == K - 1) if reverse else (t % K == 0 and t > 0)):
            ckpt[t // K] = val.copy()
        e = np.exp2((x[t].astype(dtype) * dtype(1.4426950408889634) - dtype(m[t]) * dtype(1.4426950408889634))).astype(dtype)
        e = e * dtype(2.0) ** k
        E = e.reshape(S, S)
        e0 = E[:, 0].copy() if ctx else np.full(S, e[0], dtype)
        if reverse:
            val = E * (R[None, :] + mask * val)       # F'[c][l] = e (Q[l] + m F)
            val[:, 0] = 0
            R = val.sum(axis=1) + e0 * R              # Q'[c]
        else:
            val = E * (R[:, None] + mask * val)       # A'[c][l] = e (P[c] + m A)
            val[:, 0] = 0
            R = val.sum(axis=0) + e0 * R              # P'[l]
        Kcum += k
    return dict(spill=spill, k=kexp, cum=cum, ckpt=ckpt, final=R, Kfinal=Kcum, m=m)


def combine(x, S, loops, ctx, K, fw, bw, dtype=np.float32):
    """Posterior rows from the two sweeps' spills, segment by segment (K frames), the way the
    combine kernel walks them: forward through the segment from the A checkpoint, backward
    from the F checkpoint."""
    n, V = x.shape
    mass = float(fw["final"].sum())
    gamma = np.zeros((n, V), np.float64)
    mask = np.zeros((S, S), dtype)
    if loops:
        for c in range(S):
            for l in range(1, S):
                if l != c:
                    mask[c, l] = 1
    L2E = dtype(1.4426950408889634)
    for j in range((n + K - 1) // K):
        t0, t1 = j * K, min((j + 1) * K, n)
        A = fw["ckpt"][j].copy() if (loops and j > 0) else np.zeros((S, S), dtype)
        Ahat = {}
        es = {}
        for t in range(t0, t1):
            e = np.exp2(x[t].astype(dtype) * L2E - dtype(fw["m"][t]) * L2E).astype(dtype)
            es[t] = e
            E = (e * dtype(2.0) ** int(fw["k"][t])).reshape(S, S)
            Ahat[t] = E * (fw["spill"][t][:, None] + mask * A)      # a_{t+1}[c][l], l >= 1
            A = Ahat[t].copy()
            A[:, 0] = 0
        Fv = bw["ckpt"][j].copy() if (loops and t1 < n) else np.zeros((S, S), dtype)
        for t in range(t1 - 1, t0 - 1, -1):
            e = es[t]
            Q = bw["spill"][t]
            Bhat = Q[None, :] + mask * Fv                           # b_{t+1}[c][l], l >= 1
            # exponent: a_{t+1} carries Ka_{t+1} = cum_f[t] + k_f[t]; b_{t+1} carries cum_b[t]
            Eexp = fw["Kfinal"] - (int(fw["cum"][t]) + int(fw["k"][t])) - int(bw["cum"][t])
            g = Ahat[t].astype(np.float64) * Bhat.astype(np.float64) * (2.0 ** Eexp) / mass
            g[:, 0] = 0
            row = g.reshape(-1).copy()
            e0 = (e.reshape(S, S)[:, 0] if ctx else np.full(S, e[0], dtype)) * dtype(2.0) ** int(fw["k"][t])
            gb = e0.astype(np.float64) * fw["spill"][t] * Q * (2.0 ** Eexp) / mass
            if ctx:
                row[np.arange(S) * S] += gb
            else:
                row[0] += gb.sum()
            gamma[t] = row
            Fv = (e * dtype(2.0) ** int(bw["k"][t])).reshape(S, S) * Bhat
            Fv[:, 0] = 0
    return gamma


def denominator(x, S, K=8, dtype=np.float32, **flags):
    """cost (natural log) and posterior rows of one utterance."""
    loops, ctx = _flags(**flags)
    fw = sweep(x, S, loops, ctx, K, False, dtype)
    bw = sweep(x, S, loops, ctx, K, True, dtype)
    mass = float(fw["final"].astype(np.float64).sum())
    cost = np.log(mass) - fw["Kfinal"] * np.log(2.0) + float(fw["m"].astype(np.float64).sum())
    # the backward sweep ends in Q_0; the start state is (0, 0): b_0[(0,0)] = Q_0[0]
    cost_b = np.log(float(bw["final"][0])) - bw["Kfinal"] * np.log(2.0) + float(bw["m"].astype(np.float64).sum())
    return cost, combine(x, S, loops, ctx, K, fw, bw, dtype), cost_b
