"""Torch-CPU restatement of the reference's path-sum engine and loss glue.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Used as the parity
checker (run in float64) and as the timed CPU baseline (float32, all host
threads) -- never by the product.

Follows, in ``/root/reference``:

* ``att_speech/fst_utils.py:400-488``  ``PathLogSumExp`` -> :func:`path_logsumexp`
* ``att_speech/fst_utils.py:322-397``  ``path_reduction`` (viterbi branch)
  + ``advanced_decoder.py:546-554``    -> :func:`path_viterbi`
* ``att_speech/modules/ctc_losses.py:29-43`` -> :func:`normalized_acts`
* ``att_speech/modules/decoders/advanced_decoder.py:479-517`` -> :func:`fst_loss`
* SURVEY A11 (log #paths == path_reduction on zero log-probs) -> :func:`path_count`

The padded graph format is the reference's: a list of 8 arrays
``[pred, lab_in, w_in, final, succ, lab_out, w_out, final]`` with leading
dimension B or 1, padding ``(state 0, label 0, weight -1e20)``.
"""
from __future__ import annotations

import itertools
import math

import numpy as np
import torch

NEG = -1e20


def _as_torch(mats, float_dtype):
    out = []
    for m in mats:
        t = torch.as_tensor(np.asarray(m)) if not torch.is_tensor(m) else m
        if t.is_floating_point():
            t = t.to(float_dtype)
        else:
            t = t.to(torch.int64)
        out.append(t)
    return out


def _active_counts(act_lens, T):
    """bitem(t) = number of utterances still running at frame t (lens sorted desc)."""
    lens = [int(x) for x in act_lens]
    assert lens == sorted(lens, reverse=True), "lengths must be sorted descending"
    return [sum(1 for l in lens if l > t) for t in range(T)]


def path_logsumexp(log_probs, act_lens, mats, dtype=torch.float32, neg_inf=NEG):
    """Forward-backward over a padded graph: returns ``(cost (B,), grad (T,B,L))``.

    ``cost_b = log sum over accepting length-len_b paths`` and
    ``grad = d cost / d log_probs`` (fst_utils.py:424-480).  Same tensor-op
    structure as the reference (gather / logsumexp / scatter_add on the
    shrinking active prefix), so timing it is timing the reference's algorithm.
    """
    lp = torch.as_tensor(log_probs).detach().to(dtype)
    T, bs, _ = lp.shape
    mats = _as_torch(mats, dtype)
    if mats[0].shape[0] == 1 and bs > 1:
        mats = [m.expand(bs, *m.shape[1:]) for m in mats]
    pred, lab_in, w_in, term, succ, lab_out, w_out, _ = mats
    term = term.squeeze(-1)
    n = pred.shape[1]
    counts = _active_counts(act_lens, T)

    def pick(src, idx):
        return torch.gather(src, 1, idx.reshape(idx.shape[0], -1)).reshape(idx.shape)

    alpha = torch.full((bs, n), neg_inf, dtype=dtype)
    alpha[:, 0] = 0
    alpha0 = alpha.clone()
    alphas = torch.full((T, bs, n), neg_inf, dtype=dtype)
    for t in range(T):
        k = counts[t]
        if k == 0:
            break
        alphas[t] = alpha
        tok = w_in[:k] + pick(alpha[:k], pred[:k]) + pick(lp[t, :k], lab_in[:k])
        alpha[:k] = torch.logsumexp(tok, dim=-1)
    cost = torch.logsumexp(alpha + term, dim=-1)

    beta = term.clone()
    grad = torch.zeros_like(lp)
    for t in range(T - 1, -1, -1):
        k = counts[t]
        if k == 0:
            continue
        tok = w_out[:k] + pick(beta[:k], succ[:k]) + pick(lp[t, :k], lab_out[:k])
        beta[:k] = torch.logsumexp(tok, dim=-1)
        post = (tok + (alphas[t, :k] - cost[:k, None])[:, :, None]).exp()
        grad[t, :k].scatter_add_(1, lab_out[:k].reshape(k, -1), post.reshape(k, -1))
    cost0 = torch.logsumexp(alpha0 + beta, dim=1)
    return cost, grad, cost0


def path_viterbi(log_probs, act_lens, mats, dtype=torch.float64, neg_inf=NEG):
    """Tropical path reduction with an explicit back-trace.

    Returns ``(score (B,), frame_labels (T,B) int64)``: the best accepting
    path's log-weight and the label consumed at every frame (0 for frames at or
    beyond ``len_b``, which is what ``logits.grad.min(-1)[1]`` yields there,
    advanced_decoder.py:554).  Ties resolve to the first maximal entry of the
    per-state arc list, the order ``torch.max`` follows (fst_utils.py:369-370).
    """
    lp = torch.as_tensor(log_probs).detach().to(dtype)
    T, bs, _ = lp.shape
    mats = _as_torch(mats[:4], dtype)
    if mats[0].shape[0] == 1 and bs > 1:
        mats = [m.expand(bs, *m.shape[1:]) for m in mats]
    pred, lab_in, w_in, term = mats
    term = term.squeeze(-1)
    n = pred.shape[1]
    counts = _active_counts(act_lens, T)
    alpha = torch.full((bs, n), neg_inf, dtype=dtype)
    alpha[:, 0] = 0
    back = torch.zeros((T, bs, n), dtype=torch.int64)
    for t in range(T):
        k = counts[t]
        if k == 0:
            break
        tok = (w_in[:k] + torch.gather(alpha[:k], 1, pred[:k].reshape(k, -1)).reshape(pred[:k].shape)
               + torch.gather(lp[t, :k], 1, lab_in[:k].reshape(k, -1)).reshape(pred[:k].shape))
        val, idx = _first_max(tok)
        alpha[:k] = val
        back[t, :k] = idx
    fin = alpha + term
    score, state = _first_max(fin)
    labels = torch.zeros((T, bs), dtype=torch.int64)
    for b in range(bs):
        s = int(state[b])
        for t in range(int(act_lens[b]) - 1, -1, -1):
            kk = int(back[t, b, s])
            labels[t, b] = lab_in[b, s, kk]
            s = int(pred[b, s, kk])
    return score, labels


def _first_max(x):
    val = x.max(dim=-1)[0]
    is_max = (x == val.unsqueeze(-1))
    idx = torch.argmax(is_max.to(torch.int8), dim=-1)  # first True
    return val, idx


def padded_to_arcs(mats, b=0, neg_inf=NEG):
    """Arc list ``(src, dst, label, weight)`` + final weights of graph ``b``
    from the in-edge half of a padded tuple; padding (weight <= neg_inf/2) dropped."""
    pred = np.asarray(mats[0])[b]
    lab = np.asarray(mats[1])[b]
    w = np.asarray(mats[2])[b].astype(np.float64)
    fin = np.asarray(mats[3])[b].astype(np.float64).reshape(-1)
    arcs = []
    n, k = pred.shape
    for d in range(n):
        for j in range(k):
            if w[d, j] > neg_inf / 2:
                arcs.append((int(pred[d, j]), d, int(lab[d, j]), float(w[d, j])))
    fin = np.where(fin > neg_inf / 2, fin, -np.inf)
    return n, arcs, fin


def path_count(n, arcs, final, length):
    """Number of accepting paths of exactly ``length`` arcs, modulo 2**64."""
    mask = (1 << 64) - 1
    cnt = [0] * n
    cnt[0] = 1
    for _ in range(length):
        new = [0] * n
        for (s, d, _l, _w) in arcs:
            new[d] = (new[d] + cnt[s]) & mask
        cnt = new
    tot = 0
    for s in range(n):
        if np.isfinite(final[s]):
            tot = (tot + cnt[s]) & mask
    return tot


def normalized_acts(acts, num_symbols, context_order, normalize_by_dim):
    """ctc_losses.py:29-43 (``normalize_logits=True``)."""
    if normalize_by_dim:
        shape = acts.shape
        v = acts.reshape(shape[0], shape[1], *((num_symbols,) * context_order))
        v = torch.log_softmax(v, dim=normalize_by_dim + 2)
        return v.reshape(shape)
    return torch.log_softmax(acts, dim=-1)


def fst_loss(logits, act_lens, num_mats, den_mats=None, normalize_by_dim=None,
             num_symbols=None, context_order=1, dtype=torch.float32):
    """``FSTDecoder.logits`` + ``get_fst_loss`` (advanced_decoder.py:444-517).

    ``normalize_by_dim`` None -> raw logits (CTC-G/GB), 0 -> softmax over V,
    k>0 -> softmax within context blocks.  ``den_mats`` None corresponds to
    ``denominator_red: none``.  Returns ``(loss (B,), dloss.sum()/dlogits)``.
    """
    x = torch.as_tensor(logits).detach().to(dtype).requires_grad_(True)
    T, bs, _ = x.shape
    lens = torch.as_tensor([int(v) for v in act_lens])
    acts = x
    if normalize_by_dim is not None:
        acts = normalized_acts(x, num_symbols, context_order, normalize_by_dim)
    mask = (torch.arange(T)[:, None] < lens[None, :]).to(dtype)[:, :, None]
    mx = acts.max(-1, keepdim=True)[0].detach()
    logits_sum = (mx * mask).sum(0).squeeze(-1)
    shifted = acts - mx
    ncost, ngrad, _ = path_logsumexp(shifted, act_lens, num_mats, dtype=dtype)
    loss = -ncost
    g = -ngrad
    if den_mats is not None:
        dcost, dgrad, _ = path_logsumexp(shifted, act_lens, den_mats, dtype=dtype)
        loss = loss + dcost
        g = g + dgrad
    else:
        loss = loss - logits_sum
    # chain through (acts - max) and the optional normalisation
    shifted.backward(g)
    return loss.detach(), x.grad.detach()


def brute_force(lp, n, arcs, final):
    """Exhaustive enumeration on one utterance: ``lp`` is ``(T, V)`` float64.

    Returns ``dict(cost, grad (T,V), best_score, best_labels, count)``.
    Only for tiny cases (paths grow as K**T).
    """
    lp = np.asarray(lp, dtype=np.float64)
    T, V = lp.shape
    out = [[] for _ in range(n)]
    for a in arcs:
        out[a[0]].append(a)
    paths = []  # (logweight, labels)

    def rec(state, t, w, labs):
        if t == T:
            if np.isfinite(final[state]):
                paths.append((w + final[state], tuple(labs)))
            return
        for (_s, d, lab, aw) in out[state]:
            labs.append(lab)
            rec(d, t + 1, w + aw + lp[t, lab], labs)
            labs.pop()

    rec(0, 0, 0.0, [])
    if not paths:
        return dict(cost=-np.inf, grad=np.zeros((T, V)), best_score=-np.inf,
                    best_labels=None, count=0)
    ws = np.array([p[0] for p in paths])
    m = ws.max()
    cost = m + math.log(np.exp(ws - m).sum())
    grad = np.zeros((T, V))
    for w, labs in paths:
        p = math.exp(w - cost)
        for t, lab in enumerate(labs):
            grad[t, lab] += p
    best = int(np.argmax(ws))
    return dict(cost=cost, grad=grad, best_score=ws[best],
                best_labels=np.array(paths[best][1]), count=len(paths))
