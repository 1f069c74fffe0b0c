"""Generate ``tests/golden/*.npz`` by running the REFERENCE's own code.

TEST INFRASTRUCTURE ONLY; runs only in the build container (needs
``/root/reference``).  Usage: ``python -m oracle.make_golden``.

Every fixture holds seeded inputs plus what the unmodified reference returns:

* graph matrices from ``att_speech.fst_utils.CTCGraphGen`` (its own topology
  builders + ``fst_to_matrices`` + ``batch_training_graph_matrices``, executed
  on ``oracle/minifst.py``);
* ``path_reduction`` (``PathLogSumExp``) costs and autograd gradients, in fp32
  and -- under ``torch.set_default_dtype(float64)`` -- in fp64, composed with
  ``get_normalized_acts`` and the max-shift / numerator-minus-denominator glue
  of ``FSTDecoder.get_fst_loss`` (advanced_decoder.py:479-517, restated inline
  because the decoder class needs the whole model stack to instantiate);
* Viterbi scores and the per-frame arg-min labels of ``FSTDecoder.decode``
  (advanced_decoder.py:546-554);
* ``F.ctc_loss`` and ``RawGenericCTC`` (ctc_losses.py:268-324, numpy fp64)
  cross-values for the mono / contextual-blank cases.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from . import ref_loader

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

CASES = [
    # name, context_order, S, graph_build_args, normalize_by_dim, use_den, T, lens, labels
    ("mono_ctc", 1, 7, {}, 0, False, 18, [18, 15, 11],
     [[1, 2, 2, 3], [4, 4], [6, 1, 5]]),
    ("mono_ctcg", 1, 7, {}, None, True, 16, [16, 16, 9],
     [[3, 3, 1], [2, 5, 6, 1], [4]]),
    ("bi_ctc", 2, 5, {}, 0, False, 20, [20, 17, 12],
     [[1, 2, 2, 3], [4, 4, 4], [2, 1]]),
    ("bi_ctcg", 2, 5, {}, None, True, 20, [20, 16, 13],
     [[1, 2, 2, 3, 1], [4, 4, 4], [3]]),
    ("bi_ctcgb", 2, 5, {"use_contextual_blanks": True}, None, True, 22, [22, 22, 10],
     [[2, 2, 2, 1], [1, 3], [4, 2, 4]]),
    ("bi_ctcg_noloops", 2, 4, {"allow_nonblank_selfloops": False}, None, True, 14, [14, 12],
     [[1, 2, 3], [3, 3]]),
    ("bi_ctcg_looprep", 2, 4, {"loop_using_symbol_repetitions": True}, None, True, 14, [14, 11],
     [[1, 1, 2], [3, 2]]),
    ("bi_ctc_context_norm", 2, 5, {"use_contextual_blanks": True}, 1, False, 18, [18, 14],
     [[1, 2, 3], [4, 4]]),
    # tri-char (build_ctc_trigram_decoding_fst_v2); contextual blanks: the variant whose decoding graph is
    # input-deterministic, i.e. the one the reference can also use as a denominator
    ("tri_ctc", 3, 4, {"use_context_blanks": True}, 0, False, 16, [16, 13],
     [[1, 2, 2, 3], [3, 1]]),
    ("tri_ctcg", 3, 4, {"use_context_blanks": True}, None, True, 16, [16, 12],
     [[2, 1, 3], [1, 1, 2]]),
]

ONLY = None  # set to a list of case names to (re)generate just those


def _bigram_ids(labels, S):
    # egs/wsj/data.py:146-161: id = prev * S + cur
    out, last = [], 0
    for p in labels:
        out.append(last * S + p)
        last = p
    return out


def _run_reference(fu, cl, gen, logits, lens, num_mats, den_mats, normalize_by_dim, S, order, dtype):
    old = torch.get_default_dtype()
    torch.set_default_dtype(dtype)
    try:
        x = logits.to(dtype).clone().requires_grad_(True)
        cast = lambda ms: [m.to(dtype) if m.is_floating_point() else m for m in ms]  # noqa: E731
        acts = x
        if normalize_by_dim is not None:
            acts = cl.get_normalized_acts(x, None, S, order, normalize_by_dim, normalize_logits=True)
        lens_t = torch.tensor(lens)
        T = x.size(0)
        mask = (torch.arange(T, dtype=lens_t.dtype).unsqueeze(1) < lens_t).unsqueeze(-1).to(dtype)
        mx = acts.max(-1, keepdim=True)[0].detach()
        logits_sum = (mx * mask).sum(0).squeeze(-1)
        shifted = acts - mx
        num = -fu.path_reduction(shifted, lens_t, cast(num_mats), red_kind="logsumexp", neg_inf=gen.nc_weight)
        if den_mats is not None:
            den = -fu.path_reduction(shifted, lens_t, cast(den_mats), red_kind="logsumexp",
                                     neg_inf=gen.nc_weight)
        else:
            den = logits_sum
        loss = num - den
        loss.sum().backward()
        return (loss.detach().double().numpy(), x.grad.detach().double().numpy(),
                (num + logits_sum).detach().double().numpy(),
                ((den + logits_sum) if den_mats is not None else torch.zeros_like(num)).detach().double().numpy())
    finally:
        torch.set_default_dtype(old)


def make_case(fu, cl, name, order, S, args, normalize_by_dim, use_den, T, lens, labels, seed):
    V = S ** order
    g = torch.Generator().manual_seed(seed)
    logits = (torch.randn(T, len(lens), V, generator=g) * 3).float()
    gen = fu.CTCGraphGen(context_order=order, graph_build_args=dict(args), num_symbols=S)
    B = len(lens)
    U = max(len(l) for l in labels)
    texts = torch.zeros(B, U, dtype=torch.long)
    for i, l in enumerate(labels):
        ids = _bigram_ids(l, S) if order == 2 else l  # bi-char datasets feed prev*S+cur
        texts[i, :len(l)] = torch.tensor(ids)
    text_lens = torch.tensor([len(l) for l in labels])
    num_mats = gen.get_training_matrices_batch(texts, text_lens)
    den_mats = gen.get_decoding_matrices("cpu")
    out = {"logits": logits.numpy(), "lens": np.array(lens), "S": S, "order": order,
           "normalize_by_dim": -1 if normalize_by_dim is None else normalize_by_dim,
           "use_den": int(use_den), "texts": texts.numpy(), "text_lens": text_lens.numpy(),
           "symbols": np.array([x for l in labels for x in l]),
           "args": np.array(["%s=%d" % (k, int(v)) for k, v in sorted(args.items())], dtype="U60")}
    for i, m in enumerate(num_mats):
        out["num%d" % i] = m.numpy()
    for i, m in enumerate(den_mats):
        out["den%d" % i] = m.numpy()
    for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
        loss, grad, ncost, dcost = _run_reference(fu, cl, gen, logits, lens, num_mats,
                                                  den_mats if use_den else None,
                                                  normalize_by_dim, S, order, dt)
        out["loss_" + tag] = loss
        out["grad_" + tag] = grad
        out["numloss_" + tag] = ncost
        out["denloss_" + tag] = dcost
    # Viterbi decode read-out on the decoding graph (advanced_decoder.py:546-554), fp64
    old = torch.get_default_dtype()
    torch.set_default_dtype(torch.float64)
    try:
        x = logits.double().clone().requires_grad_(True)
        dm = [m.double() if m.is_floating_point() else m for m in den_mats]
        score = fu.path_reduction(x, torch.tensor(lens), dm, red_kind="viterbi", neg_inf=gen.nc_weight)
        (-score.sum()).backward()
        out["vit_score"] = score.detach().numpy()
        out["vit_labels"] = x.grad.min(-1)[1].numpy()
        # log #paths: path_reduction on all-zero log-probs
        z = torch.zeros_like(x)
        nm = [m.double() if m.is_floating_point() else m for m in num_mats]
        out["log_count_num"] = fu.path_reduction(z, torch.tensor(lens), nm).detach().numpy()
        out["log_count_den"] = fu.path_reduction(z, torch.tensor(lens), dm).detach().numpy()
    finally:
        torch.set_default_dtype(old)
    # cross-values from the reference's OpenFst-free definitions
    if order == 1 and normalize_by_dim == 0:
        lp = torch.log_softmax(logits.double(), -1)
        tgt = torch.tensor([x for l in labels for x in l])
        out["F_ctc_loss"] = torch.nn.functional.ctc_loss(
            lp, tgt, torch.tensor(lens), text_lens, reduction="none").numpy()
        raw = []
        for b in range(B):
            A, Bm = cl.get_CTC_matrices_mono(torch.tensor(labels[b]), S)
            raw.append(float(cl.raw_generic_ctc(lp[:lens[b], b], A, Bm)))
        out["raw_generic_ctc"] = np.array(raw)
    if order == 2 and normalize_by_dim == 1:
        lp = cl.get_normalized_acts(logits.double(), None, S, 2, 1)
        raw = []
        for b in range(B):
            A, Bm = cl.get_CTC_matrices_bicontext(torch.tensor(labels[b]), S)
            raw.append(float(cl.raw_generic_ctc(lp[:lens[b], b], A, Bm)))
        out["raw_generic_ctc"] = np.array(raw)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    return out


DENSE_VARIANTS = {          # keyword arguments of bi_CTC_loss (ctc_losses.py:492-502) -> fixture key
    "plain": {},
    "noloops": {"allow_nonblank_selfloops": False},
    "repctx": {"eval_repeats_in_context": True},
    "looprepeats": {"loop_using_symbol_repetitions": True},
}


def make_dense_variants(cl, name="dense_variants", seed=2024):
    """The dense route's bi-context state machines (get_CTC_matrices_bicontext with its flags, and
    _bicontext_looprepeats, ctc_losses.py:111-226) run by the reference's own RawGenericCTC (numpy fp64,
    :268-324, through bi_CTC_loss :492-502) on per-context-normalised scores: loss values per variant."""
    g = torch.Generator().manual_seed(seed)
    S, T, B = 5, 40, 4
    lens = [40, 36, 31, 22]
    labels = [[1, 1, 3, 2, 2, 2, 4], [4, 3, 3, 1], [2, 4, 4, 4, 1, 3], [3]]
    logits = torch.randn(T, B, S * S, generator=g) * 2
    lp = cl.get_normalized_acts(logits.double(), None, S, 2, 1)
    out = dict(logits=logits.numpy(), lens=np.array(lens), S=S, order=2, normalize_by_dim=1,
               labels_cat=np.array([v for l in labels for v in l]), text_lens=np.array([len(l) for l in labels]))
    for key, kw in DENSE_VARIANTS.items():
        out["raw_" + key] = np.array([float(cl.bi_CTC_loss(lp[:lens[b], b], torch.tensor(labels[b]), **kw))
                                      for b in range(B)])
    os.makedirs(os.path.join(OUT, "dense"), exist_ok=True)     # a directory of its own: not one of the FST-route cases
    np.savez_compressed(os.path.join(OUT, "dense", name + ".npz"), **out)
    return out


def main():
    fu, cl = ref_loader.load()
    os.makedirs(OUT, exist_ok=True)
    import sys
    only = sys.argv[1:] or ONLY
    if not only or "dense_variants" in only:
        out = make_dense_variants(cl)
        print("dense_variants", {k: np.round(v, 4) for k, v in out.items() if k.startswith("raw_")})
    for i, case in enumerate(CASES):
        if only and case[0] not in only:
            continue
        out = make_case(fu, cl, *case, seed=1000 + i)
        print(case[0], "loss_f64", np.round(out["loss_f64"], 4),
              "f32-f64 grad diff %.2e" % np.abs(out["grad_f32"] - out["grad_f64"]).max(),
              {k: np.round(v, 4) for k, v in out.items() if k in ("F_ctc_loss", "raw_generic_ctc")})


if __name__ == "__main__":
    main()
