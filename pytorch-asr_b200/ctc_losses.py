"""Host-side mirror of the FST-based entries of ``att_speech/modules/ctc_losses.py``
and of the loss glue of ``FSTDecoder`` (``modules/decoders/advanced_decoder.py``).

* :func:`get_normalized_acts`  <- ctc_losses.py:29-43
* :func:`ctc_fst_loss`         <- ctc_losses.py:617-657   (``ctc_loss_fn`` plugin signature)
* :func:`ctc_fst_global_loss`  <- ctc_losses.py:663-717
* :func:`fst_loss`             <- advanced_decoder.py:444-517 fused into one library call
* :class:`FSTLossMixin`        <- ``FSTDecoder.get_fst_loss`` / ``forward`` (:454-534)

The fused path never materialises ``log_softmax(logits)``, the max-shifted copy,
``lalphas`` or the dense numerator gradient: the library reads the logits and
writes d(loss)/d(logits) directly.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from .fst_utils import CTCGraphGen, PackedGraphs, _as_graph, _check_lens, path_reduction


def get_normalized_acts(acts, acts_lens, num_symbols, context_order, normalize_by_dim,
                        normalize_logits=True):
    """ctc_losses.py:29-43 (plain torch: only used on the non-fused entry points)."""
    assert context_order == 1 or num_symbols
    del acts_lens
    if normalize_by_dim:
        act_size = acts.size()
        new_act_size = (act_size[:2] + (num_symbols,) * context_order)
        acts = acts.view(*new_act_size)
        acts = torch.nn.functional.log_softmax(acts, normalize_by_dim + 2)
        acts = acts.view(*act_size)
    elif normalize_logits:
        acts = torch.nn.functional.log_softmax(acts, dim=-1)
    return acts


def _norm_mode(normalize_by_dim):
    if normalize_by_dim is None:
        return _lib.NORM_NONE
    if normalize_by_dim == 0:
        return _lib.NORM_SOFTMAX
    return _lib.NORM_CONTEXT


class _FstLoss(torch.autograd.Function):
    """loss_b = -cost_num_b(lp) [+ cost_den_b(lp)], lp = normalise(logits); one call."""

    @staticmethod
    def forward(ctx, logits, act_lens, num, den, normalize, num_symbols):
        loss, grad = _lib.loss_fwd_bwd(logits, act_lens, num, den, normalize, num_symbols,
                                       need_grad=logits.requires_grad)
        ctx.grads = grad
        return loss

    @staticmethod
    def backward(ctx, grad_output):
        g = ctx.grads
        ctx.grads = None  # single backward, like the reference's ctx.grads (fst_utils.py:473,484)
        # grad_output scales the rows on the device; the usual .sum().backward() (all ones) touches nothing
        return _lib.scale_grad(g, grad_output), None, None, None, None, None


def fst_loss(logits, act_lens, numerator, denominator=None, normalize_by_dim=None,
             num_symbols=0):
    """Per-utterance loss of ``FSTDecoder.get_fst_loss`` on raw ``fc`` outputs.

    ``numerator`` / ``denominator``: native graphs or the reference's padded
    matrices; ``denominator=None`` is ``denominator_red: none``.
    """
    lens = _check_lens(act_lens)
    num = _as_graph(numerator)
    den = _as_graph(denominator) if denominator is not None else None
    assert num.num_graphs in (1, logits.size(1))
    return _fst_loss(logits, lens, num, den, _norm_mode(normalize_by_dim), int(num_symbols or 0))


def _fst_loss(logits, lens, num, den, mode, num_symbols):
    # (per-context softmax together with a denominator, ctc_losses.py:29-43 feeding advanced_decoder.py:497-500,
    # is handled inside ctcg_loss_fwd_bwd since round 2: normalised copy, loss, gradient pulled back in place)
    return _FstLoss.apply(logits, lens, num, den, mode, num_symbols)


def _labels_to_batch(labels, label_lens):
    seqs, ls = [], 0
    labels = labels.cpu().numpy() if torch.is_tensor(labels) else np.asarray(labels)
    for b in range(len(label_lens)):
        le = ls + int(label_lens[b])
        seqs.append(labels[ls:le])
        ls = le
    return seqs


def _numerator_from_batch(other_data_in_batch, labels, label_lens, S, context_order, gg_kwargs):
    if other_data_in_batch and "graph_matrices" in other_data_in_batch:
        return _as_graph(other_data_in_batch["graph_matrices"])
    graph_gen = CTCGraphGen(context_order=context_order, graph_build_args=gg_kwargs,
                            num_symbols=S, for_forward_only=False)
    return graph_gen.get_training_graphs(_labels_to_batch(labels, label_lens))


def ctc_fst_loss(acts, labels, act_lens, label_lens, num_symbols=0, context_order=1,
                 normalize_by_dim=None, allow_nonblank_selfloops=True,
                 loop_using_symbol_repetitions=False, other_data_in_batch=None,
                 eval_repeats_in_context=False, neg_inf=-1e20):
    """ctc_losses.py:617-657: normalise, build/fetch the training graphs, -path sum."""
    assert not (eval_repeats_in_context and loop_using_symbol_repetitions)
    S = acts.size(2)
    gg_kwargs = {}
    if context_order == 2:
        S = int(round(np.sqrt(S)))
        gg_kwargs = dict(allow_nonblank_selfloops=allow_nonblank_selfloops,
                         loop_using_symbol_repetitions=loop_using_symbol_repetitions,
                         eval_repeats_in_context=eval_repeats_in_context)
    num = _numerator_from_batch(other_data_in_batch, labels, label_lens, S, context_order, gg_kwargs)
    # the reference always normalises here (normalize_logits=True): None -> softmax over V
    mode = _lib.NORM_SOFTMAX if not normalize_by_dim else _lib.NORM_CONTEXT
    # per-context softmax (normalize_by_dim=1) is fused too: block LSE in the gather, chain rule in
    # the row pass; the block size is the number of symbols
    return _FstLoss.apply(acts, _check_lens(act_lens), num, None, mode, int(num_symbols or S))


_ctc_fst_cache = {}


def ctc_fst_global_loss(acts, labels, act_lens, label_lens, num_symbols=0, context_order=1,
                        normalize_by_dim=None, allow_nonblank_selfloops=True,
                        loop_using_symbol_repetitions=False, other_data_in_batch=None,
                        eval_repeats_in_context=False, neg_inf=-1e20):
    """ctc_losses.py:663-717: numerator minus the decoding-graph denominator on raw acts."""
    S = acts.size(2)
    key = (context_order, S, allow_nonblank_selfloops, loop_using_symbol_repetitions,
           eval_repeats_in_context)
    gg_kwargs = {}
    if context_order == 2:
        S = int(round(np.sqrt(S)))
        gg_kwargs = dict(allow_nonblank_selfloops=allow_nonblank_selfloops,
                         loop_using_symbol_repetitions=loop_using_symbol_repetitions,
                         eval_repeats_in_context=eval_repeats_in_context)
    if key not in _ctc_fst_cache:
        _ctc_fst_cache[key] = _lib.Graph.decoding(context_order, S, **gg_kwargs)
    den = _ctc_fst_cache[key]
    num = _numerator_from_batch(other_data_in_batch, labels, label_lens, S, context_order, gg_kwargs)
    return _FstLoss.apply(acts, _check_lens(act_lens), num, den, _lib.NORM_NONE, 0)


def dense_ctc_graph(labels, S, context_order, allow_nonblank_selfloops=True, eval_repeats_in_context=False,
                    loop_using_symbol_repetitions=False):
    """The state machine of ``get_CTC_matrices_mono`` / ``get_CTC_matrices_bicontext`` /
    ``get_CTC_matrices_bicontext_looprepeats`` (ctc_losses.py:67-94, :111-166, :169-226) run by ``RawGenericCTC``
    (:268-324) as an arc graph for the path-sum kernels.  The dense states emit on the STATE, so state j becomes
    node j+1 and every transition i -> j an arc labelled with j's emission; node 0 is a start node with arcs to the
    two entry states (``enter_first_num=2``); the last two states are final (``terminal_last_num=2``; three for
    the loop-repeats machine, :498-501).  States: [blank_0, tok_0, blank_1, ..., tok_{U-1}, blank_U], or for
    loop-repeats [blank, (entry_i, looping_i, blank_i) ...] where a symbol is entered in the old context, held
    as the class (l, l) and followed by the blank of its own context.
    Returns ``(n_nodes, arcs [(src, dst, label, 0.0)], final)``.  This is NOT the FST topology for
    bi-char transcripts with a repeated letter (the dense definition forbids tok -> tok there)."""
    L = [int(v) % S for v in labels]
    U = len(L)
    trans = []
    n_final = 2
    if context_order == 2 and loop_using_symbol_repetitions:
        assert not eval_repeats_in_context                      # :123
        N = 3 * U + 1
        emit = [0] * N
        n_final = 3
        trans += [(0, 0), (0, 1)] if U else [(0, 0)]
        context = 0
        for i, label in enumerate(L):
            entry = 3 * i + 1
            looping, blank, nxt = entry + 1, entry + 2, entry + 3
            emit[entry] = context * S + label
            context = label
            emit[looping] = context * S + label
            emit[blank] = context * S
            can_next = i + 1 < U and label != L[i + 1]
            trans += [(entry, looping), (entry, blank)] + ([(entry, nxt)] if can_next else [])
            trans += [(looping, looping), (looping, blank)] + ([(looping, nxt)] if can_next else [])
            trans.append((blank, blank))
            if i + 1 < U:
                trans.append((blank, nxt))
    elif context_order == 1:
        N = 2 * U + 1
        emit = [0] * N
        trans += [(0, 0), (0, 1)] if U else [(0, 0)]
        for i, l in enumerate(L):
            tok = 2 * i + 1
            emit[tok] = l
            trans += [(tok, tok), (tok, tok + 1), (tok + 1, tok + 1)]
            if tok + 2 < N:
                trans.append((tok + 1, tok + 2))
                if L[i + 1] != l:
                    trans.append((tok, tok + 2))
    else:
        N = 2 * U + 1
        emit = [0] * N
        context = 0
        for i, li in enumerate(L):
            blank, tok = 2 * i, 2 * i + 1
            emit[blank] = context * S
            emit[tok] = context * S + li
            trans += [(blank, blank), (blank, tok)]
            if allow_nonblank_selfloops:
                trans.append((tok, tok))
            trans.append((tok, tok + 1))
            if tok + 2 < N:
                rep = ((context * S + li) == (li * S + L[i + 1])) if eval_repeats_in_context else (li == L[i + 1])
                if not rep:
                    trans.append((tok, tok + 2))
            context = li
        emit[N - 1] = context * S
        trans.append((N - 1, N - 1))
    arcs = [(0, j + 1, emit[j], 0.0) for j in range(min(2, N))]
    arcs += [(i + 1, j + 1, emit[j], 0.0) for (i, j) in trans]
    final = np.full(N + 1, -np.inf, dtype=np.float32)
    final[max(1, N + 1 - n_final):] = 0.0
    return N + 1, arcs, final


def ctc_raw_loss(acts, labels, act_lens, label_lens, num_symbols=0, context_order=1, normalize_by_dim=None,
                 allow_nonblank_selfloops=True, loop_using_symbol_repetitions=False,
                 eval_repeats_in_context=False, other_data_in_batch=None, **kwargs):
    """ctc_losses.py:521-560, the plugin driver of the dense route: ``get_normalized_acts``, then per
    utterance ``RawGenericCTC`` on the matrices of ``get_CTC_matrices_mono`` / ``_bicontext`` -- whose
    mean over frames of ``-logsum(lPB + alpha + beta)`` is the negative log-likelihood of that state
    machine.  Same signature and assertions; the state machines are compiled to arc graphs on the host
    (:func:`dense_ctc_graph`) and summed by the fused kernels instead of O(T N^2) numpy in fp64."""
    assert not other_data_in_batch
    S = acts.size(2)
    if context_order == 1:
        assert allow_nonblank_selfloops and not loop_using_symbol_repetitions and not eval_repeats_in_context
        mode, ns = (_lib.NORM_SOFTMAX, 0) if not normalize_by_dim else (_lib.NORM_CONTEXT, int(num_symbols or S))
    else:
        assert context_order == 2
        assert normalize_by_dim == 1
        assert num_symbols > 0
        S = int(round(np.sqrt(S)))
        mode, ns = _lib.NORM_CONTEXT, int(num_symbols)
    seqs = _labels_to_batch(labels, label_lens)
    graphs = [dense_ctc_graph(q, S, context_order, allow_nonblank_selfloops, eval_repeats_in_context,
                              loop_using_symbol_repetitions) for q in seqs]
    num = _lib.Graph.from_arcs(graphs)
    return _FstLoss.apply(acts, _check_lens(act_lens), num, None, mode, ns)


def ctc_raw_loss_batch(acts, labels, act_lens, label_lens, num_symbols=0, context_order=1,
                       normalize_by_dim=None, allow_nonblank_selfloops=True,
                       loop_using_symbol_repetitions=False, eval_repeats_in_context=False,
                       other_data_in_batch=None, **kwargs):
    """ctc_losses.py:563-609 (``RawGenericCTCBatch``): the batched form of :func:`ctc_raw_loss`."""
    assert not (eval_repeats_in_context and loop_using_symbol_repetitions)
    if context_order == 1:
        assert not loop_using_symbol_repetitions
    return ctc_raw_loss(acts, labels, act_lens, label_lens, num_symbols, context_order, normalize_by_dim,
                        allow_nonblank_selfloops, loop_using_symbol_repetitions, eval_repeats_in_context,
                        other_data_in_batch, **kwargs)


class FSTLossMixin(object):
    """Drop-in replacement of ``FSTDecoder.get_fst_loss`` / ``forward`` / ``logits``.

    Combine with the reference decoder, e.g.::

        class FSTDecoder(FSTLossMixin, att_speech...advanced_decoder.FSTDecoder): pass

    or use :class:`FSTLoss` stand-alone on the output of any ``fc`` layer.  Expects
    ``self.graph_generator`` (a :class:`CTCGraphGen` of this package),
    ``self.normalize_by_dim``, ``self.numerator_red``, ``self.denominator_red``,
    ``self.num_symbols`` as the reference's ``__init__`` sets them (:403-411).
    """

    def _raw_logits(self, encoded, extra_ret=None):
        logits = self.fc(encoded)
        if extra_ret is not None:
            extra_ret["unnormed_logits"] = logits
        return logits

    def logits(self, encoded, encoded_lens=None, extra_ret=None):
        """advanced_decoder.py:444-452: normalised scores whenever ``normalize_by_dim`` is set (what
        ctc_forward.py:93 dumps and ``decode`` returns).  The training path does not come through
        here: ``forward`` hands the raw ``fc`` output to the loss kernel, which fuses the
        normalisation."""
        logits = self._raw_logits(encoded, extra_ret)
        if self.normalize_by_dim is not None:
            logits = get_normalized_acts(logits, encoded_lens, self.num_symbols, self.context_order,
                                         self.normalize_by_dim, normalize_logits=True)
        return logits

    def get_fst_loss(self, logits, encoded_lens, texts, text_lens, other_data_in_batch):
        gg = self.graph_generator
        if other_data_in_batch and "graph_matrices" in other_data_in_batch:
            num = _as_graph(other_data_in_batch["graph_matrices"], gg.nc_weight)
        else:
            num = gg.get_training_matrices_batch(texts, text_lens, native=True).graph
        assert self.numerator_red in ("logsumexp", "logsumexp_fwb", "logsumexp_autodiff")
        den = None
        if self.denominator_red != "none":
            assert self.denominator_red in ("logsumexp", "logsumexp_fwb", "logsumexp_autodiff")
            den = gg.get_decoding_graph()
        # (the normalisations are idempotent: logits that went through self.logits() give the same loss)
        return _fst_loss(logits, _check_lens(encoded_lens), num, den,
                         _norm_mode(self.normalize_by_dim), int(self.num_symbols or 0))

    def forward(self, encoded, encoded_lens, texts, text_lens, spkids=None, **other_data_in_batch):
        logits = self._raw_logits(encoded)
        fst_loss_ = self.get_fst_loss(logits, encoded_lens, texts, text_lens, other_data_in_batch).sum()
        return {"fst_loss": fst_loss_, "loss": fst_loss_}


    def decode(self, encoded, encoded_lens, texts=None, text_lens=None,
               return_texts_and_generated_loss=False, return_logits_text_diff=False, spkids=None,
               **other_data_in_batch):
        """``FSTDecoder.decode`` (advanced_decoder.py:536-593): best path through the decoding graph
        (tropical reduction on the device), then the collapse of its frame labels to symbols by
        walking the topology rule (:556-571 does it with OpenFst compose + shortestpath)."""
        from . import _lib
        logits = self.logits(encoded, encoded_lens)      # normalised, as the reference decodes and returns them
        gg = self.graph_generator
        lens = _check_lens(encoded_lens)
        _, frame_labels = _lib.viterbi(logits.detach().contiguous(), lens, gg.get_decoding_graph())
        frame_labels = frame_labels.cpu()
        decoded_texts = [gg.collapse(frame_labels[:lens[i], i]) for i in range(logits.size(1))]
        ret = {"decoded": decoded_texts, "logits": logits}
        if texts is not None and text_lens is not None:
            fst_text_losses = self.get_fst_loss(logits, encoded_lens, texts, text_lens, other_data_in_batch)
            fst_text_loss = fst_text_losses.sum()
            ret["loss"] = dict(fst_loss=fst_text_loss, loss=fst_text_loss)
        if return_texts_and_generated_loss:
            decoded_lens = torch.IntTensor([len(x) for x in decoded_texts])
            width = max(1, int(decoded_lens.max())) if len(decoded_texts) else 1
            padded = torch.zeros((len(decoded_texts), width), dtype=torch.int64)
            for i, x in enumerate(decoded_texts):
                padded[i, :len(x)] = torch.as_tensor(x, dtype=torch.int64)
            fst_generated_losses = self.get_fst_loss(logits, encoded_lens, padded, decoded_lens, None)
            ret["text_loss"] = fst_text_losses.tolist()
            ret["generated_loss"] = fst_generated_losses.tolist()
        if return_logits_text_diff:
            ret["logits_text_diff"] = (torch.as_tensor(encoded_lens) - torch.as_tensor(text_lens)).tolist()
        return ret


class FSTLoss(FSTLossMixin, torch.nn.Module):
    """Stand-alone loss head with ``FSTDecoder``'s constructor arguments that
    matter for the loss (advanced_decoder.py:395-411); ``fc`` defaults to identity
    so it can be applied to logits directly."""

    def __init__(self, graph_generator, num_symbols=None, normalize_by_dim=None,
                 numerator_red="logsumexp", denominator_red="logsumexp", fc=None):
        super(FSTLoss, self).__init__()
        if isinstance(graph_generator, dict):
            kw = dict(graph_generator)
            kw.pop("class_name", None)
            kw.setdefault("num_symbols", num_symbols)
            graph_generator = CTCGraphGen(**kw)
        self.graph_generator = graph_generator
        self.num_symbols = graph_generator.num_symbols
        self.context_order = graph_generator.context_order
        self.normalize_by_dim = normalize_by_dim
        self.numerator_red = numerator_red
        self.denominator_red = denominator_red
        self.fc = fc if fc is not None else torch.nn.Identity()
        if self.normalize_by_dim not in [None, 0]:
            assert graph_generator.num_classes == graph_generator.num_symbols ** self.context_order
