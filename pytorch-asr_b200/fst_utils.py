"""Host-side mirror of the hot-path half of ``att_speech/fst_utils.py``.

Same names, argument meaning and error behaviour as the reference, so that
``att_speech.fst_utils.path_reduction`` can be rebound to this module's and
``graph_generator.class_name: CTCGraphGen`` can point here:

* :func:`path_reduction`                 <- fst_utils.py:322-397
* :func:`path_logsumexp`                 <- fst_utils.py:400-488 (``PathLogSumExp``)
* :func:`batch_training_graph_matrices`  <- fst_utils.py:491-521
* :func:`make_full_ngram_table`          <- fst_utils.py:524-534
* :class:`CTCGraphGen`                   <- fst_utils.py:546-676, :1053-1068

Everything numeric runs in the CUDA library behind ``_lib``; graph construction
runs in its C++ graph compiler (no OpenFst).  No CPU fallback.
"""
from __future__ import annotations

import bisect
import collections
import itertools
import weakref

import numpy as np
import torch

from . import _lib
from . import grammar
from ._lib import Decoder, Graph


class PackedGraphs(list):
    """What :meth:`CTCGraphGen.get_training_matrices_batch` returns natively.

    Behaves as the reference's list of padded matrices only on demand (they are
    materialised lazily through ``ctcg_graph_export_padded``); ``.graph`` is the
    CSR handle the kernels consume, so no ``(B,N,K)`` int64 tensors are built or
    copied host-to-device on the training path.
    """

    def __init__(self, graph, nc_weight=-1e20):
        super(PackedGraphs, self).__init__()
        self.graph = graph
        self.nc_weight = nc_weight

    def to_padded(self, device="cpu"):
        mats = [padded_matrices(self.graph, i, self.nc_weight) for i in range(self.graph.num_graphs)]
        return batch_training_graph_matrices(mats, self.nc_weight, device=device)


class TaggedMatrices(list):
    """The reference's 4/8 padded matrices of a graph, carrying the CSR handle they were exported
    from: ``path_reduction`` / ``fst_loss`` on them use the handle (no re-import, no upload, and
    the bi-char decoding graph keeps its factorised kernels)."""

    def __init__(self, mats, graph=None):
        super(TaggedMatrices, self).__init__(mats)
        self._ctcg_graph = graph

    def __reduce__(self):
        # DataLoader workers are processes (num_workers=4, ctc.yaml:62): the matrices travel, the native
        # handle does not -- the receiving side converts them like any of the reference's own matrices
        return (list, (list(self),))


def padded_matrices(graph, index=0, nc_weight=-1e20, for_forward_only=False, device="cpu"):
    """The reference's 4/8 matrices for one graph (``fst_to_matrices`` encoding)."""
    mats = [torch.from_numpy(m) for m in graph.export_padded(index, False, nc_weight)]
    if not for_forward_only:
        mats += [torch.from_numpy(m) for m in graph.export_padded(index, True, nc_weight)]
    if device is not None and str(device) != "cpu":
        mats = [m.to(device) for m in mats]
    return TaggedMatrices(mats, graph if graph.num_graphs == 1 else None)


def batch_training_graph_matrices(matrices, nc_weight=-1e20, device="cpu"):
    """fst_utils.py:491-521: pad per-utterance tuples to ``(B, Nmax, Kmax_i)``."""
    bs = len(matrices)
    max_n = max(m[0].size(0) for m in matrices)
    max_ks = [max(m[i].size(1) for m in matrices) for i in range(len(matrices[0]))]
    out = []
    for i, m in enumerate(matrices[0]):
        out.append(torch.full((bs, max_n, max_ks[i]),
                              0 if m.dtype == torch.int64 else nc_weight,
                              dtype=m.dtype, device=device))
    for b, ms in enumerate(matrices):
        for i, m in enumerate(ms):
            out[i][b, :m.size(0), :m.size(1)] = m
    return out


def make_full_ngram_table(context_order, num_symbols, num_classes):
    """fst_utils.py:524-534."""
    if num_symbols is None:
        num_symbols = int(round(num_classes ** (1.0 / context_order)))
    num_classes = num_symbols ** context_order
    idx = np.arange(num_classes)[:, None]
    pw = num_symbols ** np.arange(context_order)[::-1][None, :]
    ngram_to_class = torch.from_numpy((idx // pw) % num_symbols).long()
    return num_symbols, num_classes, ngram_to_class


# Shared graphs in the reference's padded format (leading dimension 1: the decoding matrices a graph
# generator builds once and hands out every step, advanced_decoder.py:497-500) are converted once and
# found again by the identity of their tensors -- also after the caller's ``[m.to(device) for m in ...]``
# dropped a TaggedMatrices tag, because ``.to`` returns the same tensor when nothing changes.
_SHARED_CACHE = collections.OrderedDict()
_SHARED_CACHE_SIZE = 8
import_count = 0  # imports of padded matrices into CSR (tests assert that a shared graph is imported once)


def _as_graph(graph_matrices, nc_weight=-1e20):
    global import_count
    if isinstance(graph_matrices, Graph):
        return graph_matrices
    if isinstance(graph_matrices, PackedGraphs):
        return graph_matrices.graph
    g = getattr(graph_matrices, "_ctcg_graph", None)
    if g is not None:
        return g
    assert len(graph_matrices) in (4, 8)
    assert all(sm.size(0) == graph_matrices[0].size(0) for sm in graph_matrices)
    shared = torch.is_tensor(graph_matrices[0]) and graph_matrices[0].size(0) == 1
    if shared:
        key = tuple(id(m) for m in graph_matrices[:4])
        ent = _SHARED_CACHE.get(key)
        if ent is not None and all(r() is m for r, m in zip(ent[0], graph_matrices[:4])):
            _SHARED_CACHE.move_to_end(key)
            return ent[1]
    g = Graph.from_padded(graph_matrices, nc_weight)
    import_count += 1
    if shared:
        _SHARED_CACHE[key] = ([weakref.ref(m) for m in graph_matrices[:4]], g)
        while len(_SHARED_CACHE) > _SHARED_CACHE_SIZE:
            _SHARED_CACHE.popitem(last=False)
    return g


class _PathLogSumExp(torch.autograd.Function):
    """``PathLogSumExp`` (fst_utils.py:400-488): gradients are produced by the
    forward call and scaled by ``grad_output`` in backward, as in the reference."""

    @staticmethod
    def forward(ctx, log_probs, act_lens, graph):
        cost, grads = _lib.path_sum(log_probs, act_lens, graph, need_grad=log_probs.requires_grad)
        ctx.grads = grads
        return cost

    @staticmethod
    def backward(ctx, grad_output):
        g = ctx.grads
        ctx.grads = None  # single backward, like the reference's ctx.grads (fst_utils.py:473,484)
        return _lib.scale_grad(g, grad_output), None, None


class _PathViterbi(torch.autograd.Function):
    """Tropical reduction; its gradient is the indicator of the best path's
    labels, which is what ``FSTDecoder.decode`` reads (advanced_decoder.py:546-554)."""

    @staticmethod
    def forward(ctx, log_probs, act_lens, graph):
        score, labels = _lib.viterbi(log_probs, act_lens, graph)
        ctx.labels = labels
        ctx.lens = torch.as_tensor(np.asarray(act_lens, dtype=np.int64))
        ctx.shape = log_probs.shape
        return score

    @staticmethod
    def backward(ctx, grad_output):
        T, B, V = ctx.shape
        grad = torch.zeros(ctx.shape, dtype=grad_output.dtype, device=grad_output.device)
        mask = (torch.arange(T, device=grad.device)[:, None] < ctx.lens.to(grad.device)[None, :])
        src = (mask.to(grad.dtype) * grad_output[None, :]).unsqueeze(-1)
        grad.scatter_(2, ctx.labels.long().unsqueeze(-1), src)
        return grad, None, None


def _check_lens(act_lens):
    lens = act_lens.tolist() if torch.is_tensor(act_lens) else [int(x) for x in act_lens]
    # fst_utils.py:382,432 -- the reference requires lengths sorted descending
    assert lens == sorted(lens, reverse=True)
    return lens


def path_logsumexp(log_probs, act_lens, graph_matrices, neg_inf=-1e20):
    lens = _check_lens(act_lens)
    _, bs, _ = log_probs.size()
    graph = _as_graph(graph_matrices, neg_inf)
    assert graph.num_graphs in (1, bs)
    return _PathLogSumExp.apply(log_probs, lens, graph)


def path_reduction(log_probs, act_lens, graph_matrices, red_kind="logsumexp", neg_inf=-1e20):
    """Drop-in for ``att_speech.fst_utils.path_reduction`` (fst_utils.py:322-397).

    ``graph_matrices`` may be the reference's list of 4/8 padded tensors (leading
    dimension B or 1) or a native :class:`Graph` / :class:`PackedGraphs`.
    Returns a ``(B,)`` tensor differentiable in ``log_probs``.
    """
    lens = _check_lens(act_lens)
    _, bs, _ = log_probs.size()
    graph = _as_graph(graph_matrices, neg_inf)
    assert graph.num_graphs in (1, bs)
    if red_kind in ("logsumexp", "logsumexp_fwb", "logsumexp_autodiff"):
        return _PathLogSumExp.apply(log_probs, lens, graph)
    assert red_kind in ("viterbi", "viterbi_autodiff")
    return _PathViterbi.apply(log_probs, lens, graph)


def collapse_frame_labels(frame_labels, context_order, num_symbols, allow_nonblank_selfloops=True,
                          use_contextual_blanks=False, loop_using_symbol_repetitions=False):
    """Frame-level class ids of a best path -> emitted symbols.

    Replaces ``shortestpath(compose(chain(idx), dec_fst))`` followed by the read-out of the
    non-zero olabels in ``FSTDecoder.decode`` (advanced_decoder.py:556-571): the decoding graphs
    of fst_utils.py:694-712 / 765-817 are input-deterministic, so the path of a class sequence is
    unique and can be walked with the topology rule itself -- no OpenFst, no arc table.
    Raises ``ValueError`` when the sequence is not a path of the graph (the composition would be
    empty)."""
    S = int(num_symbols)
    labels = [int(v) for v in (frame_labels.tolist() if hasattr(frame_labels, "tolist") else frame_labels)]
    out = []
    if context_order == 1:
        state = 0
        for v in labels:
            if not 0 <= v < S:
                raise ValueError("class %d is not a label of the mono-char decoding graph" % v)
            if v != state:
                if v != 0:
                    out.append(v)
                state = v
        return out
    assert context_order == 2
    loops, ctx = bool(allow_nonblank_selfloops), bool(use_contextual_blanks)
    rep = bool(loop_using_symbol_repetitions)
    if rep:
        loops, ctx = False, True
    c1, l1 = 0, 0
    for v in labels:
        if not 0 <= v < S * S:
            raise ValueError("class %d is not a label of the bi-char decoding graph" % v)
        cls1 = c1 * S + l1 if (l1 != 0 or ctx) else 0
        if loops and l1 != 0 and c1 != l1 and v == cls1:
            continue                                  # explicit self-loop, emits nothing
        c2 = c1 if l1 == 0 else l1
        if v == 0 and not ctx:
            l2 = 0
        else:
            cv, l2 = divmod(v, S)
            if cv != c2 or (l2 == 0 and not ctx):
                raise ValueError("class %d cannot follow state (%d, %d)" % (v, c1, l1))
        if not (l2 == 0 or (c1, l1) == (c2, l2) or (rep and l1 == l2)):
            out.append(l2)
        c1, l1 = c2, l2
    return out


def collapse_by_arcs(frame_labels, arcs, final):
    """Frame-level class ids -> emitted symbols by walking an arc table (``arcs``: rows ``src, dst, class,
    emitted symbol, _``; ``final``: log-weights, -inf for non-final states; start state 0): the read-out of
    ``shortestpath(compose(chain(idx), dec_fst))`` in ``FSTDecoder.decode`` (advanced_decoder.py:556-571) for
    graphs without a closed-form rule (tri-char).  The graph need not be input-deterministic (the shared blank
    class of the tri-char topology is read by several arcs of a state): the walk keeps every state the prefix can
    be in, with one back-pointer each, and reads the symbols back from a final state -- all weights are One, so
    every accepting path is a shortest path; among several the one through the smallest state ids is taken.
    ``ValueError`` when no accepting path reads the sequence (the composition would be empty)."""
    labels = [int(v) for v in (frame_labels.tolist() if hasattr(frame_labels, "tolist") else frame_labels)]
    table = getattr(arcs, "_ctcg_table", None) if not isinstance(arcs, np.ndarray) else None
    if table is None:
        table = {}
        for src, dst, cls, out, _ in np.asarray(arcs).tolist():
            table.setdefault((src, cls), []).append((dst, out))
    frontier = {0: None}
    back = []
    for v in labels:
        nxt = {}
        for st in sorted(frontier):
            for dst, out in table.get((st, v), ()):
                if dst not in nxt:
                    nxt[dst] = (st, out)
        if not nxt:
            raise ValueError("class %d cannot follow the frames before it" % v)
        back.append(nxt)
        frontier = nxt
    ends = [st for st in sorted(frontier) if final[st] > -1e30]      # final log-weight: 0, or the grammar's
    if not ends:
        raise ValueError("the frame labels do not end in a final state")
    st, out = ends[0], []
    for nxt in reversed(back):
        st, o = nxt[st]
        if o > 0:
            out.append(o)
    return out[::-1]


def setattr_matched(obj, var, val):
    val_orig = getattr(obj, var)
    if val_orig is not None:
        assert val_orig == val, ("The value of %s is %s but should be %s" % (var, val_orig, val))
    else:
        setattr(obj, var, val)


def trigram_decoding_arcs(trigrams, use_context_blanks=False, prevent_epsilons=True):
    """The arcs of ``build_ctc_trigram_decoding_fst_v2`` (fst_utils.py:926-1050) before its ``rmepsilon``
    and ``arcsort``: states are the sorted trigrams (c0, c1, c2), final when c2 is the blank, class of a
    trigram = its state index (the shared blank class of (0,0,0) for every trigram with a blank in the
    middle unless ``use_context_blanks``), emitted symbol = the middle letter of the target.
    Returns ``(n_states, arcs (m, 5) int64 [src, dst, class, olabel, 0], final (n,) float32)``.

    The arc ORDER decides how ``compose`` numbers the states of a training graph; the reference adds the
    successors of a state by iterating a Python ``set`` of tuples (:994), so the same construct is used here
    -- in one interpreter the two orders are the same, which is what makes the exported matrices bit-exact."""
    trigrams = sorted(tuple(int(c) for c in tg) for tg in trigrams)
    assert trigrams[0] == (0, 0, 0)
    tg2state = {tg: i for i, tg in enumerate(trigrams)}

    def in_class(tg):
        if not use_context_blanks and tg[1] == 0:
            return tg2state[(0, 0, 0)]
        return tg2state[tg]

    inf = float("inf")
    arcs = []
    for i1, s1 in enumerate(trigrams):
        if s1 != (0, 0, 0):
            arcs.append((i1, i1, in_class(s1), 0, 0))          # self loop, emits nothing
        base = (s1[0], s1[2]) if s1[1] == 0 else (s1[1], s1[2])
        lo = bisect.bisect_left(trigrams, base + (0,))
        hi = bisect.bisect_left(trigrams, base + (inf,))
        for s2 in set(itertools.chain(trigrams[lo:hi], [(s1[1], 0, s1[2])])):
            if s1 == s2 or s2 == (0, 0, 0):
                continue
            if s1 != (0, 0, 0):
                if s1[-1] == 0 and s2[-1] != 0:                 # after the final blank: terminate
                    continue
                if s2[0] == 0 and s1[0] != 0:                   # no starting blank except at the start
                    continue
            if s2 not in tg2state:
                continue
            if s1 == (0, 0, 0) and not prevent_epsilons:
                raise NotImplementedError("prevent_epsilons=False needs OpenFst's rmepsilon")
            arcs.append((i1, tg2state[s2], in_class(s2), int(s2[1]), 0))
    final = np.array([0.0 if tg[-1] == 0 else -np.inf for tg in trigrams], dtype=np.float32)
    return len(trigrams), np.asarray(arcs, dtype=np.int64).reshape(-1, 5), final


def build_chain_fst(labels, arc_type="log", vocab=None):
    """fst_utils.py:195-219 for callers that want the OpenFst object (``FSTDecoder.decode`` of the reference builds
    one per utterance, advanced_decoder.py:559-563; this package's ``decode`` does not).  Needs pywrapfst."""
    try:
        import pywrapfst as fst
    except ImportError:
        raise ImportError("build_chain_fst returns an OpenFst object and needs pywrapfst; the loss and decode paths "
                          "of this package do not use it")
    f = fst.Fst(arc_type=arc_type)
    one = fst.Weight.One(f.weight_type())
    s = f.add_state()
    f.set_start(s)
    for l in labels:
        nxt = f.add_state()
        f.add_arc(s, fst.Arc(int(l), int(l), one, nxt))
        s = nxt
    f.set_final(s)
    f.arcsort("ilabel")
    return f


class BaseGraphGen(object):
    """fst_utils.py:546-676 without OpenFst; ``grammar_fst`` (an OpenFst binary file) is read and composed by
    :mod:`grammar`."""

    _decoder = None

    def __init__(self, num_symbols=None, num_classes=None, ngram_to_class_file=None,
                 context_order=None, nc_weight=-1e20, for_forward_only=False,
                 grammar_fst=None, vocabulary=None, **kwargs):
        super(BaseGraphGen, self).__init__(**kwargs)
        self.num_classes = num_classes
        self.num_symbols = num_symbols
        self.context_order = context_order
        if ngram_to_class_file is None:
            self.context_order = 1 if context_order is None else context_order
            num_symbols, num_classes, ngram_to_class = make_full_ngram_table(
                self.context_order, num_symbols, num_classes)
            setattr_matched(self, "num_symbols", num_symbols)
            setattr_matched(self, "num_classes", num_classes)
            self.ngram_to_class = ngram_to_class
            self.ngrams = ngram_to_class.tolist()
        else:
            # fst_utils.py:567-577: a pruned n-gram table, one n-gram of symbol ids per line, class = line number
            # (num_symbols is not checked against it: the file may hold fewer symbols)
            with open(ngram_to_class_file, "r") as f:
                self.ngrams = [[int(n) for n in line.split()] for line in f]
            if self.context_order is None:
                self.context_order = len(self.ngrams[0])
            assert all(len(ngr) == self.context_order for ngr in self.ngrams)
            setattr_matched(self, "num_classes", len(self.ngrams))
            self.ngram_to_class = torch.LongTensor(self.ngrams)
            if self.context_order < 3:
                # the mono- and bi-char builders number their classes by the full table whatever the file says
                # (fst_utils.py:694-712, 765-817): only a file that IS the full table is consistent with them
                full = make_full_ngram_table(self.context_order, self.num_symbols, None)[2].tolist()
                if self.ngrams != full:
                    raise NotImplementedError("a pruned n-gram table with context_order < 3 (the reference's "
                                              "builders ignore it and index classes by the full table)")
        self.grammar_fst_path = grammar_fst     # fst_utils.py:580-584
        self.vocabulary = None if grammar_fst is None else grammar.read_vocabulary(vocabulary)
        self.nc_weight = nc_weight
        self._grammar = None                    # (arcs, final) of HC o G when a grammar is given
        self._dec = self.get_decoding_fst()    # native CSR handle
        self._dec_fst = None                   # pywrapfst view of it, built on first use (decoding_fst)
        self.decoding_mats = {}
        self.for_forward_only = for_forward_only

    # -- graphs ----------------------------------------------------------
    @property
    def decoding_fst(self):
        """What the reference keeps here is an OpenFst object (``FSTDecoder.__init__`` arc-maps and
        arc-sorts it, advanced_decoder.py:420-424).  Where ``pywrapfst`` is importable -- any
        environment the reference itself runs in -- this is a real FST with the same states, arcs
        (ilabel = class + 1, olabel = emitted symbol, weight One) and arc order as
        ``build_ctc_*_decoding_fst`` (fst_utils.py:679-835), built from the native graph on first
        use; without OpenFst it is the native handle.  The loss never looks at it."""
        if self._dec_fst is None:
            self._dec_fst = self._to_openfst()
            if self._dec_fst is None:
                self._dec_fst = self.get_decoding_graph()
        return self._dec_fst

    def _to_openfst(self, arc_type="log"):
        try:
            import pywrapfst as fst
        except ImportError:
            return None
        if self._grammar is not None:
            return self._grammar_to_openfst(fst, arc_type)
        S, order = self.num_symbols, self.context_order
        args = self._build_args()
        rep = bool(args.get("loop_using_symbol_repetitions", False))
        succ, labels, weights, _ = self._dec.export_padded(0, True, self.nc_weight)
        f = fst.Fst(arc_type=arc_type)
        one = fst.Weight.One(f.weight_type())
        n = succ.shape[0]
        for st in range(n):
            assert f.add_state() == st
            f.set_final(st)
        f.set_start(0)
        for src in range(n):
            for k in range(succ.shape[1]):
                if not weights[src, k] > 0.5 * self.nc_weight:
                    continue
                dst, cls = int(succ[src, k]), int(labels[src, k])
                if order == 1:
                    out = 0 if dst == src else dst
                else:
                    l1, l2 = src % S, dst % S
                    out = 0 if (l2 == 0 or dst == src or (rep and l1 == l2)) else l2
                f.add_arc(src, fst.Arc(cls + 1, out, one, dst))
        f.arcsort("olabel")
        return f

    def _grammar_to_openfst(self, fst, arc_type):
        arcs, final, w = self._grammar
        f = fst.Fst(arc_type=arc_type)
        for st in range(final.shape[0]):
            assert f.add_state() == st
            if np.isfinite(final[st]):
                f.set_final(st, fst.Weight(f.weight_type(), -float(final[st])))
        f.set_start(0)
        for (src, dst, cls, out, _), wt in zip(arcs.tolist(), w.tolist()):
            f.add_arc(src, fst.Arc(cls + 1, out, fst.Weight(f.weight_type(), -wt), dst))
        if hasattr(f, "connect"):
            f.connect()
        return f

    def get_decoding_fst(self):
        """fst_utils.py:634-641: H o C, composed with the grammar when there is one.  The reference's
        ``rmepsilon(determinize(compose(hc, G)))`` is built as ``compose(hc, rmepsilon(G))`` (same weighted
        relation, see grammar.py); the result lives in a native decoder that serves the denominator graph and
        the transcripts' numerator graphs, LM weights included."""
        hc = self.get_hc_fst()
        if not self.grammar_fst_path:
            return hc
        n, arcs, final = self._hc_arcs(hc)
        g = grammar.remove_epsilons(*self.get_grammar_fst())
        m, (src, dst, cls, emit, w), fin = grammar.compose(n, arcs, final, *g)
        self._grammar = (np.stack([src, dst, cls, emit, np.zeros_like(src)], 1).astype(np.int64), fin, w)
        self._decoder = Decoder(m, (src, dst, cls, emit, w), fin, connect=True)
        return None                              # made on demand: ValueError if it is not input-deterministic

    def get_grammar_fst(self):
        """fst_utils.py:615-632 -> ``(n_states, start, arcs, final)`` in the vocabulary's symbol ids."""
        return grammar.grammar_acceptor(self.grammar_fst_path, self.vocabulary)

    def _hc_arcs(self, hc):
        """H o C as ``(n_states, rows (src, dst, class, emitted symbol), final)``: the built-in topologies emit
        the target's letter unless the arc is a loop or enters a blank (fst_utils.py:694-712, :765-817)."""
        S, order = self.num_symbols, self.context_order
        rep = bool(self._build_args().get("loop_using_symbol_repetitions", False))
        succ, labels, weights, _ = hc.export_padded(0, True, self.nc_weight)
        rows = []
        for src in range(succ.shape[0]):
            for k in range(succ.shape[1]):
                if not weights[src, k] > 0.5 * self.nc_weight:
                    continue
                dst, cls = int(succ[src, k]), int(labels[src, k])
                if order == 1:
                    out = 0 if dst == src else dst
                else:
                    l1, l2 = src % S, dst % S
                    out = 0 if (l2 == 0 or dst == src or (rep and l1 == l2)) else l2
                rows.append((src, dst, cls, out))
        return succ.shape[0], np.asarray(rows, dtype=np.int64), np.zeros(succ.shape[0], dtype=np.float32)

    def get_hc_fst(self):
        raise NotImplementedError()

    def _build_args(self):
        return {}

    def get_training_graphs(self, label_seqs):
        """Native batch of numerator graphs (CSR), one per transcript."""
        if self._grammar is not None:
            return self._decoder.training(label_seqs, modulo=self.num_symbols)
        return Graph.training(self.context_order, self.num_symbols, label_seqs, **self._build_args())

    def get_transcript_fst(self, labels):
        """fst_utils.py:592-601: the chain acceptor of a transcript (labels taken modulo ``num_symbols`` when the
        dataset hands over bi-char ids) as an OpenFst object -- only where pywrapfst is importable; nothing on the
        loss path uses it (the native compiler composes transcripts itself)."""
        if torch.is_tensor(labels):
            labels = labels.cpu().numpy()
        labels = np.asarray(labels)
        if np.any(labels > self.num_symbols):
            labels = labels % self.num_symbols
        return build_chain_fst(labels)

    def get_training_fst(self, labels):
        if torch.is_tensor(labels):
            labels = labels.cpu().numpy()
        return self.get_training_graphs([np.asarray(labels)])

    def get_training_matrices(self, labels, device="cpu"):
        """The reference's 4/8-tuple of padded matrices for one transcript
        (fst_utils.py:647-660); tagged with the CSR handle they came from."""
        g = self.get_training_fst(labels)
        return padded_matrices(g, 0, self.nc_weight, self.for_forward_only, device)

    def get_training_matrices_batch(self, labels, label_lens, device="cpu", native=False):
        seqs = []
        for i in range(len(labels)):
            li = labels[i]
            if torch.is_tensor(li):
                li = li.cpu().numpy()
            seqs.append(np.asarray(li)[:int(label_lens[i])])
        g = self.get_training_graphs(seqs)
        if native:
            return PackedGraphs(g, self.nc_weight)
        return PackedGraphs(g, self.nc_weight).to_padded(device)

    def get_decoding_matrices(self, device="cpu", out_edges=False):
        key = str(device)
        ret = self.decoding_mats.get(key)
        if ret is None:
            dec = self.get_decoding_graph()
            ret = TaggedMatrices([m.unsqueeze(0) for m in padded_matrices(
                dec, 0, self.nc_weight, self.for_forward_only, device)], dec)
            self.decoding_mats[key] = ret
        return ret

    def get_decoding_graph(self):
        if self._dec is None and self._decoder is not None:
            # ValueError("FST is not deterministic") where the reference's get_decoding_matrices raises it
            # (fst_utils.py:250-252): the shared-blank tri-char graph, a trigram grammar with double back-off
            self._dec = self._decoder.graph()
        return self._dec

    def collapse(self, frame_labels):
        """Emitted symbols of the path that reads ``frame_labels`` (class ids) in the decoding graph."""
        if self._grammar is not None:
            return collapse_by_arcs(frame_labels, self._grammar[0], self._grammar[1])
        return collapse_frame_labels(frame_labels, self.context_order, self.num_symbols, **self._build_args())


class CTCGraphGen(BaseGraphGen):
    """fst_utils.py:1053-1068: mono-char, bi-char and tri-char CTC topologies."""

    def __init__(self, context_order=None, graph_build_args=None, **kwargs):
        assert context_order in (1, 2, 3)
        self.graph_build_args = graph_build_args or {}
        self._decoder = None
        self._tri = None
        super(CTCGraphGen, self).__init__(context_order=context_order, **kwargs)

    def _build_args(self):
        return self.graph_build_args if self.context_order == 2 else {}

    def get_hc_fst(self):
        if self.context_order < 3:
            return Graph.decoding(self.context_order, self.num_symbols, **self._build_args())
        # build_ctc_trigram_decoding_fst_v2(S, self.ngrams, **graph_build_args), :926-1050
        args = dict(self.graph_build_args)
        if args.pop("determinize", False):
            raise NotImplementedError("determinize=True needs OpenFst's determinize")
        args.pop("add_syms", None)
        n, arcs, final = trigram_decoding_arcs(self.ngrams, **args)
        self._tri = (arcs, final)
        self._decoder = Decoder(n, arcs, final, connect=True)   # rmepsilon() on a graph without eps:eps arcs
        return None                                              # the decoding graph is made on demand (may not exist)

    def _hc_arcs(self, hc):
        if self.context_order < 3:
            return super(CTCGraphGen, self)._hc_arcs(hc)
        arcs, final = self._tri
        return final.shape[0], arcs[:, :4], final

    def get_training_graphs(self, label_seqs):
        if self.context_order < 3 or self._grammar is not None:
            return super(CTCGraphGen, self).get_training_graphs(label_seqs)
        return self._decoder.training(label_seqs, modulo=self.num_symbols)

    def collapse(self, frame_labels):
        if self.context_order < 3 or self._grammar is not None:
            return super(CTCGraphGen, self).collapse(frame_labels)
        arcs, final = self._tri                  # the arcs as built (dead ends never reach a final state)
        return collapse_by_arcs(frame_labels, arcs, final)

    def _to_openfst(self, arc_type="log"):
        if self.context_order < 3 or self._grammar is not None:
            return super(CTCGraphGen, self)._to_openfst(arc_type)
        try:
            import pywrapfst as fst
        except ImportError:
            return None
        arcs, final = self._tri
        f = fst.Fst(arc_type=arc_type)
        one = fst.Weight.One(f.weight_type())
        for st in range(final.shape[0]):
            assert f.add_state() == st
            if final[st] == 0:
                f.set_final(st)
        f.set_start(0)
        for src, dst, cls, out, _ in arcs.tolist():
            f.add_arc(src, fst.Arc(cls + 1, out, one, dst))
        f = f.rmepsilon()
        f.arcsort("olabel")
        return f
