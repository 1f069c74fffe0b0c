"""Numpy restatement of the reference's graph construction (no OpenFst).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Follows, in ``/root/reference/att_speech/fst_utils.py``:

* ``build_ctc_mono_decoding_fst``      :679-726  -> :func:`mono_decoding_graph`
* ``build_ctc_bigram_decoding_fst``    :729-835  -> :func:`bigram_decoding_graph`
* ``build_chain_fst`` + ``fst.compose``:195-219, :592-605 -> :func:`compose_with_chain`
* ``fst_to_matrices``                  :222-294  -> :func:`to_matrices`
* ``batch_training_graph_matrices``    :491-521  -> :func:`batch_matrices`
* ``BaseGraphGen`` / ``CTCGraphGen``   :546-676, :1053-1068 -> :class:`GraphGen`

A graph is a :class:`Graph`: ``n`` states (start is state 0), a list of arcs
``(src, dst, label, olabel, weight)`` where ``label`` is the 0-based class id
(the FST ilabel minus one), ``olabel`` the emitted symbol (0 = nothing) and
``weight`` a log-probability (minus the FST cost), and ``final`` log-weights
(``-inf`` for non-final states).  Per-state arc order is the order OpenFst
would hold after ``arcsort('olabel')`` (stable sort of insertion order).
"""
from __future__ import annotations

import itertools
from collections import deque

import numpy as np

NC_WEIGHT = -1e20


class Graph(object):
    def __init__(self, n, arcs, final):
        self.n = int(n)
        self.arcs = list(arcs)          # (src, dst, label, olabel, weight)
        self.final = np.asarray(final, dtype=np.float64)

    def out_lists(self):
        out = [[] for _ in range(self.n)]
        for a in self.arcs:
            out[a[0]].append(a)
        return out

    def num_arcs(self):
        return len(self.arcs)


def _sorted_by_olabel(n, arcs):
    out = [[] for _ in range(n)]
    for a in arcs:
        out[a[0]].append(a)
    res = []
    for lst in out:
        lst.sort(key=lambda a: a[3])  # stable, like ArcSort(olabel)
        res.extend(lst)
    return res


def mono_decoding_graph(S):
    """fst_utils.py:694-712: S states, all final; self-loop consumes the state's
    own class and emits nothing, s->s' consumes class s' and emits s'."""
    arcs = []
    for s in range(S):
        arcs.append((s, s, s, 0, 0.0))
        for s_next in range(S):
            if s_next == s:
                continue
            arcs.append((s, s_next, s_next, s_next, 0.0))
    return Graph(S, _sorted_by_olabel(S, arcs), np.zeros(S))


def bigram_decoding_graph(S, allow_nonblank_selfloops=True,
                          use_contextual_blanks=False,
                          loop_using_symbol_repetitions=False):
    """fst_utils.py:759-817: states (context, last letter), all final."""
    if loop_using_symbol_repetitions:
        allow_nonblank_selfloops = False
        use_contextual_blanks = True

    def in_class(c, let):
        if let != 0 or use_contextual_blanks:
            return c * S + let
        return 0

    arcs = []
    for s1 in range(S * S):
        c1, l1 = divmod(s1, S)
        if allow_nonblank_selfloops and l1 != 0 and c1 != l1:
            arcs.append((s1, s1, in_class(c1, l1), 0, 0.0))
        c2 = c1 if l1 == 0 else l1
        for l2 in range(S):
            s2 = c2 * S + l2
            if l2 == 0 or s1 == s2 or (loop_using_symbol_repetitions and l1 == l2):
                out_s = 0
            else:
                out_s = l2
            arcs.append((s1, s2, in_class(c2, l2), out_s, 0.0))
    return Graph(S * S, _sorted_by_olabel(S * S, arcs), np.zeros(S * S))


def _connected(n, arcs, final):
    """OpenFst Connect: states accessible from 0 and co-accessible, renumbered in order (what
    ``Fst.rmepsilon()`` leaves of itself on a graph without eps:eps arcs, fst_utils.py:1025)."""
    out = [[] for _ in range(n)]
    rev = [[] for _ in range(n)]
    for a in arcs:
        out[a[0]].append(a[1])
        rev[a[1]].append(a[0])
    acc = [False] * n
    acc[0] = True
    dq = deque([0])
    while dq:
        s = dq.popleft()
        for d in out[s]:
            if not acc[d]:
                acc[d] = True
                dq.append(d)
    co = [bool(np.isfinite(final[s])) for s in range(n)]
    dq = deque([s for s in range(n) if co[s]])
    while dq:
        s = dq.popleft()
        for p in rev[s]:
            if not co[p]:
                co[p] = True
                dq.append(p)
    keep = [s for s in range(n) if acc[s] and co[s]]
    remap = {s: i for i, s in enumerate(keep)}
    arcs = [(remap[a[0]], remap[a[1]], a[2], a[3], a[4]) for a in arcs if a[0] in remap and a[1] in remap]
    return len(keep), arcs, np.asarray([final[s] for s in keep])


def trigram_decoding_graph(S, use_context_blanks=False):
    """build_ctc_trigram_decoding_fst_v2 on the full table of S^3 trigrams (fst_utils.py:926-1050,
    ``prevent_epsilons=True``, ``determinize=False``): state = sorted trigram (c0, c1, c2), final when
    c2 is the blank; a self loop on every state but (0,0,0); from s1 to every s2 that continues it --
    (b0, b1, x) with (b0, b1) = (s1[0], s1[2]) if the middle of s1 is a blank else (s1[1], s1[2]), plus
    (s1[1], 0, s1[2]) -- unless s2 is s1 or (0,0,0), s1 ends in a blank and s2 does not, or s2 starts
    with a blank and s1 does not; class = index of the target trigram (of (0,0,0) when its middle is a
    blank and blanks are not contextual), emitted symbol = the target's middle letter."""
    tgs = [(a, b, c) for a in range(S) for b in range(S) for c in range(S)]
    idx = {tg: i for i, tg in enumerate(tgs)}

    def cls(tg):
        return idx[(0, 0, 0)] if (tg[1] == 0 and not use_context_blanks) else idx[tg]

    arcs = []
    for i1, s1 in enumerate(tgs):
        if s1 != (0, 0, 0):
            arcs.append((i1, i1, cls(s1), 0, 0.0))
        base = (s1[0], s1[2]) if s1[1] == 0 else (s1[1], s1[2])
        # the reference's own iteration (:994): a set of tuples, whose order decides how compose numbers the
        # states of a training graph (same interpreter => same order => bit-equal matrices)
        for s2 in set(itertools.chain([base + (x,) for x in range(S)], [(s1[1], 0, s1[2])])):
            if s2 == s1 or s2 == (0, 0, 0):
                continue
            if s1 != (0, 0, 0):
                if s1[2] == 0 and s2[2] != 0:
                    continue
                if s2[0] == 0 and s1[0] != 0:
                    continue
            arcs.append((i1, idx[s2], cls(s2), s2[1], 0.0))
    final = np.array([0.0 if tg[2] == 0 else -np.inf for tg in tgs])
    n, arcs, final = _connected(len(tgs), arcs, final)
    return Graph(n, _sorted_by_olabel(n, arcs), final)


def compose_with_chain(dec, labels):
    """compose(decoding graph, chain acceptor of ``labels``) then connect.

    Product construction over (decoding state, transcript position): an arc
    emitting nothing keeps the position, an arc emitting ``labels[pos]``
    advances it, any other arc is dropped (fst_utils.py:195-219, :603-605).
    Breadth-first numbering from (0, 0); finals are the pairs at position
    ``len(labels)`` (every decoding state is final); states that cannot reach
    a final state are removed and the rest renumbered in order.
    """
    labels = [int(x) for x in labels]
    U = len(labels)
    out = dec.out_lists()
    ids = {(0, 0): 0}
    order = [(0, 0)]
    arcs = []
    dq = deque([(0, 0)])
    while dq:
        pair = dq.popleft()
        s, pos = pair
        src = ids[pair]
        for (_, dst, lab, ol, w) in out[s]:
            if ol == 0:
                npair = (dst, pos)
            elif pos < U and ol == labels[pos]:
                npair = (dst, pos + 1)
            else:
                continue
            nid = ids.get(npair)
            if nid is None:
                nid = len(order)
                ids[npair] = nid
                order.append(npair)
                dq.append(npair)
            arcs.append((src, nid, lab, ol, w))
    n = len(order)
    final = np.full(n, -np.inf)
    for (s, pos), i in ids.items():
        if pos == U and np.isfinite(dec.final[s]):
            final[i] = dec.final[s]
    # connect: co-accessible states only (all are accessible by construction)
    rev = [[] for _ in range(n)]
    for a in arcs:
        rev[a[1]].append(a[0])
    co = np.isfinite(final)
    dq = deque(np.nonzero(co)[0].tolist())
    while dq:
        s = dq.popleft()
        for p in rev[s]:
            if not co[p]:
                co[p] = True
                dq.append(p)
    remap = -np.ones(n, dtype=np.int64)
    remap[co] = np.arange(int(co.sum()))
    arcs = [(int(remap[a[0]]), int(remap[a[1]]), a[2], a[3], a[4])
            for a in arcs if co[a[0]] and co[a[1]]]
    return Graph(int(co.sum()), arcs, final[co])


def to_matrices(g, out_edges=True, nc_weight=NC_WEIGHT):
    """fst_utils.py:222-294 on a :class:`Graph` (labels already 0-based).

    Returns ``states (N,K) int64, labels (N,K) int64, weights (N,K) float32,
    terminal (N,1) float32``; per-state arc lists sorted as tuples
    ``(other_state, label, weight)``; padding ``(0, 0, nc_weight)``.
    """
    seen = [set() for _ in range(g.n)]
    edges = [[] for _ in range(g.n)]
    for (src, dst, lab, _ol, w) in g.arcs:
        if lab < 0:
            raise ValueError("FST has eps-transitions (state=%d)" % (src,))
        if lab in seen[src]:
            raise ValueError("FST is not deterministic")
        seen[src].add(lab)
        if out_edges:
            edges[src].append((dst, lab, float(w)))
        else:
            edges[dst].append((src, lab, float(w)))
    k = max(len(e) for e in edges)
    states = np.zeros((g.n, k), dtype=np.int64)
    labels = np.zeros((g.n, k), dtype=np.int64)
    weights = np.full((g.n, k), nc_weight, dtype=np.float32)
    terminal = np.full((g.n, 1), nc_weight, dtype=np.float32)
    fin = np.isfinite(g.final)
    terminal[fin, 0] = g.final[fin]
    for s1, lst in enumerate(edges):
        for i, (s2, lab, w) in enumerate(sorted(lst)):
            states[s1, i] = s2
            labels[s1, i] = lab
            weights[s1, i] = w
    return states, labels, weights, terminal


def batch_matrices(matrices, nc_weight=NC_WEIGHT):
    """fst_utils.py:491-521: pad per-utterance tuples to (B, Nmax, Kmax_i)."""
    bs = len(matrices)
    max_n = max(m[0].shape[0] for m in matrices)
    out = []
    for i, m0 in enumerate(matrices[0]):
        max_k = max(m[i].shape[1] for m in matrices)
        fill = 0 if m0.dtype == np.int64 else nc_weight
        out.append(np.full((bs, max_n, max_k), fill, dtype=m0.dtype))
    for b, ms in enumerate(matrices):
        for i, m in enumerate(ms):
            out[i][b, :m.shape[0], :m.shape[1]] = m
    return out


class GraphGen(object):
    """Restated ``CTCGraphGen`` (fst_utils.py:546-676, :1053-1068), orders 1-3."""

    def __init__(self, context_order, num_symbols, graph_build_args=None,
                 nc_weight=NC_WEIGHT, for_forward_only=False):
        assert context_order in (1, 2, 3)
        self.context_order = context_order
        self.num_symbols = num_symbols
        self.num_classes = num_symbols ** context_order
        self.nc_weight = nc_weight
        self.for_forward_only = for_forward_only
        self.graph_build_args = dict(graph_build_args or {})
        if context_order == 1:
            self.decoding_graph = mono_decoding_graph(num_symbols)
        elif context_order == 2:
            self.decoding_graph = bigram_decoding_graph(
                num_symbols, **self.graph_build_args)
        else:
            self.decoding_graph = trigram_decoding_graph(
                num_symbols, bool(self.graph_build_args.get("use_context_blanks", False)))
        self._dec = None

    def transcript(self, labels):
        labels = np.asarray(labels)
        if np.any(labels > self.num_symbols):  # fst_utils.py:595-596
            labels = labels % self.num_symbols
        return labels

    def training_graph(self, labels):
        return compose_with_chain(self.decoding_graph, self.transcript(labels))

    def get_training_matrices(self, labels):
        g = self.training_graph(labels)
        mats = to_matrices(g, out_edges=False, nc_weight=self.nc_weight)
        if not self.for_forward_only:
            mats += to_matrices(g, out_edges=True, nc_weight=self.nc_weight)
        return mats

    def get_training_matrices_batch(self, labels, label_lens):
        mats = [self.get_training_matrices(labels[i][:int(label_lens[i])])
                for i in range(len(labels))]
        return batch_matrices(mats, self.nc_weight)

    def get_decoding_matrices(self):
        if self._dec is None:
            mats = to_matrices(self.decoding_graph, out_edges=False,
                               nc_weight=self.nc_weight)
            if not self.for_forward_only:
                mats += to_matrices(self.decoding_graph, out_edges=True,
                                    nc_weight=self.nc_weight)
            self._dec = [m[None] for m in mats]
        return self._dec


def collapse_by_arc_table(g, frame_labels):
    """Emitted symbols of the (unique) path that reads ``frame_labels`` in the input-deterministic
    graph ``g``: what ``shortestpath(compose(chain, dec_fst))`` + the olabel read-out of
    advanced_decoder.py:556-571 returns.  Walks the arc table; ``None`` if the sequence is no path."""
    out_arcs = g.out_lists()
    state, out = 0, []
    for v in frame_labels:
        nxt = [a for a in out_arcs[state] if a[2] == int(v)]
        if len(nxt) != 1:
            return None
        a = nxt[0]
        if a[3] > 0:
            out.append(a[3])
        state = a[1]
    return out



def collapse_all_paths(g, frame_labels, limit=10000):
    """Every symbol sequence an accepting path of ``g`` emits while reading ``frame_labels`` -- the candidates
    ``shortestpath(compose(chain, dec_fst))`` (advanced_decoder.py:556-571) chooses from when all weights are
    One and the graph is not input-deterministic (tri-char with a shared blank class).  Exhaustive walk;
    a set of tuples (empty when the composition is empty)."""
    out_arcs = g.out_lists()
    final = g.final
    layer = {(0, ()): True}
    for v in frame_labels:
        nxt = {}
        for (st, emitted) in layer:
            for a in out_arcs[st]:
                if a[2] == int(v):
                    nxt[(a[1], emitted + ((a[3],) if a[3] > 0 else ()))] = True
        if len(nxt) > limit:
            raise RuntimeError("too many partial paths")
        layer = nxt
    return {emitted for (st, emitted) in layer if final[st] == 0}
