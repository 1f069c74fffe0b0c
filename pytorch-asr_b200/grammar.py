"""Grammar-composed decoding graphs without OpenFst (the rest of row f4).

The reference (att_speech/fst_utils.py:615-640) reads a character n-gram grammar G written by OpenFst
(``egs/wsj/train_lm/G_char_{bg,tg}_syms.fst.gz``: ``arpa2fst`` output, back-off arcs with an epsilon label),
maps it to the log semiring, relabels it to the network's vocabulary and builds

    decoding_fst = rmepsilon(determinize(compose(HC, G)))

with pywrapfst.  None of that library is on this path's machines, and only the *weighted relation* of the
result matters to the loss (state numbers do not), so the same relation is built directly:

1. :func:`read_openfst` -- reader of OpenFst's binary ``vector`` container (header, optional symbol tables,
   states with final weight and arc list), the published file format of OpenFst 1.6/1.7;
2. :func:`grammar_acceptor` -- ``get_grammar_fst`` (:615-632): symbols re-keyed by name to the vocabulary's
   indices (``' '`` is ``<spc>``, ``<pad>`` is ``<eps>``), costs kept as they are (``arcmap to_log``);
3. :func:`remove_epsilons` -- epsilon removal in the log semiring: every state gets the non-epsilon arcs and
   the final weight of its epsilon closure, arcs that agree in (source, label, target) are summed -- what
   OpenFst's RmEpsilon does; on a back-off n-gram model direct and backed-off arcs for a symbol end in the
   same state, so the result is deterministic;
4. :func:`compose` -- product of HC (arcs that read a class and may emit a symbol) with the epsilon-free
   grammar: an arc that emits nothing moves HC alone, an arc that emits c moves the grammar over c and adds
   its weight.  Removing the grammar's epsilons first counts every (HC path, grammar path) pair once, which
   is what OpenFst's epsilon-sequencing compose filter guarantees on the other order of operations.

Weights here are log-probabilities (minus the FST cost), as everywhere in this package.  The result goes to
``ctcg_decoder_from_arcs`` (Connect + arc sort), which serves both the denominator graph and the numerator
graphs of transcripts.  Parity with pywrapfst itself is unpinned (no OpenFst here); the semantics are pinned
by brute-force enumeration in ``tests/test_grammar.py``.
"""
from __future__ import annotations

import gzip
import struct
from collections import deque, namedtuple

import numpy as np

OpenFst = namedtuple("OpenFst", "fst_type arc_type start final src ilabel olabel weight dst isymbols osymbols")

_FST_MAGIC = 2125659606       # 0x7eb2fdd6
_SYM_MAGIC = 2125658996       # 0x7eb2fb74


def read_openfst(source):
    """Parse an OpenFst binary ``vector`` FST with ``standard`` or ``log`` arcs (float weights) from a path
    (``.gz`` or plain) or a bytes object.  Weights are costs as stored; a non-final state has final cost +inf.
    ``ValueError`` for anything that is not such a file (wrong magic, other container, truncated)."""
    try:
        return _read_openfst(source)
    except (struct.error, UnicodeDecodeError, MemoryError, OverflowError) as e:
        raise ValueError("truncated or corrupt OpenFst file: %s" % e)


def _read_openfst(source):
    if isinstance(source, (bytes, bytearray)):
        data = bytes(source)
    else:
        with open(source, "rb") as f:
            data = f.read()
    if data[:2] == b"\x1f\x8b":
        data = gzip.decompress(data)
    pos = [0]

    def take(fmt):
        v = struct.unpack_from("<" + fmt, data, pos[0])
        pos[0] += struct.calcsize("<" + fmt)
        return v[0]

    def string():
        n = take("i")
        s = data[pos[0]:pos[0] + n].decode("utf-8")
        pos[0] += n
        return s

    if take("i") != _FST_MAGIC:
        raise ValueError("not an OpenFst binary file")
    fst_type, arc_type = string(), string()
    version, flags = take("i"), take("i")
    take("Q")                                   # properties
    start, n_states, n_arcs = take("q"), take("q"), take("q")
    if fst_type != "vector" or arc_type not in ("standard", "log"):
        raise ValueError("unsupported FST container %s/%s (vector with standard or log arcs expected)" % (fst_type, arc_type))
    if version < 2:
        raise ValueError("unsupported vector FST version %d" % version)

    def symbols():
        if take("i") != _SYM_MAGIC:
            raise ValueError("bad symbol table")
        string()                                # name
        take("q")                               # next available key
        table = {}
        for _ in range(take("q")):
            sym = string()
            table[take("q")] = sym
        return table

    isyms = symbols() if flags & 1 else None
    osyms = symbols() if flags & 2 else None
    final = np.empty(n_states, dtype=np.float64)
    src, il, ol, w, dst = [], [], [], [], []
    arc = struct.Struct("<iifi")
    for s in range(n_states):
        final[s] = take("f")
        k = take("q")
        for _ in range(k):
            a, b, c, d = arc.unpack_from(data, pos[0])
            pos[0] += arc.size
            src.append(s), il.append(a), ol.append(b), w.append(c), dst.append(d)
    if pos[0] != len(data):
        raise ValueError("trailing bytes after the last state (%d of %d read)" % (pos[0], len(data)))
    if n_arcs not in (0, len(src)):
        raise ValueError("arc count in the header does not match the states")
    i32 = lambda x: np.asarray(x, dtype=np.int32)  # noqa: E731
    return OpenFst(fst_type, arc_type, int(start), final, i32(src), i32(il), i32(ol), np.asarray(w, dtype=np.float64),
                   i32(dst), isyms, osyms)


def read_vocabulary(vocabulary):
    """``egs.vocab.Vocab.from_file`` (egs/vocab.py:14-21): one symbol per line, index = line number."""
    if isinstance(vocabulary, (list, tuple)):
        return list(vocabulary)
    with open(vocabulary, "r") as f:
        return [line[:-1] if line.endswith("\n") else line for line in f]


def grammar_acceptor(source, vocabulary):
    """``BaseGraphGen.get_grammar_fst`` (fst_utils.py:615-632): the grammar with its symbols re-keyed by name to
    the network's vocabulary.  Returns ``(n_states, start, arcs, final)`` with ``arcs`` a list of
    ``(src, dst, label, logweight)`` (label 0 = epsilon) and ``final`` log-weights (-inf = not final)."""
    g = source if isinstance(source, OpenFst) else read_openfst(source)
    if g.isymbols is None:
        raise ValueError("the grammar carries no symbol table to relabel by")
    itos = read_vocabulary(vocabulary)
    new = {}
    for i, s in enumerate(itos):
        new[{" ": "<spc>", "<pad>": "<eps>"}.get(s, s)] = i
    new.setdefault("<s>", len(itos))
    new.setdefault("</s>", len(itos) + 1)
    if new.get("<eps>", 0) != 0:
        raise ValueError("the vocabulary's index 0 must be the blank / epsilon (<pad>)")
    if not np.array_equal(g.ilabel, g.olabel):
        raise ValueError("the grammar is not an acceptor")
    remap = {}
    for key in np.unique(g.ilabel).tolist():
        name = g.isymbols.get(key)
        if name is None or name not in new:
            raise ValueError("grammar symbol %r (key %d) is not in the vocabulary" % (name, key))
        remap[key] = new[name]
    arcs = [(int(s), int(d), remap[int(l)], -float(w)) for s, d, l, w in zip(g.src, g.dst, g.ilabel, g.weight)]
    final = np.where(np.isfinite(g.final), -g.final, -np.inf)
    return len(final), g.start, arcs, final


def _logadd(a, b):
    if a == -np.inf:
        return b
    if b == -np.inf:
        return a
    m = max(a, b)
    return m + np.log1p(np.exp(-abs(a - b)))


def remove_epsilons(n, start, arcs, final):
    """Epsilon removal in the log semiring (OpenFst RmEpsilon without the final Connect): state q gets, for every
    state r of its epsilon closure reached with total weight e(q, r), the non-epsilon arcs of r with e(q, r) added
    and final(q) = logsum_r e(q, r) + final(r); arcs with the same (source, label, target) are summed.  The
    epsilon arcs must be acyclic (back-off arcs are).  Same return convention as :func:`grammar_acceptor`."""
    eps = [[] for _ in range(n)]
    real = [[] for _ in range(n)]
    for s, d, l, w in arcs:
        (eps if l == 0 else real)[s].append((d, l, w))
    closure = [None] * n
    state = [0] * n                                   # 0 new, 1 on the stack, 2 done

    def close(q):
        if state[q] == 2:
            return closure[q]
        if state[q] == 1:
            raise ValueError("the grammar has an epsilon cycle")
        state[q] = 1
        c = {q: 0.0}
        for d, _, w in eps[q]:
            for r, e in close(d).items():
                c[r] = _logadd(c.get(r, -np.inf), w + e)
        state[q] = 2
        closure[q] = c
        return c

    import sys
    limit = sys.getrecursionlimit()
    sys.setrecursionlimit(max(limit, n + 1000))
    try:
        for q in range(n):
            close(q)
    finally:
        sys.setrecursionlimit(limit)
    out, new_final = [], np.full(n, -np.inf)
    for q in range(n):
        merged = {}
        for r, e in closure[q].items():
            new_final[q] = _logadd(new_final[q], e + final[r])
            for d, l, w in real[r]:
                merged[(l, d)] = _logadd(merged.get((l, d), -np.inf), e + w)
        out += [(q, d, l, w) for (l, d), w in sorted(merged.items())]
    return n, start, out, new_final


def compose(hc_n, hc_arcs, hc_final, g_n, g_start, g_arcs, g_final):
    """HC o G for an epsilon-free grammar: ``hc_arcs`` rows ``(src, dst, class, emitted symbol)`` (HC start is
    state 0, weights One), grammar as returned by :func:`remove_epsilons`.  States are pairs numbered in
    breadth-first order from (0, g_start).  Returns ``(n, (src, dst, class, emitted, logweight) arrays, final)``."""
    hc_out = [[] for _ in range(hc_n)]
    for s, d, c, o in np.asarray(hc_arcs).reshape(-1, 4).tolist():
        hc_out[s].append((d, c, o))
    by_label = [{} for _ in range(g_n)]
    for s, d, l, w in g_arcs:
        if l == 0:
            raise ValueError("compose needs an epsilon-free grammar (remove_epsilons first)")
        by_label[s].setdefault(l, []).append((d, w))
    ids = {(0, g_start): 0}
    queue = deque([(0, g_start)])
    src, dst, cls, emit, wt, final = [], [], [], [], [], []
    while queue:
        h, q = queue.popleft()
        me = ids[(h, q)]
        final.append(hc_final[h] + g_final[q] if (np.isfinite(hc_final[h]) and np.isfinite(g_final[q])) else -np.inf)
        for d, c, o in hc_out[h]:
            for q2, w in (((q, 0.0),) if o == 0 else by_label[q].get(o, ())):
                key = (d, q2)
                j = ids.get(key)
                if j is None:
                    j = ids[key] = len(ids)
                    queue.append(key)
                src.append(me), dst.append(j), cls.append(c), emit.append(o), wt.append(w)
    i32 = lambda x: np.asarray(x, dtype=np.int32)  # noqa: E731
    return len(ids), (i32(src), i32(dst), i32(cls), i32(emit), np.asarray(wt, dtype=np.float32)), \
        np.asarray(final, dtype=np.float32)
