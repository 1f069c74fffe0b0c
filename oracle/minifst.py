"""A tiny pure-Python stand-in for the slice of ``pywrapfst`` the hot path uses.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

The reference builds its graphs through OpenFst's Python wrapper
(``pywrapfst``, pinned ``openfst-python==1.7.2`` in ``environment.yml:150``),
which is not present in this image and not vendored under ``/root/reference``.
The call sites on the path (``att_speech/fst_utils.py:209-219, 246-279, 604,
694-712, 765-817``) use only: ``Fst``, ``Arc``, ``Weight.One/Zero``,
``add_state/add_arc/set_start/set_final``, ``arcsort``, ``states/arcs/final/
num_states/num_arcs/start``, ``properties(I_DETERMINISTIC)`` and ``compose``.
This module restates the published semantics of exactly those calls for the
log/tropical semirings (weights are costs, One = 0, Zero = +inf, Times = +) so
that the reference's *own* builders and ``fst_to_matrices`` can be executed
unmodified when generating golden vectors.

``compose`` follows OpenFst's definition for the case that occurs on the path
(left operand may emit epsilon on its output side, right operand has no
epsilon input labels): states are pairs, an arc of A with olabel 0 advances A
alone, an arc with olabel x pairs with every arc of B reading x; result is
trimmed to accessible and co-accessible states (``connect=True`` is the
pywrapfst default).  State numbers are assigned in breadth-first discovery
order; OpenFst's own numbering may differ, so results are only comparable up
to graph isomorphism (losses/gradients do not depend on state numbering).
"""
from __future__ import annotations

import math
from collections import deque

I_DETERMINISTIC = 0x0000000000010000  # kIDeterministic in OpenFst's properties.h
NO_STATE_ID = -1


class Weight(object):
    """A cost in the log or tropical semiring."""

    __slots__ = ("_type", "_value")

    def __init__(self, weight_type, value):
        self._type = weight_type
        self._value = float(value)

    @staticmethod
    def One(weight_type):
        return Weight(weight_type, 0.0)

    @staticmethod
    def Zero(weight_type):
        return Weight(weight_type, math.inf)

    def type(self):
        return self._type

    def __float__(self):
        return self._value

    def __neg__(self):
        return -self._value

    def __repr__(self):
        return "Weight(%s, %r)" % (self._type, self._value)


def _w(x):
    return float(x)


class Arc(object):
    __slots__ = ("ilabel", "olabel", "weight", "nextstate")

    def __init__(self, ilabel, olabel, weight, nextstate):
        self.ilabel = int(ilabel)
        self.olabel = int(olabel)
        self.weight = weight
        self.nextstate = int(nextstate)

    def __repr__(self):
        return "Arc(%d,%d,%r,%d)" % (self.ilabel, self.olabel,
                                     _w(self.weight), self.nextstate)


class SymbolTable(object):
    def __init__(self):
        self._syms = {}

    def add_symbol(self, sym, key=None):
        if key is None:
            key = len(self._syms)
        self._syms[key] = sym
        return key


class Fst(object):
    def __init__(self, arc_type="standard"):
        self._arc_type = arc_type
        self._arcs = []
        self._final = []
        self._start = NO_STATE_ID

    # -- types ---------------------------------------------------------
    def arc_type(self):
        return self._arc_type

    def weight_type(self):
        return "tropical" if self._arc_type == "standard" else self._arc_type

    # -- construction --------------------------------------------------
    def add_state(self):
        self._arcs.append([])
        self._final.append(math.inf)
        return len(self._arcs) - 1

    def set_start(self, s):
        self._start = int(s)
        return self

    def set_final(self, s, weight=None):
        self._final[s] = 0.0 if weight is None else _w(weight)
        return self

    def add_arc(self, s, arc):
        self._arcs[s].append(arc)
        return self

    def arcsort(self, sort_type="ilabel"):
        if sort_type == "ilabel":
            key = lambda a: a.ilabel  # noqa: E731
        else:
            key = lambda a: a.olabel  # noqa: E731
        for arcs in self._arcs:
            arcs.sort(key=key)  # stable, like OpenFst's ArcSort
        return self

    def set_input_symbols(self, syms):
        return self

    def set_output_symbols(self, syms):
        return self

    # -- inspection ----------------------------------------------------
    def start(self):
        return self._start

    def num_states(self):
        return len(self._arcs)

    def num_arcs(self, s):
        return len(self._arcs[s])

    def states(self):
        return iter(range(len(self._arcs)))

    def arcs(self, s):
        return iter(self._arcs[s])

    def final(self, s):
        return Weight(self.weight_type(), self._final[s])

    def rmepsilon(self, connect=True):
        """``Fst.rmepsilon()`` (in place, returns self) for the one case the reference reaches
        (fst_utils.py:1025, ``prevent_epsilons=True``): no arc with an epsilon input AND output, so
        nothing is removed and what remains of OpenFst's RmEpsilon is its default ``connect``
        (inaccessible / non-coaccessible states dropped, the others renumbered in order)."""
        for arcs in self._arcs:
            for a in arcs:
                if a.ilabel == 0 and a.olabel == 0:
                    raise NotImplementedError("rmepsilon with eps:eps arcs is not restated")
        if connect:
            t = _connect(self)
            self._arcs, self._final, self._start = t._arcs, t._final, t._start
        return self

    def properties(self, mask, test):
        props = 0
        if mask & I_DETERMINISTIC:
            det = True
            for arcs in self._arcs:
                seen = set()
                for a in arcs:
                    if a.ilabel in seen:
                        det = False
                        break
                    seen.add(a.ilabel)
                if not det:
                    break
            if det:
                props |= I_DETERMINISTIC
        return props & mask


def _connect(f):
    """Keep states that are both accessible and co-accessible (OpenFst Connect)."""
    n = f.num_states()
    if n == 0 or f.start() == NO_STATE_ID:
        return Fst(f.arc_type())
    acc = [False] * n
    acc[f.start()] = True
    dq = deque([f.start()])
    while dq:
        s = dq.popleft()
        for a in f._arcs[s]:
            if not acc[a.nextstate]:
                acc[a.nextstate] = True
                dq.append(a.nextstate)
    rev = [[] for _ in range(n)]
    for s in range(n):
        for a in f._arcs[s]:
            rev[a.nextstate].append(s)
    coacc = [math.isfinite(f._final[s]) for s in range(n)]
    dq = deque([s for s in range(n) if coacc[s]])
    while dq:
        s = dq.popleft()
        for p in rev[s]:
            if not coacc[p]:
                coacc[p] = True
                dq.append(p)
    keep = [acc[s] and coacc[s] for s in range(n)]
    remap = {}
    out = Fst(f.arc_type())
    for s in range(n):
        if keep[s]:
            remap[s] = out.add_state()
    if f.start() in remap:
        out.set_start(remap[f.start()])
    for s in range(n):
        if not keep[s]:
            continue
        if math.isfinite(f._final[s]):
            out.set_final(remap[s], f._final[s])
        for a in f._arcs[s]:
            if keep[a.nextstate]:
                out.add_arc(remap[s], Arc(a.ilabel, a.olabel, a.weight,
                                          remap[a.nextstate]))
    return out


def arcmap(ifst, delta=0.0009765625, map_type="identity", power=1.0, weight=None):
    """``fst.arcmap(f, 0, 'to_standard', 0)`` (advanced_decoder.py:422-424): the same states and arcs
    re-typed to the tropical semiring (the numeric weight values carry over).  Only the
    ``to_standard`` / ``to_log`` / ``identity`` maps the reference uses are restated."""
    arc_type = {"to_standard": "standard", "to_std": "standard", "to_log": "log",
                "identity": ifst.arc_type()}[map_type]
    out = Fst(arc_type=arc_type)
    for s in ifst.states():
        out.add_state()
    if ifst.start() != NO_STATE_ID:
        out.set_start(ifst.start())
    for s in ifst.states():
        out._final[s] = ifst._final[s]
        for a in ifst.arcs(s):
            out.add_arc(s, Arc(a.ilabel, a.olabel, Weight(out.weight_type(), float(a.weight)), a.nextstate))
    return out


def compose(ifst1, ifst2, connect=True):
    """``fst.compose`` for (A with output epsilons) o (B without input epsilons)."""
    for s in ifst2.states():
        for a in ifst2.arcs(s):
            if a.ilabel == 0:
                raise NotImplementedError(
                    "minifst.compose: epsilon input labels on the right operand")
    wtype = ifst1.weight_type()
    out = Fst(ifst1.arc_type())
    if ifst1.start() == NO_STATE_ID or ifst2.start() == NO_STATE_ID:
        return out
    # index B's arcs by input label
    b_by_label = []
    for s in ifst2.states():
        d = {}
        for a in ifst2.arcs(s):
            d.setdefault(a.ilabel, []).append(a)
        b_by_label.append(d)
    ids = {}
    dq = deque()

    def state_id(pair):
        sid = ids.get(pair)
        if sid is None:
            sid = out.add_state()
            ids[pair] = sid
            dq.append(pair)
        return sid

    out.set_start(state_id((ifst1.start(), ifst2.start())))
    while dq:
        pair = dq.popleft()
        s1, s2 = pair
        src = ids[pair]
        f1 = _w(ifst1.final(s1))
        f2 = _w(ifst2.final(s2))
        if math.isfinite(f1) and math.isfinite(f2):
            out.set_final(src, f1 + f2)
        for a in ifst1.arcs(s1):
            if a.olabel == 0:
                dst = state_id((a.nextstate, s2))
                out.add_arc(src, Arc(a.ilabel, 0, Weight(wtype, _w(a.weight)), dst))
            else:
                for b in b_by_label[s2].get(a.olabel, ()):
                    dst = state_id((a.nextstate, b.nextstate))
                    out.add_arc(src, Arc(a.ilabel, b.olabel,
                                         Weight(wtype, _w(a.weight) + _w(b.weight)),
                                         dst))
    return _connect(out) if connect else out
