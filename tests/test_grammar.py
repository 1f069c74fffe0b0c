"""Grammar-composed decoding graphs (row f4; fst_utils.py:615-640) without OpenFst.

pywrapfst is not available anywhere this runs, so the reference's ``rmepsilon(determinize(compose(HC, G)))``
cannot be executed: parity at that boundary is unpinned.  What is pinned here is the semantics, by brute force:
the OpenFst file reader against files written byte by byte in this test (and the reference's own grammar files
where the tree is present), epsilon removal against an enumeration of every epsilon path, and the composed
graph against a sum over every frame-label sequence."""
import gzip
import importlib
import itertools
import math
import os
import struct

import numpy as np
import pytest
import torch

from oracle import engine, graphs

REF_LM = "/root/reference/egs/wsj/train_lm"
REF_VOCAB = "/root/reference/egs/wsj/vocabulary.txt"
needs_ref = pytest.mark.skipif(not os.path.exists(REF_LM), reason="reference tree not present")


def _gr(P):
    return importlib.import_module(P.CTCGraphGen.__module__.rsplit(".", 1)[0] + ".grammar")


def _write_openfst(states, start, symbols, arc_type="standard"):
    """OpenFst binary ``vector`` file (version 2, both symbol tables): ``states`` = [(final cost, [(ilabel,
    olabel, cost, next)...])]."""
    def s(x):
        b = x.encode()
        return struct.pack("<i", len(b)) + b

    def table():
        out = struct.pack("<i", 2125658996) + s("syms") + struct.pack("<q", len(symbols)) + struct.pack("<q", len(symbols))
        for key, sym in symbols.items():
            out += s(sym) + struct.pack("<q", key)
        return out

    out = struct.pack("<i", 2125659606) + s("vector") + s(arc_type) + struct.pack("<iiQ", 2, 3, 0)
    out += struct.pack("<qqq", start, len(states), 0) + table() + table()
    for fin, arcs in states:
        out += struct.pack("<fq", fin, len(arcs))
        for a in arcs:
            out += struct.pack("<iifi", *a)
    return out


# a back-off bigram model over {a, b}: state 0 = start, 1 = unigram (back-off) state, 2 = after a, 3 = after b
SYMS = {0: "<eps>", 1: "<s>", 2: "</s>", 3: "b", 4: "a", 5: "<spc>"}
INF = float("inf")
TOY = [
    (INF, [(0, 0, 0.0, 1), (4, 4, 0.9, 2)]),                       # start: eps to unigram, a seen after <s>
    (1.2, [(4, 4, 1.1, 2), (3, 3, 0.7, 3), (5, 5, 2.0, 1)]),       # unigrams (final = </s>)
    (0.8, [(0, 0, 0.5, 1), (3, 3, 0.4, 3)]),                       # after a: back-off, explicit a b
    (INF, [(0, 0, 0.3, 1), (4, 4, 0.6, 2), (3, 3, 1.5, 3)]),       # after b
]
VOCAB = ["<pad>", " ", "a", "b"]                                    # network symbols: 0 blank, 1 space, 2 a, 3 b


def _toy(P, tmp_path=None):
    data = _write_openfst(TOY, 0, SYMS)
    return _gr(P).read_openfst(gzip.compress(data) if tmp_path is None else data)


def test_reader_reads_what_the_format_says(lib_built, tmp_path):
    P = lib_built
    gr = _gr(P)
    raw = _write_openfst(TOY, 0, SYMS)
    g = gr.read_openfst(raw)
    assert (g.fst_type, g.arc_type, g.start) == ("vector", "standard", 0)
    assert g.isymbols == SYMS and g.osymbols == SYMS
    assert g.src.tolist() == [0, 0, 1, 1, 1, 2, 2, 3, 3, 3]
    assert g.ilabel.tolist() == [0, 4, 4, 3, 5, 0, 3, 0, 4, 3] and np.array_equal(g.ilabel, g.olabel)
    assert np.allclose(g.weight, [0.0, 0.9, 1.1, 0.7, 2.0, 0.5, 0.4, 0.3, 0.6, 1.5])
    assert g.dst.tolist() == [1, 2, 2, 3, 1, 1, 3, 1, 2, 3]
    assert np.isinf(g.final[0]) and np.isclose(g.final[1], 1.2) and np.isclose(g.final[2], 0.8)
    path = tmp_path / "g.fst.gz"
    path.write_bytes(gzip.compress(raw))
    g2 = gr.read_openfst(str(path))                                 # gzip file, as the reference ships them
    assert g2.dst.tolist() == g.dst.tolist()
    with pytest.raises(ValueError):
        gr.read_openfst(raw[:-3])
    with pytest.raises(ValueError):
        gr.read_openfst(b"\x00" * 64)
    # relabelling by symbol NAME (get_grammar_fst, fst_utils.py:620-632): a -> 2, b -> 3, <spc> -> 1
    n, start, arcs, final = gr.grammar_acceptor(g, VOCAB)
    assert sorted({l for _, _, l, _ in arcs}) == [0, 1, 2, 3]
    assert (0, 2, 2, pytest.approx(-0.9)) in [(s, d, l, w) for s, d, l, w in arcs]
    with pytest.raises(ValueError):
        gr.grammar_acceptor(g, ["<pad>", "a"])                      # b is not in the vocabulary


def _all_paths_weight(n, start, arcs, final, string):
    """log of the total weight of every path (epsilon arcs included) that spells ``string`` and ends final."""
    out = [[] for _ in range(n)]
    for s, d, l, w in arcs:
        out[s].append((d, l, w))
    total = [-math.inf]

    def walk(q, i, acc):
        if i == len(string) and final[q] > -math.inf:
            total[0] = np.logaddexp(total[0], acc + final[q])
        for d, l, w in out[q]:
            if l == 0:
                walk(d, i, acc + w)
            elif i < len(string) and l == string[i]:
                walk(d, i + 1, acc + w)

    walk(start, 0, 0.0)
    return float(total[0])


def test_epsilon_removal_keeps_every_strings_weight(lib_built):
    P = lib_built
    gr = _gr(P)
    g = gr.grammar_acceptor(_toy(P), VOCAB)
    ge = gr.remove_epsilons(*g)
    assert all(l != 0 for _, _, l, _ in ge[2])
    keys = [(s, l, d) for s, d, l, _ in ge[2]]
    assert len(keys) == len(set(keys))                               # merged: one arc per (source, label, target)
    assert len({(s, l) for s, l, _ in keys}) == len(keys)            # and, for a bigram model, deterministic
    for length in range(0, 5):
        for string in itertools.product([1, 2, 3], repeat=length):
            a = _all_paths_weight(*g, string)
            b = _all_paths_weight(*ge, string)
            assert (a == b == -math.inf) or abs(a - b) < 1e-12, (string, a, b)
    cyc = (2, 0, [(0, 1, 0, -0.1), (1, 0, 0, -0.1), (0, 1, 2, -1.0)], np.array([-np.inf, 0.0]))
    with pytest.raises(ValueError):
        gr.remove_epsilons(*cyc)


@pytest.mark.parametrize("order", [1, 2])
def test_composed_graphs_sum_what_brute_force_sums(lib_built, tmp_path, order):
    """Denominator = log sum over EVERY frame-label sequence of exp(sum of its logits) * G(its collapse);
    numerator of a transcript = the same sum restricted to the sequences that collapse to it."""
    P = lib_built
    gr = _gr(P)
    path = tmp_path / "toy.fst.gz"
    path.write_bytes(gzip.compress(_write_openfst(TOY, 0, SYMS)))
    S = len(VOCAB)
    V = S ** order
    gen = P.CTCGraphGen(context_order=order, graph_build_args={}, num_symbols=S, grammar_fst=str(path), vocabulary=VOCAB)
    plain = graphs.mono_decoding_graph(S) if order == 1 else graphs.bigram_decoding_graph(S)
    gacc = gr.grammar_acceptor(str(path), VOCAB)
    T = 5 if order == 1 else 4
    rng = np.random.default_rng(order)
    x = rng.standard_normal((T, 1, V))
    by_text = {}
    for pi in itertools.product(range(V), repeat=T):
        text = graphs.collapse_by_arc_table(plain, pi)
        if text is None:
            continue
        score = sum(x[t, 0, c] for t, c in enumerate(pi))
        by_text[tuple(text)] = np.logaddexp(by_text.get(tuple(text), -math.inf), score)
    want_den = -math.inf
    lm = {}
    for text, acoustic in by_text.items():
        lm[text] = _all_paths_weight(*gacc, text)
        want_den = np.logaddexp(want_den, acoustic + lm[text])
    den = gen.get_decoding_graph()
    dm = [m.numpy() for m in gen.get_decoding_matrices()]
    got_den, _, _ = engine.path_logsumexp(torch.as_tensor(x), [T], dm, dtype=torch.float64)
    assert abs(float(got_den[0]) - want_den) < 1e-5                 # (the graph's weights are fp32)
    assert den.info().structured == 0                                # weighted: never the factorised kernels
    texts = [t for t in by_text if len(t) > 0 and lm[t] > -math.inf][:6]
    for text in texts:
        nm = gen.get_training_matrices(torch.as_tensor(list(text)))
        got, _, _ = engine.path_logsumexp(torch.as_tensor(x), [T], [m.unsqueeze(0).numpy() for m in nm], dtype=torch.float64)
        assert abs(float(got[0]) - (by_text[text] + lm[text])) < 1e-5, text
    # a transcript the grammar gives no weight to has no numerator path
    dead = [t for t in by_text if lm[t] == -math.inf]
    assert dead == [] or gen.get_training_graphs([list(dead[0])]) is not None
    # decode collapse walks the composed arcs
    some = next(pi for pi in itertools.product(range(V), repeat=T)
                if graphs.collapse_by_arc_table(plain, pi) == list(texts[0]))
    assert gen.collapse(list(some)) == list(texts[0])


@needs_ref
def test_reference_grammar_files(lib_built):
    """The reference's own character LMs (egs/wsj/train_lm, arpa2fst output): sizes, symbols, and what composing
    them gives -- the bigram model leaves the bi-char topology's 2401 states in place and only adds weights; the
    trigram model's double back-off makes the epsilon-free graph non-deterministic, which the reference's
    fst_to_matrices refuses too (fst_utils.py:250-252)."""
    P = lib_built
    gr = _gr(P)
    bg = gr.read_openfst(os.path.join(REF_LM, "G_char_bg_syms.fst.gz"))
    tg = gr.read_openfst(os.path.join(REF_LM, "G_char_tg_syms.fst.gz"))
    assert (len(bg.final), len(bg.src), int((bg.ilabel == 0).sum())) == (52, 1057, 80)
    assert (len(tg.final), len(tg.src), int((tg.ilabel == 0).sum())) == (981, 9743, 1253)
    assert bg.isymbols[0] == "<eps>" and bg.isymbols[8] == "<spc>" and len(bg.isymbols) == 51
    vocab = gr.read_vocabulary(REF_VOCAB)
    assert len(vocab) == 49 and vocab[0] == "<pad>" and vocab[2] == " "
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=49,
                        grammar_fst=os.path.join(REF_LM, "G_char_bg_syms.fst.gz"), vocabulary=REF_VOCAB)
    info = gen.get_decoding_graph().info()
    assert info.max_states == 2401 and info.structured == 0
    mono = P.CTCGraphGen(context_order=1, graph_build_args={}, num_symbols=49,
                         grammar_fst=os.path.join(REF_LM, "G_char_tg_syms.fst.gz"), vocabulary=REF_VOCAB)
    with pytest.raises(ValueError):
        mono.get_decoding_graph()
    with pytest.raises(ValueError):                                   # and so are the transcripts' graphs
        mono.get_training_graphs([[3, 4, 5, 3]])
    ok = P.CTCGraphGen(context_order=1, graph_build_args={}, num_symbols=49,
                       grammar_fst=os.path.join(REF_LM, "G_char_bg_syms.fst.gz"), vocabulary=REF_VOCAB)
    assert ok.get_decoding_graph().info().max_states == 97 and ok.get_training_graphs([[3, 4, 5, 3]]).num_graphs == 1


@pytest.mark.gpu
@pytest.mark.parametrize("order,S_extra", [(1, 0), (2, 0)])
def test_grammar_loss_on_the_gpu_matches_the_fp64_oracle(lib_built, tmp_path, order, S_extra):
    """CTC-G with a grammar in both graphs (weighted arcs, generic CSR kernels) against the fp64 oracle engine
    on the matrices exported from the same graphs."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    P = lib_built
    path = tmp_path / "toy.fst.gz"
    path.write_bytes(gzip.compress(_write_openfst(TOY, 0, SYMS)))
    S = len(VOCAB)
    V = S ** order
    gen = P.CTCGraphGen(context_order=order, graph_build_args={}, num_symbols=S, grammar_fst=str(path), vocabulary=VOCAB)
    rng = np.random.default_rng(9 + order)
    T, B = 33, 4
    lens = [33, 30, 24, 17]
    seqs = [[2, 3, 3, 2, 1, 2], [3, 2, 3], [2, 2, 3, 1, 3, 2, 3], [3]]
    x = (rng.standard_normal((T, B, V)) * 2).astype(np.float32)
    num, den = gen.get_training_graphs(seqs), gen.get_decoding_graph()
    loss, grad = P._lib.loss_fwd_bwd(torch.from_numpy(x).cuda(), lens, num, den, P.NORM_NONE, S)
    nm = P.batch_training_graph_matrices([P.padded_matrices(num, i) for i in range(B)])
    dm = gen.get_decoding_matrices()
    want_loss, want_grad = engine.fst_loss(x, lens, [m.numpy() for m in nm], [m.numpy() for m in dm], None, S, order,
                                           dtype=torch.float64)
    assert np.abs(loss.cpu().double().numpy() - want_loss.numpy()).max() / np.abs(want_loss.numpy()).max() < 1e-5
    assert np.abs(grad.cpu().double().numpy() - want_grad.numpy()).max() < 1e-5
    assert (want_loss.numpy() > 0).all()                             # numerator paths are a subset, weights included


def _random_backoff_bigram(S, rng, keep=0.6):
    """A back-off bigram model over the symbols 1..S-1 in arpa2fst's layout (start, unigram state, one state per
    history), with a fraction ``keep`` of the bigrams explicit: states, symbol table, vocabulary."""
    syms = {0: "<eps>", 1: "<s>", 2: "</s>"}
    vocab = ["<pad>"] + ["s%d" % i for i in range(1, S)]
    for i in range(1, S):
        syms[2 + i] = "s%d" % i
    hist = lambda i: 1 + i                                           # noqa: E731  state after symbol i
    states = [(INF, [(0, 0, 0.0, 1)])]
    states.append((float(rng.uniform(2, 4)), [(2 + i, 2 + i, float(rng.uniform(1, 5)), hist(i)) for i in range(1, S)]))
    for i in range(1, S):
        arcs = [(0, 0, float(rng.uniform(0.1, 1.5)), 1)]
        arcs += [(2 + j, 2 + j, float(rng.uniform(0.5, 4)), hist(j)) for j in range(1, S) if rng.random() < keep]
        states.append((float(rng.uniform(2, 5)) if rng.random() < 0.5 else INF, arcs))
    return states, syms, vocab


@pytest.mark.gpu
def test_bigram_grammar_at_the_full_vocabulary_matches_the_fp64_oracle(lib_built, tmp_path):
    """A weighted denominator at WSJ size: bi-char H o C (S=49, 2401 states, 119 953 arcs) composed with a random
    back-off bigram model over 48 symbols -- same states, every emitting arc weighted -- on the generic CSR
    kernels, numerator graphs carrying the grammar's weights too; loss and gradient against the fp64 oracle."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    P = lib_built
    S = 49
    rng = np.random.default_rng(49)
    states, syms, vocab = _random_backoff_bigram(S, rng)
    path = tmp_path / "big.fst.gz"
    path.write_bytes(gzip.compress(_write_openfst(states, 0, syms)))
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S, grammar_fst=str(path), vocabulary=vocab)
    den = gen.get_decoding_graph()
    info = den.info()
    assert info.max_states == S * S and info.structured == 0 and info.total_arcs > 100000
    T, B, V = 48, 2, S * S
    lens = [48, 37]
    seqs = [rng.integers(1, S, size=11).tolist(), rng.integers(1, S, size=8).tolist()]
    x = (rng.standard_normal((T, B, V)) * 3).astype(np.float32)
    num = gen.get_training_graphs(seqs)
    loss, grad = P._lib.loss_fwd_bwd(torch.from_numpy(x).cuda(), lens, num, den, P.NORM_NONE, S)
    nm = P.batch_training_graph_matrices([P.padded_matrices(num, i) for i in range(B)])
    dm = gen.get_decoding_matrices()
    want_loss, want_grad = engine.fst_loss(x, lens, [m.numpy() for m in nm], [m.numpy() for m in dm], None, S, 2,
                                           dtype=torch.float64)
    assert np.abs(loss.cpu().double().numpy() - want_loss.numpy()).max() / np.abs(want_loss.numpy()).max() < 1e-5
    assert np.abs(grad.cpu().double().numpy() - want_grad.numpy()).max() < 1e-5
    # the same topology without the grammar takes the factorised kernels and gives a different (unweighted) loss
    plain = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S)
    l0, _ = P._lib.loss_fwd_bwd(torch.from_numpy(x).cuda(), lens, plain.get_training_graphs(seqs), plain.get_decoding_graph(),
                                P.NORM_NONE, S)
    assert (l0 - loss).abs().min() > 1e-3


def test_grammar_composes_with_the_trichar_topology(lib_built, tmp_path):
    """Order 3 (fst_utils.py:926-1050) under a bigram grammar: the tri-char state already knows the last symbol, so
    the composed graph keeps the topology's states and arcs and only gains weights; every transcript's numerator
    is below the denominator (fp64 oracle engine on the exported matrices), and the decode collapse walks the
    composed arcs."""
    P = lib_built
    path = tmp_path / "toy.fst.gz"
    path.write_bytes(gzip.compress(_write_openfst(TOY, 0, SYMS)))
    S = len(VOCAB)
    args = {"use_context_blanks": True}
    gen = P.CTCGraphGen(context_order=3, graph_build_args=dict(args), num_symbols=S, grammar_fst=str(path), vocabulary=VOCAB)
    plain = P.CTCGraphGen(context_order=3, graph_build_args=dict(args), num_symbols=S)
    gi, pi = gen.get_decoding_graph().info(), plain.get_decoding_graph().info()
    assert (gi.max_states, gi.total_arcs) == (pi.max_states, pi.total_arcs) and gi.structured == 0
    rng = np.random.default_rng(0)
    T = 12
    x = torch.as_tensor(rng.standard_normal((T, 1, S ** 3)))
    den, _, _ = engine.path_logsumexp(x, [T], [m.numpy() for m in gen.get_decoding_matrices()], dtype=torch.float64)
    den0, _, _ = engine.path_logsumexp(x, [T], [m.numpy() for m in plain.get_decoding_matrices()], dtype=torch.float64)
    assert float(den[0]) < float(den0[0])                      # probabilities below one on every emitting arc
    for text in ([2, 3, 2], [3], [2, 2, 3, 1, 2]):
        nm = gen.get_training_matrices(torch.as_tensor(text))
        num, _, _ = engine.path_logsumexp(x, [T], [m.unsqueeze(0).numpy() for m in nm], dtype=torch.float64)
        assert float(num[0]) < float(den[0])
        nm0 = plain.get_training_matrices(torch.as_tensor(text))
        num0, _, _ = engine.path_logsumexp(x, [T], [m.unsqueeze(0).numpy() for m in nm0], dtype=torch.float64)
        gr = _gr(P)
        lm = _all_paths_weight(*gr.grammar_acceptor(str(path), VOCAB), tuple(text))
        assert abs((float(num[0]) - float(num0[0])) - lm) < 1e-5   # the numerator carries exactly the grammar's weight


def test_epsilon_removal_on_random_automata(lib_built):
    """Random weighted automata with acyclic epsilon arcs (chains of back-offs, epsilons into and out of final
    states, parallel arcs): the epsilon-free automaton gives every string up to length 4 the same total weight."""
    P = lib_built
    gr = _gr(P)
    rng = np.random.default_rng(77)
    for trial in range(25):
        n = int(rng.integers(2, 7))
        arcs = []
        for s in range(n):
            for _ in range(int(rng.integers(1, 5))):
                arcs.append((s, int(rng.integers(n)), int(rng.integers(1, 4)), float(-rng.uniform(0, 2))))
            for d in range(s + 1, n):                       # epsilons only forwards: acyclic
                if rng.random() < 0.4:
                    arcs.append((s, d, 0, float(-rng.uniform(0, 2))))
        final = np.where(rng.random(n) < 0.5, -rng.uniform(0, 2, size=n), -np.inf)
        final[n - 1] = 0.0
        g = (n, 0, arcs, final)
        ge = gr.remove_epsilons(*g)
        assert all(l != 0 for _, _, l, _ in ge[2])
        for length in range(0, 5):
            for string in itertools.product([1, 2, 3], repeat=length):
                a, b = _all_paths_weight(*g, string), _all_paths_weight(*ge, string)
                assert (a == b == -math.inf) or abs(a - b) < 1e-9, (trial, string, a, b)
