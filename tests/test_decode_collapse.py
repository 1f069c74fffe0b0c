"""Collapse of a best path's frame labels to symbols (FSTDecoder.decode, advanced_decoder.py:556-571).

The reference composes the chain of frame classes with the decoding FST and reads the non-zero
olabels of the shortest path (OpenFst).  The product walks the topology rule directly
(pytorch-asr_b200/fst_utils.py:collapse_frame_labels); the checker walks the arc table of the
oracle's restatement of build_ctc_mono_decoding_fst / build_ctc_bigram_decoding_fst."""
import numpy as np
import pytest
import torch

from oracle import graphs

CASES = [
    (1, 7, {}),
    (2, 6, {}),
    (2, 6, {"allow_nonblank_selfloops": False}),
    (2, 5, {"use_contextual_blanks": True}),
    (2, 5, {"loop_using_symbol_repetitions": True}),
]


def _oracle_graph(order, S, args):
    return graphs.mono_decoding_graph(S) if order == 1 else graphs.bigram_decoding_graph(S, **args)


def _random_path_labels(g, rng, n):
    """Class ids along a random path of the graph, biased towards staying (repeats)."""
    out_arcs = g.out_lists()
    state, labels = 0, []
    for _ in range(n):
        arcs = out_arcs[state]
        stay = [a for a in arcs if a[1] == state]
        a = stay[0] if (stay and rng.random() < 0.4) else arcs[int(rng.integers(len(arcs)))]
        labels.append(a[2])
        state = a[1]
    return labels


@pytest.mark.parametrize("order,S,args", CASES)
def test_collapse_equals_arc_table_walk_on_random_paths(lib_built, order, S, args):
    P = lib_built
    g = _oracle_graph(order, S, args)
    gen = P.CTCGraphGen(context_order=order, graph_build_args=args, num_symbols=S)
    rng = np.random.default_rng(17 * S + order)
    for n in (0, 1, 2, 9, 60):
        for _ in range(20):
            labels = _random_path_labels(g, rng, n)
            want = graphs.collapse_by_arc_table(g, labels)
            assert want is not None
            assert gen.collapse(labels) == want
            assert gen.collapse(torch.tensor(labels, dtype=torch.int32)) == want


@pytest.mark.parametrize("order,S,args", CASES)
def test_collapse_rejects_exactly_what_the_graph_rejects(lib_built, order, S, args):
    P = lib_built
    g = _oracle_graph(order, S, args)
    gen = P.CTCGraphGen(context_order=order, graph_build_args=args, num_symbols=S)
    rng = np.random.default_rng(5)
    V = S ** order
    seen_reject = False
    for _ in range(300):
        labels = rng.integers(0, V, size=int(rng.integers(1, 5))).tolist()
        want = graphs.collapse_by_arc_table(g, labels)
        if want is None:
            seen_reject = True
            with pytest.raises(ValueError):
                gen.collapse(labels)
        else:
            assert gen.collapse(labels) == want
    assert seen_reject or order == 1      # every label sequence is a path of the mono-char graph


def test_plain_ctc_collapse_is_the_textbook_rule(lib_built):
    gen = lib_built.CTCGraphGen(context_order=1, graph_build_args={}, num_symbols=5)
    assert gen.collapse([0, 1, 1, 0, 1, 2, 2, 2, 0, 0, 3]) == [1, 1, 2, 3]


@pytest.mark.parametrize("S,args", [(4, {}), (4, {"use_context_blanks": True}), (5, {})])
def test_trichar_collapse_walks_the_arc_table(lib_built, S, args):
    """Tri-char graphs (fst_utils.py:926-1050) have no closed-form rule and, with the shared blank class, are
    not input-deterministic: the product keeps every state a prefix can be in.  Checked against the oracle's
    exhaustive walk of its own restatement of the graph: the transducer is functional on accepted sequences (one
    emitted string whatever path is taken), and that string is what the product returns; sequences that end in a
    non-final state or leave the graph are rejected by both."""
    P = lib_built
    g = graphs.trigram_decoding_graph(S, **args)
    gen = P.CTCGraphGen(context_order=3, graph_build_args=dict(args), num_symbols=S)
    out_arcs = g.out_lists()
    rng = np.random.default_rng(S + len(args))
    accepted = rejected = 0
    for n in (1, 2, 3, 8, 25, 60):
        for _ in range(25):
            state, labels, emitted = 0, [], []
            for _ in range(n):
                arcs = out_arcs[state]
                stay = [a for a in arcs if a[1] == state]
                a = stay[0] if (stay and rng.random() < 0.4) else arcs[int(rng.integers(len(arcs)))]
                labels.append(a[2])
                if a[3] > 0:
                    emitted.append(a[3])
                state = a[1]
            want = graphs.collapse_all_paths(g, labels)
            assert len(want) <= 1                                  # functional
            if g.final[state] == 0:
                assert want == {tuple(emitted)}
            if want:
                accepted += 1
                assert gen.collapse(labels) == list(next(iter(want)))
                assert gen.collapse(torch.tensor(labels, dtype=torch.int32)) == list(next(iter(want)))
            else:
                rejected += 1
                with pytest.raises(ValueError):
                    gen.collapse(labels)
    assert accepted > 20 and rejected > 20
    with pytest.raises(ValueError):
        gen.collapse([S ** 3 - 1, 1])                              # not a path at all
