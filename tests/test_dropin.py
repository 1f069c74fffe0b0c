"""The drop-in routes of INTEGRATION.md against the reference's own classes (CPU; needs
``/root/reference``, skipped where it is absent -- the GPU box).

Route 1: ``class FSTDecoderB200(P.FSTLossMixin, advanced_decoder.FSTDecoder)`` must construct with the
reference's ``__init__`` (which arc-maps ``graph_generator.decoding_fst`` through pywrapfst), resolve
``forward / logits / get_fst_loss / decode`` to the mixin, return the reference's normalised
``logits()`` and fail loudly -- not fall back -- when there is no CUDA device.
Route 2: padded matrices handed back to ``path_reduction`` keep or regain their native graph.
"""
import importlib
import sys
import types

import numpy as np
import pytest
import torch

from oracle import graphs, ref_loader

needs_ref = pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")


def _reference_decoder_module():
    ref_loader.load()
    if "tensorboardX" not in sys.modules:      # logger dependency of the decoder module, off-path
        tb = types.ModuleType("tensorboardX")
        tb.SummaryWriter = type("SummaryWriter", (), {})
        sys.modules["tensorboardX"] = tb
    import builtins
    if not hasattr(builtins, "unicode"):       # py2-era builtin used by att_speech/utils.py:74
        builtins.unicode = str
    return importlib.import_module("att_speech.modules.decoders.advanced_decoder")


@needs_ref
@pytest.mark.parametrize("order,args", [(1, {}), (2, {}), (2, {"use_contextual_blanks": True}),
                                        (2, {"loop_using_symbol_repetitions": True})])
def test_decoding_fst_is_the_reference_fst(lib_built, order, args):
    """``CTCGraphGen.decoding_fst`` rebuilt through pywrapfst (here: the oracle's stand-in) state for
    state and arc for arc equal to what the reference's builder makes."""
    P = lib_built
    fu, _ = ref_loader.load()
    S = 5
    gen = P.CTCGraphGen(context_order=order, graph_build_args=args, num_symbols=S)
    mine = gen.decoding_fst
    want = (fu.build_ctc_mono_decoding_fst(S) if order == 1 else fu.build_ctc_bigram_decoding_fst(S, **args))
    assert mine.num_states() == want.num_states() and mine.start() == want.start() == 0

    def arcs(f):
        return sorted((s, a.ilabel, a.olabel, a.nextstate) for s in f.states() for a in f.arcs(s))
    assert arcs(mine) == arcs(want)


@needs_ref
def test_mixin_composes_with_the_reference_decoder(lib_built):
    P = lib_built
    ad = _reference_decoder_module()
    fu, _ = ref_loader.load()
    fu.CTCGraphGenB200 = P.CTCGraphGen          # INTEGRATION.md route 1: the one line added to att_speech/fst_utils.py

    class FSTDecoderB200(P.FSTLossMixin, ad.FSTDecoder):
        pass

    S, H = 6, 8
    vocab = [str(i) for i in range(S)]
    kw = dict(sample_batch={"features": torch.zeros(3, 2, H)}, num_classes=S * S,
              graph_generator=dict(class_name="CTCGraphGenB200", context_order=2, graph_build_args={}),
              normalize_by_dim=0, vocabulary=vocab)
    torch.manual_seed(3)
    ad.Globals.cuda = False                     # the reference's global device switch (no GPU in this container)
    dec = FSTDecoderB200(**kw).eval()           # (in training mode the reference's LutLinear writes to its global logger)
    for name in ("forward", "logits", "get_fst_loss", "decode"):
        assert getattr(FSTDecoderB200, name) is getattr(P.FSTLossMixin, name)
    assert isinstance(dec.graph_generator, P.CTCGraphGen) and dec.dec_fst is not None
    # logits(): the reference's semantics (normalised) with the reference's own fc
    enc = torch.randn(7, 2, H)
    extra = {}
    got = dec.logits(enc, torch.tensor([7, 5]), extra)
    want = ad.FSTDecoder.logits(dec, enc, torch.tensor([7, 5]))
    assert torch.allclose(got, want, atol=1e-6)
    assert torch.allclose(got.exp().sum(-1), torch.ones(7, 2), atol=1e-5)
    assert torch.equal(extra["unnormed_logits"], dec.fc(enc))
    # no CPU path: the loss refuses CPU tensors instead of computing something else
    texts = torch.tensor([[1, 2], [3, 0]])
    with pytest.raises(RuntimeError):
        dec.forward(enc, torch.tensor([7, 5]), texts, torch.tensor([2, 1]))


def test_padded_matrices_carry_their_graph_and_shared_graphs_are_imported_once(lib_built):
    P = lib_built
    fu = P.fst_utils if hasattr(P, "fst_utils") else importlib.import_module(P.CTCGraphGen.__module__)
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=7)
    mats = gen.get_decoding_matrices()
    assert fu._as_graph(mats) is gen.get_decoding_graph()            # tagged: no import at all
    before = fu.import_count
    plain = [m.to("cpu") for m in mats]                              # what FSTDecoder does (:458): the tag is gone
    g1 = fu._as_graph(plain)
    g2 = fu._as_graph([m.to("cpu") for m in mats])                   # next step: same tensors, new list
    assert g1 is g2 and fu.import_count == before + 1
    assert g1.info().structured == 1                                 # recognised: factorised kernels
    # a different graph in tensors of the same shape is not confused with it
    other = [m.clone() for m in P.CTCGraphGen(context_order=2, graph_build_args={"use_contextual_blanks": True},
                                              num_symbols=7).get_decoding_matrices()]
    g3 = fu._as_graph(other)
    assert g3 is not g1 and fu.import_count == before + 2
    # per-utterance numerator matrices (leading dimension B) are converted every time, never cached
    num = gen.get_training_matrices_batch(torch.tensor([[1, 2, 3], [4, 4, 0]]), torch.tensor([3, 2]))
    assert fu._as_graph(num) is not fu._as_graph(num)
    # a weighted graph is not mistaken for the bi-char topology
    ogen = graphs.GraphGen(2, 7, {})
    dm = [np.array(m, copy=True) for m in ogen.get_decoding_matrices()]
    dm[2][0, 3, 0] = -0.5
    assert P.Graph.from_padded([torch.as_tensor(m) for m in dm]).info().structured == 0


@needs_ref
@pytest.mark.parametrize("S,args", [(4, {}), (5, {"use_context_blanks": True}), (6, {})])
def test_trichar_graphs_equal_the_reference(lib_built, S, args):
    """``CTCGraphGen(context_order=3)`` (build_ctc_trigram_decoding_fst_v2, fst_utils.py:926-1050): decoding FST,
    training matrices and -- where the graph is input-deterministic -- decoding matrices, bit-exact against the
    reference's own generator run on the oracle's OpenFst stand-in."""
    import contextlib
    import io
    P = lib_built
    fu, _ = ref_loader.load()
    with contextlib.redirect_stdout(io.StringIO()):          # the reference prints its progress
        rgen = fu.CTCGraphGen(context_order=3, graph_build_args=dict(args), num_symbols=S)
    gen = P.CTCGraphGen(context_order=3, graph_build_args=dict(args), num_symbols=S)
    assert gen.num_classes == rgen.num_classes == S ** 3

    def arcs(f):
        return [(s, a.ilabel, a.olabel, a.nextstate) for s in f.states() for a in f.arcs(s)]
    mine, want = gen.decoding_fst, rgen.decoding_fst
    assert mine.num_states() == want.num_states()
    assert arcs(mine) == arcs(want)                          # same arcs in the same order
    assert [mine.final(s) == want.final(s) for s in mine.states()]
    rng = np.random.default_rng(S)
    for _ in range(4):
        labels = rng.integers(1, S, size=int(rng.integers(1, 7)))
        want_m = rgen.get_training_matrices(torch.as_tensor(labels))
        got_m = gen.get_training_matrices(torch.as_tensor(labels))
        assert len(got_m) == len(want_m) == 8
        for a, b in zip(got_m, want_m):
            assert a.dtype == b.dtype and torch.equal(a, b)
    if args.get("use_context_blanks"):
        for a, b in zip(gen.get_decoding_matrices(), rgen.get_decoding_matrices("cpu")):
            assert torch.equal(a, b)
    else:                                                    # shared blank class: numerator only, in both
        with pytest.raises(ValueError):
            rgen.get_decoding_matrices("cpu")
        with pytest.raises(ValueError):
            gen.get_decoding_matrices()


@needs_ref
@pytest.mark.parametrize("order,kw", [(1, {}), (2, {}), (2, {"allow_nonblank_selfloops": False}),
                                      (2, {"eval_repeats_in_context": True}),
                                      (2, {"loop_using_symbol_repetitions": True})])
def test_dense_state_machines_equal_the_reference_matrices(lib_built, order, kw):
    """``dense_ctc_graph`` against ``get_CTC_matrices_mono`` / ``_bicontext`` / ``_bicontext_looprepeats``
    (ctc_losses.py:67-94, :111-226) of the live reference: transition i -> j exactly where A[i, j] = 1, labelled
    with the one class B makes state j emit; entry and terminal states as ``mono_CTC_loss`` / ``bi_CTC_loss``
    pass them to ``RawGenericCTC`` (:486-502)."""
    P = lib_built
    _, cl = ref_loader.load()
    dense = importlib.import_module(P.FSTLoss.__module__).dense_ctc_graph
    S = 5
    rng = np.random.default_rng(order * 10 + len(kw))
    for U in (1, 2, 3, 7):
        for _ in range(6):
            lab = rng.integers(1, S, size=U)
            if U >= 3:
                lab[1] = lab[0]                                   # a repeated letter
            if order == 1:
                A, Bm = cl.get_CTC_matrices_mono(torch.as_tensor(lab), S)
            else:
                A, Bm = cl.get_CTC_matrices_bicontext(torch.as_tensor(lab), S, **kw)
            n_term = 3 if kw.get("loop_using_symbol_repetitions") else 2
            n, arcs, final = dense(lab.tolist(), S, order, **kw)
            N = A.shape[0]
            assert n == N + 1
            emit = Bm.argmax(0).tolist()
            assert (Bm.sum(0) == 1).all()
            want = {(0, j + 1, emit[j]) for j in range(min(2, N))}
            want |= {(i + 1, j + 1, emit[j]) for i in range(N) for j in range(N) if A[i, j] != 0}
            assert {(a, b, c) for a, b, c, _ in arcs} == want and len(arcs) == len(want)
            assert [i for i in range(n) if final[i] == 0] == list(range(N + 1 - n_term, N + 1))


@needs_ref
@pytest.mark.parametrize("args", [{}, {"use_context_blanks": True}])
def test_pruned_trigram_table_equals_the_reference(lib_built, tmp_path, args):
    """``ngram_to_class_file`` (fst_utils.py:567-577): a pruned tri-char table, class = line number.  Same decoding
    FST (arcs in the same order) and training matrices as the reference's generator reading the same file."""
    import contextlib
    import io
    import itertools
    P = lib_built
    fu, _ = ref_loader.load()
    S = 5
    keep = [tg for tg in itertools.product(range(S), repeat=3)
            if not (tg[0] == 3 and tg[1] == 4) and not (tg[1] == 2 and tg[2] == 2 and tg[0] != 0)]
    path = tmp_path / "trigrams.txt"
    path.write_text("".join("%d %d %d\n" % tg for tg in keep))
    with contextlib.redirect_stdout(io.StringIO()):
        rgen = fu.CTCGraphGen(context_order=3, graph_build_args=dict(args), num_symbols=S, ngram_to_class_file=str(path))
    gen = P.CTCGraphGen(context_order=3, graph_build_args=dict(args), num_symbols=S, ngram_to_class_file=str(path))
    assert gen.num_classes == rgen.num_classes == len(keep) < S ** 3
    assert torch.equal(gen.ngram_to_class, rgen.ngram_to_class)

    def arcs(f):
        return [(s, a.ilabel, a.olabel, a.nextstate) for s in f.states() for a in f.arcs(s)]
    assert gen.decoding_fst.num_states() == rgen.decoding_fst.num_states()
    assert arcs(gen.decoding_fst) == arcs(rgen.decoding_fst)
    for labels in ([1, 2, 3], [4, 1, 1, 3, 2], [2], [1, 4, 3, 1]):          # none needs a pruned trigram
        want_m = rgen.get_training_matrices(torch.as_tensor(labels))
        got_m = gen.get_training_matrices(torch.as_tensor(labels))
        for a, b in zip(got_m, want_m):
            assert a.dtype == b.dtype and torch.equal(a, b)
    # orders 1 and 2 index their classes by the full table whatever the file says: only the full table is accepted
    full = tmp_path / "bigrams.txt"
    full.write_text("".join("%d %d\n" % bg for bg in itertools.product(range(S), repeat=2)))
    assert P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S, ngram_to_class_file=str(full)).num_classes == S * S
    pruned = tmp_path / "bigrams_pruned.txt"
    pruned.write_text("".join("%d %d\n" % bg for bg in itertools.product(range(S), repeat=2) if bg != (3, 4)))
    with pytest.raises(NotImplementedError):
        P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S, ngram_to_class_file=str(pruned))


@needs_ref
def test_transcript_fst_is_the_reference_chain(lib_built):
    """``get_transcript_fst`` / ``build_chain_fst`` (fst_utils.py:195-219, :592-601) on the OpenFst stand-in."""
    P = lib_built
    fu, _ = ref_loader.load()
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=7)
    rgen = fu.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=7)
    for labels in ([3, 1, 1, 6], [8, 22, 45]):                # the second: bi-char dataset ids, taken modulo S
        mine = gen.get_transcript_fst(torch.as_tensor(labels))
        want = rgen.get_transcript_fst(torch.as_tensor(labels))
        arcs = lambda f: [(s, a.ilabel, a.olabel, a.nextstate) for s in f.states() for a in f.arcs(s)]  # noqa: E731
        assert mine.num_states() == want.num_states() and arcs(mine) == arcs(want)
