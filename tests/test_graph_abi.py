"""CPU tests of the host side: the C-ABI library loads and exports every symbol of
``include/ctcg_b200.h``; the C++ graph compiler reproduces the reference's
matrices bit-exactly; error behaviour mirrors the reference's exceptions."""
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT, golden_names, load_golden
from oracle import graphs


def test_library_exports_every_declared_symbol(lib_built):
    P = lib_built
    header = open(os.path.join(ROOT, "include", "ctcg_b200.h")).read()
    declared = set(re.findall(r"^(?:int|int32_t|void|size_t|uint64_t|const char\*)\s+(ctcg_[a-z_0-9]+)\s*\(", header, re.M))
    assert declared == set(P._lib.EXPORTS)
    L = P._lib.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert P._lib.lib().ctcg_strerror(0) == b"ok"


@pytest.mark.parametrize("name", golden_names())
def test_graph_compiler_reproduces_reference_matrices(lib_built, name):
    P = lib_built
    d = load_golden(name)
    gen = P.CTCGraphGen(context_order=d["order"], graph_build_args=d["args"], num_symbols=d["S"])
    assert gen.num_classes == d["S"] ** d["order"]
    mats = gen.get_training_matrices_batch(torch.as_tensor(d["texts"]), torch.as_tensor(d["text_lens"]))
    assert len(mats) == 8
    for got, want in zip(mats, d["num_mats"]):
        assert got.numpy().dtype == want.dtype and np.array_equal(got.numpy(), want)
    for got, want in zip(gen.get_decoding_matrices("cpu"), d["den_mats"]):
        assert np.array_equal(got.numpy(), want)
    # single transcript, bi-char ids (prev*S+cur) are reduced modulo S like fst_utils.py:595-596
    one = gen.get_training_matrices(d["texts"][0][:d["text_lens"][0]])
    n0 = int((d["num_mats"][3][0, :, 0] > -1e19).nonzero()[0].max()) + 1
    assert one[0].shape[0] == n0


def test_padded_round_trip(lib_built):
    P = lib_built
    d = load_golden("bi_ctcgb")
    g = P.Graph.from_padded(d["num_mats"])
    assert g.num_graphs == len(d["lens"])
    for b in range(g.num_graphs):
        n_real = int((d["num_mats"][2][b] > -1e19).any(-1).nonzero()[0].max()) + 1
        for out_edges, off in ((False, 0), (True, 4)):
            st, lab, w, term = g.export_padded(b, out_edges)
            k = st.shape[1]
            assert np.array_equal(st[:n_real], d["num_mats"][off][b, :n_real, :k])
            assert np.array_equal(lab[:n_real], d["num_mats"][off + 1][b, :n_real, :k])
            assert np.array_equal(w[:n_real], d["num_mats"][off + 2][b, :n_real, :k])
            assert np.array_equal(term[:n_real, 0], d["num_mats"][3][b, :n_real, 0])


def test_generic_arc_import_and_errors(lib_built):
    P = lib_built
    g = graphs.GraphGen(2, 4, {}).training_graph([1, 2, 2])
    arcs = [(a[0], a[1], a[2], a[4]) for a in g.arcs]
    h = P.Graph.from_arcs([(g.n, arcs, g.final)])
    want = graphs.to_matrices(g, out_edges=False)
    got = h.export_padded(0, False)
    assert all(np.array_equal(a, b) for a, b in zip(got, want))
    with pytest.raises(ValueError, match="eps"):       # fst_utils.py:272-274
        P.Graph.from_arcs([(2, [(0, 1, -1, 0.0)], [-np.inf, 0.0])])
    with pytest.raises(ValueError, match="deterministic"):  # fst_utils.py:250-252
        P.Graph.from_arcs([(2, [(0, 1, 3, 0.0), (0, 0, 3, 0.0)], [-np.inf, 0.0])])


def test_bichar_topology_sizes(lib_built):
    P = lib_built
    info = P.Graph.decoding(2, 49).info()
    assert (info.total_states, info.total_arcs, info.max_in_degree, info.max_out_degree) == (2401, 119953, 51, 50)
    assert info.max_labels == 2353 and info.structured == 1
    info = P.Graph.decoding(2, 49, use_contextual_blanks=True).info()
    assert info.max_labels == 2401
    info = P.Graph.decoding(1, 49).info()
    assert (info.total_states, info.total_arcs) == (49, 2401)
    rng = np.random.default_rng(0)
    tr = P.Graph.training(2, 49, [rng.integers(1, 49, size=100) for _ in range(4)])
    ti = tr.info()
    assert ti.num_graphs == 4 and ti.max_states == 201 and ti.max_in_degree == 3


def test_compute_entry_points_fail_loudly_without_cuda(lib_built):
    P = lib_built
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    gen = P.CTCGraphGen(context_order=1, graph_build_args={}, num_symbols=5)
    num = gen.get_training_graphs([[1, 2]])
    x = torch.randn(6, 1, 5)
    with pytest.raises(RuntimeError, match="no CPU path"):
        P.fst_loss(x, [6], num, None, normalize_by_dim=0)
    with pytest.raises(RuntimeError):
        P._lib.loss_host(x, [6], num, None, P.NORM_SOFTMAX)


def test_unsorted_lengths_are_rejected_like_the_reference(lib_built):
    P = lib_built
    gen = P.CTCGraphGen(context_order=1, graph_build_args={}, num_symbols=5)
    num = gen.get_training_graphs([[1], [2]])
    with pytest.raises(AssertionError):                 # fst_utils.py:382,432
        P.path_reduction(torch.randn(6, 2, 5), torch.tensor([4, 6]), num)


def test_training_graph_cache_and_pickling_of_tagged_matrices(lib_built):
    """Rest of SURVEY f1: transcripts seen before are not composed again (the reference recomputes them every
    epoch, kaldi_dataset.py:63-66, 230-232), and matrices made in a DataLoader worker survive the trip."""
    import pickle
    P = lib_built
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=9)
    rng = np.random.default_rng(4)
    seqs = [rng.integers(1, 9, size=n).tolist() for n in (7, 5, 5, 3)]
    P._lib.graph_cache_clear()
    g1 = gen.get_training_graphs(seqs)
    h, m, n = P._lib.graph_cache_info()
    assert (h, m, n) == (0, 4, 4)
    g2 = gen.get_training_graphs(seqs[::-1])                       # another epoch, another order
    assert P._lib.graph_cache_info()[:2] == (4, 4)
    for i in range(4):                                             # the cached graphs are the same graphs
        for a, b in zip(g1.export_padded(i), g2.export_padded(3 - i)):
            assert np.array_equal(a, b)
    other = P.CTCGraphGen(context_order=2, graph_build_args={"use_contextual_blanks": True}, num_symbols=9)
    other.get_training_graphs(seqs[:1])                            # same transcript, other topology: a miss
    assert P._lib.graph_cache_info()[:2] == (4, 5)
    mats = gen.get_training_matrices(seqs[0])
    back = pickle.loads(pickle.dumps(mats))
    assert type(back) is list and all(torch.equal(a, b) for a, b in zip(mats, back))
