"""CPU tests: the oracle against the reference's golden vectors, brute force and
(when the reference tree is present) the reference itself."""
import numpy as np
import pytest
import torch

from conftest import golden_names, load_golden
from oracle import engine, graphs, ref_loader


@pytest.mark.parametrize("name", golden_names())
def test_oracle_graphs_match_reference_matrices(name):
    d = load_golden(name)
    gen = graphs.GraphGen(d["order"], d["S"], d["args"])
    mats = gen.get_training_matrices_batch(d["texts"], d["text_lens"])
    for got, want in zip(mats, d["num_mats"]):
        assert got.dtype == want.dtype and np.array_equal(got, want)
    for got, want in zip(gen.get_decoding_matrices(), d["den_mats"]):
        assert np.array_equal(got, want)


@pytest.mark.parametrize("name", golden_names())
def test_oracle_engine_matches_reference_values(name):
    d = load_golden(name)
    den = d["den_mats"] if d["use_den"] else None
    for tag, dt, tol in (("f32", torch.float32, 0.0), ("f64", torch.float64, 1e-12)):
        loss, grad = engine.fst_loss(d["logits"], d["lens"], d["num_mats"], den,
                                     d["normalize_by_dim"], d["S"], d["order"], dtype=dt)
        # fp32: same op order as the reference -> bit-exact; fp64: to rounding
        assert np.abs(loss.double().numpy() - d["loss_" + tag]).max() <= tol * 1e3 + (0 if tol else 1e-4)
        assert np.abs(grad.double().numpy() - d["grad_" + tag]).max() <= (tol if tol else 2e-6)


@pytest.mark.parametrize("name", golden_names())
def test_oracle_viterbi_and_counts(name):
    d = load_golden(name)
    score, labels = engine.path_viterbi(d["logits"], d["lens"], d["den_mats"], dtype=torch.float64)
    assert np.allclose(score.numpy(), d["vit_score"], atol=1e-9)
    assert np.array_equal(labels.numpy(), d["vit_labels"])
    for key, mats in (("log_count_num", d["num_mats"]), ("log_count_den", d["den_mats"])):
        for b, ln in enumerate(d["lens"]):
            gi = b if mats[0].shape[0] > 1 else 0
            n, arcs, fin = engine.padded_to_arcs(mats, gi)
            # counts here are far below 2**53, so the log comparison is exact enough
            cnt = engine.path_count(n, arcs, fin, int(ln))
            assert cnt > 0 and abs(np.log(float(cnt)) - d[key][b]) < 1e-6


def test_mono_fst_route_equals_ctc_loss_and_dense_definition():
    d = load_golden("mono_ctc")
    assert np.allclose(d["loss_f64"], d["F_ctc_loss"], atol=1e-9)
    assert np.allclose(d["loss_f64"], d["raw_generic_ctc"], atol=1e-5)
    c = load_golden("bi_ctc_context_norm")
    # dense bi-context definition agrees where the transcript has no repeated letter
    assert abs(c["loss_f64"][0] - c["raw_generic_ctc"][0]) < 1e-5


def test_brute_force_agrees_with_engine():
    rng = np.random.default_rng(3)
    for order, S, args in [(1, 3, {}), (2, 3, {}), (2, 3, {"use_contextual_blanks": True})]:
        gen = graphs.GraphGen(order, S, args)
        V = S ** order
        T = 5
        lp = rng.normal(size=(T, 1, V)) * 2
        for g, name in ((gen.training_graph([1, 2]), "num"), (gen.decoding_graph, "den")):
            mats = [m[None] for m in graphs.to_matrices(g, False)] + [m[None] for m in graphs.to_matrices(g, True)]
            cost, grad, cost0 = engine.path_logsumexp(lp, [T], mats, dtype=torch.float64)
            arcs = [(a[0], a[1], a[2], a[4]) for a in g.arcs]
            bf = engine.brute_force(lp[:, 0], g.n, arcs, g.final)
            assert abs(float(cost[0]) - bf["cost"]) < 1e-10, name
            assert abs(float(cost0[0]) - bf["cost"]) < 1e-10
            assert np.abs(grad[:, 0].numpy() - bf["grad"]).max() < 1e-10
            assert engine.path_count(g.n, arcs, g.final, T) == bf["count"]
            sc, labs = engine.path_viterbi(lp, [T], mats)
            assert abs(float(sc[0]) - bf["best_score"]) < 1e-10
            assert np.array_equal(labs[:, 0].numpy(), bf["best_labels"])


def test_oracle_invariants():
    d = load_golden("bi_ctcg")
    loss, grad = engine.fst_loss(d["logits"], d["lens"], d["num_mats"], d["den_mats"], None, dtype=torch.float64)
    assert (loss >= 0).all()          # numerator paths are a subset of denominator paths
    g = grad.numpy()
    for b, ln in enumerate(d["lens"]):
        assert np.abs(g[ln:, b]).max(initial=0.0) == 0.0          # rows past the end are zero
        assert np.abs(g[:ln, b].sum(-1)).max() < 1e-9              # -gamma_num + gamma_den sums to 0
    shift = torch.randn(d["logits"].shape[0], d["logits"].shape[1], 1, dtype=torch.float64)
    loss2, _ = engine.fst_loss(torch.as_tensor(d["logits"]).double() + shift, d["lens"], d["num_mats"],
                               d["den_mats"], None, dtype=torch.float64)
    assert np.allclose(loss.numpy(), loss2.numpy(), atol=1e-8)    # per-frame shifts cancel


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")
def test_oracle_equals_reference_live():
    fu, _ = ref_loader.load()
    rng = np.random.default_rng(11)
    S, order, args = 6, 2, {"use_contextual_blanks": True}
    ref = fu.CTCGraphGen(context_order=order, graph_build_args=args, num_symbols=S)
    mine = graphs.GraphGen(order, S, args)
    labs = [rng.integers(1, S, size=n) for n in (5, 3, 1)]
    texts = np.zeros((3, 5), dtype=np.int64)
    for i, l in enumerate(labs):
        texts[i, :len(l)] = l
    tl = np.array([5, 3, 1])
    rm = ref.get_training_matrices_batch(torch.as_tensor(texts), torch.as_tensor(tl))
    mm = mine.get_training_matrices_batch(texts, tl)
    assert all(np.array_equal(a.numpy(), b) for a, b in zip(rm, mm))
    lp = torch.randn(15, 3, S * S).log_softmax(-1).requires_grad_(True)
    lens = [15, 12, 7]
    c = fu.path_reduction(lp, torch.tensor(lens), rm)
    c.sum().backward()
    c2, g2, _ = engine.path_logsumexp(lp, lens, mm)
    assert torch.equal(c.detach(), c2) and torch.equal(lp.grad, g2)


def test_postprocess_oracle_properties():
    """oracle/postprocess.py (ctc_forward.py:96-128): normalisations sum to one, the '#' mass
    moves to the blank, marginals average the contexts; float32 (the reference's dtype) agrees
    with float64."""
    from oracle import postprocess as pp
    rng = np.random.default_rng(3)
    S = 5
    lp = np.log(rng.dirichlet(np.ones(S * S), size=(6, 2))).astype(np.float32)
    o = pp.postprocess(lp, block_normalize=True)
    assert np.allclose(np.exp(o).reshape(6, 2, S, S).sum(-1), 1.0, atol=1e-12)
    o = pp.postprocess(lp, block_marginalize=True)
    assert o.shape == (6, 2, S) and np.allclose(np.exp(o).sum(-1), 1.0 / S, atol=1e-12)
    mono = np.log(rng.dirichlet(np.ones(S), size=(6, 2)))
    o = pp.postprocess(mono, imitate_biphones=True)
    assert o.shape == (6, 2, S * S) and np.allclose(np.exp(o).sum(-1), 1.0, atol=1e-12)
    assert np.allclose(pp.postprocess(mono, imitate_biphones=True, block_normalize=True), np.tile(mono, (1, 1, S)))
    o = pp.postprocess(lp, transfer_hash_prob=True)
    assert np.allclose(np.exp(o[:, :, 0]), np.exp(lp[:, :, 0].astype(np.float64)) + np.exp(lp[:, :, 3].astype(np.float64)),
                       rtol=1e-6)
    assert np.allclose(o[:, :, 3], np.log(1e-30))
    for kw in ({"block_normalize": True}, {"block_marginalize": True}, {"transfer_hash_prob": True, "block_normalize": True}):
        assert np.abs(pp.postprocess(lp, dtype=np.float32, **kw) - pp.postprocess(lp, **kw)).max() < 1e-5
