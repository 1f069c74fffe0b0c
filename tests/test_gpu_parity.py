"""GPU parity tests: the CUDA path (through the C ABI) against the oracle and the
reference's golden vectors.

Tolerances are BASELINE.json's: loss within 1e-5 relative, gradients within 1e-5
absolute, both against fp64; Viterbi alignments and path counts bit-exact.
"""
import numpy as np
import pytest
import torch

from conftest import golden_names, load_golden
from oracle import engine, graphs

pytestmark = pytest.mark.gpu

LOSS_RTOL = 1e-5
GRAD_ATOL = 1e-5


@pytest.fixture(scope="module")
def P(lib_built):
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    return lib_built


def _rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float((np.abs(a - b) / np.maximum(np.abs(b), 1.0)).max())


def _run_fused(P, logits, lens, num, den, normalize_by_dim, S=0):
    x = torch.as_tensor(logits).float().cuda().requires_grad_(True)
    loss = P.fst_loss(x, lens, num, den, normalize_by_dim=normalize_by_dim, num_symbols=S)
    loss.sum().backward()
    torch.cuda.synchronize()
    return loss.detach().cpu().double().numpy(), x.grad.detach().cpu().double().numpy()


@pytest.mark.parametrize("name", golden_names())
def test_fused_loss_matches_reference_golden(P, name):
    d = load_golden(name)
    gen = P.CTCGraphGen(context_order=d["order"], graph_build_args=d["args"], num_symbols=d["S"])
    num = gen.get_training_graphs(d["label_seqs"])
    den = gen.get_decoding_graph() if d["use_den"] else None
    loss, grad = _run_fused(P, d["logits"], d["lens"].tolist(), num, den, d["normalize_by_dim"], d["S"])
    assert _rel(loss, d["loss_f64"]) < LOSS_RTOL
    assert np.abs(grad - d["grad_f64"]).max() < GRAD_ATOL


@pytest.mark.parametrize("name", golden_names())
def test_path_reduction_dropin_on_reference_matrices(P, name):
    """path_reduction fed with the reference's own padded matrices (fst_utils.py:322)."""
    d = load_golden(name)
    lens = torch.tensor(d["lens"])
    lp = torch.log_softmax(torch.as_tensor(d["logits"]).double(), -1)
    for mats in (d["num_mats"], d["den_mats"]):
        x = lp.float().cuda().requires_grad_(True)
        tm = [torch.as_tensor(m) for m in mats]
        cost = P.path_reduction(x, lens, tm)
        cost.sum().backward()
        want, wgrad, _ = engine.path_logsumexp(lp, d["lens"], mats, dtype=torch.float64)
        assert _rel(cost.detach().cpu(), want) < LOSS_RTOL
        assert (x.grad.cpu().double() - wgrad).abs().max().item() < GRAD_ATOL


def test_context_normalised_loss_through_plugin_signature(P):
    """ctc_fst_loss with normalize_by_dim=1 (ctc_losses.py:617-657)."""
    d = load_golden("bi_ctc_context_norm")
    x = torch.as_tensor(d["logits"]).cuda().requires_grad_(True)
    labels_cat = torch.as_tensor(np.concatenate([np.asarray(s) for s in d["label_seqs"]]))
    loss = P.ctc_fst_loss(x, labels_cat, torch.tensor(d["lens"]), torch.tensor(d["text_lens"]),
                          d["S"], 2, 1, True, False,
                          {"graph_matrices": [torch.as_tensor(m) for m in d["num_mats"]]})
    loss.sum().backward()
    assert _rel(loss.detach().cpu(), d["loss_f64"]) < LOSS_RTOL
    assert np.abs(x.grad.cpu().double().numpy() - d["grad_f64"]).max() < GRAD_ATOL


@pytest.mark.parametrize("name", golden_names())
def test_viterbi_alignment_bit_exact(P, name):
    d = load_golden(name)
    gen = P.CTCGraphGen(context_order=d["order"], graph_build_args=d["args"], num_symbols=d["S"])
    x = torch.as_tensor(d["logits"]).cuda().requires_grad_(True)
    score = P.path_reduction(x, torch.tensor(d["lens"]), gen.get_decoding_graph(), red_kind="viterbi")
    (-score.sum()).backward()
    sel = x.grad.min(-1)[1].cpu().numpy()          # advanced_decoder.py:554
    assert np.array_equal(sel, d["vit_labels"])
    assert _rel(score.detach().cpu(), d["vit_score"]) < LOSS_RTOL
    _, labels = P._lib.viterbi(x, d["lens"].tolist(), gen.get_decoding_graph())
    assert np.array_equal(labels.cpu().numpy(), d["vit_labels"])


def test_path_counts_bit_exact_including_wraparound(P):
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=6)
    ogen = graphs.GraphGen(2, 6, {})
    lens = [150, 40, 7, 0]
    got = P._lib.path_count(lens, gen.get_decoding_graph())
    g = ogen.decoding_graph
    arcs = [(a[0], a[1], a[2], a[4]) for a in g.arcs]
    want = [engine.path_count(g.n, arcs, g.final, n) for n in lens]
    assert [int(v) for v in got] == want
    assert want[0] != 0 and 36 ** 150 > 2 ** 64          # the long one wrapped
    seqs = [[1, 2, 2, 3], [4, 4, 4], [5], [2, 1]]
    num = gen.get_training_graphs(seqs)
    lens = [30, 12, 3, 1]
    got = P._lib.path_count(lens, num)
    for b, s in enumerate(seqs):
        tg = ogen.training_graph(s)
        arcs = [(a[0], a[1], a[2], a[4]) for a in tg.arcs]
        assert int(got[b]) == engine.path_count(tg.n, arcs, tg.final, lens[b])
    d = load_golden("bi_ctcg")
    cnt = P._lib.path_count(d["lens"].tolist(), P.Graph.from_padded(d["num_mats"]))
    assert np.allclose(np.log(cnt.astype(np.float64)), d["log_count_num"], atol=1e-9)


def _random_case(order, S, args, T, B, seed, peaky=False):
    rng = np.random.default_rng(seed)
    V = S ** order
    lens = sorted([T] + rng.integers(int(0.6 * T), T + 1, size=B - 1).tolist(), reverse=True)
    seqs = [rng.integers(1, S, size=max(1, ln // 4)).tolist() for ln in lens]
    logits = (rng.standard_normal((T, B, V)) * 3).astype(np.float32)
    if peaky:
        for b in range(B):
            pos = np.sort(rng.choice(lens[b], size=len(seqs[b]), replace=False))
            prev = 0
            for p, s in zip(pos, seqs[b]):
                logits[p, b, (prev * S + s) if order == 2 else s] += 8
                prev = s
    return logits, lens, seqs


def _oracle(order, S, args, logits, lens, seqs, use_den, normalize_by_dim):
    ogen = graphs.GraphGen(order, S, args)
    U = max(len(s) for s in seqs)
    lab = np.zeros((len(seqs), U), dtype=np.int64)
    for i, s in enumerate(seqs):
        lab[i, :len(s)] = s
    nm = ogen.get_training_matrices_batch(lab, [len(s) for s in seqs])
    dm = ogen.get_decoding_matrices() if use_den else None
    return engine.fst_loss(logits, lens, nm, dm, normalize_by_dim, S, order, dtype=torch.float64)


@pytest.mark.parametrize("order,S,args,use_den,nbd,T,B,peaky", [
    (1, 49, {}, False, 0, 400, 4, False),                         # cfg1 shape, mono CTC
    (1, 49, {}, True, None, 120, 3, True),                        # mono CTC-G
    (2, 12, {}, False, 0, 100, 6, False),                         # bi-char CTC
    (2, 12, {}, True, None, 100, 6, True),                        # CTC-G
    (2, 12, {"use_contextual_blanks": True}, True, None, 90, 5, False),   # CTC-GB
    (2, 10, {"loop_using_symbol_repetitions": True}, True, None, 60, 3, False),
    (2, 10, {"allow_nonblank_selfloops": False}, True, 0, 60, 3, True),   # softmax + den: LSE cancels
    (2, 12, {"use_contextual_blanks": True}, False, 1, 80, 4, False),     # per-context softmax (normalize_by_dim=1)
    (2, 49, {"use_contextual_blanks": True}, False, 1, 50, 2, True),
])
def test_fused_loss_matches_fp64_oracle_random(P, order, S, args, use_den, nbd, T, B, peaky):
    logits, lens, seqs = _random_case(order, S, args, T, B, seed=1000 + T + B, peaky=peaky)
    want_loss, want_grad = _oracle(order, S, args, logits, lens, seqs, use_den, nbd)
    gen = P.CTCGraphGen(context_order=order, graph_build_args=args, num_symbols=S)
    num = gen.get_training_graphs(seqs)
    den = gen.get_decoding_graph() if use_den else None
    loss, grad = _run_fused(P, logits, lens, num, den, nbd, S)
    assert _rel(loss, want_loss.numpy()) < LOSS_RTOL
    assert np.abs(grad - want_grad.numpy()).max() < GRAD_ATOL


@pytest.mark.parametrize("variant", ["ctc", "ctcg", "ctcgb"])
def test_full_vocabulary_slice_matches_fp64_oracle(P, variant):
    """S=49 (V=2401), T=400: two utterances of the BASELINE batch against the fp64 oracle."""
    args = {"use_contextual_blanks": True} if variant == "ctcgb" else {}
    use_den, nbd = (False, 0) if variant == "ctc" else (True, None)
    T = 400 if variant == "ctc" else 120
    logits, lens, seqs = _random_case(2, 49, args, T, 2, seed=1002)
    want_loss, want_grad = _oracle(2, 49, args, logits, lens, seqs, use_den, nbd)
    gen = P.CTCGraphGen(context_order=2, graph_build_args=args, num_symbols=49)
    loss, grad = _run_fused(P, logits, lens, gen.get_training_graphs(seqs),
                            gen.get_decoding_graph() if use_den else None, nbd, 49)
    assert _rel(loss, want_loss.numpy()) < LOSS_RTOL
    assert np.abs(grad - want_grad.numpy()).max() < GRAD_ATOL


def _record(name, value):
    """Largest deviations seen by the bench-size parity tests, kept for DESIGN.md (best effort)."""
    import json
    import os
    d = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if not os.path.isdir(d):
        return
    path = os.path.join(d, "parity_maxerr.json")
    try:
        data = json.load(open(path)) if os.path.exists(path) else {}
        data[name] = value
        json.dump(data, open(path, "w"), indent=1, sort_keys=True)
    except (OSError, ValueError):
        pass


@pytest.mark.parametrize("variant,S,B,T,peaky", [
    ("ctcg", 49, 64, 400, False),     # cfg3, the bench default: CTC-G, V=2401
    ("ctcg", 49, 64, 400, True),
    ("ctcgb", 49, 32, 800, False),    # cfg4 / cfg5 shape: CTC-GB, T=800 (an utterance does not depend on B)
    ("ctcgb", 49, 32, 800, True),
    ("ctcg", 36, 64, 400, False),     # BASELINE's nominal V=1296
    ("ctc", 36, 64, 400, False),
    ("ctc", 49, 64, 400, True),       # cfg2
    ("ctcg", 49, 3, 1400, False),     # > 256 distinct labels per utterance: the combine's read-modify-write tail
])
def test_bench_configurations_match_fp64_oracle(P, variant, S, B, T, peaky):
    """The configurations bench.py times (seed-1003 batch at full B and T) against the fp64 oracle:
    the CUDA path runs the whole batch, the oracle three utterances of it (first, middle, last)."""
    args = {"use_contextual_blanks": True} if variant == "ctcgb" else {}
    use_den, nbd = (False, 0) if variant == "ctc" else (True, None)
    logits, lens, seqs = _random_case(2, S, args, T, B, seed=1003, peaky=peaky)
    gen = P.CTCGraphGen(context_order=2, graph_build_args=args, num_symbols=S)
    loss, grad = _run_fused(P, logits, lens, gen.get_training_graphs(seqs),
                            gen.get_decoding_graph() if use_den else None, nbd, S)
    sub = [0, B // 2, B - 1]
    want_loss, want_grad = _oracle(2, S, args, np.ascontiguousarray(logits[:, sub]), [lens[i] for i in sub],
                                   [seqs[i] for i in sub], use_den, nbd)
    dl = _rel(loss[sub], want_loss.numpy())
    dg = float(np.abs(grad[:, sub] - want_grad.numpy()).max())
    _record("%s_S%d_B%d_T%d%s" % (variant, S, B, T, "_peaky" if peaky else ""), {"loss_rel": dl, "grad_abs": dg})
    assert dl < LOSS_RTOL
    assert dg < GRAD_ATOL


@pytest.mark.parametrize("S,T,args", [(49, 400, {}), (49, 200, {"use_contextual_blanks": True}), (36, 300, {})])
def test_viterbi_and_collapse_at_full_vocabulary(P, S, T, args):
    """Best path over the full bi-char decoding graph (fst_utils.py:366-397, advanced_decoder.py:546-571)
    at production size: score and per-frame labels against the oracle's tropical recursion, and the
    collapse of the labels against the oracle's walk over the arc table."""
    logits, lens, seqs = _random_case(2, S, args, T, 2, seed=515 + S, peaky=True)
    gen = P.CTCGraphGen(context_order=2, graph_build_args=args, num_symbols=S)
    ogen = graphs.GraphGen(2, S, args)
    lp = torch.as_tensor(logits)
    want_score, want_labels = engine.path_viterbi(lp, lens, ogen.get_decoding_matrices(), dtype=torch.float64)
    score, labels = P._lib.viterbi(lp.cuda(), lens, gen.get_decoding_graph())
    assert _rel(score.cpu().numpy(), want_score.numpy()) < LOSS_RTOL
    labels = labels.cpu().numpy()
    want_labels = np.asarray(want_labels)
    for b, n in enumerate(lens):
        got = P.collapse_frame_labels(labels[:n, b].tolist(), 2, S, **args)
        assert got == graphs.collapse_by_arc_table(ogen.decoding_graph, labels[:n, b].tolist())
        if not np.array_equal(labels[:n, b], want_labels[:n, b]):
            # fp32 recursion vs fp64 oracle: a different path is acceptable only when it is a path of the
            # graph (checked by the collapse above) whose own fp64 score ties with the optimum
            mine = float(sum(np.float64(logits[t, b, labels[t, b]]) for t in range(n)))
            assert abs(mine - float(want_score[b])) < 2e-3


@pytest.mark.parametrize("S,args", [(12, {}), (49, {"use_contextual_blanks": True})])
def test_wide_logit_spread_exercises_the_power_of_two_rescale(P, S, args):
    """Logits ~ 12 * N(0,1): the scaled-linear sweeps of the bi-char denominator leave their
    exponent band every few frames, so the lagged power-of-two rescale (sweeps) and its replay in
    the combine kernel are on the path (oracle/bichar_model.py shows it firing at this spread)."""
    T, B = 96, 3
    logits, lens, seqs = _random_case(2, S, args, T, B, seed=977 + S)
    logits = (logits * 4).astype(np.float32)
    want_loss, want_grad = _oracle(2, S, args, logits, lens, seqs, True, None)
    gen = P.CTCGraphGen(context_order=2, graph_build_args=args, num_symbols=S)
    loss, grad = _run_fused(P, logits, lens, gen.get_training_graphs(seqs), gen.get_decoding_graph(), None, S)
    assert _rel(loss, want_loss.numpy()) < LOSS_RTOL
    assert np.abs(grad - want_grad.numpy()).max() < GRAD_ATOL


def test_all_row_alignments_and_vector_paths_match_fp64_oracle(P):
    """B = 4 with V = 2401: the logits rows of the four utterances start at the four possible
    offsets inside a 16-byte window (the sweeps' producers work on the aligned window) and
    B*V is a multiple of 4, which is the condition for the 16-byte path of the combine kernel;
    T = 43 leaves a partial last segment (scalar path) next to full ones."""
    T, B = 43, 4
    logits, lens, seqs = _random_case(2, 49, {}, T, B, seed=4242)
    want_loss, want_grad = _oracle(2, 49, {}, logits, lens, seqs, True, None)
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=49)
    loss, grad = _run_fused(P, logits, lens, gen.get_training_graphs(seqs), gen.get_decoding_graph(), None, 49)
    assert _rel(loss, want_loss.numpy()) < LOSS_RTOL
    assert np.abs(grad - want_grad.numpy()).max() < GRAD_ATOL


@pytest.mark.parametrize("S,args", [(49, {}), (49, {"use_contextual_blanks": True}), (36, {}),
                                    (11, {"allow_nonblank_selfloops": False})])
def test_fused_bichar_kernel_equals_generic_csr_kernel(P, S, args):
    """The factorised bi-char denominator kernel against the generic CSR kernel on the same
    decoding graph (imported arc by arc through the reference's padded matrices; the import recognises
    the topology, ``set_generic`` drops the tag and with it the fused route)."""
    T, B = 64, 3
    logits, lens, seqs = _random_case(2, S, args, T, B, seed=77 + S)
    gen = P.CTCGraphGen(context_order=2, graph_build_args=args, num_symbols=S)
    den = gen.get_decoding_graph()
    mats = den.export_padded(0, out_edges=False) + den.export_padded(0, out_edges=True)
    generic = P.Graph.from_padded([torch.as_tensor(m)[None] for m in mats])
    assert generic.info().structured == 1          # the re-imported decoding graph is recognised ...
    generic.set_generic()                          # ... and is sent to the generic kernel on request
    assert generic.info().structured == 0
    x = torch.from_numpy(logits).cuda()
    c_fused, g_fused = P._lib.path_sum(x, lens, den)
    c_gen, g_gen = P._lib.path_sum(x, lens, generic)
    assert _rel(c_fused.cpu().numpy(), c_gen.cpu().numpy()) < 2e-6
    assert (g_fused - g_gen).abs().max().item() < 5e-6
    assert (g_fused[lens[-1]:, -1] == 0).all()


@pytest.mark.parametrize("variant,B,T", [("ctc", 64, 400), ("ctcg", 64, 400), ("ctcgb", 32, 800)])
def test_full_size_properties(P, variant, B, T):
    """BASELINE-size batches: size-independent properties of the posteriors."""
    args = {"use_contextual_blanks": True} if variant == "ctcgb" else {}
    use_den, nbd = (False, 0) if variant == "ctc" else (True, None)
    logits, lens, seqs = _random_case(2, 49, args, T, B, seed=1003)
    gen = P.CTCGraphGen(context_order=2, graph_build_args=args, num_symbols=49)
    num = gen.get_training_graphs(seqs)
    den = gen.get_decoding_graph() if use_den else None
    loss, grad = _run_fused(P, logits, lens, num, den, nbd, 49)
    assert np.isfinite(loss).all() and np.isfinite(grad).all()
    assert (loss > 0).all()
    for b in range(0, B, 7):
        assert np.abs(grad[lens[b]:, b]).max(initial=0.0) == 0.0           # pad rows stay zero
        assert np.abs(grad[:lens[b], b].sum(-1)).max() < 2e-5               # posteriors sum to one
    # per-frame shifts of the logits change nothing (softmax / num-den invariance)
    shift = np.random.default_rng(5).standard_normal((T, B, 1)).astype(np.float32) * 4
    loss2, grad2 = _run_fused(P, logits + shift, lens, num, den, nbd, 49)
    assert _rel(loss2, loss) < 2e-5
    # logits + shift is rounded to fp32 (ulp ~2e-6 at |x| ~ 20), so the inputs themselves differ
    assert np.abs(grad2 - grad).max() < 1e-4
    # an utterance's result does not depend on its batch neighbours
    sub = [0, B // 2, B - 1]
    gsub = gen.get_training_graphs([seqs[i] for i in sub])
    loss3, grad3 = _run_fused(P, logits[:, sub], [lens[i] for i in sub], gsub, den, nbd, 49)
    assert _rel(loss3, loss[sub]) < 1e-6
    assert np.abs(grad3 - grad[:, sub]).max() < 1e-6


def test_edge_cases(P):
    S = 5
    gen = P.CTCGraphGen(context_order=1, graph_build_args={}, num_symbols=S)
    ogen = graphs.GraphGen(1, S, {})
    rng = np.random.default_rng(9)
    # (a) exactly one path: T == U (no repeats); (b) a repeat needs T >= U+1; (c) B == 1
    seqs = [[1, 2, 3, 4]]
    x = (rng.standard_normal((4, 1, S)) * 2).astype(np.float32)
    loss, grad = _run_fused(P, x, [4], gen.get_training_graphs(seqs), None, 0)
    lp = torch.log_softmax(torch.as_tensor(x).double(), -1)
    want = -sum(lp[t, 0, s] for t, s in enumerate(seqs[0]))
    assert abs(loss[0] - float(want)) < 1e-5
    # (d) infeasible: T < U.  The reference returns a finite ~1e20-scale loss, never NaN/inf
    loss, grad = _run_fused(P, x[:3], [3], gen.get_training_graphs(seqs), None, 0)
    assert np.isfinite(loss).all() and loss[0] > 1e19 and np.isfinite(grad).all()
    # (e) ragged batch with an empty utterance and a 1-frame utterance
    seqs = [[1, 1, 2], [3], [4]]
    lens = [9, 1, 0]
    x = (rng.standard_normal((9, 3, S)) * 2).astype(np.float32)
    loss, grad = _run_fused(P, x, lens, gen.get_training_graphs(seqs), gen.get_decoding_graph(), None)
    U = 3
    lab = np.zeros((3, U), dtype=np.int64)
    for i, s in enumerate(seqs):
        lab[i, :len(s)] = s
    nm = ogen.get_training_matrices_batch(lab, [3, 1, 1])
    wl, wg = engine.fst_loss(x, lens, nm, ogen.get_decoding_matrices(), None, dtype=torch.float64)
    assert _rel(loss[:2], wl.numpy()[:2]) < LOSS_RTOL
    assert loss[2] > 1e19                      # empty utterance with a non-empty transcript: infeasible
    assert np.abs(grad[:, :2] - wg.numpy()[:, :2]).max() < GRAD_ATOL
    assert np.abs(grad[:, 2]).max() == 0.0
    # (f) the C ABI itself does not need sorted lengths
    seqs = [[1], [2, 3], [4, 4]]
    lens_u = [4, 9, 6]
    num = gen.get_training_graphs(seqs)
    xc = torch.as_tensor(x).cuda()
    l_u, g_u = P._lib.loss_fwd_bwd(xc, lens_u, num, None, P.NORM_SOFTMAX)
    order = [1, 2, 0]
    l_s, g_s = P._lib.loss_fwd_bwd(xc[:, order].contiguous(), [lens_u[i] for i in order],
                                   gen.get_training_graphs([seqs[i] for i in order]), None, P.NORM_SOFTMAX)
    assert torch.allclose(l_u[order], l_s, rtol=1e-6) and torch.allclose(g_u[:, order], g_s, atol=1e-7)


def test_host_buffer_entry_point_equals_device_path(P):
    logits, lens, seqs = _random_case(2, 8, {}, 50, 4, seed=21)
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=8)
    num, den = gen.get_training_graphs(seqs), gen.get_decoding_graph()
    xh = torch.as_tensor(logits).pin_memory()
    lh, gh = P._lib.loss_host(xh, lens, num, den, P.NORM_NONE)
    ld, gd = P._lib.loss_fwd_bwd(xh.cuda(), lens, num, den, P.NORM_NONE)
    assert torch.equal(lh, ld.cpu()) and torch.equal(gh, gd.cpu())
    lh2, _ = P._lib.loss_host(xh, lens, num, None, P.NORM_SOFTMAX, want_grad=False)
    ld2, _ = P._lib.loss_fwd_bwd(xh.cuda(), lens, num, None, P.NORM_SOFTMAX, need_grad=False)
    assert torch.equal(lh2, ld2.cpu())


def test_decoder_module_and_grad_scale(P):
    logits, lens, seqs = _random_case(2, 7, {}, 40, 3, seed=33)
    head = P.FSTLoss({"class_name": "CTCGraphGen", "context_order": 2, "graph_build_args": {}},
                     num_symbols=7)
    x = torch.as_tensor(logits).cuda().requires_grad_(True)
    U = max(len(s) for s in seqs)
    texts = torch.zeros(3, U, dtype=torch.long)
    for i, s in enumerate(seqs):
        prev = 0
        for j, v in enumerate(s):          # bi-char dataset ids: prev*S+cur (egs/wsj/data.py:146-161)
            texts[i, j] = prev * 7 + v
            prev = v
    out = head(x, torch.tensor(lens), texts, torch.tensor([len(s) for s in seqs]))
    assert set(out) == {"fst_loss", "loss"}
    (2.5 * out["loss"]).backward()
    wl, wg = _oracle(2, 7, {}, logits, lens, seqs, True, None)
    assert abs(out["loss"].item() - wl.sum().item()) / wl.sum().item() < LOSS_RTOL
    assert np.abs(x.grad.cpu().double().numpy() - 2.5 * wg.numpy()).max() < 2.5 * GRAD_ATOL
    # per-utterance weights through the C ABI's grad_scale
    w = torch.tensor([0.5, 2.0, -1.0])
    gen = head.graph_generator
    _, g = P._lib.loss_fwd_bwd(x.detach(), lens, gen.get_training_graphs(seqs), gen.get_decoding_graph(),
                               P.NORM_NONE, grad_scale=w)
    assert np.abs(g.cpu().double().numpy() - wg.numpy() * w.numpy()[None, :, None]).max() < 2 * GRAD_ATOL


@pytest.mark.gpu
@pytest.mark.parametrize("flags", [
    {}, {"transfer_hash_prob": True}, {"block_normalize": True}, {"block_marginalize": True},
    {"transfer_hash_prob": True, "block_normalize": True}, {"transfer_hash_prob": True, "block_marginalize": True},
    {"imitate_biphones": True}, {"imitate_biphones": True, "block_normalize": True},
    {"imitate_biphones": True, "transfer_hash_prob": True},
])
def test_ctc_forward_postprocessing(P, flags):
    """postprocess_logprobs against the numpy statements of ctc_forward.py:96-128 (fp64 oracle; the
    reference itself computes in float32, tolerance 1e-5 absolute on log-probabilities)."""
    from oracle import postprocess as pp
    rng = np.random.default_rng(17)
    mono = flags.get("imitate_biphones", False)
    S, T, B = (49, 37, 3)
    V = S if mono else S * S
    lp = torch.log_softmax(torch.from_numpy(rng.standard_normal((T, B, V)).astype(np.float32) * 3), -1)
    got = P.postprocess_logprobs(lp.cuda(), **flags).cpu().double().numpy()
    want = pp.postprocess(lp.numpy(), **flags)
    assert got.shape == want.shape
    assert np.abs(got - want).max() < 1e-5
    ref32 = pp.postprocess(lp.numpy(), dtype=np.float32, **flags)        # what the reference would write
    assert np.abs(got - ref32).max() < 2e-5


@pytest.mark.gpu
@pytest.mark.parametrize("order", [1, 2])
def test_decode_returns_the_transcript_on_peaked_logits(P, order):
    """FSTDecoder.decode (advanced_decoder.py:536-593) through the stand-alone head: Viterbi frame
    labels on the device, collapse by the topology rule; logits peaked on an alignment of the
    transcript (symbol held one extra frame where the graph has a self-loop, blank in between)."""
    S, B = 9, 3
    rng = np.random.default_rng(31 + order)
    seqs = [rng.integers(1, S, size=n).tolist() for n in (7, 5, 4)]
    aligns = []
    for seq in seqs:
        frames, ctx_sym = [], 0
        for cur in seq:
            cls = cur if order == 1 else ctx_sym * S + cur
            frames.append(cls)
            if order == 1 or ctx_sym != cur:
                frames.append(cls)          # hold: explicit self-loop
            frames.append(0)                # blank (global)
            ctx_sym = cur
        aligns.append(frames)
    lens = sorted((len(a) for a in aligns), reverse=True)
    order_idx = sorted(range(B), key=lambda i: -len(aligns[i]))
    aligns = [aligns[i] for i in order_idx]
    seqs = [seqs[i] for i in order_idx]
    T, V = lens[0], S ** order
    logits = (rng.standard_normal((T, B, V)) * 0.5).astype(np.float32)
    for b, a in enumerate(aligns):
        for t, c in enumerate(a):
            logits[t, b, c] += 12
    gen = P.CTCGraphGen(context_order=order, graph_build_args={}, num_symbols=S)
    head = P.FSTLoss(gen, normalize_by_dim=None)
    U = max(len(q) for q in seqs)
    texts = torch.zeros((B, U), dtype=torch.int64)
    for i, q in enumerate(seqs):
        texts[i, :len(q)] = torch.tensor(q)
    ret = head.decode(torch.from_numpy(logits).cuda(), torch.tensor(lens), texts,
                      torch.tensor([len(q) for q in seqs]))
    assert ret["decoded"] == seqs
    assert torch.isfinite(ret["loss"]["loss"]).item()



def test_path_reduction_on_decoding_matrices_takes_the_factorised_kernels_and_imports_once(P):
    """Boundary route 2: ``path_reduction`` rebound into the reference, fed with
    ``graph_generator.get_decoding_matrices(device)`` every step (advanced_decoder.py:497-500)."""
    import importlib
    fu = importlib.import_module(P.CTCGraphGen.__module__)
    S, T, B = 12, 40, 3
    logits, lens, seqs = _random_case(2, S, {}, T, B, seed=61)
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S)
    x = torch.from_numpy(logits).cuda()
    mats = gen.get_decoding_matrices(x.device)
    plain = [m.to(x.device) for m in mats]               # the reference's own copy loop (:458): plain list, same tensors
    before = fu.import_count
    want, wgrad, _ = engine.path_logsumexp(torch.from_numpy(logits), lens, graphs.GraphGen(2, S, {}).get_decoding_matrices(),
                                           dtype=torch.float64)
    for step in range(2):
        xs = x.clone().requires_grad_(True)
        P._lib.profile(True)
        cost = P.path_reduction(xs, torch.tensor(lens), [m.to(x.device) for m in plain])
        cost.sum().backward()
        torch.cuda.synchronize()
        names = [n for n, _ in P._lib.profile_collect()]
        P._lib.profile(False)
        assert "bisweep" in names and not any(n.startswith("path_fwd_bwd") for n in names)
        assert _rel(cost.detach().cpu(), want) < LOSS_RTOL
        assert (xs.grad.cpu().double() - wgrad).abs().max().item() < GRAD_ATOL
    assert fu.import_count == before + 1                  # converted once, found again by tensor identity
    xs = x.clone().requires_grad_(True)                   # the tagged list needs no import at all
    P.path_reduction(xs, torch.tensor(lens), mats).sum().backward()
    assert fu.import_count == before + 1


@pytest.mark.parametrize("nbd", [None, 0, 1])
def test_logits_and_decode_follow_the_reference_normalisation(P, nbd):
    """``FSTDecoder.logits`` returns normalised scores when ``normalize_by_dim`` is set
    (advanced_decoder.py:444-452) and ``decode`` finds the best path on those (:536-554)."""
    S, T, B = 8, 30, 2
    args = {"use_contextual_blanks": True}
    logits, lens, seqs = _random_case(2, S, args, T, B, seed=88)
    gen = P.CTCGraphGen(context_order=2, graph_build_args=args, num_symbols=S)
    head = P.FSTLoss(gen, normalize_by_dim=nbd, denominator_red="none" if nbd is not None else "logsumexp")
    x = torch.from_numpy(logits).cuda()
    got = head.logits(x, torch.tensor(lens))
    want = engine.normalized_acts(torch.from_numpy(logits).double(), S, 2, nbd) if nbd is not None \
        else torch.from_numpy(logits).double()
    assert (got.cpu().double() - want).abs().max().item() < 1e-5
    ret = head.decode(x, torch.tensor(lens))
    assert (ret["logits"].cpu().double() - want).abs().max().item() < 1e-5
    ogen = graphs.GraphGen(2, S, args)
    _, want_labels = engine.path_viterbi(want, lens, ogen.get_decoding_matrices(), dtype=torch.float64)
    for b, n in enumerate(lens):
        assert ret["decoded"][b] == graphs.collapse_by_arc_table(ogen.decoding_graph, want_labels[:n, b].tolist())


def test_context_normalisation_together_with_a_denominator(P):
    """normalize_by_dim=1 with denominator_red='logsumexp' (the reference allows any combination:
    ctc_losses.py:29-43 feeding advanced_decoder.py:497-500)."""
    S, T, B = 9, 50, 3
    args = {"use_contextual_blanks": True}
    logits, lens, seqs = _random_case(2, S, args, T, B, seed=123)
    want_loss, want_grad = _oracle(2, S, args, logits, lens, seqs, True, 1)
    gen = P.CTCGraphGen(context_order=2, graph_build_args=args, num_symbols=S)
    loss, grad = _run_fused(P, logits, lens, gen.get_training_graphs(seqs), gen.get_decoding_graph(), 1, S)
    assert _rel(loss, want_loss.numpy()) < LOSS_RTOL
    assert np.abs(grad - want_grad.numpy()).max() < GRAD_ATOL
    # straight through the C entry point (normalised copy, loss, gradient pulled back in place), with per-utterance
    # weights; the plain workspace size is refused, the host-buffer entry point sizes its own
    num, den = gen.get_training_graphs(seqs), gen.get_decoding_graph()
    w = torch.tensor([0.5, -1.5, 2.0])
    l2, g2 = P._lib.loss_fwd_bwd(torch.from_numpy(logits).cuda(), lens, num, den, P.NORM_CONTEXT, S, grad_scale=w)
    assert _rel(l2.cpu().numpy(), want_loss.numpy()) < LOSS_RTOL
    assert np.abs(g2.cpu().double().numpy() - want_grad.numpy() * w.numpy()[None, :, None]).max() < 2 * GRAD_ATOL
    L = P._lib.lib()
    V = S * S
    small = L.ctcg_workspace_bytes(T, B, V, num.ptr, den.ptr)
    assert L.ctcg_workspace_bytes_norm(T, B, V, num.ptr, den.ptr, P.NORM_CONTEXT) >= small + T * B * V * 4
    assert L.ctcg_workspace_bytes_norm(T, B, V, num.ptr, den.ptr, P.NORM_NONE) - small < 256
    ws = torch.empty(small, dtype=torch.uint8, device="cuda")
    x = torch.from_numpy(logits).cuda()
    lo, gr = torch.empty(B, device="cuda"), torch.empty(T, B, V, device="cuda")
    lens_np = np.asarray(lens, dtype=np.int32)
    st = L.ctcg_loss_fwd_bwd(x.data_ptr(), T, B, V, lens_np.ctypes.data, num.ptr, den.ptr, P.NORM_CONTEXT, S, None,
                             lo.data_ptr(), gr.data_ptr(), ws.data_ptr(), small, None)
    assert st == 7                                                   # CTCG_EWORKSPACE
    lh, gh = P._lib.loss_host(torch.from_numpy(logits), lens, num, den, P.NORM_CONTEXT, S)
    assert _rel(lh.numpy(), want_loss.numpy()) < LOSS_RTOL and np.abs(gh.double().numpy() - want_grad.numpy()).max() < GRAD_ATOL


def test_per_utterance_grad_output_scales_rows_on_the_device(P):
    """``PathLogSumExp.backward`` (fst_utils.py:482-485): grad_output[b] * grads; the all-ones case of
    ``.sum().backward()`` must not touch the gradient (ctcg_scale_grad skips rows with factor 1)."""
    S, T, B = 10, 40, 4
    logits, lens, seqs = _random_case(2, S, {}, T, B, seed=321)
    wl, wg = _oracle(2, S, {}, logits, lens, seqs, True, None)
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S)
    num, den = gen.get_training_graphs(seqs), gen.get_decoding_graph()
    w = torch.tensor([1.0, -2.0, 0.25, 1.0])
    x = torch.from_numpy(logits).cuda().requires_grad_(True)
    (P.fst_loss(x, lens, num, den) * w.cuda()).sum().backward()
    assert np.abs(x.grad.cpu().double().numpy() - wg.numpy() * w.numpy()[None, :, None]).max() < 2 * GRAD_ATOL
    g = torch.from_numpy(wg.numpy()).float().cuda()
    keep = g.clone()
    P._lib.scale_grad(g, torch.ones(B, device="cuda").sum().expand(B) / B)      # broadcast scalar 1.0 (stride 0)
    assert torch.equal(g, keep)
    P._lib.scale_grad(g, torch.tensor(3.0, device="cuda").expand(B))
    assert torch.allclose(g, keep * 3)


@pytest.mark.parametrize("S,args,T,B", [(49, {}, 400, 4), (49, {"use_contextual_blanks": True}, 200, 3),
                                        (36, {"allow_nonblank_selfloops": False}, 150, 3), (7, {}, 60, 5)])
def test_factorised_viterbi_equals_the_generic_kernel(P, S, args, T, B):
    """The tropical counterpart of the factorised denominator (biviterbi_kernel) against the generic CSR
    Viterbi on the same decoding graph (fst_utils.py:366-397 on the arcs of :765-817): same frame labels,
    same score; the timings of both are kept for DESIGN.md."""
    import time
    logits, lens, seqs = _random_case(2, S, args, T, B, seed=900 + S + T, peaky=True)
    gen = P.CTCGraphGen(context_order=2, graph_build_args=args, num_symbols=S)
    fact = gen.get_decoding_graph()
    generic = P.Graph.decoding(2, S, **args).set_generic()
    assert fact.info().structured == 1 and generic.info().structured == 0
    x = torch.from_numpy(logits).cuda()
    out = {}
    for name, g in (("factorised", fact), ("generic", generic)):
        P._lib.viterbi(x, lens, g)                      # warm-up (graph upload, workspace)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        score, labels = P._lib.viterbi(x, lens, g)
        torch.cuda.synchronize()
        out[name] = (score.cpu().numpy(), labels.cpu().numpy(), (time.perf_counter() - t0) * 1e3)
    assert np.array_equal(out["factorised"][1], out["generic"][1])
    assert _rel(out["factorised"][0], out["generic"][0]) < 1e-6
    _record("viterbi_S%d_T%d_B%d_ms" % (S, T, B), {"factorised": out["factorised"][2], "generic": out["generic"][2]})


@pytest.mark.parametrize("name,fn", [("mono_ctc", "ctc_raw_loss"), ("mono_ctc", "ctc_raw_loss_batch"),
                                     ("bi_ctc_context_norm", "ctc_raw_loss"), ("bi_ctc_context_norm", "ctc_raw_loss_batch")])
def test_dense_route_plugin_drivers_match_raw_generic_ctc(P, name, fn):
    """``ctc_loss_fn: ctc_raw_loss`` / ``ctc_raw_loss_batch`` (ctc_losses.py:521-609, resolved by name at
    advanced_decoder.py:238): same signature, and the values the reference's own ``RawGenericCTC`` (numpy
    fp64, :268-324) returned on these inputs (fixture key ``raw_generic_ctc``)."""
    d = load_golden(name)
    x = torch.as_tensor(d["logits"]).cuda().requires_grad_(True)
    labels_cat = torch.as_tensor(np.concatenate([np.asarray(s) for s in d["label_seqs"]]))
    loss = getattr(P, fn)(x, labels_cat, torch.tensor(d["lens"]), torch.tensor(d["text_lens"]),
                          d["S"], d["order"], d["normalize_by_dim"], True, False)
    loss.sum().backward()
    assert _rel(loss.detach().cpu(), d["raw_generic_ctc"]) < LOSS_RTOL
    # gradient: the fp64 oracle on the same state machines (the dense bi-context topology is not the FST one
    # when a letter repeats, so the FST-route fixture gradient does not apply)
    import importlib
    cl = importlib.import_module(P.FSTLoss.__module__)
    g = P.Graph.from_arcs([cl.dense_ctc_graph(q, d["S"], d["order"]) for q in d["label_seqs"]])
    nm = P.batch_training_graph_matrices([P.padded_matrices(g, i) for i in range(g.num_graphs)])
    want_loss, want_grad = engine.fst_loss(d["logits"], d["lens"].tolist(), [m.numpy() for m in nm], None,
                                           d["normalize_by_dim"], d["S"], d["order"], dtype=torch.float64)
    assert _rel(want_loss.numpy(), d["raw_generic_ctc"]) < 1e-6      # the reference returns it rounded to fp32
    assert np.abs(x.grad.cpu().double().numpy() - want_grad.numpy()).max() < GRAD_ATOL


@pytest.mark.parametrize("order,S,args,use_den,nbd,T,B", [
    (2, 12, {}, True, None, 37, 5),                                  # CTC-G: sweeps, combine with the fold, groups off (B < 64)
    (2, 49, {"use_contextual_blanks": True}, True, None, 45, 64),    # CTC-GB at B=64: 8 length groups
    (2, 12, {}, False, 0, 37, 5),                                    # plain CTC: rows kernel
    (2, 9, {"use_contextual_blanks": True}, False, 1, 30, 4),        # per-context softmax
    (3, 4, {"use_context_blanks": True}, True, None, 33, 3),         # tri-char: generic CSR denominator
    (1, 7, {}, True, None, 29, 4),                                   # mono-char CTC-G: generic denominator
])
def test_workspace_and_outputs_stay_inside_their_bounds(P, order, S, args, use_den, nbd, T, B):
    """compute-sanitizer is closed on this pool, so the bounds are checked directly: the workspace
    (exactly ``ctcg_workspace_bytes``), the loss and the gradient sit between guard bands of a known
    pattern, the C entry point is called on them, and no guard byte may change; the results must equal
    those of the normal path."""
    rng = np.random.default_rng(T * 31 + B)
    V = S ** order
    lens = sorted([T] + rng.integers(int(0.6 * T), T + 1, size=B - 1).tolist(), reverse=True)
    seqs = [rng.integers(1, S, size=max(1, ln // 4)).tolist() for ln in lens]
    x = torch.from_numpy((rng.standard_normal((T, B, V)) * 3).astype(np.float32)).cuda()
    gen = P.CTCGraphGen(context_order=order, graph_build_args=args, num_symbols=S)
    num = gen.get_training_graphs(seqs)
    den = gen.get_decoding_graph() if use_den else None
    mode = {None: P.NORM_NONE, 0: P.NORM_SOFTMAX, 1: P.NORM_CONTEXT}[nbd]
    want_loss, want_grad = P._lib.loss_fwd_bwd(x, lens, num, den, mode, S)      # uploads the graphs too
    L = P._lib.lib()
    need = L.ctcg_workspace_bytes(T, B, V, num.ptr, den.ptr if den is not None else None)
    G = 1 << 16                                                                  # guard bytes (keeps 16-byte alignment)
    ws_all = torch.full((need + 2 * G,), 0xA5, dtype=torch.uint8, device="cuda")
    out_all = torch.full((2 * G // 4 + T * B * V,), 12345.0, dtype=torch.float32, device="cuda")
    loss_all = torch.full((2 * 64 + B,), 12345.0, dtype=torch.float32, device="cuda")
    ws, grad, loss = ws_all[G:G + need], out_all[G // 4:G // 4 + T * B * V], loss_all[64:64 + B]
    lens_np = np.ascontiguousarray(np.asarray(lens, dtype=np.int32))
    for _ in range(2):                                                           # twice: stale workspace contents
        st = L.ctcg_loss_fwd_bwd(x.data_ptr(), T, B, V, lens_np.ctypes.data, num.ptr, den.ptr if den is not None else None,
                                 int(mode), int(S), None, loss.data_ptr(), grad.data_ptr(), ws.data_ptr(), need, None)
        assert st == 0
        torch.cuda.synchronize()
        assert bool((ws_all[:G] == 0xA5).all()) and bool((ws_all[G + need:] == 0xA5).all())
        assert bool((out_all[:G // 4] == 12345.0).all()) and bool((out_all[G // 4 + T * B * V:] == 12345.0).all())
        assert bool((loss_all[:64] == 12345.0).all()) and bool((loss_all[64 + B:] == 12345.0).all())
        assert torch.equal(loss, want_loss)
        if order == 2:                                                           # factorised / rows kernels: run-to-run bit-identical
            assert torch.equal(grad.view(T, B, V), want_grad)
        else:                                                                    # generic denominator sums arcs with atomics
            assert torch.allclose(grad.view(T, B, V), want_grad, rtol=0, atol=2e-6)
    # one byte less of workspace is refused, nothing is launched
    launches = P._lib.launch_count()
    st = L.ctcg_loss_fwd_bwd(x.data_ptr(), T, B, V, lens_np.ctypes.data, num.ptr, den.ptr if den is not None else None,
                             int(mode), int(S), None, loss.data_ptr(), grad.data_ptr(), ws.data_ptr(), need - 1, None)
    assert st == 7 and P._lib.launch_count() == launches                         # CTCG_EWORKSPACE


def test_upload_logits_moves_the_live_frames_only(P):
    """``ctcg_upload_logits``: the device copy equals the host tensor on every frame t < len_b and is zero
    elsewhere; an unsorted batch and a batch without padding are copied whole; the loss computed on the trimmed
    copy is bit-identical to the loss on the full copy (no kernel reads the padding)."""
    rng = np.random.default_rng(5)
    T, B, S = 41, 9, 7
    V = S * S
    host = torch.from_numpy(rng.standard_normal((T, B, V)).astype(np.float32)).pin_memory()
    lens = [41, 41, 40, 33, 33, 33, 20, 7, 0]
    dev = P.upload_logits(host, lens)
    torch.cuda.synchronize()
    want = host.clone()
    for b, ln in enumerate(lens):
        want[ln:, b] = 0
    assert torch.equal(dev.cpu(), want)
    # out= re-uses a buffer (stale contents must not survive in the padding)
    buf = torch.full((T, B, V), 7.0, device="cuda")
    assert P.upload_logits(host, lens, out=buf) is buf and torch.equal(buf.cpu(), want)
    # not sorted: whole copy;  no padding: whole copy
    assert torch.equal(P.upload_logits(host, lens[::-1]).cpu(), host)
    assert torch.equal(P.upload_logits(host, [T] * B).cpu(), host)
    with pytest.raises(Exception):
        P.upload_logits(host, [T + 1] + [T] * (B - 1))
    # the loss does not see the difference
    lens2 = [41, 40, 33, 33, 29, 25, 20, 18, 15]
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S)
    seqs = [rng.integers(1, S, size=max(1, ln // 4)).tolist() for ln in lens2]
    num, den = gen.get_training_graphs(seqs), gen.get_decoding_graph()
    la, ga = P._lib.loss_fwd_bwd(host.cuda(), lens2, num, den, P.NORM_NONE, S)
    lb, gb = P._lib.loss_fwd_bwd(P.upload_logits(host, lens2), lens2, num, den, P.NORM_NONE, S)
    assert torch.equal(la, lb) and torch.equal(ga, gb)
    lh, gh = P._lib.loss_host(host, lens2, num, den, P.NORM_NONE, S)       # the host entry point goes through it
    assert torch.equal(lh, la.cpu()) and torch.equal(gh, ga.cpu())


@pytest.mark.gpu
def test_trichar_decode_follows_a_peaked_alignment(P):
    """FSTDecoder.decode on the tri-char graph with contextual blanks (the input-deterministic tri-char
    topology, fst_utils.py:926-1050): Viterbi on the generic kernel, collapse by the arc-table walk.  The logits
    are peaked on random accepted paths of the oracle's restatement of the graph; the decoded symbols must be
    what those paths emit."""
    from oracle import graphs
    S, B = 4, 3
    V = S ** 3
    g = graphs.trigram_decoding_graph(S, use_context_blanks=True)
    out_arcs = g.out_lists()
    rng = np.random.default_rng(77)
    aligns, emitted = [], []
    for n in (34, 27, 19):
        while True:
            state, labels, out = 0, [], []
            for _ in range(n):
                arcs = out_arcs[state]
                stay = [a for a in arcs if a[1] == state]
                a = stay[0] if (stay and rng.random() < 0.4) else arcs[int(rng.integers(len(arcs)))]
                labels.append(a[2])
                if a[3] > 0:
                    out.append(a[3])
                state = a[1]
            if g.final[state] == 0:
                break
        aligns.append(labels)
        emitted.append(out)
    lens = [len(a) for a in aligns]
    T = lens[0]
    logits = (rng.standard_normal((T, B, V)) * 0.5).astype(np.float32)
    for b, a in enumerate(aligns):
        for t, c in enumerate(a):
            logits[t, b, c] += 12
    gen = P.CTCGraphGen(context_order=3, graph_build_args={"use_context_blanks": True}, num_symbols=S)
    head = P.FSTLoss(gen, normalize_by_dim=None)
    ret = head.decode(torch.from_numpy(logits).cuda(), torch.tensor(lens))
    assert ret["decoded"] == emitted


@pytest.mark.gpu
@pytest.mark.parametrize("variant,kw", [("plain", {}), ("noloops", {"allow_nonblank_selfloops": False}),
                                        ("repctx", {"eval_repeats_in_context": True}),
                                        ("looprepeats", {"loop_using_symbol_repetitions": True})])
def test_dense_route_bicontext_variants_match_the_reference_values(P, variant, kw):
    """``ctc_raw_loss`` with the flags of ``get_CTC_matrices_bicontext`` and the loop-repeats machine
    (ctc_losses.py:111-226): the values the reference's RawGenericCTC returned (``tests/golden/dense/dense_variants.npz``,
    made by oracle/make_golden.py from the reference's own bi_CTC_loss), transcripts with doubled and tripled
    letters; gradient against the fp64 oracle on the same state machines."""
    import importlib
    import os
    d = dict(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "dense", "dense_variants.npz")))
    S = int(d["S"])
    x = torch.as_tensor(d["logits"]).cuda().requires_grad_(True)
    labels_cat = torch.as_tensor(d["labels_cat"].astype(np.int64))
    ends = np.concatenate([[0], np.cumsum(d["text_lens"])])
    d["label_seqs"] = [d["labels_cat"][ends[i]:ends[i + 1]].tolist() for i in range(len(d["text_lens"]))]
    args = dict(allow_nonblank_selfloops=True, loop_using_symbol_repetitions=False, eval_repeats_in_context=False)
    args.update(kw)
    loss = P.ctc_raw_loss(x, labels_cat, torch.tensor(d["lens"]), torch.tensor(d["text_lens"]), S, 2, 1, **args)
    loss.sum().backward()
    assert _rel(loss.detach().cpu(), d["raw_" + variant]) < LOSS_RTOL
    cl = importlib.import_module(P.FSTLoss.__module__)
    g = P.Graph.from_arcs([cl.dense_ctc_graph(q, S, 2, args["allow_nonblank_selfloops"], args["eval_repeats_in_context"],
                                              args["loop_using_symbol_repetitions"]) for q in d["label_seqs"]])
    nm = P.batch_training_graph_matrices([P.padded_matrices(g, i) for i in range(g.num_graphs)])
    want_loss, want_grad = engine.fst_loss(d["logits"], d["lens"].tolist(), [m.numpy() for m in nm], None, 1, S, 2,
                                           dtype=torch.float64)
    assert _rel(want_loss.numpy(), d["raw_" + variant]) < 1e-6
    assert np.abs(x.grad.cpu().double().numpy() - want_grad.numpy()).max() < GRAD_ATOL
    got_b = P.ctc_raw_loss_batch(x.detach(), labels_cat, torch.tensor(d["lens"]), torch.tensor(d["text_lens"]), S, 2, 1,
                                 **args)
    assert torch.equal(got_b, loss.detach())


def test_misaligned_logits_and_error_detail(P):
    """A logits view that starts off a 16-byte boundary: the Python wrapper re-aligns it (same result as the
    aligned copy); the C entry point refuses it with CTCG_EUNSUPPORTED and says why in
    ``ctcg_last_error_detail`` (the status code alone lost the reason)."""
    rng = np.random.default_rng(8)
    T, B, S = 30, 3, 7
    V = S * S
    lens = [30, 26, 19]
    seqs = [rng.integers(1, S, size=max(1, ln // 4)).tolist() for ln in lens]
    flat = torch.from_numpy(rng.standard_normal(T * B * V + 1).astype(np.float32)).cuda()
    x_off = flat[1:].view(T, B, V)                       # data_ptr % 16 == 4
    assert x_off.data_ptr() % 16 != 0
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S)
    num, den = gen.get_training_graphs(seqs), gen.get_decoding_graph()
    la, ga = P._lib.loss_fwd_bwd(x_off, lens, num, den, P.NORM_NONE, S)
    lb, gb = P._lib.loss_fwd_bwd(x_off.clone(), lens, num, den, P.NORM_NONE, S)
    assert torch.equal(la, lb) and torch.equal(ga, gb)
    L = P._lib.lib()
    need = L.ctcg_workspace_bytes(T, B, V, num.ptr, den.ptr)
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    loss, grad = torch.empty(B, device="cuda"), torch.empty(T, B, V, device="cuda")
    lens_np = np.asarray(lens, dtype=np.int32)
    st = L.ctcg_loss_fwd_bwd(x_off.data_ptr(), T, B, V, lens_np.ctypes.data, num.ptr, den.ptr, 0, S, None,
                             loss.data_ptr(), grad.data_ptr(), ws.data_ptr(), need, None)
    assert st == 8                                        # CTCG_EUNSUPPORTED
    assert b"16-byte aligned" in L.ctcg_last_error_detail()


def test_repeated_identical_calls_replay_a_captured_step(P):
    """``ctcg_loss_fwd_bwd`` replays a CUDA graph for a call whose every argument equals an earlier one's (captured
    on the third sighting).  The logits' CONTENTS are read anew, the kernel count is the same as on the ordinary
    path, other lengths / graphs / buffers never match, and the results are bit-identical to the ordinary path
    (the numerator's posteriors carry no atomics; B >= 64 so that the length groups and their streams are in it)."""
    rng = np.random.default_rng(11)
    S, T, B = 49, 60, 64
    V = S * S
    lens = sorted([T] + rng.integers(int(0.6 * T), T + 1, size=B - 1).tolist(), reverse=True)
    seqs = [rng.integers(1, S, size=max(1, ln // 4)).tolist() for ln in lens]
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S)
    num, den = gen.get_training_graphs(seqs), gen.get_decoding_graph()
    a = torch.from_numpy((rng.standard_normal((T, B, V)) * 3).astype(np.float32)).cuda()
    b = torch.from_numpy((rng.standard_normal((T, B, V)) * 3).astype(np.float32)).cuda()
    L = P._lib.lib()
    num.upload(a.device), den.upload(a.device)
    need = L.ctcg_workspace_bytes(T, B, V, num.ptr, den.ptr)
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    x = a.clone()
    loss, grad = torch.empty(B, device="cuda"), torch.empty(T, B, V, device="cuda")
    lens_np = np.asarray(lens, dtype=np.int32)

    def call(xt, lens_arr, lo, gr):
        st = L.ctcg_loss_fwd_bwd(xt.data_ptr(), T, B, V, lens_arr.ctypes.data, num.ptr, den.ptr, 0, S, None,
                                 lo.data_ptr(), gr.data_ptr(), ws.data_ptr(), need, None)
        assert st == 0
        torch.cuda.synchronize()
        return lo.clone(), gr.clone()

    n0 = P._lib.launch_count()
    i0 = P._lib.step_graph_info()
    la1, ga1 = call(x, lens_np, loss, grad)              # first sighting: ordinary path
    per_call = P._lib.launch_count() - n0
    call(x, lens_np, loss, grad)                         # second: ordinary
    la2, ga2 = call(x, lens_np, loss, grad)              # third: captured, then replayed
    la3, ga3 = call(x, lens_np, loss, grad)              # fourth: replay
    assert P._lib.launch_count() - n0 == 4 * per_call
    i1 = P._lib.step_graph_info()
    assert (i1[0] - i0[0], i1[1] - i0[1], i1[2] - i0[2]) == (2, 1, 2)
    assert torch.equal(la1, la2) and torch.equal(la1, la3) and torch.equal(ga1, ga2) and torch.equal(ga1, ga3)
    x.copy_(b)                                           # same buffers, new contents
    lb_replay, gb_replay = call(x, lens_np, loss, grad)
    other_loss, other_grad = torch.empty(B, device="cuda"), torch.empty(T, B, V, device="cuda")
    lb, gb = call(b, lens_np, other_loss, other_grad)    # other buffers: never seen, ordinary path
    assert torch.equal(lb_replay, lb) and torch.equal(gb_replay, gb)
    assert not torch.equal(lb, la1)
    # other lengths in the same buffers: a new signature, and the old one still replays afterwards
    lens2 = np.asarray([T] + [max(1, v - 1) for v in lens[1:]], dtype=np.int32)
    lc, gc = call(x, lens2, loss, grad)
    want_c = P._lib.loss_fwd_bwd(b, lens2.tolist(), num, den, P.NORM_NONE, S)
    assert torch.equal(lc, want_c[0]) and torch.equal(gc, want_c[1])
    ld, gd = call(x, lens_np, loss, grad)
    assert torch.equal(ld, lb) and torch.equal(gd, gb)


def test_captured_step_cache_under_many_interleaved_signatures(P):
    """More live signatures than the cache has entries, visited in a random order often enough for captures and
    evictions to happen: every call must return what the ordinary path returns for the same arguments
    (reference values computed once per signature with the cache bypassed through fresh output buffers)."""
    rng = np.random.default_rng(2024)
    S, T, B = 12, 24, 16
    V = S * S
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S)
    den = gen.get_decoding_graph()
    L = P._lib.lib()
    sigs = []
    for k in range(20):                                   # 20 signatures > 16 entries
        lens = sorted([T] + rng.integers(T // 2, T + 1, size=B - 1).tolist(), reverse=True)
        seqs = [rng.integers(1, S, size=max(1, ln // 4)).tolist() for ln in lens]
        num = gen.get_training_graphs(seqs)
        x = torch.from_numpy((rng.standard_normal((T, B, V)) * 2).astype(np.float32)).cuda()
        num.upload(x.device), den.upload(x.device)
        need = L.ctcg_workspace_bytes(T, B, V, num.ptr, den.ptr)
        ws = torch.empty(need, dtype=torch.uint8, device="cuda")
        loss, grad = torch.empty(B, device="cuda"), torch.empty(T, B, V, device="cuda")
        want = P._lib.loss_fwd_bwd(x, lens, num, den, P.NORM_NONE, S)          # wrapper: its own buffers, first sighting
        sigs.append(dict(x=x, lens=np.asarray(lens, dtype=np.int32), num=num, ws=ws, need=need, loss=loss, grad=grad,
                         want=want))
    i0 = P._lib.step_graph_info()
    order = rng.integers(0, len(sigs), size=160).tolist() + [3] * 6 + [17] * 6
    for k in order:
        s = sigs[k]
        s["loss"].fill_(-7.0), s["grad"].fill_(-7.0)
        st = L.ctcg_loss_fwd_bwd(s["x"].data_ptr(), T, B, V, s["lens"].ctypes.data, s["num"].ptr, den.ptr, 0, S, None,
                                 s["loss"].data_ptr(), s["grad"].data_ptr(), s["ws"].data_ptr(), s["need"], None)
        assert st == 0
        assert torch.equal(s["loss"], s["want"][0]) and torch.equal(s["grad"], s["want"][1]), k
    i1 = P._lib.step_graph_info()
    assert i1[1] - i0[1] >= 2 and i1[0] - i0[0] >= 8      # captures and replays did happen


def _sweep_cases():
    """24 seeded random configurations over the flags, sizes and corner shapes (B=1, T=1, S at the limits of the
    factorised kernels' templates, odd V, ragged lengths down to one frame, per-utterance weights)."""
    rng = np.random.default_rng(20260)
    flagsets = [{}, {"use_contextual_blanks": True}, {"allow_nonblank_selfloops": False},
                {"loop_using_symbol_repetitions": True}, {"use_contextual_blanks": True, "allow_nonblank_selfloops": False}]
    cases = []
    for i in range(24):
        order = 2 if i % 4 else 1
        S = int(rng.choice([4, 5, 7, 9, 13, 16, 23, 36, 49] if order == 2 else [3, 8, 30, 49]))
        T = int(rng.choice([1, 2, 7, 8, 9, 16, 17, 31, 64, 65, 100]))
        B = int(rng.choice([1, 2, 3, 8, 17]))
        args = dict(flagsets[int(rng.integers(len(flagsets)))]) if order == 2 else {}
        use_den = bool(rng.integers(2))
        nbd = None if use_den and rng.random() < 0.7 else [None, 0, 1][int(rng.integers(3))]
        if order == 1 and nbd == 1:
            nbd = 0
        cases.append((i, order, S, args, use_den, nbd, T, B))
    return cases


@pytest.mark.parametrize("i,order,S,args,use_den,nbd,T,B", _sweep_cases())
def test_random_sweep_of_shapes_and_flags_matches_fp64_oracle(P, i, order, S, args, use_den, nbd, T, B):
    rng = np.random.default_rng(5000 + i)
    V = S ** order
    lens = sorted([T] + rng.integers(1, T + 1, size=B - 1).tolist(), reverse=True)      # ragged, down to one frame
    seqs = [rng.integers(1, S, size=max(1, min(ln, rng.integers(1, max(2, ln // 2 + 1))))).tolist() for ln in lens]
    logits = (rng.standard_normal((T, B, V)) * 3).astype(np.float32)
    want_loss, want_grad = _oracle(order, S, args, logits, lens, seqs, use_den, nbd)
    gen = P.CTCGraphGen(context_order=order, graph_build_args=args, num_symbols=S)
    num = gen.get_training_graphs(seqs)
    den = gen.get_decoding_graph() if use_den else None
    w = torch.from_numpy(rng.uniform(-2, 2, size=B).astype(np.float32))
    x = torch.from_numpy(logits).cuda().requires_grad_(True)
    loss = P.fst_loss(x, lens, num, den, normalize_by_dim=nbd, num_symbols=S)
    (loss * w.cuda()).sum().backward()
    got, wl = loss.detach().cpu().double().numpy(), want_loss.numpy()
    feasible = np.abs(wl) < 1e19                           # T too short for the transcript: ~1e20 in both, zero gradient
    assert (np.abs(got[~feasible]) > 1e19).all()
    if feasible.any():
        assert np.abs(got[feasible] - wl[feasible]).max() <= LOSS_RTOL * max(1.0, np.abs(wl[feasible]).max())
    wg = want_grad.numpy() * w.numpy()[None, :, None]
    wg[:, ~feasible] = 0
    assert np.abs(x.grad.cpu().double().numpy() - wg).max() < 2 * GRAD_ATOL * max(1.0, float(w.abs().max()))


@pytest.mark.parametrize("n_states,deg,uniform", [(600, 20, False), (700, 18, True), (300, 24, False)])
def test_generic_kernels_on_wide_irregular_weighted_graphs(P, n_states, deg, uniform):
    """The generic CSR kernels on a random wide graph with weights and irregular in-degree, imported arc by arc:
    >= 512 states take the transposed arc tables (per-arc posteriors when the arcs into a state carry different
    labels, per-state occupancy when they share one), fewer the warp-per-state shape.  Path sum, gradient, Viterbi
    score and frame labels against the fp64 oracle engine on the exported matrices; the loss with this graph as
    the denominator of a small CTC numerator as well."""
    rng = np.random.default_rng(n_states + deg)
    V = 64
    arcs = []
    state_label = rng.integers(0, V, size=n_states)
    for s in range(n_states):
        if uniform:                                   # label = the target's: arcs into a state share it; targets with
            dst, seen = [], set()                     # distinct labels keep the graph input-deterministic
            for d in rng.permutation(n_states).tolist():
                if state_label[d] not in seen:
                    seen.add(int(state_label[d]))
                    dst.append(d)
                if len(dst) == deg:
                    break
            labs = [int(state_label[d]) for d in dst]
        else:
            dst = rng.choice(n_states, size=deg, replace=False).tolist()
            labs = rng.choice(V, size=deg, replace=False).tolist()
        for d, l in zip(dst, labs):
            arcs.append((s, int(d), int(l), float(-rng.uniform(0, 3))))
    final = np.where(rng.random(n_states) < 0.5, -rng.uniform(0, 2, size=n_states), -np.inf).astype(np.float32)
    final[0] = 0.0
    g = P.Graph.from_arcs([(n_states, arcs, final)])
    info = g.info()
    assert info.total_arcs >= 16 * n_states and info.structured == 0
    T, B = 23, 3
    lens = [23, 17, 9]
    x = (rng.standard_normal((T, B, V)) * 2).astype(np.float32)
    mats = [m.unsqueeze(0).numpy() for m in P.padded_matrices(g, 0)]
    want_cost, want_grad, _ = engine.path_logsumexp(torch.as_tensor(x), lens, mats, dtype=torch.float64)
    cost, grad = P._lib.path_sum(torch.from_numpy(x).cuda(), lens, g)
    assert _rel(cost.cpu().numpy(), want_cost.numpy()) < LOSS_RTOL
    assert np.abs(grad.cpu().double().numpy() - want_grad.numpy()).max() < GRAD_ATOL
    want_score, want_labels = engine.path_viterbi(torch.as_tensor(x), lens, mats)
    score, labels = P._lib.viterbi(torch.from_numpy(x).cuda(), lens, g)
    assert _rel(score.cpu().numpy(), np.asarray(want_score)) < 1e-6
    for b, ln in enumerate(lens):
        assert np.array_equal(labels.cpu().numpy()[:ln, b], np.asarray(want_labels)[:ln, b])
