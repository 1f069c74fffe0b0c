"""``ctcg_viterbi_static`` -- the entry point that allocates nothing (its scratch is a device array
built into the library; no counterpart in the reference)
-- against the existing ``ctcg_viterbi``, through the C ABI, on two calls in a row.

Kept in a file of its own that pytest collects last.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

LOSS_RTOL = 1e-5   # the tolerance of tests/test_gpu_parity.py; frame labels are compared exactly


def _case(S, T, B, seed):
    rng = np.random.default_rng(seed)
    lens = sorted([T] + rng.integers(int(0.6 * T), T + 1, size=B - 1).tolist(), reverse=True)
    logits = (rng.standard_normal((T, B, S * S)) * 3).astype(np.float32)
    return torch.from_numpy(logits).cuda(), lens


def _rel(a, b):
    a, b = a.double().cpu().numpy(), b.double().cpu().numpy()
    return float((np.abs(a - b) / np.maximum(np.abs(b), 1.0)).max())


def test_viterbi_static_matches_viterbi_on_two_calls_in_a_row(lib_built):
    P = lib_built
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    dev = torch.device("cuda", torch.cuda.current_device())
    # two different problems back to back: the second one is smaller and finds the scratch full of the
    # first one's back-pointers; then the first again (nothing may depend on what the array held)
    gens = {S: P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S) for S in (12, 7)}
    for S, T, B, seed in [(12, 90, 5, 11), (7, 40, 3, 12), (12, 90, 5, 11)]:
        graph = gens[S].get_decoding_graph()
        x, lens = _case(S, T, B, seed)
        want_score, want_labels = P._lib.viterbi(x, lens, graph)       # uploads the graph (set-up)
        # everything the call needs exists before it: inputs, outputs, lengths on the device
        lens_dev = torch.tensor(lens, dtype=torch.int32, device=dev)
        score = torch.full((B,), float("nan"), dtype=torch.float32, device=dev)
        labels = torch.full((T, B), -7, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        P._lib.viterbi_static(x, lens_dev, graph, score, labels)
        assert torch.cuda.memory_allocated(dev) == before              # the wrapper made no torch allocation either
        torch.cuda.synchronize()
        assert _rel(score, want_score) < LOSS_RTOL
        assert torch.equal(labels, want_labels)


def test_viterbi_static_clamps_lengths_and_rejects_oversize(lib_built):
    P = lib_built
    dev = torch.device("cuda", torch.cuda.current_device())
    S, T, B = 7, 30, 2
    graph = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S).get_decoding_graph()
    x, lens = _case(S, T, B, 5)
    want_score, want_labels = P._lib.viterbi(x, [T, 0], graph)
    score = torch.empty(B, dtype=torch.float32, device=dev)
    labels = torch.empty((T, B), dtype=torch.int32, device=dev)
    # lengths outside [0, T] are clamped on the device (the host never sees them)
    P._lib.viterbi_static(x, torch.tensor([T + 9, -3], dtype=torch.int32, device=dev), graph, score, labels)
    torch.cuda.synchronize()
    assert _rel(score, want_score) < LOSS_RTOL and torch.equal(labels, want_labels)
    # beyond the built-in scratch: a clear error and nothing launched.  The size check comes before any
    # access, so a small tensor viewed with a huge T is enough (expand: no memory behind it)
    row = (S * S + 3) // 4 * 4
    T_big = P._lib.STATIC_SCRATCH_BYTES // (4 * B * row) + 1
    launches = P._lib.launch_count()
    status = P._lib.lib().ctcg_viterbi_static(x.data_ptr(), T_big, B, S * S, torch.zeros(B, dtype=torch.int32, device=dev).data_ptr(),
                                              graph.ptr, score.data_ptr(), labels.data_ptr(), None)
    assert status == 10                                                 # CTCG_ESCRATCH
    assert b"scratch" in P._lib.lib().ctcg_strerror(status)
    assert P._lib.launch_count() == launches
