"""The factorised bi-char recurrences the CUDA sweeps implement (csrc/kernels_bisweep.cu), stated in
numpy (oracle/bichar_model.py), against the reference algorithm (PathLogSumExp on the matrices of
build_ctc_bigram_decoding_fst, fst_utils.py:400-488, 759-817) as restated in oracle/engine.py.
Runs on the CPU: it pins the mathematics of the split-sweep design (column sums P, row sums Q, folded
blank states, lagged power-of-two rescaling, K-frame checkpoints, combine formula) independently of
any GPU."""
import numpy as np
import pytest
import torch

from oracle import bichar_model as bm
from oracle import engine, graphs


@pytest.mark.parametrize("S,args,T,scale", [
    (6, {}, 40, 3.0),
    (7, {"use_contextual_blanks": True}, 33, 3.0),
    (5, {"allow_nonblank_selfloops": False}, 20, 3.0),
    (6, {"loop_using_symbol_repetitions": True}, 25, 3.0),
    (6, {}, 120, 12.0),          # wide logit spread: the power-of-two rescale fires
])
@pytest.mark.parametrize("K", [4, 8])
def test_split_sweep_model_matches_reference_algorithm(S, args, T, scale, K):
    rng = np.random.default_rng(S * 100 + T)
    x = (rng.standard_normal((T, 1, S * S)) * scale).astype(np.float32)
    dm = graphs.GraphGen(2, S, args).get_decoding_matrices()
    cost, grad, _ = engine.path_logsumexp(torch.from_numpy(x), [T], dm, dtype=torch.float64)
    c_fwd, gamma, c_bwd = bm.denominator(x[:, 0], S, K=K, **args)
    want = float(cost[0])
    assert abs(c_fwd - want) <= 1e-5 * abs(want)          # loss tolerance of the GPU tests
    assert abs(c_bwd - want) <= 1e-5 * abs(want)          # the backward sweep ends in the same total
    assert np.abs(gamma - grad[:, 0].numpy()).max() < 1e-5
    assert np.abs(gamma.sum(-1) - 1).max() < 1e-5         # one unit of posterior mass per frame


def test_rescale_is_exercised():
    rng = np.random.default_rng(3)
    x = (rng.standard_normal((150, 36)) * 12).astype(np.float32)
    fw = bm.sweep(x, 6, True, False, 8, False)
    bw = bm.sweep(x, 6, True, False, 8, True)
    assert np.count_nonzero(fw["k"]) > 0 and np.count_nonzero(bw["k"]) > 0
