"""Which half of the CTC-G(B) gradient carries the error at long T: numerator-only (plain CTC through the
same numerator kernels) and denominator-only (path_sum on the decoding graph) against the fp64 oracle."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import pytorch_asr_b200 as P  # noqa: E402
from oracle import engine, graphs  # noqa: E402
from test_gpu_parity import _oracle, _random_case  # noqa: E402

for T in (400, 800, 1600):
    for peaky in (False, True):
        args = {"use_contextual_blanks": True}
        S, B = 49, 2
        logits, lens, seqs = _random_case(2, S, args, T, B, seed=1003, peaky=peaky)
        gen = P.CTCGraphGen(context_order=2, graph_build_args=args, num_symbols=S)
        ogen = graphs.GraphGen(2, S, args)
        x = torch.from_numpy(logits).cuda()
        # denominator only
        c, g = P._lib.path_sum(x, lens, gen.get_decoding_graph())
        wc, wg, _ = engine.path_logsumexp(torch.from_numpy(logits), lens, ogen.get_decoding_matrices(), dtype=torch.float64)
        den_err = (g.cpu().double() - wg).abs().max().item()
        den_rel = ((c.cpu().double() - wc).abs() / wc.abs()).max().item()
        # numerator only: plain CTC = softmax - gamma_num
        l, gr = P._lib.loss_fwd_bwd(x, lens, gen.get_training_graphs(seqs), None, P.NORM_SOFTMAX)
        wl, wgr = _oracle(2, S, args, logits, lens, seqs, False, 0)
        num_err = (gr.cpu().double() - wgr).abs().max().item()
        num_rel = ((l.cpu().double() - wl).abs() / wl.abs()).max().item()
        print("T=%d peaky=%d  den: grad %.2e cost rel %.2e   num(ctc): grad %.2e loss rel %.2e"
              % (T, peaky, den_err, den_rel, num_err, num_rel), flush=True)
