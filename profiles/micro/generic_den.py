"""The generic CSR kernel on wide shared graphs (tens of arcs per state): bi-char decoding graph forced off the
factorised kernels, at the cfg3 shape.  Run twice: CTCG_PATH_NARROW=1 (one state per thread, round 1) and default
(one state per warp, 1024 threads).   python profiles/micro/generic_den.py > gpurun_out/generic_den_<tag>.json"""
import importlib
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
P = importlib.import_module("pytorch-asr_b200")

out = {"narrow": os.environ.get("CTCG_PATH_NARROW", "0")}
for name, order, S, T, B in (("bichar_S49_cfg3", 2, 49, 400, 64), ("bichar_S49_B8", 2, 49, 400, 8), ("mono_S49", 1, 49, 400, 64)):
    rng = np.random.default_rng(1)
    V = S ** order
    lens = sorted([T] + rng.integers(int(0.6 * T), T + 1, size=B - 1).tolist(), reverse=True)
    seqs = [rng.integers(1, S, size=max(1, ln // 4)).tolist() for ln in lens]
    x = torch.from_numpy((rng.standard_normal((T, B, V)) * 3).astype(np.float32)).cuda()
    gen = P.CTCGraphGen(context_order=order, graph_build_args={}, num_symbols=S)
    num = gen.get_training_graphs(seqs)
    den = P.Graph.decoding(order, S)
    ref = P._lib.loss_fwd_bwd(x, lens, num, den, P.NORM_NONE, S) if order == 2 else None
    den.set_generic()
    for _ in range(2):
        got = P._lib.loss_fwd_bwd(x, lens, num, den, P.NORM_NONE, S)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 3
    e0.record()
    for _ in range(n):
        got = P._lib.loss_fwd_bwd(x, lens, num, den, P.NORM_NONE, S)
    e1.record()
    torch.cuda.synchronize()
    rec = {"ms_per_step": e0.elapsed_time(e1) / n, "frames": int(sum(lens))}
    if ref is not None:
        rec["max_rel_loss_diff_vs_factorised"] = float(((got[0] - ref[0]).abs() / ref[0].abs()).max())
        rec["max_abs_grad_diff_vs_factorised"] = float((got[1] - ref[1]).abs().max())
    out[name] = rec
print(json.dumps(out))
