"""Row f2 (projection fused into the loss), the measurement behind DESIGN section 8: what the output projection
fc(encoded) of cfg3 costs on this GPU at the reference's precision (fp32 SIMT) and on the tensor cores (TF32,
BF16), and what those precisions do to the logits and to the CTC-G gradient (bar: 1e-5 absolute vs fp64).
Library GEMMs through torch (cuBLAS); nothing here is on the product path.
    python profiles/micro/f2_gemm.py > gpurun_out/f2_gemm.json"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import importlib
P = importlib.import_module("pytorch-asr_b200")

T, B, H, S = 400, 64, 640, 49
V = S * S
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(1)
enc = torch.randn(T * B, H, generator=g, device=dev)
W = torch.randn(V, H, generator=g, device=dev) * (3.0 / H ** 0.5)      # logits ~ N(0, 3^2) like the bench input
bias = torch.randn(V, generator=g, device=dev) * 0.1


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


out = {"shape": {"T": T, "B": B, "H": H, "V": V}, "gflop_forward": 2 * T * B * H * V / 1e9}
ref64 = torch.addmm(bias.double(), enc.double(), W.double().t())
res = {}
torch.backends.cuda.matmul.allow_tf32 = False
res["fp32"] = (timed(lambda: torch.addmm(bias, enc, W.t())), torch.addmm(bias, enc, W.t()))
torch.backends.cuda.matmul.allow_tf32 = True
res["tf32"] = (timed(lambda: torch.addmm(bias, enc, W.t())), torch.addmm(bias, enc, W.t()))
torch.backends.cuda.matmul.allow_tf32 = False
eb, Wb = enc.bfloat16(), W.bfloat16()
res["bf16"] = (timed(lambda: torch.addmm(bias.bfloat16(), eb, Wb.t())), torch.addmm(bias.bfloat16(), eb, Wb.t()).float())
# 3xTF32: hi/lo split of both operands, three tensor-core products (lo*lo dropped), fp32 accumulate
torch.backends.cuda.matmul.allow_tf32 = True


def split(x):
    hi = (x.view(torch.int32) & -8192).view(torch.float32)      # keep 10 mantissa bits (TF32)
    return hi, x - hi


eh, el = split(enc)
Wh, Wl = split(W)
f3 = lambda: torch.addmm(bias, eh, Wh.t()) + (eh @ Wl.t() + el @ Wh.t())   # noqa: E731
res["3xtf32"] = (timed(f3), f3())
torch.backends.cuda.matmul.allow_tf32 = False

# the CTC-G gradient on each version of the logits, against the gradient on the fp64 logits rounded to fp32
rng = np.random.default_rng(3)
lens = sorted([T] + rng.integers(int(0.6 * T), T + 1, size=B - 1).tolist(), reverse=True)
seqs = [rng.integers(1, S, size=max(1, ln // 4)).tolist() for ln in lens]
gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S)
num, den = gen.get_training_graphs(seqs), gen.get_decoding_graph()
l0, g0 = P._lib.loss_fwd_bwd(ref64.float().view(T, B, V).contiguous(), lens, num, den, P.NORM_NONE, S)
for k, (ms, y) in res.items():
    l, gr = P._lib.loss_fwd_bwd(y.view(T, B, V).contiguous(), lens, num, den, P.NORM_NONE, S)
    out[k] = {"ms_forward": ms, "tflops": out["gflop_forward"] / ms,
              "max_abs_logit_err_vs_fp64": float((y.double() - ref64).abs().max()),
              "max_abs_grad_diff": float((gr - g0).abs().max()),
              "max_rel_loss_diff": float(((l - l0).abs() / l0.abs()).max())}
out["loss_step_ms_for_scale"] = timed(lambda: P._lib.loss_fwd_bwd(res["fp32"][1].view(T, B, V), lens, num, den, P.NORM_NONE, S))
print(json.dumps(out))
