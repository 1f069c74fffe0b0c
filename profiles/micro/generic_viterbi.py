"""One generic-CSR Viterbi on the bi-char decoding graph (S=49, T=400, B=4) forced off the factorised kernel:
the launch profiled by `ncu -k regex:viterbi_kernel` to see what bounds the wide generic kernels."""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
P = importlib.import_module("pytorch-asr_b200")
S, T, B = 49, 400, 4
rng = np.random.default_rng(0)
x = torch.from_numpy((rng.standard_normal((T, B, S * S)) * 3).astype(np.float32)).cuda()
g = P.Graph.decoding(2, S).set_generic()
for _ in range(2):
    P._lib.viterbi(x, [T] * B, g)
torch.cuda.synchronize()
