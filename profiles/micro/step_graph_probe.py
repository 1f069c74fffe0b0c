"""How the captured-step cache behaves under the bench's two loops (public API with autograd, C entry point direct):
replays / captures / ordinary calls per phase, and the GPU time per step once every call is a replay."""
import importlib, json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
P = importlib.import_module("pytorch-asr_b200")
S, T, B = 49, 400, 64
V = S * S
rng = np.random.default_rng(1)
lens = sorted([T] + rng.integers(int(0.6 * T), T + 1, size=B - 1).tolist(), reverse=True)
seqs = [rng.integers(1, S, size=max(1, ln // 4)).tolist() for ln in lens]
gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S)
num, den = gen.get_training_graphs(seqs), gen.get_decoding_graph()
xs = [(torch.randn(T, B, V, device="cuda") * 3).requires_grad_(True) for _ in range(3)]
out = {}


def api(i):
    x = xs[i % 3]
    x.grad = None
    P.fst_loss(x, lens, num, den, num_symbols=S).sum().backward()


def cabi(i):
    P._lib.loss_fwd_bwd(xs[i % 3].detach(), lens, num, den, P.NORM_NONE, S)


for name, fn in (("public_api", api), ("c_abi", cabi)):
    rec = []
    i = 0
    for phase in range(4):
        a0 = P._lib.step_graph_info()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            fn(i)
            i += 1
        e1.record()
        torch.cuda.synchronize()
        a1 = P._lib.step_graph_info()
        rec.append({"ms_per_step": e0.elapsed_time(e1) / 20, "replays": a1[0] - a0[0], "captures": a1[1] - a0[1], "ordinary": a1[2] - a0[2]})
    out[name] = rec
print(json.dumps(out))
