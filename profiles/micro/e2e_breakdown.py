"""Where the end-to-end step of bench.py goes (cfg3): H2D, graph compile+upload, loss, backward, D2H."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import bench
import pytorch_asr_b200 as P

batch = bench.make_batch("cfg3")
T, B, V, S = batch["T"], batch["B"], batch["V"], batch["S"]
dev = torch.device("cuda", 0)
gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S)
den = gen.get_decoding_graph()
host = (torch.randn(T, B, V) * 3).pin_memory()
lens = batch["lens"]


def timed(fn, n=10):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        r = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3, r


ms, x = timed(lambda: host.to(dev, non_blocking=True))
print("H2D 246 MB pinned           %.2f ms  (%.1f GB/s)" % (ms, host.numel() * 4 / ms / 1e6))
ms, num = timed(lambda: gen.get_training_graphs(batch["seqs"]))
print("graph compile (host)        %.2f ms" % ms)
ms, _ = timed(lambda: gen.get_training_graphs(batch["seqs"]).upload(dev))
print("graph compile + upload      %.2f ms" % ms)
xg = x.clone().requires_grad_(True)
def fwd():
    return P.fst_loss(xg, lens, num, den, normalize_by_dim=None, num_symbols=S)
ms, loss = timed(fwd)
print("fst_loss forward (fused)    %.2f ms" % ms)
def fb():
    xg.grad = None
    l = P.fst_loss(xg, lens, num, den, normalize_by_dim=None, num_symbols=S).sum()
    l.backward()
    return l
ms, l = timed(fb)
print("forward + backward          %.2f ms" % ms)
ms, _ = timed(lambda: float(fb().item()))
print("forward + backward + item   %.2f ms" % ms)
ms, _ = timed(lambda: torch.empty_like(x))
print("empty_like(x)               %.3f ms" % ms)
