"""Forward-only (cost) and full (cost + grad) time of the bi-char kernel under CTCG_BICHAR_DEBUG masks."""
import os, sys, subprocess
here = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1:
    sys.path.insert(0, here)
    import torch, bench
    import pytorch_asr_b200 as P
    batch = bench.make_batch("cfg3")
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=batch["S"])
    den = gen.get_decoding_graph()
    x = (torch.randn(batch["T"], batch["B"], batch["V"]) * 3).cuda()
    out = []
    for need_grad in (False, True):
        for _ in range(3):
            P._lib.path_sum(x, batch["lens"], den, need_grad=need_grad)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            P._lib.path_sum(x, batch["lens"], den, need_grad=need_grad)
        e1.record(); torch.cuda.synchronize()
        out.append(e0.elapsed_time(e1) / 20)
    print("fwd-only %.4f ms   fwd+bwd %.4f ms" % tuple(out))
else:
    for mask in (0, 8, 4, 12, 32, 44):  # never mask the TMA issue: its consumers would wait for ever
        env = dict(os.environ, CTCG_BICHAR_DEBUG=str(mask))
        r = subprocess.run([sys.executable, __file__, "child"], capture_output=True, text=True, env=env)
        print("mask %3d  %s" % (mask, r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-200:]), flush=True)
