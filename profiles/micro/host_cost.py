"""Host time to ENQUEUE one step (no synchronisation inside the loop; the GPU queue is kept short by syncing every
16 steps outside the timed sections): C entry point alone, with 8 / 1 length groups, and the public API call."""
import importlib, json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
P = importlib.import_module("pytorch-asr_b200")
S, T, B = 49, 400, 64
V = S * S
rng = np.random.default_rng(1)
lens = sorted([T] + rng.integers(int(0.6 * T), T + 1, size=B - 1).tolist(), reverse=True)
seqs = [rng.integers(1, S, size=max(1, ln // 4)).tolist() for ln in lens]
x = torch.from_numpy((rng.standard_normal((T, B, V)) * 3).astype(np.float32)).cuda()
gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=S)
num, den = gen.get_training_graphs(seqs), gen.get_decoding_graph()
SYNC_EVERY = int(os.environ.get("SYNC_EVERY", "8"))
out = {"sync_every": SYNC_EVERY, "groups_env": os.environ.get("CTCG_GROUPS"), "serial_env": os.environ.get("CTCG_SERIAL")}


def host_ms(fn, n=64):
    for _ in range(8):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for i in range(n):
        t0 = time.perf_counter()
        fn()
        tot += time.perf_counter() - t0
        if i % SYNC_EVERY == SYNC_EVERY - 1:
            torch.cuda.synchronize()
    return 1e3 * tot / n


out["c_abi_host_ms"] = host_ms(lambda: P._lib.loss_fwd_bwd(x, lens, num, den, P.NORM_NONE, S))
xr = x.clone().requires_grad_(True)


def api():
    xr.grad = None
    P.fst_loss(xr, lens, num, den, num_symbols=S).sum().backward()


out["public_api_host_ms"] = host_ms(api)
L = P._lib.lib()
need = L.ctcg_workspace_bytes(T, B, V, num.ptr, den.ptr)
ws = torch.empty(need, dtype=torch.uint8, device="cuda")
lo, gr = torch.empty(B, device="cuda"), torch.empty(T, B, V, device="cuda")
lens_np = np.asarray(lens, dtype=np.int32)
args = (x.data_ptr(), T, B, V, lens_np.ctypes.data, num.ptr, den.ptr, 0, S, None, lo.data_ptr(), gr.data_ptr(), ws.data_ptr(), need, None)
out["raw_ctypes_call_host_ms"] = host_ms(lambda: L.ctcg_loss_fwd_bwd(*args))
print(json.dumps(out))
