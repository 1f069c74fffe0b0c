"""One cfg3 loss+grad call with CTCG_BICHAR_TRACE set (needs a CTCG_TRACE_BUILD=1 library)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import bench
import pytorch_asr_b200 as P

out = sys.argv[1]
batch = bench.make_batch("cfg3")
gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=batch["S"])
num = gen.get_training_graphs(batch["seqs"]); den = gen.get_decoding_graph()
x = (torch.randn(batch["T"], batch["B"], batch["V"]) * 3).cuda()
for _ in range(3):
    P._lib.loss_fwd_bwd(x, batch["lens"], num, den, P.NORM_NONE, batch["S"])
torch.cuda.synchronize()
os.environ["CTCG_BICHAR_TRACE"] = out
P._lib.loss_fwd_bwd(x, batch["lens"], num, den, P.NORM_NONE, batch["S"])
torch.cuda.synchronize()
print("trace written", os.path.getsize(out))
