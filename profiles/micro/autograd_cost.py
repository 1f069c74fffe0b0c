"""Host cost of `f(x).sum().backward()` for a trivial custom autograd Function on CUDA tensors: what the public-API
call pays on top of the C entry point regardless of what the Function does."""
import json, time, torch


class F(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, loss, grad):
        ctx.g = grad
        return loss.clone()

    @staticmethod
    def backward(ctx, go):
        return ctx.g, None, None


x = torch.zeros(400, 64, 2401, device="cuda", requires_grad=True)
loss = torch.zeros(64, device="cuda")
grad = torch.zeros_like(x)


def call():
    x.grad = None
    F.apply(x, loss, grad).sum().backward()


def fwd_only():
    F.apply(x, loss, grad).sum()


out = {}
for name, fn in (("apply_sum_backward", call), ("apply_sum", fwd_only)):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    n = 200
    for i in range(n):
        t0 = time.perf_counter()
        fn()
        tot += time.perf_counter() - t0
        if i % 4 == 3:
            torch.cuda.synchronize()
    out[name + "_host_ms"] = 1e3 * tot / n
print(json.dumps(out))
