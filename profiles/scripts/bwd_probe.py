"""Where the backward of the fused loss spends its time: the multiply, the autograd accumulation. Measurement script."""
import torch
dev = torch.device("cuda", 0)
n = 384 * 10 * 232 * 400
g = torch.rand(n, device=dev); s = torch.ones((), device=dev)
def ev(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
print("g * s (new tensor)      ", ev(lambda: g * s))
print("g.mul_(s) in place      ", ev(lambda: g.mul_(s)))
print("g.clone()               ", ev(lambda: g.clone()))
class F(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(g); return x.sum() * 0
    @staticmethod
    def backward(ctx, go):
        (gg,) = ctx.saved_tensors
        return gg * go
class F2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(g); return x.sum() * 0
    @staticmethod
    def backward(ctx, go):
        (gg,) = ctx.saved_tensors
        return gg.mul_(go)
x = torch.rand(n, device=dev, requires_grad=True)
for Fn in (F, F2):
    def run():
        x.grad = None
        y = Fn.apply(x)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); y.backward(); b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b)
    ts = [run() for _ in range(6)]
    print(Fn.__name__, "backward ms", ts[-3:])
