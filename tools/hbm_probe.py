"""HBM probe (GPU box): achievable pure-read, pure-write and copy bandwidth with library kernels, to put the
roofline fractions of the streaming kernels in context."""
import torch
n = 1 << 29   # 512 Mi bf16 = 1 GiB
x = torch.ones(n, dtype=torch.bfloat16, device="cuda")
y = torch.empty_like(x)
def t(f, reps=10):
    f(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
gb = n * 2 / 1e9
print("copy  (r+w) %.0f GB/s" % (2 * gb / t(lambda: y.copy_(x)) * 1e3))
print("write (fill) %.0f GB/s" % (gb / t(lambda: y.fill_(1.0)) * 1e3))
print("read  (sum)  %.0f GB/s" % (gb / t(lambda: x.view(torch.int16).sum()) * 1e3))
xf = x.view(torch.float32)
print("read  (f32 max) %.0f GB/s" % (gb / t(lambda: xf.max()) * 1e3))
z = torch.empty(n // 4, dtype=torch.bfloat16, device="cuda")
print("r1:w4 (repeat_interleave-like expand copy) %.0f GB/s" % ((gb + gb / 4) / t(lambda: y.view(4, -1).copy_(z.view(1, -1).expand(4, -1))) * 1e3))
