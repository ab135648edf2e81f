"""HBM micro-benchmarks on the GPU box: fill (write-only), copy, read-only reduce."""
import torch
n = 1 << 28   # 1 GiB of f32
a = torch.empty(n, dtype=torch.float32, device='cuda'); b = torch.empty_like(a)
def t(f, reps=10):
    f(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
ms = t(lambda: a.fill_(1.0)); print('fill  %.1f GB/s' % (n * 4 / ms / 1e6))
ms = t(lambda: a.zero_()); print('zero  %.1f GB/s' % (n * 4 / ms / 1e6))
ms = t(lambda: b.copy_(a)); print('copy  %.1f GB/s (r+w)' % (2 * n * 4 / ms / 1e6))
ms = t(lambda: a.sum()); print('read  %.1f GB/s' % (n * 4 / ms / 1e6))
