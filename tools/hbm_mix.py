"""HBM ceilings for the bicubic traffic mix (1 byte read : 4 bytes written): write-only fill, copy, and a 1:4 torch op."""
import torch
def t(fn, reps=10):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(3):
        torch.cuda.synchronize(); e0.record()
        for _ in range(reps): fn()
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps)
    return best
n = 300 * 1080 * 1920 * 3
src = torch.randint(0, 255, (n,), dtype=torch.uint8, device="cuda")
dst = torch.empty(4 * n, dtype=torch.uint8, device="cuda")
ms = t(lambda: dst.fill_(7)); print(f"fill  {4*n/ms/1e6:8.0f} GB/s write-only ({ms:.3f} ms)")
a = torch.empty(4 * n, dtype=torch.uint8, device="cuda")
ms = t(lambda: dst.copy_(a)); print(f"copy  {8*n/ms/1e6:8.0f} GB/s 1:1 ({ms:.3f} ms)")
# 1 read : 4 written -- expand (broadcast) each 16-byte source chunk into 4 destination chunks
s4 = src.view(-1, 1, 16); d4 = dst.view(-1, 4, 16)
ms = t(lambda: d4.copy_(s4.expand(-1, 4, 16))); print(f"1:4   {5*n/ms/1e6:8.0f} GB/s ({ms:.3f} ms)")
s8 = src.view(torch.int64).view(-1, 1); d8 = dst.view(torch.int64).view(-1, 4)
ms = t(lambda: d8.copy_(s8.expand(-1, 4))); print(f"1:4 i64 {5*n/ms/1e6:8.0f} GB/s ({ms:.3f} ms)")
