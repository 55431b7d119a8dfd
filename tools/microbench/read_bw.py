"""Read-only and copy bandwidth of library kernels on this box (context for the SpMV streaming numbers)."""
import torch
dev = torch.device("cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for mb in (134, 335, 671, 2048):
    n = mb * (1 << 20) // 4
    a = torch.ones(n, dtype=torch.float32, device=dev)
    b = torch.empty_like(a)
    for name, fn, bytes_ in (("sum (read only)", lambda: a.sum(), 4 * n), ("copy (read + write)", lambda: b.copy_(a), 8 * n)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(10):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        print(f"{mb} MB {name}: {best * 1e3:.1f} us  {bytes_ / best / 1e6:.0f} GB/s")
